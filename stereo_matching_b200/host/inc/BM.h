// BM.h — drop-in replacement for the reference's inc/BM.h (inc/BM.h:1-20): block matching (census + recursive cost
// filter + winner-take-all, no aggregation), left view only.
#pragma once

#include "Solver.h"

class BM : public Solver
{
public:
    explicit BM(int h, int w, int s, int d);
    virtual ~BM();

    BM(const BM&) = delete;
    BM& operator=(const BM&) = delete;

    virtual void process(Mat &img_l, Mat &img_r);
    virtual void process(Mat &img_l, Mat &img_r, Mat &sky_mask, Mat &sky_mask_beta);

protected:
    // BM decimates with source row i, not i*scale (src/BM.cpp:23-24); the debug view follows it.
    virtual int debug_source_row(int i) const { return i; }
};

typedef std::shared_ptr<BM> BMSolverPtr;
