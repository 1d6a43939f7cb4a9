// SGM.h — drop-in replacement for the reference's inc/SGM.h (inc/SGM.h:1-29): semi-global matching solver.
// The eight path volumes L1..L8 and their per-pixel minima, which the reference allocates on the host
// (src/SGM.cpp:7-24), live in device memory inside the engine; this class only carries the parameters.
#pragma once

#include "Solver.h"
#include "utils.h"

const bool USE_8_PATH = true;  // inc/SGM.h:7: 8 aggregation paths (false: the 4 axis-aligned ones)

class SGM : public Solver
{
public:
    explicit SGM(int h, int w, int s, int d);
    virtual ~SGM();

    SGM(const SGM&) = delete;
    SGM& operator=(const SGM&) = delete;

    virtual void process(Mat &img_l, Mat &img_r);
    virtual void process(Mat &img_l, Mat &img_r, Mat &sky_mask, Mat &sky_mask_beta);

private:
    int P1, P2;  // src/SGM.cpp:27-28
};

typedef std::shared_ptr<SGM> SGMSolverPtr;
