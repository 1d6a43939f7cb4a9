// BM.cpp — host shell of the block-matching solver (reference: src/BM.cpp:9-105): census + recursive cost filter +
// winner-take-all with the |dd| > 2 uniqueness gap, left view only, then the post filter.
//
// One documented deviation: the reference's BM::process never writes filtered_disp before post-filtering it
// (src/BM.cpp:52-96 fills `disp` only), so its get_disp() returns a stale buffer.  The engine defines the float result
// as compute_subpixel(disp) followed by the post filter, which is what SGM does (src/SGM.cpp:443); `disp` itself is
// bit-identical to the reference's.
#include "../inc/BM.h"

BM::BM(int h, int w, int s, int d) : Solver(h, w, s, d)
{
    open_engine(SGM_ALGO_BM, 8, h, w);
}

BM::~BM() {}

void BM::process(Mat &img_l, Mat &img_r)
{
    run_engine(img_l, img_r);
}

void BM::process(Mat &img_l, Mat &img_r, Mat &sky_mask, Mat &sky_mask_beta)
{
    this->sky_mask = sky_mask;            // src/BM.cpp:102-104
    this->sky_mask_beta = sky_mask_beta;
    process(img_l, img_r);
}
