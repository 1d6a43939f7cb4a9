// SGM.cpp — host shell of the SGM solver (reference: src/SGM.cpp).  The reference's process() is ~800 lines of
// OpenMP loops (cost volume, 8 path recurrences, sum, WTA, the whole thing again for the right view, LR check, post
// filter: src/SGM.cpp:32-826); here it is one call into the device engine.
#include "../inc/SGM.h"

SGM::SGM(int h, int w, int s, int d) : Solver(h, w, s, d), P1(10), P2(100)  // src/SGM.cpp:27-28
{
    open_engine(SGM_ALGO_SGM, USE_8_PATH ? 8 : 4, h, w);
}

SGM::~SGM() {}

void SGM::process(Mat &img_l, Mat &img_r)
{
    run_engine(img_l, img_r);  // also refreshes disp_beta / filtered_disp_beta (src/SGM.cpp:801)
}

void SGM::process(Mat &img_l, Mat &img_r, Mat &sky_mask, Mat &sky_mask_beta)
{
    this->sky_mask = sky_mask;            // src/SGM.cpp:831-833
    this->sky_mask_beta = sky_mask_beta;
    process(img_l, img_r);
}
