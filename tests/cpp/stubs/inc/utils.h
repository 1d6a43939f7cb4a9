// Stand-in for the reference's inc/utils.h (only what the solver shells could touch).  TEST INFRASTRUCTURE.
#pragma once
#include "global.h"
double get_cur_ms();

struct CamIntrinsics  // inc/utils.h:14-20
{
    float fx;
    float fy;
    float cx;
    float cy;
};
