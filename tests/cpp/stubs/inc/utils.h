// Stand-in for the reference's inc/utils.h (only what the solver shells could touch).  TEST INFRASTRUCTURE.
#pragma once
#include "global.h"
double get_cur_ms();
