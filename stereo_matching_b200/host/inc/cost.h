// cost.h — drop-in replacement for the reference's inc/cost.h (inc/cost.h:1-14).
//
// The per-pixel probes keep their signatures.  They are single-pixel host utilities (one window, one disparity per
// call) and are NOT on the process() path: the dense forms of the same costs run on the device
// (sgm_cost_volume_sad in include/sgm_b200.h, and the census / Hamming volume inside the engine).
#pragma once

#include "global.h"
#include "Solver.h"

// Mean absolute / squared difference of a win_h x win_w window at l_pt and the window shifted by disp / scale
// (src/cost.cpp:4-30, :33-59).  `weight` is accepted and ignored, as in the reference build (WEIGHTED_COST = 0).
float SAD(const Mat &img_l, const Mat &img_r, const Point &l_pt, int disp, int win_h, int win_w, float* weight, int scale=1);
float SSD(const Mat &img_l, const Mat &img_r, const Point &l_pt, int disp, int win_h, int win_w, float* weight, int scale=1);

// Census transform of both windows and their Hamming distance (src/cost.cpp:62-96).
int CT(const Mat &img_l, const Mat &img_r, const Point &l_pt, int disp, int win_h, int win_w, float* weight, int scale=1);
int hamming_cost(uint64_t ct_l, uint64_t ct_r);

// Census bit strings of pixel (u, v) in both images, written to cost_table_x[v * cols + u] (src/cost.cpp:99-129).
void CT_pts(const Mat &img_l, const Mat &img_r, int u, int v, int win_h, int win_w, float* weight,
            uint64_t *cost_table_l, uint64_t *cost_table_r);
