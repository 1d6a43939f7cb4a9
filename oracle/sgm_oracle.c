/* oracle/sgm_oracle.c — CPU restatement of the reference's stereo hot path.  TEST INFRASTRUCTURE.
 * See sgm_oracle.h for scope and the parity pin.  All float arithmetic is IEEE binary32 in the
 * reference's operation order (compile with -ffp-contract=off, never -ffast-math).
 * OpenMP is used only across units the reference itself treats as independent.
 */
#include "sgm_oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define IDX3(i, j, k, W, D) ((((size_t)(i)) * (size_t)(W) + (size_t)(j)) * (size_t)(D) + (size_t)(k))

static int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

void orc_default_params(orc_params* p, int h, int w, int scale, int max_disp, int algo) {
    memset(p, 0, sizeof(*p));
    p->height = h;
    p->width = w;
    p->scale = scale;
    p->max_disp = max_disp;
    p->algo = algo;
    p->paths = 8;            /* inc/SGM.h:7 */
    p->P1 = 10;              /* src/SGM.cpp:27 */
    p->P2 = 100;             /* src/SGM.cpp:28 */
    p->census_h = 7;         /* inc/Solver.h:10 */
    p->census_w = 9;         /* inc/Solver.h:11 */
    p->box_h = 3;            /* inc/Solver.h:12 */
    p->box_w = 5;            /* inc/Solver.h:13 */
    p->unique_ratio = 0.7f;  /* inc/Solver.h:14 */
    p->lr_thresh = 1.0f;     /* inc/Solver.h:16 */
    p->median_h = 5;         /* inc/Solver.h:17 */
    p->median_w = 5;         /* inc/Solver.h:18 */
    p->speckle_size = 1000;  /* inc/Solver.h:19 */
    p->speckle_dis = 2;      /* inc/Solver.h:20 */
    p->sad_h = 9;            /* BASELINE.json config 2 */
    p->sad_w = 9;
}

/* src/SGM.cpp:40-61 (rows i*scale) and src/BM.cpp:17-38 (rows i: BM does not scale the row index). */
void orc_subsample(const uint8_t* src, int h, int w, int scale, int bm_rows, uint8_t* dst) {
    const int ih = h / scale, iw = w / scale;
    for (int i = 0; i < ih; ++i) {
        const uint8_t* row = src + (size_t)(bm_rows ? i : i * scale) * w;
        for (int j = 0; j < iw; ++j) dst[(size_t)i * iw + j] = row[j * scale];
    }
}

/* src/Solver.cpp:124-125: cv::GaussianBlur(img, out, Size(3,3), 2, 1).  OpenCV's CV_8U path is fixed
 * point: kx = [82,92,82]/256 (sigma 2), ky = [70,116,70]/256 (sigma 1), horizontal pass in 8.8,
 * vertical in 16.16, (v + 2^15) >> 16, BORDER_REFLECT_101.  Pinned against cv2 4.13 in the tests. */
static int reflect101(int p, int n) {
    if (n == 1) return 0;
    if (p < 0) return -p;
    if (p >= n) return 2 * n - 2 - p;
    return p;
}
void orc_gaussian_blur3(const uint8_t* src, int h, int w, uint8_t* dst) {
    static const int kx[3] = {82, 92, 82}, ky[3] = {70, 116, 70};
    int* hb = (int*)malloc(sizeof(int) * (size_t)h * w);
#pragma omp parallel for
    for (int i = 0; i < h; ++i) {
        const uint8_t* s = src + (size_t)i * w;
        for (int j = 0; j < w; ++j)
            hb[(size_t)i * w + j] = kx[0] * s[reflect101(j - 1, w)] + kx[1] * s[j] + kx[2] * s[reflect101(j + 1, w)];
    }
#pragma omp parallel for
    for (int i = 0; i < h; ++i) {
        const int* a = hb + (size_t)reflect101(i - 1, h) * w;
        const int* b = hb + (size_t)i * w;
        const int* c = hb + (size_t)reflect101(i + 1, h) * w;
        for (int j = 0; j < w; ++j) dst[(size_t)i * w + j] = (uint8_t)((ky[0] * a[j] + ky[1] * b[j] + ky[2] * c[j] + 32768) >> 16);
    }
    free(hb);
}

/* src/cost.cpp:99-129 (CT_pts): census over a (2*(win_h/2)+1) x (2*(win_w/2)+1) window, replicate-clamped
 * coordinates, centre skipped, bit = neighbour > centre, appended as (v << 1) | bit in row-major order. */
void orc_census(const uint8_t* img, int h, int w, int win_h, int win_w, uint64_t* table) {
    const int rh = win_h / 2, rw = win_w / 2;
#pragma omp parallel for
    for (int v = 0; v < h; ++v)
        for (int u = 0; u < w; ++u) {
            const uint8_t ctr = img[(size_t)v * w + u];
            uint64_t val = 0;
            for (int a = -rh; a <= rh; ++a) {
                const int y = clampi(v + a, 0, h - 1);
                for (int b = -rw; b <= rw; ++b) {
                    if (a == 0 && b == 0) continue;
                    const int x = clampi(u + b, 0, w - 1);
                    val = (val << 1) | (uint64_t)(img[(size_t)y * w + x] > ctr);
                }
            }
            table[(size_t)v * w + u] = val;
        }
}

/* src/cost.cpp:132-144 counts set bits of the xor with a shift loop == popcount. */
static int popcount64(uint64_t x) {
    int c = 0;
    while (x) {
        c += (int)(x & 1u);
        x >>= 1;
    }
    return c;
}

/* src/Solver.cpp:143-194 (left: right column max(j - d/scale, 0)) and :197-248 (beta: left column
 * min(j + d/scale, W-1)); sky pixels (mask == 255) get cost 0 at d == 0 and 999999 elsewhere. */
void orc_hamming_volume(const uint64_t* ct_l, const uint64_t* ct_r, const uint8_t* sky, int h, int w, int d,
                        int scale, int beta, float* cost) {
#pragma omp parallel for
    for (int i = 0; i < h; ++i)
        for (int j = 0; j < w; ++j) {
            float* c = cost + IDX3(i, j, 0, w, d);
            if (sky && sky[(size_t)i * w + j] == 255) {
                for (int k = 0; k < d; ++k) c[k] = (k == 0) ? 0.0f : 999999.0f;
                continue;
            }
            for (int k = 0; k < d; ++k) {
                uint64_t a, b;
                if (!beta) {
                    int jr = j - k / scale;
                    if (jr < 0) jr = 0;
                    a = ct_l[(size_t)i * w + j];
                    b = ct_r[(size_t)i * w + jr];
                } else {
                    int jl = j + k / scale;
                    if (jl > w - 1) jl = w - 1;
                    a = ct_l[(size_t)i * w + jl];
                    b = ct_r[(size_t)i * w + j];
                }
                c[k] = (float)popcount64(a ^ b);
            }
        }
}

/* src/cost.cpp:4-30 (SAD) evaluated in a build_dsi-shaped loop (src/Solver.cpp:96-117).  The sum of
 * absolute differences is accumulated in float (exact: <= 81*255) and divided by win_h then win_w.
 * beta != 0: the right-view volume by symmetry with src/Solver.cpp:239-243 (x_l = min(x_r + d/scale, W-1)). */
void orc_sad_volume(const uint8_t* img_l, const uint8_t* img_r, int h, int w, int d, int win_h, int win_w, int scale,
                    int beta, float* cost) {
#pragma omp parallel for
    for (int i = 0; i < h; ++i)
        for (int j = 0; j < w; ++j)
            for (int k = 0; k < d; ++k) {
                float acc = 0;
                for (int a = -win_h / 2; a <= win_h / 2; ++a) {
                    const int y = clampi(i + a, 0, h - 1);
                    const uint8_t* pl = img_l + (size_t)y * w;
                    const uint8_t* pr = img_r + (size_t)y * w;
                    for (int b = -win_w / 2; b <= win_w / 2; ++b) {
                        int xl, xr;
                        if (!beta) {
                            xl = clampi(j + b, 0, w - 1);
                            xr = xl - k / scale;
                            if (xr < 0) xr = 0;
                        } else {
                            xr = clampi(j + b, 0, w - 1);
                            xl = xr + k / scale;
                            if (xl > w - 1) xl = w - 1;
                        }
                        acc += (float)abs((int)pl[xl] - (int)pr[xr]);
                    }
                }
                cost[IDX3(i, j, k, w, d)] = acc / win_h / win_w;
            }
}

/* src/Solver.cpp:296-330.  In place: from the (win - win/2)-th output on, the value subtracted from the
 * running sum has already been overwritten with its filtered value, so this is a float recurrence along
 * j per (row, d), not a box mean.  Rows are independent (the reference runs them serially). */
void orc_box_filter_h(float* cost, int h, int w, int d, int win) {
    const size_t back = (size_t)(win / 2 + 1) * d;
#pragma omp parallel for
    for (int i = 0; i < h; ++i)
        for (int k = 0; k < d; ++k) {
            float sum = 0;
            size_t idx = IDX3(i, 0, k, w, d);
            for (int j = 0; j < win; ++j) {
                sum += cost[idx];
                idx += d;
            }
            for (int j = win / 2; j < w - win / 2; ++j) {
                cost[idx - back] = sum / win;
                if (j == w - win / 2 - 1) break;
                sum += cost[idx];
                sum -= cost[idx - (size_t)win * d];
                idx += d;
            }
        }
}

/* src/Solver.cpp:333-368: the same recurrence along i per (column, d). */
void orc_box_filter_v(float* cost, int h, int w, int d, int win) {
    const size_t step = (size_t)w * d;
    const size_t back = (size_t)(win / 2 + 1) * step;
#pragma omp parallel for
    for (int j = 0; j < w; ++j)
        for (int k = 0; k < d; ++k) {
            float sum = 0;
            size_t idx = (size_t)j * d + k;
            for (int i = 0; i < win; ++i) {
                sum += cost[idx];
                idx += step;
            }
            for (int i = win / 2; i < h - win / 2; ++i) {
                cost[idx - back] = sum / win;
                if (i == h - win / 2 - 1) break;
                sum += cost[idx];
                sum -= cost[idx - (size_t)win * step];
                idx += step;
            }
        }
}

/* One SGM path, src/SGM.cpp:82-369: predecessor pixel (i - di, j - dj); L = C on pixels whose predecessor
 * lies outside the image (:99,139,179,219,266,282,330,346), else
 *   L = MIN(MIN(MIN(Lp[d], Lp[d-1]+P1), Lp[d+1]+P1), minLp+P2) + (C - minLp)      (:105-108)
 * with d-1 / d+1 clamped to [0, D-1] (:96-97); minL = min_d L stored per pixel (:110-117). */
static void path_pixel(const float* C, const float* Lp, int has_prev, float min_prev, int d, int P1, int P2, float* L,
                       float* min_out) {
    float m = FLT_MAX;
    for (int k = 0; k < d; ++k) {
        float v;
        if (!has_prev) {
            v = C[k];
        } else {
            const int km = k - 1 < 0 ? 0 : k - 1;
            const int kp = k + 1 > d - 1 ? d - 1 : k + 1;
            float a = Lp[k], b = Lp[km] + P1;
            v = a > b ? b : a;
            b = Lp[kp] + P1;
            v = v > b ? b : v;
            b = min_prev + P2;
            v = v > b ? b : v;
            v += (C[k] - min_prev);
        }
        L[k] = v;
        if (v < m) m = v;
    }
    *min_out = m;
}

static void run_path(const float* cost, int h, int w, int d, int P1, int P2, int di, int dj, float* L, float* minL) {
    if (di == 0) {
        /* horizontal: rows independent, sequential along j in the direction of travel (:82-159) */
#pragma omp parallel for
        for (int i = 0; i < h; ++i)
            for (int s = 0; s < w; ++s) {
                const int j = dj > 0 ? s : w - 1 - s;
                const int jp = j - dj;
                const int has_prev = (jp >= 0 && jp < w);
                path_pixel(cost + IDX3(i, j, 0, w, d), has_prev ? L + IDX3(i, jp, 0, w, d) : NULL, has_prev,
                           has_prev ? minL[(size_t)i * w + jp] : 0.f, d, P1, P2, L + IDX3(i, j, 0, w, d),
                           &minL[(size_t)i * w + j]);
            }
        return;
    }
    /* vertical and diagonal: rows in the direction of travel, all columns of a row independent (:162-369) */
    for (int s = 0; s < h; ++s) {
        const int i = di > 0 ? s : h - 1 - s;
        const int ip = i - di;
#pragma omp parallel for
        for (int j = 0; j < w; ++j) {
            const int jp = j - dj;
            const int has_prev = (ip >= 0 && ip < h && jp >= 0 && jp < w);
            path_pixel(cost + IDX3(i, j, 0, w, d), has_prev ? L + IDX3(ip, jp, 0, w, d) : NULL, has_prev,
                       has_prev ? minL[(size_t)ip * w + jp] : 0.f, d, P1, P2, L + IDX3(i, j, 0, w, d),
                       &minL[(size_t)i * w + j]);
        }
    }
}

/* src/SGM.cpp:82-390: the eight paths and their sum, in the reference's association order
 *   S = (((L1+L2)+L3)+L4) + (((L5+L6)+L7)+L8)                                     (:386-390)
 * L1 ->, L2 <-, L3 down, L4 up, L5 down-right, L6 down-left, L7 up-right, L8 up-left. */
void orc_aggregate(const float* cost, int h, int w, int d, int P1, int P2, int paths, float* S) {
    static const int DI[8] = {0, 0, 1, -1, 1, 1, -1, -1};
    static const int DJ[8] = {1, -1, 0, 0, 1, -1, 1, -1};
    const size_t n = (size_t)h * w, nd = n * d;
    float* L = (float*)malloc(nd * sizeof(float));
    float* B = paths == 8 ? (float*)malloc(nd * sizeof(float)) : NULL;
    float* minL = (float*)malloc(n * sizeof(float));
    for (int p = 0; p < paths; ++p) {
        run_path(cost, h, w, d, P1, P2, DI[p], DJ[p], L, minL);
        float* acc = p < 4 ? S : B;
        if (p == 0 || p == 4) {
            memcpy(acc, L, nd * sizeof(float));
        } else {
#pragma omp parallel for
            for (size_t t = 0; t < nd; ++t) acc[t] = acc[t] + L[t];
        }
    }
    if (paths == 8) {
#pragma omp parallel for
        for (size_t t = 0; t < nd; ++t) S[t] = S[t] + B[t];
    }
    free(L);
    free(B);
    free(minL);
}

/* Integer restatement of orc_aggregate (the engine's SGM_FLAG_INTEGER mode): the same eight paths
 * (src/SGM.cpp:82-369) and their sum (:386-390) in 16-bit integer arithmetic.  It applies when the costs are small
 * non-negative integers (raw Hamming costs: cost filter window 1x1, no sky fill); the float loops of the reference then
 * only ever produce integers < 2^24, every float operation is exact, and this function returns the same numbers as
 * orc_aggregate (tests/test_oracle_cpu.py checks that), which is what makes the integer mode comparable with the
 * reference at all. */
static void path_pixel_u16(const uint16_t* C, const uint16_t* Lp, int has_prev, int min_prev, int d, int P1, int P2,
                           uint16_t* L, uint16_t* min_out) {
    int m = 0x7fffffff;
    for (int k = 0; k < d; ++k) {
        int v;
        if (!has_prev) {
            v = C[k];
        } else {
            const int km = k - 1 < 0 ? 0 : k - 1;
            const int kp = k + 1 > d - 1 ? d - 1 : k + 1;
            v = Lp[k];
            if (Lp[km] + P1 < v) v = Lp[km] + P1;
            if (Lp[kp] + P1 < v) v = Lp[kp] + P1;
            if (min_prev + P2 < v) v = min_prev + P2;
            v += C[k] - min_prev;
        }
        L[k] = (uint16_t)v;
        if (v < m) m = v;
    }
    *min_out = (uint16_t)m;
}

void orc_aggregate_u16(const uint16_t* cost, int h, int w, int d, int P1, int P2, int paths, uint16_t* S) {
    static const int DI[8] = {0, 0, 1, -1, 1, 1, -1, -1};
    static const int DJ[8] = {1, -1, 0, 0, 1, -1, 1, -1};
    const size_t n = (size_t)h * w, nd = n * d;
    uint16_t* L = (uint16_t*)malloc(nd * sizeof(uint16_t));
    uint16_t* minL = (uint16_t*)malloc(n * sizeof(uint16_t));
    memset(S, 0, nd * sizeof(uint16_t));
    for (int p = 0; p < paths; ++p) {
        const int di = DI[p], dj = DJ[p];
        /* pixels in the direction of travel: the predecessor (i - di, j - dj) is always finished first */
        for (int si = 0; si < h; ++si) {
            const int i = di >= 0 ? si : h - 1 - si;
            for (int sj = 0; sj < w; ++sj) {
                const int j = dj >= 0 ? sj : w - 1 - sj;
                const int ip = i - di, jp = j - dj;
                const int has_prev = ip >= 0 && ip < h && jp >= 0 && jp < w;
                path_pixel_u16(cost + IDX3(i, j, 0, w, d), has_prev ? L + IDX3(ip, jp, 0, w, d) : NULL, has_prev,
                               has_prev ? minL[(size_t)ip * w + jp] : 0, d, P1, P2, L + IDX3(i, j, 0, w, d),
                               &minL[(size_t)i * w + j]);
            }
        }
        for (size_t t = 0; t < nd; ++t) S[t] = (uint16_t)(S[t] + L[t]);
    }
    free(L);
    free(minL);
}

/* src/SGM.cpp:373-418 (min_gap = 1) and src/BM.cpp:52-91 (min_gap = 2): first strict minimum, first
 * occurrence of the smallest value different from it, invalid (= D+1) iff
 * min/sec > ratio && |min_d - sec_d| > min_gap.  sec_d is deliberately carried across pixels in raster
 * order like the reference's variable (it only matters when a second value exists, which resets it). */
void orc_wta(const float* S, int h, int w, int d, float ratio, int min_gap, uint16_t* disp) {
    int min_d = d + 1, sec_d = d + 1;
    for (int i = 0; i < h; ++i)
        for (int j = 0; j < w; ++j) {
            const float* c = S + IDX3(i, j, 0, w, d);
            float mn = FLT_MAX, sec = FLT_MAX;
            for (int k = 0; k < d; ++k)
                if (c[k] < mn) {
                    mn = c[k];
                    min_d = k;
                }
            for (int k = 0; k < d; ++k)
                if (c[k] < sec && c[k] != mn) {
                    sec = c[k];
                    sec_d = k;
                }
            if (mn / sec > ratio && abs(min_d - sec_d) > min_gap)
                disp[(size_t)i * w + j] = (uint16_t)(d + 1);
            else
                disp[(size_t)i * w + j] = (uint16_t)min_d;
        }
}

/* src/Solver.cpp:569-597: parabola through S[d-1], S[d], S[d+1], clamped to D-1. */
void orc_subpixel(const uint16_t* disp, const float* S, int h, int w, int d, float* out) {
#pragma omp parallel for
    for (int i = 0; i < h; ++i)
        for (int j = 0; j < w; ++j) {
            const int k = disp[(size_t)i * w + j];
            float* o = out + (size_t)i * w + j;
            if (k > d - 1) {
                *o = (float)(d + 1);
            } else if (!k || k == d - 1) {
                *o = (float)k;
            } else {
                const float* c = S + IDX3(i, j, k, w, d);
                const float c0 = c[0], cm = c[-1], cp = c[1];
                const float v = k + (cm - cp) / (2 * (cm + cp - 2 * c0));
                const float cap = (d - 1) * 1.f;
                *o = cap < v ? cap : v; /* std::min(v, cap) */
            }
        }
}

/* src/SGM.cpp:803-818: for j >= dl, dr = right(i, int(j - dl/scale)); |dl - dr| > thresh -> invalid. */
void orc_lr_check(float* fl, const float* fr, int h, int w, int d, int scale, float thresh) {
#pragma omp parallel for
    for (int i = 0; i < h; ++i)
        for (int j = 0; j < w; ++j) {
            float* dl = fl + (size_t)i * w + j;
            if (j >= *dl) {
                const int jr = (int)(j - *dl / scale);
                const float dr = fr[(size_t)i * w + jr];
                if (fabs(*dl - dr) > thresh) *dl = (float)(d + 1);
            }
        }
}

static int cmp_int(const void* a, const void* b) {
    const int x = *(const int*)a, y = *(const int*)b;
    return (x > y) - (x < y);
}

/* src/Solver.cpp:604-630: in place, raster order, interior pixels only; an invalid pixel (> D-1) takes
 * the upper median of the valid window values truncated to int when more than half the window is valid. */
void orc_median_fill(float* f, int h, int w, int d, int win_h, int win_w) {
    const int rh = win_h / 2, rw = win_w / 2;
    int* v = (int*)malloc(sizeof(int) * (size_t)win_h * win_w);
    for (int i = rh; i < h - rh; ++i)
        for (int j = rw; j < w - rw; ++j) {
            if (f[(size_t)i * w + j] <= d - 1) continue;
            int cnt = 0;
            for (int m = i - rh; m <= i + rh; ++m)
                for (int n = j - rw; n <= j + rw; ++n)
                    if (f[(size_t)m * w + n] <= d - 1) v[cnt++] = (int)f[(size_t)m * w + n];
            if (cnt > win_w * win_h / 2) {
                qsort(v, (size_t)cnt, sizeof(int), cmp_int);
                f[(size_t)i * w + j] = (float)v[cnt / 2];
            }
        }
    free(v);
}

/* src/Solver.cpp:493-566 (speckle_filter_new) with serial semantics: 4-connected components under
 * |a - b| < max_dis, components of area <= max_size are set to `value`. */
static int uf_find(int i, const int* label) {
    while (i != label[i]) i = label[i];
    return i;
}
static void uf_union(int i, int j, int* label) {
    const int a = uf_find(i, label), b = uf_find(j, label);
    if (a != b) label[a] = b;
}
void orc_speckle(float* f, int h, int w, float value, int max_size, int max_dis) {
    const int n = h * w;
    int* label = (int*)malloc(sizeof(int) * (size_t)n);
    int* area = (int*)calloc((size_t)n, sizeof(int));
    for (int t = 0; t < n; ++t) label[t] = t;
    for (int t = 0; t < n; ++t) {
        const int r = t / w, c = t % w;
        if (r > 0 && fabs(f[t] - f[t - w]) < max_dis) uf_union(t - w, t, label);
        if (r < h - 1 && fabs(f[t] - f[t + w]) < max_dis) uf_union(t + w, t, label);
        if (c > 0 && fabs(f[t] - f[t - 1]) < max_dis) uf_union(t - 1, t, label);
        if (c < w - 1 && fabs(f[t] - f[t + 1]) < max_dis) uf_union(t + 1, t, label);
    }
    for (int t = 0; t < n; ++t) {
        label[t] = uf_find(t, label);
        area[label[t]]++;
    }
    for (int t = 0; t < n; ++t)
        if (area[label[t]] <= max_size) f[t] = value;
    free(label);
    free(area);
}

uint64_t orc_fnv1a64(const void* buf, size_t bytes) {
    const uint8_t* p = (const uint8_t*)buf;
    uint64_t hsh = 1469598103934665603ULL;
    for (size_t i = 0; i < bytes; ++i) hsh = (hsh ^ p[i]) * 1099511628211ULL;
    return hsh;
}

#define TAP(dst, src, bytes)                  \
    do {                                      \
        if (dst) memcpy((dst), (src), (bytes)); \
    } while (0)

/* src/SGM.cpp:32-826 (SGM), src/BM.cpp:9-97 (BM) and the SAD-BM definition of sgm_oracle.h. */
int orc_process(const orc_params* p, const uint8_t* left, const uint8_t* right, const uint8_t* sky_l,
                const uint8_t* sky_r, orc_taps* taps) {
    orc_taps none;
    memset(&none, 0, sizeof(none));
    if (!taps) taps = &none;
    if (!p || !left || !right) return -1;
    const int s = p->scale, D = p->max_disp;
    if (p->height <= 0 || p->width <= 0 || s <= 0 || D <= 0) return -1;
    const int h = p->height / s, w = p->width / s;
    const size_t n = (size_t)h * w, nd = n * D;
    const int is_bm = p->algo != ORC_ALGO_SGM;
    const int views = p->algo == ORC_ALGO_BM ? 1 : 2;
    const uint8_t* sky[2] = {sky_l, sky_r};

    uint8_t* img[2];
    uint8_t* blur[2];
    uint64_t* ct[2];
    for (int v = 0; v < 2; ++v) {
        img[v] = (uint8_t*)malloc(n);
        blur[v] = (uint8_t*)malloc(n);
        ct[v] = (uint64_t*)malloc(n * sizeof(uint64_t));
    }
    orc_subsample(left, p->height, p->width, s, is_bm, img[0]);
    orc_subsample(right, p->height, p->width, s, is_bm, img[1]);
    for (int v = 0; v < 2; ++v) TAP(taps->img[v], img[v], n);

    if (p->algo != ORC_ALGO_BM_SAD) {
        /* src/Solver.cpp:120-140 */
        for (int v = 0; v < 2; ++v) {
            orc_gaussian_blur3(img[v], h, w, blur[v]);
            orc_census(blur[v], h, w, p->census_h / s, p->census_w / s, ct[v]);
            TAP(taps->blur[v], blur[v], n);
            TAP(taps->census[v], ct[v], n * sizeof(uint64_t));
        }
    }

    float* cost = (float*)malloc(nd * sizeof(float));
    float* S = p->algo == ORC_ALGO_SGM ? (float*)malloc(nd * sizeof(float)) : NULL;
    uint16_t* disp = (uint16_t*)malloc(n * sizeof(uint16_t));
    float* fsub[2] = {(float*)malloc(n * sizeof(float)), (float*)malloc(n * sizeof(float))};

    for (int v = 0; v < views; ++v) {
        const float* vol;
        if (p->algo == ORC_ALGO_BM_SAD) {
            orc_sad_volume(img[0], img[1], h, w, D, p->sad_h / s, p->sad_w / s, s, v, cost);
            TAP(taps->cost_raw[v], cost, nd * sizeof(float));
        } else {
            orc_hamming_volume(ct[0], ct[1], sky[v], h, w, D, s, v, cost);
            TAP(taps->cost_raw[v], cost, nd * sizeof(float));
            orc_box_filter_h(cost, h, w, D, p->box_w / s); /* src/SGM.cpp:71, src/BM.cpp:46 */
            TAP(taps->cost_h[v], cost, nd * sizeof(float));
            orc_box_filter_v(cost, h, w, D, p->box_h / s); /* src/SGM.cpp:72, src/BM.cpp:47 */
            TAP(taps->cost_hv[v], cost, nd * sizeof(float));
        }
        if (p->algo == ORC_ALGO_SGM) {
            orc_aggregate(cost, h, w, D, p->P1, p->P2, p->paths, S);
            TAP(taps->S[v], S, nd * sizeof(float));
            vol = S;
        } else {
            vol = cost;
        }
        orc_wta(vol, h, w, D, p->unique_ratio, is_bm ? 2 : 1, disp);
        TAP(taps->disp[v], disp, n * sizeof(uint16_t));
        if (p->algo == ORC_ALGO_BM_SAD) {
            for (size_t t = 0; t < n; ++t) fsub[v][t] = (float)disp[t];
        } else {
            /* src/SGM.cpp:443,801.  The reference's BM never writes filtered_disp (src/BM.cpp:52-96);
             * the restatement defines it as compute_subpixel(disp) on the filtered cost volume. */
            orc_subpixel(disp, vol, h, w, D, fsub[v]);
        }
        TAP(taps->fsub[v], fsub[v], n * sizeof(float));
    }

    float* f = fsub[0];
    if (views == 2) orc_lr_check(f, fsub[1], h, w, D, s, p->lr_thresh);
    TAP(taps->f_lr, f, n * sizeof(float));
    orc_median_fill(f, h, w, D, p->median_h, p->median_w); /* src/Solver.cpp:604-630 */
    TAP(taps->f_median, f, n * sizeof(float));
    orc_speckle(f, h, w, (float)(D + 1), p->speckle_size / s, p->speckle_dis); /* src/Solver.cpp:634 */
    TAP(taps->f_final, f, n * sizeof(float));

    for (int v = 0; v < 2; ++v) {
        free(img[v]);
        free(blur[v]);
        free(ct[v]);
        free(fsub[v]);
    }
    free(cost);
    free(S);
    free(disp);
    return 0;
}
