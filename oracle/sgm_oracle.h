/* oracle/sgm_oracle.h — CPU restatement of the reference's stereo hot path.  TEST INFRASTRUCTURE.
 *
 * Plain-C restatement of hilbertw/stereo_matching's SGM / BM CPU path, one function per stage, each
 * citing the reference file:line it follows.  It exists because the compiled reference (oracle/_ref)
 *   - cannot expose every intermediate (the left-pass path sum is overwritten by the beta pass),
 *   - asserts D in {32,64,128} and stores disparity in uchar (src/Solver.cpp:10,16,18), so BASELINE's
 *     D=256 config needs a *widened* restatement (uint16 disparity), and
 *   - has no SAD block-matching pipeline (src/BM.cpp uses the census volume), so BASELINE's config 2
 *     is defined here from the reference's own SAD() (src/cost.cpp:4-30).
 * Parity pin: tests/test_oracle_cpu.py checks every stage of this file bit-for-bit against
 * oracle/_ref/libsgm_ref.so (the reference itself, compiled here) and against tests/golden/.
 * Results for D>128 and for the SAD-BM mode are "widened restatement", not "reference".
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use this library.
 */
#ifndef SGM_ORACLE_H_
#define SGM_ORACLE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_ALGO_SGM = 0, ORC_ALGO_BM = 1, ORC_ALGO_BM_SAD = 2 };

typedef struct orc_params {
    int height, width;       /* full-resolution input size (ctor args h, w) */
    int scale;               /* 1 or 2 (src/Solver.cpp:8) */
    int max_disp;            /* D; the reference allows 32/64/128, the restatement any multiple of 32 */
    int algo;                /* ORC_ALGO_* */
    int paths;               /* 4 or 8 (inc/SGM.h:7 USE_8_PATH) */
    int P1, P2;              /* src/SGM.cpp:27-28 */
    int census_h, census_w;  /* WIN_H, WIN_W (inc/Solver.h:10-11), divided by scale at use */
    int box_w, box_h;        /* COST_WIN_W, COST_WIN_H (inc/Solver.h:12-13), divided by scale at use */
    float unique_ratio;      /* UNIQUE_RATIO (inc/Solver.h:14) */
    float lr_thresh;         /* LR_CHECK_DIS (inc/Solver.h:16) */
    int median_h, median_w;  /* MEDIAN_FILTER_H/W (inc/Solver.h:17-18) */
    int speckle_size;        /* SPECKLE_SIZE (inc/Solver.h:19), divided by scale at use */
    int speckle_dis;         /* SPECKLE_DIS (inc/Solver.h:20) */
    int sad_h, sad_w;        /* SAD window for ORC_ALGO_BM_SAD, divided by scale at use */
} orc_params;

/* Every pointer is nullable; non-null ones receive that stage's buffer.  Index [0] = left view,
 * [1] = right ("beta") view.  Sizes use ih = height/scale, iw = width/scale, n = ih*iw. */
typedef struct orc_taps {
    uint8_t* img[2];       /* n   : subsampled inputs */
    uint8_t* blur[2];      /* n   : after the 3x3 Gaussian */
    uint64_t* census[2];   /* n   : cost_table_l / cost_table_r */
    float* cost_raw[2];    /* n*D : Hamming (or SAD) volume */
    float* cost_h[2];      /* n*D : after cost_horizontal_filter */
    float* cost_hv[2];     /* n*D : after cost_vertical_filter */
    float* S[2];           /* n*D : summed path volume (SGM only) */
    uint16_t* disp[2];     /* n   : WTA + uniqueness result (invalid = D+1) */
    float* fsub[2];        /* n   : compute_subpixel output, before the LR-check */
    float* f_lr;           /* n   : left, after LR-check */
    float* f_median;       /* n   : left, after the hole-filling median */
    float* f_final;        /* n   : left, after the speckle filter == get_disp() */
} orc_taps;

void orc_default_params(orc_params* p, int h, int w, int scale, int max_disp, int algo);
int orc_process(const orc_params* p, const uint8_t* left, const uint8_t* right, const uint8_t* sky_l,
                const uint8_t* sky_r, orc_taps* taps);

/* Individual stages (same code orc_process runs). */
void orc_subsample(const uint8_t* src, int h, int w, int scale, int bm_rows, uint8_t* dst);
void orc_gaussian_blur3(const uint8_t* src, int h, int w, uint8_t* dst);
void orc_census(const uint8_t* img, int h, int w, int win_h, int win_w, uint64_t* table);
void orc_hamming_volume(const uint64_t* ct_l, const uint64_t* ct_r, const uint8_t* sky, int h, int w, int d,
                        int scale, int beta, float* cost);
void orc_sad_volume(const uint8_t* img_l, const uint8_t* img_r, int h, int w, int d, int win_h, int win_w, int scale,
                    int beta, float* cost);
void orc_box_filter_h(float* cost, int h, int w, int d, int win);
void orc_box_filter_v(float* cost, int h, int w, int d, int win);
void orc_aggregate(const float* cost, int h, int w, int d, int P1, int P2, int paths, float* S);
/* Integer (16-bit) form of orc_aggregate for integer-valued costs: the oracle of the engine's SGM_FLAG_INTEGER mode. */
void orc_aggregate_u16(const uint16_t* cost, int h, int w, int d, int P1, int P2, int paths, uint16_t* S);
void orc_wta(const float* S, int h, int w, int d, float ratio, int min_gap, uint16_t* disp);
void orc_subpixel(const uint16_t* disp, const float* S, int h, int w, int d, float* out);
void orc_lr_check(float* fl, const float* fr, int h, int w, int d, int scale, float thresh);
void orc_median_fill(float* f, int h, int w, int d, int win_h, int win_w);
void orc_speckle(float* f, int h, int w, float value, int max_size, int max_dis);
uint64_t orc_fnv1a64(const void* buf, size_t bytes);

#ifdef __cplusplus
}
#endif
#endif /* SGM_ORACLE_H_ */
