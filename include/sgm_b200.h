/* sgm_b200.h — C ABI of the B200-native stereo matching engine (libsgm_b200.so).
 *
 * This is the drop-in boundary for the reference's dense-stereo hot path.  The reference
 * (hilbertw/stereo_matching) runs that path on the CPU inside its Solver / SGM / BM classes
 * (inc/Solver.h:23-70, inc/SGM.h:10-27, inc/BM.h:6-18); here it runs as hand-written sm_100a
 * CUDA kernels behind an opaque handle.  The C++ shells in stereo_matching_b200/host/ keep the
 * reference class surface and forward to these entry points; INTEGRATION.md shows the binding.
 *
 * Conventions: plain pointers and sizes only; host pointers unless a name says `_device`;
 * strides are in BYTES; every call returns 0 (SGM_OK) or an sgm_status error code and never aborts;
 * sgm_last_error() gives the message.  One handle processes one frame at a time and is not
 * thread-safe; distinct handles are independent (one per GPU / per in-flight frame).
 * There is NO CPU fallback: without a CUDA device sgm_create() fails with SGM_ERR_CUDA.
 */
#ifndef SGM_B200_H_
#define SGM_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SGM_B200_ABI_VERSION 1

#if defined(__GNUC__)
#define SGM_API __attribute__((visibility("default")))
#else
#define SGM_API
#endif

typedef struct sgm_engine sgm_engine; /* opaque */

typedef enum sgm_status {
    SGM_OK = 0,
    SGM_ERR_INVALID = 1,     /* bad argument / shape (the reference would assert: src/Solver.cpp:6-10, src/SGM.cpp:34-38) */
    SGM_ERR_UNSUPPORTED = 2, /* parameter combination the kernels are not built for */
    SGM_ERR_CUDA = 3,        /* CUDA runtime error, or no device */
    SGM_ERR_NOMEM = 4,
    SGM_ERR_STATE = 5        /* result requested before any frame was processed */
} sgm_status;

typedef enum sgm_algo {
    SGM_ALGO_SGM = 0,    /* SGM::process  (src/SGM.cpp:32-826): census + recursive box + 4/8-path SGM, both views, LR, post filter */
    SGM_ALGO_BM = 1,     /* BM::process   (src/BM.cpp:9-97): census + recursive box + WTA(|dd|>2), left view, post filter */
    SGM_ALGO_BM_SAD = 2  /* SAD block matching built from SAD() (src/cost.cpp:4-30): both views, WTA(|dd|>2), LR, post filter */
} sgm_algo;

/* Replaces the compile-time constants of inc/Solver.h:10-20, inc/SGM.h:7 and src/SGM.cpp:27-28
 * and the ctor arguments Solver(h, w, s, d) (src/Solver.cpp:4-16).  sgm_default_params() fills
 * the reference's values. */
typedef struct sgm_params {
    int32_t height, width;       /* full-resolution input size (ctor h, w) */
    int32_t scale;               /* 1 or 2: nearest-neighbour decimation before matching */
    int32_t max_disp;            /* D: 32, 64, 128 (reference) or 256 (widened, uint16 disparity) */
    int32_t algo;                /* sgm_algo */
    int32_t paths;               /* 4 or 8 aggregation paths (USE_8_PATH) */
    int32_t P1, P2;              /* SGM penalties (10, 100) */
    int32_t census_h, census_w;  /* WIN_H, WIN_W = 7, 9 (divided by scale at use, like the reference) */
    int32_t box_w, box_h;        /* COST_WIN_W, COST_WIN_H = 5, 3 (divided by scale at use) */
    float unique_ratio;          /* UNIQUE_RATIO = 0.7 */
    float lr_thresh;             /* LR_CHECK_DIS = 1 */
    int32_t median_h, median_w;  /* MEDIAN_FILTER_H/W = 5, 5 */
    int32_t speckle_size;        /* SPECKLE_SIZE = 1000 (divided by scale at use) */
    int32_t speckle_dis;         /* SPECKLE_DIS = 2 */
    int32_t sad_h, sad_w;        /* SAD window (9, 9) for SGM_ALGO_BM_SAD (divided by scale at use) */
    int32_t device;              /* CUDA device ordinal */
    uint32_t flags;              /* SGM_FLAG_* */
} sgm_params;

#define SGM_FLAG_KEEP_VOLUMES 0x1u /* keep the summed path volume S of both views readable through sgm_get_tap (debug; costs 4 B per disp.px) */
#define SGM_FLAG_LINE_WALKERS 0x4u /* aggregate with the independent line-walker kernels (8 launches, 88 B/disp.px) instead of the fused sweeps */
/* Integer / DPX mode (north_star): with the cost filter switched off (box_w / scale == box_h / scale == 1) every
 * number on the reference's path (src/SGM.cpp:82-418) is an integer < 2^15 that float32 holds exactly, so the engine keeps the
 * volumes as 16-bit integers (half the bytes) and runs the min-plus recurrences on s16x2 pairs with the DPX instructions
 * VIMNMX3 / VIADDMNMX.  Results are IDENTICAL to the float path under the same parameters (tested tap by tap); sgm_create rejects
 * parameter sets for which that would not hold (SGM_ERR_UNSUPPORTED).  Sky masks are accepted (P2 <= 1365): masked pixels cost
 * 2 * P2 instead of the reference's 999999 (src/Solver.cpp:165-179), which provably changes no disparity, sub-pixel value or
 * validity anywhere; only the cost / path-sum debug taps of the masked pixels themselves read 2 * P2 + t instead of 999999 + t. */
#define SGM_FLAG_INTEGER 0x8u
/* The C++ shells keep Solver's protected result Mats (disp, disp_beta, filtered_disp_beta: inc/Solver.h:43-46) filled after every
 * process(): with this flag the 8-bit disparities of both views (narrowed on the device) and the right view's sub-pixel image are
 * copied to pinned host memory on the frame's stream together with the result, so sgm_get_disparity_u8 / sgm_get_subpixel_beta
 * are plain host copies (no synchronous cudaMemcpy, no allocation). */
#define SGM_FLAG_HOST_TAPS 0x10u
#define SGM_FLAG_STAGE_TIMING 0x2u /* record CUDA events around the pipeline stages on the engine's stream (sgm_get_stage_ms) */

/* Buffers readable after a frame, for parity tests and debugging (the reference keeps the same data in
 * Solver's protected members, inc/Solver.h:43-52).  view: 0 = left, 1 = right ("beta"). */
typedef enum sgm_tap {
    SGM_TAP_CENSUS = 0,    /* uint64 [ih*iw]       cost_table_l (view 0) / cost_table_r (view 1) */
    SGM_TAP_COST = 1,      /* float  [ih*iw*D]     cost volume after both box filters (SAD volume for BM_SAD) */
    SGM_TAP_PATHSUM = 2,   /* float  [ih*iw*D]     summed path volume S (needs SGM_FLAG_KEEP_VOLUMES) */
    SGM_TAP_DISP = 3,      /* uint16 [ih*iw]       WTA + uniqueness disparity, invalid = D+1 (Solver::disp / disp_beta) */
    SGM_TAP_SUBPIXEL = 4,  /* float  [ih*iw]       compute_subpixel output before the LR-check */
    SGM_TAP_LR = 5,        /* float  [ih*iw]       left view after the LR-check (view must be 0) */
    SGM_TAP_MEDIAN = 6,    /* float  [ih*iw]       left view after the hole-filling median (view must be 0) */
    SGM_TAP_FINAL = 7,     /* float  [ih*iw]       left view after the speckle filter == Solver::get_disp() */
    SGM_TAP_BLUR = 8       /* uint8  [ih*iw]       Gaussian-blurred input */
} sgm_tap;

typedef enum sgm_stage {
    SGM_STAGE_UPLOAD = 0,     /* H2D of the inputs */
    SGM_STAGE_COST = 1,       /* subsample, blur, census, Hamming/SAD volume, recursive box filters */
    SGM_STAGE_AGGREGATE = 2,  /* all path kernels incl. the fused sum/WTA/uniqueness/sub-pixel epilogue */
    SGM_STAGE_POST = 3,       /* LR-check, hole-filling median, speckle filter */
    SGM_STAGE_DOWNLOAD = 4,   /* D2H of the result */
    SGM_STAGE_COUNT = 5
} sgm_stage;

SGM_API void sgm_default_params(sgm_params* p, int height, int width, int scale, int max_disp, int algo);

/* Solver::Solver / SGM::SGM / BM::BM (src/Solver.cpp:4-52, src/SGM.cpp:4-29): validates, allocates all
 * device volumes, pinned staging and the stream. */
SGM_API int sgm_create(sgm_engine** out, const sgm_params* params);
/* Solver::~Solver / SGM::~SGM (src/Solver.cpp:710-719, src/SGM.cpp:837-858). */
SGM_API void sgm_destroy(sgm_engine* e);
/* Message of the last error on this handle; with e == NULL, of the last failed sgm_create on this thread. */
SGM_API const char* sgm_last_error(const sgm_engine* e);

/* Solver::process(img_l, img_r, sky_mask, sky_mask_beta) (src/SGM.cpp:829-834, src/BM.cpp:100-105): host
 * images in (8-bit, `height` x `width`, row strides in bytes), blocking; the result is then available from
 * the getters.  sky_l / sky_r may be NULL (Solver::process(img_l, img_r)); when given they are
 * (height/scale) x (width/scale) masks with 255 = sky (src/Solver.cpp:146-179). */
SGM_API int sgm_process(sgm_engine* e, const uint8_t* left, size_t left_stride, const uint8_t* right, size_t right_stride,
                const uint8_t* sky_l, size_t sky_l_stride, const uint8_t* sky_r, size_t sky_r_stride);

/* Asynchronous pair for the frame-stream runner: sgm_enqueue copies the inputs into the handle's pinned
 * staging, queues H2D + all kernels + D2H on the handle's stream and returns; sgm_wait blocks until done. */
SGM_API int sgm_enqueue(sgm_engine* e, const uint8_t* left, size_t left_stride, const uint8_t* right, size_t right_stride,
                const uint8_t* sky_l, size_t sky_l_stride, const uint8_t* sky_r, size_t sky_r_stride);
SGM_API int sgm_wait(sgm_engine* e);

/* Same pipeline with inputs already resident in device memory (no H2D, no D2H): queues the kernels on the
 * handle's stream and returns; follow with sgm_wait.  Pointers are device pointers on the handle's device. */
SGM_API int sgm_enqueue_device(sgm_engine* e, const void* left_device, size_t left_stride, const void* right_device,
                       size_t right_stride, const void* sky_l_device, size_t sky_l_stride, const void* sky_r_device,
                       size_t sky_r_stride);

/* Solver::get_disp() (inc/Solver.h:36): the final float disparity (invalid = D+1), (height/scale) x (width/scale). */
SGM_API int sgm_get_disparity_f32(sgm_engine* e, float* dst, size_t dst_stride);
/* Solver::disp / disp_beta narrowed to 8 bits like the reference's CV_8UC1 Mats (only for D <= 128). */
SGM_API int sgm_get_disparity_u8(sgm_engine* e, int view, uint8_t* dst, size_t dst_stride);
SGM_API int sgm_get_disparity_u16(sgm_engine* e, int view, uint16_t* dst, size_t dst_stride);
/* Solver::filtered_disp_beta as SGM::process leaves it (src/SGM.cpp:801): the right view's sub-pixel disparity. */
SGM_API int sgm_get_subpixel_beta(sgm_engine* e, float* dst, size_t dst_stride);
/* Device pointer of the final float disparity (valid until the next frame), for device-side consumers. */
SGM_API const float* sgm_disparity_device(sgm_engine* e);

/* Debug / parity taps: copies tap `which` of `view` into dst (bytes must equal sgm_tap_bytes). */
SGM_API size_t sgm_tap_bytes(const sgm_engine* e, int which);
SGM_API int sgm_get_tap(sgm_engine* e, int which, int view, void* dst, size_t bytes);

/* Introspection for bench.py: working size, kernel launches of the last frame, per-stage device times of the
 * last frame (needs SGM_FLAG_STAGE_TIMING), the CUDA stream (cudaStream_t) the kernels are launched on. */
SGM_API int sgm_get_size(const sgm_engine* e, int* out_h, int* out_w, int* out_d);
SGM_API int sgm_get_launch_count(const sgm_engine* e);
SGM_API int sgm_get_stage_ms(sgm_engine* e, float* ms, int count);
SGM_API void* sgm_get_stream(sgm_engine* e);
SGM_API size_t sgm_device_bytes(const sgm_engine* e);

/* Dense forms of the free functions of inc/cost.h:7-14 (SAD / census), computed on the device:
 * out is a host float volume [ih][iw][D], as Solver::build_dsi (src/Solver.cpp:96-117) would fill it. */
SGM_API int sgm_cost_volume_sad(int device, const uint8_t* left, const uint8_t* right, int height, int width, int max_disp,
                        int win_h, int win_w, int scale, float* out);
/* SSD (src/cost.cpp:33-59): sum of squared differences / win_h / win_w; windows up to 258 pixels (exact float accumulation). */
SGM_API int sgm_cost_volume_ssd(int device, const uint8_t* left, const uint8_t* right, int height, int width, int max_disp,
                        int win_h, int win_w, int scale, float* out);
/* CT (src/cost.cpp:62-96): per-disparity census + Hamming distance with the function's own conventions (right centre pixel at
 * x - d unscaled, taps with x_r < 0 dropped); windows up to 64 pixels. */
SGM_API int sgm_cost_volume_ct(int device, const uint8_t* left, const uint8_t* right, int height, int width, int max_disp,
                       int win_h, int win_w, int scale, float* out);

/* Solver::show_disp(debug_view) (src/Solver.cpp:55-93) with Solver::colormap (:652-707), rendered on the device from the
 * last frame: dst is a (2 * height/scale) x (width/scale) 8-bit BGR image (3 bytes per pixel, dst_stride bytes per row):
 * the working-resolution left image on top, the colour-mapped disparity from row height/scale - 1 on. */
SGM_API int sgm_render_debug_view(sgm_engine* e, uint8_t* dst_bgr, size_t dst_stride);

/* The disparity -> point cloud loop of node.cpp:119-143 on the device, for the last frame: every pixel, in raster order,
 * whose disparity is valid and whose depth Zc = (fx+fy)/2 * baseline / (d + 1e-6) is <= max_range yields
 * (Xc, Yc, Zc) = ((j*scale - cx) * Zc / fx, (i*scale - cy) * Zc / fy, Zc) and the grey value left(i, j).
 * xyz receives 3 doubles per point, intensity 1 byte per point (either may be NULL), at most `capacity` points;
 * *count receives the number of points found (it can exceed capacity: call again with larger buffers). */
typedef struct sgm_camera {
    float fx, fy, cx, cy;   /* CamIntrinsics of inc/utils.h:14-20 */
    double baseline;        /* node.cpp:118 (0.5) */
    float max_range;        /* node.cpp:117 (100) */
} sgm_camera;
SGM_API int sgm_point_cloud(sgm_engine* e, const sgm_camera* cam, double* xyz, uint8_t* intensity, size_t capacity, size_t* count);

/* ---- Frame-stream runner ------------------------------------------------------------------------------------------
 * The reference's caller (node.cpp:49,93,104) owns ONE Solver and calls process() frame after frame; a Solver keeps no
 * state between frames (scratch members only), so a stream of stereo pairs shards by frame.  sgm_stream is that loop for
 * 1..N GPUs of one box behind one handle: frame k goes to device k mod N, every device keeps `inflight` frames in flight
 * (one engine handle each: own stream, own pinned staging for inputs and result), one host thread per device issues the
 * launches, and results come back in submission order.  No collective and no peer traffic: inputs and results cross PCIe.
 *
 *   sgm_stream_submit  copies the pair (and masks, nullable) into pinned staging and returns; it blocks only while the slot
 *                      the frame maps to is still occupied (back-pressure: at most n_devices * inflight frames occupy slots,
 *                      so a producer whose consumer has stopped stalls there).  *frame_index (nullable) = k.
 *   sgm_stream_next    blocks until the OLDEST pending frame is done and hands out its final disparity (Solver::get_disp():
 *                      float32, (height/scale) x (width/scale), invalid = D+1): copied into dst when dst != NULL, and/or as
 *                      a pointer into the pinned result buffer (*view, dense rows).  A view keeps its slot occupied until
 *                      sgm_stream_release or the next sgm_stream_next on this stream: it is never overwritten while it can
 *                      still be read.  Returns SGM_ERR_STATE when no frame is pending (it never waits for a submit).
 *   sgm_stream_pending frames that occupy a slot: submitted and not yet handed out, plus a view not yet released.  A caller
 *                      that submits and collects on ONE thread calls sgm_stream_next whenever this equals
 *                      n_devices * inflight, and so never waits for itself.
 * One producer thread and one consumer thread may call submit and next / release concurrently; two producers (or two consumers)
 * may not.  A frame that fails reports its error from sgm_stream_next (message: sgm_stream_last_error); the stream stays usable. */
typedef struct sgm_stream sgm_stream; /* opaque */
SGM_API int sgm_stream_create(sgm_stream** out, const sgm_params* params, const int* devices, int n_devices, int inflight);
SGM_API void sgm_stream_destroy(sgm_stream* s);
SGM_API int sgm_stream_submit(sgm_stream* s, const uint8_t* left, size_t left_stride, const uint8_t* right, size_t right_stride,
                              const uint8_t* sky_l, size_t sky_l_stride, const uint8_t* sky_r, size_t sky_r_stride,
                              uint64_t* frame_index);
SGM_API int sgm_stream_next(sgm_stream* s, float* dst, size_t dst_stride, const float** view, uint64_t* frame_index);
SGM_API int sgm_stream_release(sgm_stream* s);
SGM_API int sgm_stream_pending(const sgm_stream* s);
/* Kernel launches issued so far by all devices of the stream (bench.py's gpu_launches). */
SGM_API uint64_t sgm_stream_launches(const sgm_stream* s);
/* With s == NULL: message of the last failed sgm_stream_create on this thread. */
SGM_API const char* sgm_stream_last_error(const sgm_stream* s);

/* No counterpart in the reference: added for callers that must not allocate (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The speckle filter of Solver::post_filter
 * (src/Solver.cpp:514-566 as called at :634: 4-connected components under |a - b| < max_dis; components of at most max_size
 * pixels become invalid_value) on a caller-owned DEVICE image, queued on `stream` (a cudaStream_t; NULL = the default stream)
 * of `device`; the call returns once the four kernels are queued.  in and out are height x width dense float32 device
 * buffers and may be the same buffer.  The label / area scratch is a static 16 MiB __device__ array inside the library
 * (held by every GPU that loads it), so the call performs no cudaMalloc / cudaFree / allocating library call; images of more
 * than SGM_SPECKLE_STATIC_MAX_PIXELS pixels are rejected with SGM_ERR_UNSUPPORTED.  All calls on a device SHARE that array:
 * calls that overlap on the GPU (different streams, different threads) race and give wrong results — keep them on one stream
 * or order them with events.  Nothing has to be cleared between calls (the first kernel rewrites what the later ones read).
 * The first call per device looks the array's address up (cudaGetSymbolAddress): make it outside graph capture.
 * Same result, bit for bit, as the speckle stage of sgm_process (SGM_TAP_MEDIAN -> SGM_TAP_FINAL). */
#define SGM_SPECKLE_STATIC_MAX_PIXELS (1 << 21)
SGM_API int sgm_speckle_filter_static(int device, const float* in_device, float* out_device, int height, int width, float invalid_value,
                                      int max_size, float max_dis, void* stream);

SGM_API int sgm_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* SGM_B200_H_ */
