// Solver.cpp — host shell of the Solver base class over the device engine (reference: src/Solver.cpp).
//
// Everything the reference's Solver computes on the CPU (census tables, Hamming volume, recursive cost filters,
// sub-pixel fit, hole-filling median, speckle filter: src/Solver.cpp:120-635) happens inside sgm_process(); this file
// only validates, moves images in and disparities out, and draws the debug view.
#include "../inc/Solver.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace {

// The reference reports misuse through assert() (src/Solver.cpp:6-10, src/SGM.cpp:34-38), i.e. message + abort.
// Engine errors are reported the same way: loudly, never by falling back to another path.
void die(const char *what, const sgm_engine *e)
{
    std::fprintf(stderr, "stereo_matching_b200: %s: %s\n", what, sgm_last_error(e));
    std::abort();
}

}  // namespace

Solver::Solver(int h, int w, int s, int d)
    : img_h(0), img_w(0), scale(s), max_disp(d), invalid_disp(d + 1), cost_table_l(nullptr), cost_table_r(nullptr), cost(nullptr),
      weight(nullptr), engine(nullptr), two_views(false)
{
    // same domain as the reference (src/Solver.cpp:6-10); d == 256 is the engine's widened mode (uint16 disparities
    // inside, the CV_8UC1 `disp` Mats then saturate, get_disp() stays exact)
    if (!(h > 0 && w > 0 && s > 0 && d > 0) || !(s == 1 || s == 2) || !(d == 32 || d == 64 || d == 128 || d == 256)) {
        std::fprintf(stderr, "stereo_matching_b200: Solver(h=%d, w=%d, s=%d, d=%d): need h, w > 0, s in {1, 2}, d in {32, 64, 128, 256}\n", h, w, s, d);
        std::abort();
    }
    img_h = h / s;
    img_w = w / s;
    disp.create(img_h, img_w, CV_8UC1);
    disp_beta.create(img_h, img_w, CV_8UC1);
    filtered_disp.create(img_h, img_w, CV_32FC1);
    filtered_disp_beta.create(img_h, img_w, CV_32FC1);
    colored_disp.create(img_h, img_w, CV_8UC3);
#ifdef SGM_B200_HAVE_LKREFINE
    lkptr = LKSubPixel::create(h, w, s, d);  // src/Solver.cpp:51; the refinement itself is never invoked (src/SGM.cpp:823-824)
#endif
}

Solver::~Solver()
{
    sgm_destroy(engine);
    delete[] cost_table_l;
    delete[] cost_table_r;
    delete[] cost;
}

void Solver::open_engine(int algo, int paths, int full_h, int full_w)
{
    sgm_params p;
    sgm_default_params(&p, full_h, full_w, scale, max_disp, algo);
    p.paths = paths;
    p.census_h = WIN_H;
    p.census_w = WIN_W;
    p.box_h = COST_WIN_H;
    p.box_w = COST_WIN_W;
    p.unique_ratio = UNIQUE_RATIO;
    p.lr_thresh = LR_CHECK_DIS;
    p.median_h = MEDIAN_FILTER_H;
    p.median_w = MEDIAN_FILTER_W;
    p.speckle_size = SPECKLE_SIZE;
    p.speckle_dis = SPECKLE_DIS;
    if (const char *dev = std::getenv("SGM_B200_DEVICE")) p.device = std::atoi(dev);
    // Opt-in integer / DPX mode of the engine (include/sgm_b200.h: SGM_FLAG_INTEGER): SGM without the cost filter
    // (COST_WIN_W = COST_WIN_H = 1 in the reference's terms) on 16-bit volumes.  Sky masks are accepted like in the float
    // mode (masked pixels cost 2 * P2 instead of 999999, with identical results: include/sgm_b200.h).
    if (algo == SGM_ALGO_SGM && std::getenv("SGM_B200_INTEGER")) {
        p.box_w = p.box_h = scale;
        p.flags |= SGM_FLAG_INTEGER;
    }
    p.flags |= SGM_FLAG_HOST_TAPS;  // disp / disp_beta / filtered_disp_beta follow the result to pinned host memory
    two_views = algo != SGM_ALGO_BM;
    if (sgm_create(&engine, &p) != SGM_OK) die("cannot create the device pipeline (no CPU path exists)", nullptr);
}

void Solver::run_engine(Mat &l, Mat &r)
{
    // src/SGM.cpp:34-38 / src/BM.cpp:11-15
    if (l.rows != r.rows || l.cols != r.cols || l.type() != CV_8UC1 || r.type() != CV_8UC1 || l.rows / scale != img_h ||
        l.cols / scale != img_w) {
        std::fprintf(stderr, "stereo_matching_b200: process(): images must be CV_8UC1, equal size, %d x %d at scale %d\n",
                     img_w * scale, img_h * scale, scale);
        std::abort();
    }
    img_l = l;  // header copies: the debug view reads the left image later (src/SGM.cpp:59-60)
    img_r = r;
    // src/Solver.cpp:146-152 prints per view; a mask that is present must be a CV_8UC1 image of the working size (the
    // reference would index it as one): anything else is misuse and ends like the other argument checks, loudly
    const Mat *masks[2] = {&sky_mask, &sky_mask_beta};
    bool have[2];
    for (int v = 0; v < 2; ++v) {
        have[v] = !masks[v]->empty();
        if (have[v] && (masks[v]->rows != img_h || masks[v]->cols != img_w || masks[v]->type() != CV_8UC1)) {
            std::fprintf(stderr, "stereo_matching_b200: process(): sky mask %d must be CV_8UC1, %d x %d (got %d x %d)\n", v, img_w, img_h,
                         masks[v]->cols, masks[v]->rows);
            std::abort();
        }
        std::printf(have[v] ? "sky mask found\n" : "no sky mask\n");
    }
    if (sgm_process(engine, l.ptr<uchar>(0), l.step, r.ptr<uchar>(0), r.step, have[0] ? sky_mask.ptr<uchar>(0) : nullptr,
                    have[0] ? (size_t)sky_mask.step : 0, have[1] ? sky_mask_beta.ptr<uchar>(0) : nullptr,
                    have[1] ? (size_t)sky_mask_beta.step : 0) != SGM_OK)
        die("process() failed", engine);

    // Results: all of them reached pinned host memory on the frame's stream (SGM_FLAG_HOST_TAPS), the 8-bit disparities
    // already narrowed on the device; what is left is one row-wise copy each into the Mats (which may be pitched).
    if (sgm_get_disparity_f32(engine, filtered_disp.ptr<float>(0), filtered_disp.step) != SGM_OK) die("get_disp", engine);
    if (sgm_get_disparity_u8(engine, 0, disp.ptr<uchar>(0), disp.step) != SGM_OK) die("disp", engine);
    if (two_views) {
        if (sgm_get_disparity_u8(engine, 1, disp_beta.ptr<uchar>(0), disp_beta.step) != SGM_OK) die("disp_beta", engine);
        // right-view sub-pixel disparity, which the reference leaves in filtered_disp_beta (src/SGM.cpp:801)
        if (sgm_get_subpixel_beta(engine, filtered_disp_beta.ptr<float>(0), filtered_disp_beta.step) != SGM_OK) die("filtered_disp_beta", engine);
    }
    // the masks stay set until the next 4-argument process(), as in the reference (src/SGM.cpp:831-833)
}

// ---- the reference's protected stage helpers as refreshers of the host mirrors (see inc/Solver.h) -------------------------
void Solver::build_cost_table()
{
    const size_t n = (size_t)img_h * img_w;
    if (!cost_table_l) cost_table_l = new uint64_t[n];
    if (!cost_table_r) cost_table_r = new uint64_t[n];
    if (sgm_get_tap(engine, SGM_TAP_CENSUS, 0, cost_table_l, n * sizeof(uint64_t)) != SGM_OK) die("build_cost_table", engine);
    if (sgm_get_tap(engine, SGM_TAP_CENSUS, 1, cost_table_r, n * sizeof(uint64_t)) != SGM_OK) die("build_cost_table", engine);
}

static void fetch_volume(sgm_engine *engine, float *&cost, int view, size_t count)
{
    if (!cost) cost = new float[count];
    if (sgm_get_tap(engine, SGM_TAP_COST, view, cost, count * sizeof(float)) != SGM_OK) die("build_dsi", engine);
}

void Solver::build_dsi() { fetch_volume(engine, cost, 0, (size_t)img_h * img_w * max_disp); }
void Solver::build_dsi_from_table() { fetch_volume(engine, cost, 0, (size_t)img_h * img_w * max_disp); }
void Solver::build_dsi_from_table_beta() { fetch_volume(engine, cost, 1, (size_t)img_h * img_w * max_disp); }
void Solver::cost_horizontal_filter(int) {}  // fused into the volume kernel on the device (src/Solver.cpp:296-330)
void Solver::cost_vertical_filter(int) {}    // done on the device (src/Solver.cpp:333-368)
void Solver::post_filter() {}                // done on the device (src/Solver.cpp:600-635); filtered_disp holds the result

float Solver::find_dsi_mean_max()
{
    if (!cost) build_dsi();
    double max_cost = 0, mean_cost = 0;
    const size_t count = (size_t)img_h * img_w * max_disp;
    for (size_t k = 0; k < count; ++k) {
        mean_cost += cost[k];
        if (cost[k] > max_cost) max_cost = cost[k];
    }
    mean_cost /= (double)count;
    std::printf("max_cost: %lf, mean_cost: %lf\n", max_cost, mean_cost);
    return (float)mean_cost;
}

float Solver::find_table_mean_max()
{
    if (!cost_table_r) build_cost_table();
    double max_cost = 0, mean_cost = 0;
    const size_t n = (size_t)img_h * img_w;
    for (size_t k = 0; k < n; ++k) {
        mean_cost += (double)cost_table_r[k];
        if ((double)cost_table_r[k] > max_cost) max_cost = (double)cost_table_r[k];
    }
    mean_cost /= (double)(n * max_disp);  // the reference divides by h * w * max_disp here too (src/Solver.cpp:290)
    std::printf("max_cost: %lf, mean_cost: %lf\n", max_cost, mean_cost);
    return (float)mean_cost;
}

void Solver::fetch_cost(float *p)
{
    const size_t count = (size_t)img_h * img_w * max_disp;
    if (!cost) cost = new float[count];
    std::memcpy(cost, p, count * sizeof(float));  // src/Solver.cpp:371-374
}

// src/Solver.cpp:569-597 runs inside the device pipeline; this returns its result for the disparity image given: the left
// view's for `disp`, the right view's for `disp_beta` (any other Mat: the left view's).
void Solver::compute_subpixel(const Mat &d, Mat &out)
{
    const int view = (two_views && &d == &disp_beta) ? 1 : 0;
    if (out.rows != img_h || out.cols != img_w || out.type() != CV_32FC1) out.create(img_h, img_w, CV_32FC1);
    std::vector<float> tmp((size_t)img_h * img_w);
    if (sgm_get_tap(engine, SGM_TAP_SUBPIXEL, view, tmp.data(), tmp.size() * sizeof(float)) != SGM_OK) die("compute_subpixel", engine);
    for (int i = 0; i < img_h; ++i) std::memcpy(out.ptr<float>(i), tmp.data() + (size_t)i * img_w, (size_t)img_w * sizeof(float));
}

void Solver::fetch_disparity(uchar *d)
{
    for (int i = 0; i < disp.rows; ++i) std::memcpy(disp.ptr<uchar>(i), d + (size_t)i * disp.cols, disp.cols);
}

void Solver::fetch_disparity(float *d)
{
    for (int i = 0; i < filtered_disp.rows; ++i)
        std::memcpy(filtered_disp.ptr<float>(i), d + (size_t)i * filtered_disp.cols, filtered_disp.cols * sizeof(float));
}

// Colour map of filtered_disp (blue -> cyan -> green -> yellow -> red over the disparity range, invalid = black;
// src/Solver.cpp:652-707).  Rendered on the device by sgm_render_debug_view; this keeps the reference's protected helper.
void Solver::colormap()
{
    Mat view = Mat::zeros(img_h * 2, img_w, CV_8UC3);
    if (sgm_render_debug_view(engine, view.ptr<uchar>(0), view.step) != SGM_OK) die("colormap", engine);
    Mat lower = view(Rect(0, img_h - 1, img_w, img_h));
    lower.copyTo(colored_disp);
}

// src/Solver.cpp:55-93: left image on top, colour-mapped disparity below (the lower block starts at row img_h - 1).
// Composed by one device kernel from the engine's working-resolution left image and the final disparity.
void Solver::show_disp(Mat &debug_view)
{
    debug_view = Mat::zeros(img_h * 2, img_w, CV_8UC3);
    if (sgm_render_debug_view(engine, debug_view.ptr<uchar>(0), debug_view.step) != SGM_OK) die("show_disp", engine);
}

// node.cpp:119-143 as one call: camera-frame points (raster order) and their grey values, computed on the device.
size_t Solver::get_point_cloud(const CamIntrinsics &cam, double baseline, float max_range, std::vector<cv::Point3d> &pts,
                               std::vector<uchar> &pixel)
{
    sgm_camera c;
    c.fx = cam.fx;
    c.fy = cam.fy;
    c.cx = cam.cx;
    c.cy = cam.cy;
    c.baseline = baseline;
    c.max_range = max_range;
    const size_t cap = (size_t)img_h * img_w;
    std::vector<double> xyz(cap * 3);
    pixel.resize(cap);
    size_t n = 0;
    if (sgm_point_cloud(engine, &c, xyz.data(), pixel.data(), cap, &n) != SGM_OK) die("get_point_cloud", engine);
    pixel.resize(n);
    pts.resize(n);
    for (size_t k = 0; k < n; ++k) pts[k] = cv::Point3d(xyz[3 * k], xyz[3 * k + 1], xyz[3 * k + 2]);
    return n;
}
