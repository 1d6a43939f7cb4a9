// host_shell_main.cpp — drives the C++ Solver / SGM / BM shells exactly the way the reference's node.cpp does
// (node.cpp:49 make_shared<SGM>(h, w, scale, max_disp) into a SolverPtr; :93 process(l, r, sky, sky_beta);
// :104 get_disp(); :107 show_disp(debug_view)) and dumps the results for tests/test_host_shell.py.
//
//   host_shell_main <sgm|bm> H W SCALE D in_left.raw in_right.raw [sky_l.raw sky_r.raw|- -] out_prefix
//   host_shell_main bench <sgm|bm> H W SCALE D in_left.raw in_right.raw FRAMES     (bench.py's `blocking` leg: one JSON line)
// TEST INFRASTRUCTURE.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "inc/BM.h"
#include "inc/SGM.h"

double get_cur_ms() {
    using namespace std::chrono;
    return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}

static bool read_raw(const char* path, Mat& m) {
    FILE* f = std::fopen(path, "rb");
    if (!f) return false;
    bool ok = true;
    for (int i = 0; i < m.rows && ok; ++i) ok = std::fread(m.ptr<uchar>(i), 1, m.cols, f) == (size_t)m.cols;
    std::fclose(f);
    return ok;
}

static void write_raw(const std::string& path, const Mat& m, size_t elem) {
    FILE* f = std::fopen(path.c_str(), "wb");
    for (int i = 0; i < m.rows; ++i) std::fwrite(m.ptr<uchar>(i), elem, m.cols, f);
    std::fclose(f);
}

// Blocking frame loop exactly as node.cpp:93,104 runs it: process(l, r) then get_disp(), one frame at a time.
static int bench_main(int argc, char** argv) {
    if (argc != 10) return 2;
    const bool bm = std::strcmp(argv[2], "bm") == 0;
    const int h = std::atoi(argv[3]), w = std::atoi(argv[4]), s = std::atoi(argv[5]), d = std::atoi(argv[6]);
    const int frames = std::atoi(argv[9]);
    Mat img_l(h, w, CV_8UC1), img_r(h, w, CV_8UC1);
    if (!read_raw(argv[7], img_l) || !read_raw(argv[8], img_r) || frames < 1) return 3;
    SolverPtr sv;
    if (bm)
        sv = std::make_shared<BM>(h, w, s, d);
    else
        sv = std::make_shared<SGM>(h, w, s, d);
    FILE* quiet = std::freopen("/dev/null", "w", stdout);  // the shells print "no sky mask" per frame like the reference
    (void)quiet;
    std::vector<double> ms((size_t)frames);
    double checksum = 0.0;
    for (int k = 0; k < 5; ++k) sv->process(img_l, img_r);  // warm-up
    const double t_begin = get_cur_ms();
    for (int k = 0; k < frames; ++k) {
        const double t0 = get_cur_ms();
        sv->process(img_l, img_r);
        const Mat& f = sv->get_disp();
        checksum += f.ptr<float>(h / s / 2)[w / s / 2];
        ms[(size_t)k] = get_cur_ms() - t0;
    }
    const double total = get_cur_ms() - t_begin;
    std::sort(ms.begin(), ms.end());
    std::fprintf(stderr, "{\"blocking_fps\": %.2f, \"latency_ms_median\": %.4f, \"latency_ms_p90\": %.4f, \"frames\": %d, \"checksum\": %.3f}\n",
                 1e3 * frames / total, ms[(size_t)frames / 2], ms[(size_t)(frames * 9) / 10], frames, checksum);
    return 0;
}

int main(int argc, char** argv) {
    if (argc >= 2 && std::strcmp(argv[1], "bench") == 0) return bench_main(argc, argv);
    if (argc != 11) {
        std::fprintf(stderr, "usage: %s <sgm|bm> H W SCALE D left.raw right.raw sky_l.raw|- sky_r.raw|- out_prefix\n", argv[0]);
        return 2;
    }
    const bool bm = std::strcmp(argv[1], "bm") == 0;
    const int h = std::atoi(argv[2]), w = std::atoi(argv[3]), s = std::atoi(argv[4]), d = std::atoi(argv[5]);
    // a pitched left image (extra bytes per row) exercises Mat::step like a ROI of a larger image would
    Mat big_l(h, w + 24, CV_8UC1), img_r(h, w, CV_8UC1);
    Mat img_l = big_l(Rect(0, 0, w, h));
    if (!read_raw(argv[6], img_l) || !read_raw(argv[7], img_r)) return 3;
    const bool have_sky = std::strcmp(argv[8], "-") != 0;
    Mat sky_l, sky_r;
    if (have_sky) {
        sky_l.create(h / s, w / s, CV_8UC1);
        sky_r.create(h / s, w / s, CV_8UC1);
        if (!read_raw(argv[8], sky_l) || !read_raw(argv[9], sky_r)) return 3;
    }
    const std::string out = argv[10];

    SolverPtr sv;
    if (bm)
        sv = std::make_shared<BM>(h, w, s, d);
    else
        sv = std::make_shared<SGM>(h, w, s, d);
    const double t0 = get_cur_ms();
    if (have_sky)
        sv->process(img_l, img_r, sky_l, sky_r);
    else
        sv->process(img_l, img_r);
    const double t1 = get_cur_ms();
    Mat disp = sv->get_disp();
    Mat debug_view;
    sv->show_disp(debug_view);
    std::printf("time cost: %lf ms\n", t1 - t0);
    std::printf("disp %d x %d type %d, debug view %d x %d channels %d\n", disp.rows, disp.cols, disp.type(), debug_view.rows,
                debug_view.cols, debug_view.channels());
    write_raw(out + "_disp.f32", disp, sizeof(float));
    write_raw(out + "_debug.bgr", debug_view, 3);
    // second frame on the same solver (node.cpp would loop over frames): must give the same result
    if (have_sky)
        sv->process(img_l, img_r, sky_l, sky_r);
    else
        sv->process(img_l, img_r);
    write_raw(out + "_disp2.f32", sv->get_disp(), sizeof(float));
    // node.cpp:114-143 as one call (KITTI-like intrinsics)
    CamIntrinsics cam = {718.856f, 718.856f, 607.1928f, 185.2157f};
    std::vector<cv::Point3d> pts;
    std::vector<uchar> pix;
    const size_t n = sv->get_point_cloud(cam, 0.5, 100.f, pts, pix);
    std::printf("pointcloud size: %zu, %zu\n", pts.size(), pix.size());
    FILE* f = std::fopen((out + "_cloud.f64").c_str(), "wb");
    for (size_t k = 0; k < n; ++k) {
        const double v[3] = {pts[k].x, pts[k].y, pts[k].z};
        std::fwrite(v, sizeof(double), 3, f);
    }
    std::fclose(f);
    return 0;
}
