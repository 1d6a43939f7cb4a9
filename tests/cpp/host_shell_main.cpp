// host_shell_main.cpp — drives the C++ Solver / SGM / BM shells exactly the way the reference's node.cpp does
// (node.cpp:49 make_shared<SGM>(h, w, scale, max_disp) into a SolverPtr; :93 process(l, r, sky, sky_beta);
// :104 get_disp(); :107 show_disp(debug_view)) and dumps the results for tests/test_host_shell.py.
//
//   host_shell_main <sgm|bm> H W SCALE D in_left.raw in_right.raw [sky_l.raw sky_r.raw|- -] out_prefix
//   host_shell_main bench <sgm|bm> H W SCALE D in_left.raw in_right.raw FRAMES     (bench.py's `blocking` leg: one JSON line)
//   host_shell_main stream H W SCALE D N_GPUS INFLIGHT FRAMES PAIR_DIR N_PAIRS     (BASELINE config 5: a frame stream over N GPUs
//        from ONE process through the C runner sgm_stream_*; PAIR_DIR holds l<k>.raw / r<k>.raw; one JSON line on stderr)
// TEST INFRASTRUCTURE.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
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

// The protected surface of inc/Solver.h:38-69, reached the way a subclass of the reference would reach it.
struct ProbeSGM : public SGM {
    ProbeSGM(int h, int w, int s, int d) : SGM(h, w, s, d) {}
    // census tables, the filtered left / right cost volumes and the sub-pixel images of the last frame, through the helpers
    bool dump(const std::string& out) {
        build_cost_table();
        build_dsi_from_table();
        cost_horizontal_filter(COST_WIN_W / scale);
        cost_vertical_filter(COST_WIN_H / scale);
        const size_t n = (size_t)img_h * img_w, N = n * max_disp;
        FILE* f = std::fopen((out + "_ctl.u64").c_str(), "wb");
        std::fwrite(cost_table_l, sizeof(uint64_t), n, f);
        std::fclose(f);
        f = std::fopen((out + "_cost_l.f32").c_str(), "wb");
        std::fwrite(cost, sizeof(float), N, f);
        std::fclose(f);
        const float mean_l = find_dsi_mean_max();
        build_dsi_from_table_beta();
        f = std::fopen((out + "_cost_r.f32").c_str(), "wb");
        std::fwrite(cost, sizeof(float), N, f);
        std::fclose(f);
        Mat sub_r;
        compute_subpixel(disp_beta, sub_r);
        write_raw(out + "_sub_r.f32", sub_r, sizeof(float));
        write_raw(out + "_disp8.u8", disp, 1);
        write_raw(out + "_fbeta.f32", filtered_disp_beta, sizeof(float));
        post_filter();
        std::printf("probe mean_l %f weight %p\n", mean_l, (void*)weight);
        return weight == nullptr && cost != nullptr;
    }
};

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

// FNV-1a 64 of a float image, the hash the golden vectors use.
static unsigned long long fnv1a64(const void* data, size_t bytes) {
    const unsigned char* p = static_cast<const unsigned char*>(data);
    unsigned long long h = 1469598103934665603ull;
    for (size_t k = 0; k < bytes; ++k) h = (h ^ p[k]) * 1099511628211ull;
    return h;
}

// The reference's frame loop (node.cpp:49,93,104) over a stream of pairs and N GPUs: a producer thread submits, this thread
// collects in order (zero-copy views of the pinned results), the first N_PAIRS results are hashed for the parity check.
static int stream_main(int argc, char** argv) {
    if (argc != 11) return 2;
    const int h = std::atoi(argv[2]), w = std::atoi(argv[3]), s = std::atoi(argv[4]), d = std::atoi(argv[5]);
    const int n_gpus = std::atoi(argv[6]), inflight = std::atoi(argv[7]), frames = std::atoi(argv[8]), n_pairs = std::atoi(argv[10]);
    const std::string dir = argv[9];
    if (n_gpus < 1 || inflight < 1 || frames < 1 || n_pairs < 1) return 2;
    std::vector<Mat> L, R;
    for (int k = 0; k < n_pairs; ++k) {
        L.emplace_back(h, w, CV_8UC1);
        R.emplace_back(h, w, CV_8UC1);
        if (!read_raw((dir + "/l" + std::to_string(k) + ".raw").c_str(), L[k]) || !read_raw((dir + "/r" + std::to_string(k) + ".raw").c_str(), R[k])) return 3;
    }
    sgm_params p;
    sgm_default_params(&p, h, w, s, d, SGM_ALGO_SGM);
    sgm_stream* st = nullptr;
    if (sgm_stream_create(&st, &p, nullptr, n_gpus, inflight) != SGM_OK) {
        std::fprintf(stderr, "sgm_stream_create: %s\n", sgm_stream_last_error(nullptr));
        return 4;
    }
    const int ih = h / s, iw = w / s;
    uint64_t next_index = 0;  // frame indices are counted per stream, across the two runs below
    auto run = [&](int count, std::vector<unsigned long long>* hashes) -> bool {
        std::atomic<bool> ok(true), producing(true);
        std::thread producer([&] {
            for (int k = 0; k < count && ok; ++k)
                if (sgm_stream_submit(st, L[k % n_pairs].ptr<uchar>(0), L[k % n_pairs].step, R[k % n_pairs].ptr<uchar>(0), R[k % n_pairs].step,
                                      nullptr, 0, nullptr, 0, nullptr) != SGM_OK)
                    ok = false;
            producing = false;
        });
        int got = 0;
        // collect until the producer is done and nothing is pending: also after a failure, so that a producer waiting for a
        // slot is never left alone (it would wait for ever)
        while (producing || sgm_stream_pending(st) > 0) {
            if (sgm_stream_pending(st) == 0) {
                std::this_thread::yield();  // the producer has not got there yet
                continue;
            }
            const float* view = nullptr;
            uint64_t idx = 0;
            if (sgm_stream_next(st, nullptr, 0, &view, &idx) != SGM_OK || idx != next_index) ok = false;
            ++next_index;
            if (ok && hashes && got < n_pairs) hashes->push_back(fnv1a64(view, (size_t)ih * iw * sizeof(float)));
            sgm_stream_release(st);  // done with the view: its slot may take a new frame
            ++got;
        }
        producer.join();
        return ok && got == count;
    };
    std::vector<unsigned long long> hashes;
    if (!run(std::max(n_pairs, 4 * n_gpus * inflight), &hashes)) {  // warm-up + parity frames
        std::fprintf(stderr, "stream failed: %s\n", sgm_stream_last_error(st));
        return 5;
    }
    const double t0 = get_cur_ms();
    if (!run(frames, nullptr)) {
        std::fprintf(stderr, "stream failed: %s\n", sgm_stream_last_error(st));
        return 5;
    }
    const double ms = get_cur_ms() - t0;
    std::fprintf(stderr, "{\"frames\": %d, \"n_gpus\": %d, \"inflight\": %d, \"wall_ms\": %.2f, \"frames_per_s\": %.2f, \"launches\": %llu, \"first_hashes\": [",
                 frames, n_gpus, inflight, ms, 1e3 * frames / ms, (unsigned long long)sgm_stream_launches(st));
    for (size_t k = 0; k < hashes.size(); ++k) std::fprintf(stderr, "%s\"%016llx\"", k ? ", " : "", hashes[k]);
    std::fprintf(stderr, "]}\n");
    sgm_stream_destroy(st);
    return 0;
}

int main(int argc, char** argv) {
    if (argc >= 2 && std::strcmp(argv[1], "bench") == 0) return bench_main(argc, argv);
    if (argc >= 2 && std::strcmp(argv[1], "stream") == 0) return stream_main(argc, argv);
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
    std::shared_ptr<ProbeSGM> probe;
    if (bm)
        sv = std::make_shared<BM>(h, w, s, d);
    else
        sv = probe = std::make_shared<ProbeSGM>(h, w, s, d);
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
    if (probe && !probe->dump(out)) return 6;
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
