// cost_probe_main.cpp — calls the per-pixel probes of inc/cost.h (SAD, CT_pts, hamming_cost) for every pixel of a small
// pair and dumps the results for tests/test_host_shell.py.   cost_probe_main H W D left.raw right.raw out_prefix
// TEST INFRASTRUCTURE.
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "inc/cost.h"

double get_cur_ms() { return 0.0; }

int main(int argc, char** argv) {
    if (argc != 7) return 2;
    const int h = std::atoi(argv[1]), w = std::atoi(argv[2]), d = std::atoi(argv[3]);
    Mat l(h, w, CV_8UC1), r(h, w, CV_8UC1);
    FILE* f = std::fopen(argv[4], "rb");
    if (!f || std::fread(l.data, 1, (size_t)h * w, f) != (size_t)h * w) return 3;
    std::fclose(f);
    f = std::fopen(argv[5], "rb");
    if (!f || std::fread(r.data, 1, (size_t)h * w, f) != (size_t)h * w) return 3;
    std::fclose(f);
    std::vector<float> sad((size_t)h * w * d);
    std::vector<uint64_t> ctl((size_t)h * w), ctr((size_t)h * w);
    std::vector<int> ham((size_t)h * w);
    for (int i = 0; i < h; ++i)
        for (int j = 0; j < w; ++j) {
            for (int k = 0; k < d; ++k) sad[((size_t)i * w + j) * d + k] = SAD(l, r, Point(j, i), k, 9, 9, nullptr);
            CT_pts(l, r, j, i, WIN_H, WIN_W, nullptr, ctl.data(), ctr.data());
            ham[(size_t)i * w + j] = hamming_cost(ctl[(size_t)i * w + j], ctr[(size_t)i * w + j]);
        }
    const std::string out = argv[6];
    auto dump = [&](const std::string& name, const void* p, size_t bytes) {
        FILE* o = std::fopen((out + name).c_str(), "wb");
        std::fwrite(p, 1, bytes, o);
        std::fclose(o);
    };
    dump("_sad.f32", sad.data(), sad.size() * 4);
    dump("_ctl.u64", ctl.data(), ctl.size() * 8);
    dump("_ctr.u64", ctr.data(), ctr.size() * 8);
    dump("_ham.i32", ham.data(), ham.size() * 4);
    return 0;
}
