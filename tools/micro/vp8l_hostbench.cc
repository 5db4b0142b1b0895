#include "host_vp8l.h"
#include <chrono>
#include <cstdio>
#include <vector>
using namespace webpb200;
int main(int argc, char** argv) {
    for (int a = 1; a < argc; a++) {
        FILE* f = fopen(argv[a], "rb"); std::vector<uint8_t> d(1 << 24); size_t n = fread(d.data(), 1, d.size(), f); fclose(f);
        // ALPH payload: 1 header byte then the VP8L stream (headerless)
        std::vector<uint32_t> argb(1920 * 1080 + 8);
        double best = 1e9; int st = 0;
        for (int r = 0; r < 5; r++) {
            Vp8lPlan plan;
            auto t0 = std::chrono::steady_clock::now();
            st = vp8l_decode_entropy(d.data() + 1, n - 1, 1920, 1080, argb.data(), &plan);
            double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
            if (ms < best) best = ms;
        }
        printf("%s: %zu bytes, status %d, entropy decode %.2f ms (%.0f MPix/s)\n", argv[a], n, st, best, 1920 * 1080 / best / 1e3);
    }
}
