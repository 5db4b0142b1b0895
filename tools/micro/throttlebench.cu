// throttlebench.cu -- does the SM issue rate dip in short windows under a steady integer load?
// Each block's warp 0 repeats a fixed dependent-IMAD chain (about 1 us) and stamps %globaltimer and clock64 around
// every repetition; the host prints, per configuration, the distribution of repetition times.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o throttlebench throttlebench.cu ; ./throttlebench
#include <cstdio>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>

constexpr int kReps = 3000;

__global__ void k_probe(long long* ns, long long* cyc, int chain, int* sink) {
    const bool rec = (threadIdx.x >> 5) == 0;
    int v = threadIdx.x + 1;
    for (int r = 0; r < kReps; r++) {
        long long t0, c0 = clock64();
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
        for (int i = 0; i < chain; i++) v = v * 1664525 + 1013904223;
        long long t1, c1 = clock64();
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
        if (rec && (threadIdx.x & 31) == 0) {
            ns[(size_t)blockIdx.x * kReps + r] = t1 - t0;
            cyc[(size_t)blockIdx.x * kReps + r] = c1 - c0;
        }
    }
    if (v == 12345) *sink = v;
}

int main() {
    long long *d_ns, *d_cyc;
    int* d_sink;
    const int blocks = 148;
    cudaMalloc(&d_ns, sizeof(long long) * blocks * kReps);
    cudaMalloc(&d_cyc, sizeof(long long) * blocks * kReps);
    cudaMalloc(&d_sink, 4);
    std::vector<long long> ns(blocks * kReps), cyc(blocks * kReps);
    for (int threads : {32, 256, 1024}) {
        for (int pass = 0; pass < 3; pass++) {
            k_probe<<<blocks, threads>>>(d_ns, d_cyc, 400, d_sink);
            cudaDeviceSynchronize();
        }
        cudaMemcpy(ns.data(), d_ns, sizeof(long long) * ns.size(), cudaMemcpyDeviceToHost);
        cudaMemcpy(cyc.data(), d_cyc, sizeof(long long) * cyc.size(), cudaMemcpyDeviceToHost);
        std::vector<long long> s(ns);
        std::sort(s.begin(), s.end());
        const long long med = s[s.size() / 2];
        size_t slow = 0;
        long long slow_ns = 0, total = 0;
        for (long long v : ns) { total += v; if (v > 3 * med) { slow++; slow_ns += v; } }
        std::vector<long long> sc(cyc);
        std::sort(sc.begin(), sc.end());
        printf("%4d threads/block: repetition median %lld ns (%lld cycles), p99 %lld ns, max %lld ns; slower than 3x median: %zu of %zu (%.2f %% of the time); cycles p99 %lld max %lld\n",
               threads, med, sc[sc.size() / 2], s[s.size() * 99 / 100], s.back(), slow, ns.size(), 100.0 * slow_ns / total, sc[sc.size() * 99 / 100], sc.back());
        // slow windows of SM 0 in time order
        printf("   block 0 slow repetitions (index: ns):");
        int shown = 0;
        for (int r = 0; r < kReps && shown < 16; r++)
            if (ns[r] > 3 * med) { printf(" %d:%lld", r, ns[r]); shown++; }
        printf("\n");
    }
    return 0;
}
