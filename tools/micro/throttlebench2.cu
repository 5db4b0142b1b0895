// throttlebench2.cu -- does a worker warp slow down while the rest of the chip spin-polls a flag?
// Workers (one warp per block, first `workers` blocks) repeat a ~1 us mix of shared-memory and integer work and
// stamp %globaltimer around every repetition; all other warps of the grid poll a global flag the way the
// wavefront kernels poll an edge record (ld.relaxed.gpu + nanosleep).  Prints the distribution of repetition times.
#include <cstdio>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>

constexpr int kReps = 4000;

__global__ void k_mix(long long* ns, int workers, int sleep_ns, unsigned long long* flag, int* done, int* sink) {
    __shared__ int sm[8][64];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (blockIdx.x < workers && warp == 0) {
        int v = lane;
        for (int r = 0; r < kReps; r++) {
            long long t0, t1;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
            for (int i = 0; i < 40; i++) {
                sm[warp][lane] = v;
                __syncwarp();
                v = sm[warp][(lane * 5 + i) & 31] * 3 + v;
                v += __shfl_xor_sync(0xffffffffu, v, 1);
                __syncwarp();
            }
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
            if (lane == 0) ns[(size_t)blockIdx.x * kReps + r] = t1 - t0;
        }
        if (v == 12345) *sink = v;
        __syncwarp();
        if (lane == 0 && atomicAdd(done, 1) == workers - 1) atomicExch(flag, 1ull);
        return;
    }
    if (lane < 8) {
        unsigned long long f;
        do {
            if (sleep_ns > 0) __nanosleep(sleep_ns);
            asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(f) : "l"(flag + lane * 0));
        } while (f == 0);
    }
}

int main() {
    long long* d_ns;
    unsigned long long* d_flag;
    int *d_done, *d_sink;
    const int grid = 148 * 4;
    cudaMalloc(&d_ns, sizeof(long long) * grid * kReps);
    cudaMalloc(&d_flag, 256);
    cudaMalloc(&d_done, 4);
    cudaMalloc(&d_sink, 4);
    std::vector<long long> ns;
    for (int workers : {grid, 148, 16}) {
        for (int sleep_ns : {-1, 0, 40, 1000}) {
            if (workers == grid && sleep_ns >= 0) continue;
            if (workers != grid && sleep_ns < 0) continue;
            for (int pass = 0; pass < 2; pass++) {
                cudaMemset(d_flag, 0, 256);
                cudaMemset(d_done, 0, 4);
                k_mix<<<grid, 256>>>(d_ns, workers, sleep_ns, d_flag, d_done, d_sink);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
            }
            ns.assign((size_t)workers * kReps, 0);
            cudaMemcpy(ns.data(), d_ns, sizeof(long long) * ns.size(), cudaMemcpyDeviceToHost);
            std::vector<long long> s(ns);
            std::sort(s.begin(), s.end());
            const long long med = s[s.size() / 2];
            size_t slow = 0;
            long long slow_ns = 0, total = 0;
            for (long long v : ns) { total += v; if (v > 3 * med) { slow++; slow_ns += v; } }
            printf("workers %4d (blocks), pollers %s sleep %5d ns: repetition median %lld ns, p99 %lld, max %lld; > 3x median: %zu of %zu reps = %.2f %% of the time\n",
                   workers, workers == grid ? "none," : "rest, ", sleep_ns, med, s[s.size() * 99 / 100], s.back(), slow, ns.size(), 100.0 * slow_ns / total);
        }
    }
    return 0;
}
