// placement.cu -- where does the hardware put the blocks of a grid that does not fill the GPU?
// Every block records its SM and start time, then stays resident for a while.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o placement placement.cu ; ./placement [threads] [smem_kb] [hold_us]
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <vector>
__global__ void k(unsigned* sm_of, unsigned long long* t_of, unsigned hold_us, int regs_hog) {
    extern __shared__ char sm[];
    unsigned long long t0, t;
    unsigned s;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    asm volatile("mov.u32 %0, %%smid;" : "=r"(s));
    if (threadIdx.x == 0) { sm_of[blockIdx.x] = s; t_of[blockIdx.x] = t0; sm[0] = 1; }
    do { __nanosleep(1000); asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); } while (t - t0 < 1000ull * hold_us);
}
int main(int argc, char** argv) {
    const int threads = argc > 1 ? atoi(argv[1]) : 128, smem_kb = argc > 2 ? atoi(argv[2]) : 19, hold = argc > 3 ? atoi(argv[3]) : 3000;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_kb * 1024);
    int dev_sms = 0; cudaDeviceGetAttribute(&dev_sms, cudaDevAttrMultiProcessorCount, 0);
    unsigned* d_sm; unsigned long long* d_t;
    cudaMalloc(&d_sm, 1 << 16); cudaMalloc(&d_t, 1 << 17);
    for (int n : {148, 256, 296, 512, 768, 1024, 1500}) {
        k<<<n, threads, smem_kb * 1024>>>(d_sm, d_t, hold, 0);
        cudaDeviceSynchronize();
        std::vector<unsigned> sm(n); std::vector<unsigned long long> t(n);
        cudaMemcpy(sm.data(), d_sm, 4 * n, cudaMemcpyDeviceToHost); cudaMemcpy(t.data(), d_t, 8 * n, cudaMemcpyDeviceToHost);
        unsigned long long t0 = t[0]; for (auto x : t) if (x < t0) t0 = x;
        std::map<unsigned, int> per; int early = 0;
        for (int i = 0; i < n; i++) if (t[i] - t0 < 1000ull * hold / 2) { per[sm[i]]++; early++; }
        std::map<int, int> hist; for (auto& p : per) hist[p.second]++;
        printf("grid %4d x %d thr, %d KB: %d blocks started at once on %zu of %d SMs; blocks per SM:", n, threads, smem_kb, early, per.size(), dev_sms);
        for (auto& h : hist) printf(" %dx%d", h.second, h.first);
        printf("   first 24 blocks' SMs:"); for (int i = 0; i < 24 && i < n; i++) printf(" %u", sm[i]);
        printf("\n");
    }
    return 0;
}
