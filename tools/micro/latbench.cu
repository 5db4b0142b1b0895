// Dependent-issue latencies of the instructions a serial decoder is made of (one warp, one SM).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define REP16(x) x x x x x x x x x x x x x x x x
#define BENCH(name, init, body)                                              \
    __global__ void name(int n, unsigned seed, unsigned* out, long long* cyc) { \
        __shared__ unsigned sm[256];                                          \
        for (int i = threadIdx.x; i < 256; i += 32) sm[i] = (i * 7 + 1) & 255; \
        __syncthreads();                                                       \
        unsigned a = seed, b = seed * 3 + 1; init;                             \
        long long t0 = clock64();                                              \
        for (int i = 0; i < n; i++) { REP16(body) }                           \
        long long t1 = clock64();                                              \
        out[threadIdx.x] = a + b; if (!threadIdx.x) cyc[0] = t1 - t0;          \
    }
BENCH(k_iadd, , a = a + b;)
BENCH(k_imad, , a = a * b + 5;)
BENCH(k_shf, , a = (a << (b & 31)) | 1;)
BENCH(k_flo, , a = __clz(a | 1) + b;)
BENCH(k_lop_shf, , a = ((a >> 3) ^ b) + 1;)
BENCH(k_sel, , a = (a >= b) ? a - b : a + 77;)
BENCH(k_lds, , a = sm[a & 255];)
BENCH(k_prmt, , a = __byte_perm(a, b, a & 0x7777) + 1;)
BENCH(k_popc, , a = __popc(a) + b;)
// a taken branch per step (data-dependent, always taken in practice)
__global__ void k_branch(int n, unsigned seed, unsigned* out, long long* cyc) {
    unsigned a = seed, b = seed * 3 + 1;
    long long t0 = clock64();
    for (int i = 0; i < n * 16; i++) {
        a = a + b;
        if (a == 0x12345) { a = a * 3; b++; }
    }
    long long t1 = clock64();
    out[threadIdx.x] = a + b; if (!threadIdx.x) cyc[0] = t1 - t0;
}
__global__ void k_branch2(int n, unsigned seed, unsigned* out, long long* cyc) {  // alternating taken / not taken forward branches
    unsigned a = seed, b = seed * 3 + 1;
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < n * 16; i++) {
        a = a + b;
        if (a & 1) { a = a ^ 0x55; asm volatile(""); } else { a = a + 0x33; asm volatile(""); }
        if (a & 2) { b = b ^ 0x51; asm volatile(""); } else { b = b + 0x13; asm volatile(""); }
    }
    long long t1 = clock64();
    out[threadIdx.x] = a + b; if (!threadIdx.x) cyc[0] = t1 - t0;
}
int main() {
    unsigned* out; long long* cyc; cudaMalloc(&out, 256); cudaMalloc(&cyc, 8);
    const int n = 20000;
#define RUN(k, ops) for (int r = 0; r < 2; r++) { k<<<1, 32>>>(n, 12345u, out, cyc); cudaDeviceSynchronize(); } \
    { long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost); printf("%-10s %.2f cycles per step (%s)\n", #k, (double)c / (n * 16.0), ops); }
    RUN(k_iadd, "IADD") RUN(k_imad, "IMAD") RUN(k_shf, "AND+SHF+LOP") RUN(k_flo, "LOP+FLO+IADD") RUN(k_lop_shf, "SHF+LOP+IADD")
    RUN(k_sel, "ISETP+2 ALU+SEL") RUN(k_lds, "LOP+LDS") RUN(k_prmt, "LOP+PRMT+IADD") RUN(k_popc, "POPC+IADD") RUN(k_branch, "IADD+ISETP+BRA loop") RUN(k_branch2, "loop with 2 if/else")
    return 0;
}
