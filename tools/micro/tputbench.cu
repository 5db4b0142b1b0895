// Issue throughput of a few integer instructions (many warps, independent chains): cycles per warp-instruction per SM sub-partition.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define REP8(x) x x x x x x x x
#define KERNEL(name, body)                                                         \
    __global__ void name(int n, unsigned seed, unsigned* out) {                     \
        unsigned a0 = seed + threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, a4 = a0 * 11, a5 = a0 * 13, a6 = a0 * 17, a7 = a0 * 19; \
        const unsigned b = seed * 2654435761u | 1;                                   \
        for (int i = 0; i < n; i++) { REP8(body) }                                   \
        out[blockIdx.x * blockDim.x + threadIdx.x] = a0 ^ a1 ^ a2 ^ a3 ^ a4 ^ a5 ^ a6 ^ a7; \
    }
#define ALL(op) a0 = op(a0); a1 = op(a1); a2 = op(a2); a3 = op(a3); a4 = op(a4); a5 = op(a5); a6 = op(a6); a7 = op(a7);
#define OP_IMAD(x) ((x) * b + 7)
#define OP_MULHI(x) (__umulhi((x), b) + 1)
#define OP_SHF(x) __funnelshift_l((x), b, 3)
#define OP_SAD(x) __usad((x), b, 3u)
#define OP_PRMT(x) __byte_perm((x), b, 0x2103)
#define OP_CLZ(x) (__clz((x) | 1) + b)
#define OP_MNMX(x) (unsigned)min(max((int)(x), (int)b - 7), (int)b)
#define OP_IADD3(x) ((x) + b + (unsigned)i)
#define OP_LOP(x) (((x) ^ b) & (unsigned)(i | 0xffff))
__device__ __forceinline__ unsigned sat8(unsigned x) { int r; asm("cvt.sat.u8.s32 %0, %1;" : "=r"(r) : "r"((int)x + 5)); return r; }
__device__ __forceinline__ unsigned pack16(unsigned x, unsigned y) { unsigned d; asm("cvt.pack.sat.s16.s32 %0, %1, %2;" : "=r"(d) : "r"((int)x), "r"((int)y)); return d; }
#define OP_SAT8(x) sat8(x)
#define OP_PACK(x) pack16((x), b)
KERNEL(k_imad, ALL(OP_IMAD)) KERNEL(k_mulhi, ALL(OP_MULHI)) KERNEL(k_shf, ALL(OP_SHF)) KERNEL(k_sad, ALL(OP_SAD))
KERNEL(k_prmt, ALL(OP_PRMT)) KERNEL(k_clz, ALL(OP_CLZ)) KERNEL(k_mnmx, ALL(OP_MNMX)) KERNEL(k_sat8, ALL(OP_SAT8)) KERNEL(k_pack, ALL(OP_PACK)) KERNEL(k_iadd3, ALL(OP_IADD3)) KERNEL(k_lop, ALL(OP_LOP))
int main() {
    unsigned* out; cudaMalloc(&out, 148 * 4 * 1024 * 4);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    const int n = 4000;
#define RUN(k, ops_per_body) for (int r = 0; r < 2; r++) { cudaEventRecord(a); k<<<148 * 2, 1024>>>(n, 12345u, out); cudaEventRecord(b); cudaEventSynchronize(b); } \
    { float ms; cudaEventElapsedTime(&ms, a, b); double warp_instr = 148.0 * 2 * 32 * (double)n * 64 * ops_per_body; \
      printf("%-8s %.2f cycles per warp-instruction per scheduler (assuming 1.965 GHz, %d instr per op)\n", #k, ms * 1e-3 * 1.965e9 * 148 * 4 / warp_instr, ops_per_body); }
    RUN(k_imad, 1) RUN(k_mulhi, 2) RUN(k_shf, 1) RUN(k_sad, 1) RUN(k_prmt, 1) RUN(k_clz, 2) RUN(k_mnmx, 2) RUN(k_sat8, 2) RUN(k_pack, 1) RUN(k_iadd3, 1) RUN(k_lop, 1)
    return 0;
}
