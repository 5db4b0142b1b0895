// Latency of one boolean-decoder step on one lane, for several formulations (cycles per symbol).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

struct St { uint32_t value, range; int bits; uint32_t pos; };

__device__ __forceinline__ void refill(St& s, const uint8_t* d) {
    uint32_t b0 = d[s.pos], b1 = d[s.pos + 1], b2 = d[s.pos + 2];
    s.value = (s.value << 24) | (b0 << 16) | (b1 << 8) | b2; s.pos += 3; s.bits += 24;
}
template <int V> __device__ __forceinline__ int get(St& s, const uint8_t* d, int prob, const uint8_t* lut) {
    if (s.bits < 0) refill(s, d);
    uint32_t split;
    if (V == 0) split = 1 + (((s.range - 1) * (uint32_t)prob) >> 8);
    else split = (s.range * (uint32_t)prob + (256 - prob)) >> 8;
    const uint32_t big = split << s.bits;
    const bool one = s.value >= big;
    uint32_t r = one ? s.range - split : split;
    s.value = one ? s.value - big : s.value;
    int shift;
    if (V == 0 || V == 1) shift = __clz(r) - 24;
    else if (V == 2) shift = (r < 128) + (r < 64) + (r < 32) + (r < 16) + (r < 8) + (r < 4) + (r < 2);
    else shift = lut[r];
    s.range = r << shift;
    s.bits -= shift;
    return one;
}
template <int V> __global__ void k(const uint8_t* d, int n, int* out, long long* cyc) {
    __shared__ uint8_t lut[256];
    __shared__ uint8_t probs[256];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) { lut[i] = i ? __clz(i) - 24 : 7; probs[i] = 1 + (i * 37) % 255; }
    __syncthreads();
    if (threadIdx.x) return;
    St s{0, 255, -8, 0};
    int acc = 0, p = 0;
    long long t0 = clock64();
    for (int i = 0; i < n; i++) {
        int b = get<V>(s, d, probs[p], lut);
        p = (p * 2 + b + 1) & 255;   // next probability depends on the bit, as in a token tree
        acc += b;
    }
    long long t1 = clock64();
    out[0] = acc; cyc[0] = t1 - t0;
}
// variant 4: probability already in a register (no dependent table load)
__global__ void k4(const uint8_t* d, int n, int* out, long long* cyc) {
    if (threadIdx.x) return;
    St s{0, 255, -8, 0};
    int acc = 0; uint32_t pr = 0x80402010u;
    long long t0 = clock64();
    for (int i = 0; i < n; i++) {
        int b = get<1>(s, d, 1 + ((pr >> (8 * (acc & 3))) & 0xfe), nullptr);
        acc += b;
    }
    long long t1 = clock64();
    out[0] = acc; cyc[0] = t1 - t0;
}
int main() {
    const int n = 200000;
    uint8_t* h = (uint8_t*)malloc(n + 64);
    uint32_t x = 12345;
    for (int i = 0; i < n + 64; i++) { x = x * 1664525u + 1013904223u; h[i] = x >> 24; }
    uint8_t* d; int* out; long long* cyc;
    cudaMalloc(&d, n + 64); cudaMalloc(&out, 4); cudaMalloc(&cyc, 8);
    cudaMemcpy(d, h, n + 64, cudaMemcpyHostToDevice);
    for (int v = 0; v < 5; v++) {
        for (int rep = 0; rep < 2; rep++) {
            if (v == 0) k<0><<<1, 32>>>(d, n, out, cyc);
            if (v == 1) k<1><<<1, 32>>>(d, n, out, cyc);
            if (v == 2) k<2><<<1, 32>>>(d, n, out, cyc);
            if (v == 3) k<3><<<1, 32>>>(d, n, out, cyc);
            if (v == 4) k4<<<1, 32>>>(d, n, out, cyc);
            cudaDeviceSynchronize();
        }
        int o; long long c;
        cudaMemcpy(&o, out, 4, cudaMemcpyDeviceToHost); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
        printf("variant %d: %.1f cycles/symbol (ones %d) %s\n", v, (double)c / n, o, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
