// D2H copy rates: linear vs 2-D (pitch) copies into pinned memory, alone and under a busy kernel.
#include <cstdio>
#include <cstdint>
#include <cstring>
#include <cuda_runtime.h>
__global__ void spin(long long cycles, int* sink) {
    long long t0 = clock64(); int a = 0;
    while (clock64() - t0 < cycles) a++;
    if (a == 123456789) *sink = a;
}
// one lane per warp: byte loads walking a private 50 KB stream + scattered 2-byte stores, ~100 cycles apart
__global__ void memspin(long long cycles, const uint8_t* in, uint16_t* out, int mode) {
    if (threadIdx.x & 31) return;
    const size_t w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const uint8_t* p = in + w * 51200;
    uint16_t* o = out + w * 409600;
    long long t0 = clock64(); unsigned a = 1, pos = 0;
    while (clock64() - t0 < cycles) {
        if (mode & 1) { a = a * 31 + __ldg(p + (pos % 51200)); pos += 3; }
        if (mode & 2) o[(a * 2654435761u >> 8) % 409600] = (uint16_t)a;
        if (mode & 4) __threadfence_block();
        for (int i = 0; i < 8; i++) a = a * 1664525u + 1013904223u;
    }
    if (a == 123456789) out[0] = a;
}
// 1024 warps, each adds to a host-mapped counter every `period` cycles (system-scope atomic over PCIe)
__global__ void atomspin(long long cycles, long long period, unsigned* host_counter) {
    if (threadIdx.x & 31) return;
    long long t0 = clock64(), next = t0 + period + (blockIdx.x * 977) % period;
    while (clock64() - t0 < cycles) {
        if (clock64() >= next) {
            __threadfence();
            if (period & 1) *reinterpret_cast<volatile unsigned*>(host_counter + blockIdx.x) = (unsigned)next;  // plain posted write, own slot
            else atomicAdd_system(host_counter + (blockIdx.x & 7), 1u);
            next += period;
        }
    }
}
int main() {
    const size_t pitch = 786432, n = 1024, total = pitch * n;
    uint8_t *d, *h; int* sink;
    cudaMalloc(&d, total); cudaMallocHost(&h, total); cudaMalloc(&sink, 4);
    cudaMemset(d, 1, total);
    cudaStream_t s, s2; cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking); cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    uint8_t* din; uint16_t* dout;
    cudaMalloc(&din, (size_t)2048 * 51200); cudaMalloc(&dout, (size_t)2048 * 409600 * 2);
    cudaMemset(din, 3, (size_t)2048 * 51200); cudaMemset(dout, 0, (size_t)2048 * 409600 * 2);
    int busy_mode = 0;
    auto run = [&](const char* name, size_t width, bool twod, bool busy) {
        for (int rep = 0; rep < 2; rep++) {
            if (busy && busy_mode == 0) spin<<<1024, 64, 0, s2>>>(100000000LL, sink);
            if (busy && busy_mode) memspin<<<1024, 64, 0, s2>>>(60000000LL, din, dout, busy_mode);
            cudaEventRecord(a, s);
            if (twod) cudaMemcpy2DAsync(h, pitch, d, pitch, width, n, cudaMemcpyDeviceToHost, s);
            else cudaMemcpyAsync(h, d, width * n, cudaMemcpyDeviceToHost, s);
            cudaEventRecord(b, s);
            cudaStreamSynchronize(s);
            float ms; cudaEventElapsedTime(&ms, a, b);
            if (rep) printf("%-40s %7.1f MB in %6.2f ms = %5.1f GB/s\n", name, width * n / 1e6, ms, width * n / ms / 1e6);
            cudaDeviceSynchronize();
        }
    };
    for (size_t mb : {1, 8, 52, 200}) {
        for (int rep = 0; rep < 3; rep++) {
            if (rep == 2) memset(h, rep, mb << 20);  // freshly written by the CPU
            cudaEventRecord(a, s);
            cudaMemcpyAsync(d, h, mb << 20, cudaMemcpyHostToDevice, s);
            cudaEventRecord(b, s);
            cudaStreamSynchronize(s);
            float ms; cudaEventElapsedTime(&ms, a, b);
            if (rep) printf("H2D %3zu MB%s: %6.3f ms = %5.1f GB/s\n", mb, rep == 2 ? " (just written by the CPU)" : "", ms, (mb << 20) / ms / 1e6);
        }
    }
    run("linear 805 MB", pitch, false, false);
    run("linear 201 MB", pitch / 4, false, false);
    run("2D 1024 x 196608 (pitch 786432)", 196608, true, false);
    run("2D 1024 x 98304", 98304, true, false);
    run("2D 1024 x 49152", 49152, true, false);
    run("2D 1024 x 185856 (121 rows of 1536)", 185856, true, false);
    run("linear 201 MB under busy kernel", pitch / 4, false, true);
    run("2D 1024 x 196608 under busy kernel", 196608, true, true);
    unsigned *hc, *dc;
    cudaHostAlloc((void**)&hc, 8192, cudaHostAllocMapped); cudaHostGetDevicePointer((void**)&dc, hc, 0);
    for (long long period : {2000000LL, 2000001LL, 200001LL, 20001LL}) {  // odd period = plain stores instead of atomics  // 1024 warps: one atomic per ~1 us, 0.1 us, 0.01 us overall
        for (int rep = 0; rep < 2; rep++) {
            atomspin<<<1024, 64, 0, s2>>>(60000000LL, period, dc);
            cudaEventRecord(a, s);
            cudaMemcpyAsync(h, d, pitch / 4 * n, cudaMemcpyDeviceToHost, s);
            cudaEventRecord(b, s);
            cudaStreamSynchronize(s);
            float ms; cudaEventElapsedTime(&ms, a, b);
            if (rep) printf("linear 201 MB under host atomics every %lld cycles/warp: %6.2f ms = %5.1f GB/s\n", period, ms, pitch / 4 * n / ms / 1e6);
            cudaDeviceSynchronize();
        }
    }
    // host polling an event at 50 kHz while the copy runs
    {
        cudaEvent_t never; cudaEventCreate(&never);
        for (int rep = 0; rep < 2; rep++) {
            spin<<<1024, 64, 0, s2>>>(60000000LL, sink);
            cudaEventRecord(never, s2);
            cudaEventRecord(a, s);
            cudaMemcpyAsync(h, d, pitch / 4 * n, cudaMemcpyDeviceToHost, s);
            cudaEventRecord(b, s);
            while (cudaEventQuery(b) == cudaErrorNotReady) { cudaEventQuery(never); volatile unsigned v = hc[0]; (void)v; }
            float ms; cudaEventElapsedTime(&ms, a, b);
            if (rep) printf("linear 201 MB while the host polls events: %6.2f ms = %5.1f GB/s\n", ms, pitch / 4 * n / ms / 1e6);
            cudaDeviceSynchronize();
        }
    }
    for (busy_mode = 1; busy_mode < 2; busy_mode++) {
        char name[64]; snprintf(name, sizeof name, "linear 201 MB under memspin mode %d", busy_mode);
        run(name, pitch / 4, false, true);
    }
    return 0;
}
