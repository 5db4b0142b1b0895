// kernels.cu -- hand-written sm_100a kernels for everything after VP8 entropy decoding.
//
//   k_predict     dequantisation + inverse WHT/DCT + intra prediction (16x16, chroma, B_PRED)
//   k_loopfilter  simple / normal in-loop deblocking filter
//   k_yuv2rgb     fancy (bilinear) or simple chroma upsampling + YUV420 -> RGB / RGBA
//   k_alpha_*     ALPH plane un-filter (none / horizontal / vertical / gradient) + merge
//   k_composite   animation frame compositing (clear, copy, non-premultiplied src-over blend)
//
// Arithmetic follows the reference decoder (image-webp v0.2.4) bit for bit; each device
// function cites the reference lines it reproduces.  Scheduling does not: the reference walks
// macroblocks in raster order on one thread, here one WARP owns one macroblock ROW of one
// image and rows advance as a wavefront (row y may work on macroblock x once row y-1 has
// finished macroblock x+1, i.e. the x+2y order the dependencies allow).  Rows are handed out
// by a global ticket counter in (image, row) order, so every dependency points at a lower
// ticket and the scheme cannot deadlock whatever the residency.
//
// Hand-off between rows: what row y needs from row y-1 is its bottom edge (one pixel row for the
// prediction, four for the loop filter).  The producer writes that edge, per macroblock, into an
// EDGE RECORD of 8-byte units {4 pixels, tag}, tag = producer's macroblock row + 1; the consumer
// polls the unit until the tag is there.  An aligned 8-byte access is single-copy atomic, so the
// pixels and their valid mark arrive together: no counter, no release/acquire fence on either side
// (round 1's counters cost a MEMBAR per macroblock in the producer and one per wait in the
// consumer -- 30 % of all stall samples on large frames).  The records are cleared before a pass.  Integer/byte work bound by
// instruction issue and HBM, so no tensor cores and no TMA tiles: coalesced 16-byte loads of
// the level array, shared-memory staging of the macroblock workspace, L2-coherent (ld.cg)
// reads of neighbour rows.
#include "kernels.cuh"

#include <cstdlib>
#include <cstring>

namespace webpb200 {

ScratchLayout scratch_layout(uint64_t total_rows, uint64_t total_mbs, uint64_t edge_slots, bool with_filter) {
    ScratchLayout l;
    auto up = [](uint64_t v) { return (v + 255) & ~(uint64_t)255; };
    l.tickets_off = 0;
    l.edge_pred_off = 256;
    l.edge_lf_off = up(l.edge_pred_off + kEdgePredBytes * edge_slots);
    l.zero_bytes = up(l.edge_lf_off + (with_filter ? kEdgeFilterBytes * edge_slots : 0));
    l.rowmap_off = l.zero_bytes;
    l.order_off = up(l.rowmap_off + 4 * total_rows);
    l.recbase_off = up(l.order_off + 4 * total_rows);
    l.mbinfo_off = up(l.recbase_off + 4 * total_rows);
    l.total = up(l.mbinfo_off + total_mbs);
    return l;
}

uint64_t edge_slots_of(const webpb200_frame_desc* frames, int n) {
    uint64_t slots = 0;
    for (int i = 0; i < n; i++) slots += 2 * (uint64_t)frames[i].mbw;
    return slots;
}

#ifndef WF_SLEEP_NS
#define WF_SLEEP_NS 40  // pause between two polls of an edge record
#endif
#ifdef WF_TRACE
// Debug build only (tools/wf_trace.py): global-timer stamps (ns) of four consecutive macroblock rows of the
// prediction wavefront.  [row - WF_TRACE][stamp 0 step start, 1 after the wait for the row above, 2 step end
// (edge record written), 3 smid][mbx]
__device__ long long g_wf_trace[4][4][1024];
__device__ __forceinline__ long long wf_now() { long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define WF_STAMP(i, x) do { if (rid >= (uint32_t)(WF_TRACE) && rid < (uint32_t)(WF_TRACE) + 4 && lane == 0 && (x) < 1024) g_wf_trace[rid - (WF_TRACE)][i][x] = wf_now(); } while (0)
#else
#define WF_STAMP(i, x) do { } while (0)
#endif

namespace {

#ifndef BPRED_UNROLL
#define BPRED_UNROLL 10  // 1 = the ten B_PRED steps as a rolled loop (less code to fetch)
#endif
constexpr int kBpredUnroll = BPRED_UNROLL;
#ifndef WF_WARPS
#define WF_WARPS 8
#endif
constexpr int kWarpsPerBlock = WF_WARPS;
constexpr int kThreads = kWarpsPerBlock * 32;
constexpr unsigned kFull = 0xffffffffu;

// ---------------------------------------------------------------------------------------------
// small helpers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int clamp255(int v) { return __vimin_s32_relu(v, 255); }  // max(min(v, 255), 0): one VIMNMX.RELU
// four values clamped to 0..255 and packed, a lowest: two saturating cvt.pack instead of eight min/max and the shifts
__device__ __forceinline__ uint32_t clamp_pack4(int a, int b, int c, int d) {
    uint32_t hi, r;
    asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(hi) : "r"(d), "r"(c), "r"(0));
    asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(b), "r"(a), "r"(hi));
    return r;
}

__device__ __forceinline__ uint32_t ld_cg_u32(const void* p) { return __ldcg(reinterpret_cast<const unsigned int*>(p)); }
__device__ __forceinline__ uint2 ld_cg_u64(const void* p) { return __ldcg(reinterpret_cast<const uint2*>(p)); }
__device__ __forceinline__ uint4 ld_cg_u128(const void* p) { return __ldcg(reinterpret_cast<const uint4*>(p)); }

// one unit of an edge record: {4 pixels, tag} read / written as ONE aligned 64-bit access
__device__ __forceinline__ uint2 ld_edge(const uint2* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p));
    return make_uint2((uint32_t)v, (uint32_t)(v >> 32));
}
__device__ __forceinline__ void st_edge(uint2* p, uint32_t pixels, uint32_t tag) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"((unsigned long long)pixels | ((unsigned long long)tag << 32)) : "memory");
}
// `v` is what an earlier ld_edge(p) returned; poll until the producer's tag is there and return the pixels.
// Watchdog: a wait that lasts absurdly long (a scheduling bug, never a data property) raises the abort
// word; every waiter then stops waiting, the kernel drains with garbage pixels and the host reports an
// error instead of hanging the device.
// The whole warp waits together: the loop's exit is a warp-uniform vote, so the lanes never leave it one by one (a
// loop that only some lanes run and leave at different times puts the warp on the compiler's "not converged" copy of
// the code that follows, where every warp-level operation is several times slower).  `want`: this lane's unit matters.
__device__ __forceinline__ uint32_t wait_edge(const uint2* p, uint2 v, uint32_t tag, bool want, unsigned int* abort_word) {
    unsigned spins = 0;
    bool ok = !want || v.y == tag;
    while (!__all_sync(kFull, ok)) {
#if WF_SLEEP_NS > 0
        __nanosleep(WF_SLEEP_NS);
#endif
        if (!ok) {
            v = ld_edge(p);
            ok = v.y == tag;
        }
        if ((++spins & 0x3ff) == 0) {
            if (spins > (1u << 24)) atomicExch(abort_word, 1u);
            if (*reinterpret_cast<volatile unsigned int*>(abort_word)) break;
        }
    }
    return v.x;
}

// Every lane of the warp has a unit to wait for (normal loop filter): each lane polls its own; measured faster than
// the voting loop there, and the warp reconverges at the caller's next __syncwarp.
__device__ __forceinline__ uint32_t wait_edge_each(const uint2* p, uint2 v, uint32_t tag, unsigned int* abort_word) {
    unsigned spins = 0;
    while (v.y != tag) {
#if WF_SLEEP_NS > 0
        __nanosleep(WF_SLEEP_NS);
#endif
        v = ld_edge(p);
        if ((++spins & 0x3ff) == 0) {
            if (spins > (1u << 24)) atomicExch(abort_word, 1u);
            if (*reinterpret_cast<volatile unsigned int*>(abort_word)) break;
        }
    }
    return v.x;
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// ---------------------------------------------------------------------------------------------
// B_PRED lookup table.  Every 4x4 sub-block predictor except B_DC is, per pixel, either
// (a + 2b + c + 2) >> 2 over three edge pixels (avg3; avg2(a,b) = avg3(a,b,a), copy = avg3(a,a,a))
// or clamp(a + b - c) (B_TM).  One word per (mode, pixel): three signed byte offsets from the
// sub-block origin inside the 32-byte-stride workspace, and a TM flag in the top byte.
// Reference: prediction.rs:339-499 (VE, HE, LD, RD, VR, VL, HD, HU) and 238-269 (TM).
// ---------------------------------------------------------------------------------------------
constexpr int kWsStride = 32;
struct EdgeRef {
    int off;
};
constexpr EdgeRef eA(int i) { return {-kWsStride + i}; }
constexpr EdgeRef eP() { return {-kWsStride - 1}; }
constexpr EdgeRef eL(int j) { return {kWsStride * j - 1}; }

struct BpredTable {
    uint32_t w[10 * 16];
};

uint32_t pack3(EdgeRef a, EdgeRef b, EdgeRef c, int tm) {
    return (uint32_t)(uint8_t)(int8_t)a.off | ((uint32_t)(uint8_t)(int8_t)b.off << 8) |
           ((uint32_t)(uint8_t)(int8_t)c.off << 16) | ((uint32_t)tm << 24);
}

BpredTable make_bpred_table() {
    BpredTable t;
    std::memset(&t, 0, sizeof(t));
    auto set3 = [&](int mode, int x, int y, EdgeRef a, EdgeRef b, EdgeRef c) { t.w[mode * 16 + y * 4 + x] = pack3(a, b, c, 0); };
    auto set2 = [&](int mode, int x, int y, EdgeRef a, EdgeRef b) { t.w[mode * 16 + y * 4 + x] = pack3(a, b, a, 0); };
    auto set1 = [&](int mode, int x, int y, EdgeRef a) { t.w[mode * 16 + y * 4 + x] = pack3(a, a, a, 0); };
    // top row with the corner in front: T(-1) = P, T(i) = A_i
    auto T = [&](int i) { return i < 0 ? eP() : eA(i); };
    // edge e0..e8 = L3 L2 L1 L0 P A0 A1 A2 A3 (prediction.rs:318-337)
    auto E = [&](int i) { return i < 4 ? eL(3 - i) : i == 4 ? eP() : eA(i - 5); };
    for (int y = 0; y < 4; y++)
        for (int x = 0; x < 4; x++) {
            t.w[1 * 16 + y * 4 + x] = pack3(eL(y), eA(x), eP(), 1);           // B_TM
            set3(2, x, y, T(x - 1), T(x), T(x + 1));                         // B_VE
            if (y == 0) set3(3, x, y, eP(), eL(0), eL(1));                   // B_HE
            else if (y < 3) set3(3, x, y, eL(y - 1), eL(y), eL(y + 1));
            else set3(3, x, y, eL(2), eL(3), eL(3));
            const int i = x + y;                                             // B_LD
            if (i < 6) set3(4, x, y, eA(i), eA(i + 1), eA(i + 2));
            else set3(4, x, y, eA(6), eA(7), eA(7));
            const int j = 3 - y + x;                                         // B_RD
            set3(5, x, y, E(j), E(j + 1), E(j + 2));
        }
    // B_VR (prediction.rs:417-436)
    set3(6, 0, 3, E(1), E(2), E(3));
    set3(6, 0, 2, E(2), E(3), E(4));
    set3(6, 1, 3, E(3), E(4), E(5)); set3(6, 0, 1, E(3), E(4), E(5));
    set2(6, 1, 2, E(4), E(5));       set2(6, 0, 0, E(4), E(5));
    set3(6, 2, 3, E(4), E(5), E(6)); set3(6, 1, 1, E(4), E(5), E(6));
    set2(6, 2, 2, E(5), E(6));       set2(6, 1, 0, E(5), E(6));
    set3(6, 3, 3, E(5), E(6), E(7)); set3(6, 2, 1, E(5), E(6), E(7));
    set2(6, 3, 2, E(6), E(7));       set2(6, 2, 0, E(6), E(7));
    set3(6, 3, 1, E(6), E(7), E(8));
    set2(6, 3, 0, E(7), E(8));
    // B_VL (prediction.rs:438-457)
    set2(7, 0, 0, eA(0), eA(1));
    set3(7, 0, 1, eA(0), eA(1), eA(2));
    set2(7, 0, 2, eA(1), eA(2));        set2(7, 1, 0, eA(1), eA(2));
    set3(7, 1, 1, eA(1), eA(2), eA(3)); set3(7, 0, 3, eA(1), eA(2), eA(3));
    set2(7, 1, 2, eA(2), eA(3));        set2(7, 2, 0, eA(2), eA(3));
    set3(7, 1, 3, eA(2), eA(3), eA(4)); set3(7, 2, 1, eA(2), eA(3), eA(4));
    set2(7, 2, 2, eA(3), eA(4));        set2(7, 3, 0, eA(3), eA(4));
    set3(7, 2, 3, eA(3), eA(4), eA(5)); set3(7, 3, 1, eA(3), eA(4), eA(5));
    set3(7, 3, 2, eA(4), eA(5), eA(6));
    set3(7, 3, 3, eA(5), eA(6), eA(7));
    // B_HD (prediction.rs:459-478)
    set2(8, 0, 3, E(0), E(1));
    set3(8, 1, 3, E(0), E(1), E(2));
    set2(8, 0, 2, E(1), E(2));       set2(8, 2, 3, E(1), E(2));
    set3(8, 1, 2, E(1), E(2), E(3)); set3(8, 3, 3, E(1), E(2), E(3));
    set2(8, 2, 2, E(2), E(3));       set2(8, 0, 1, E(2), E(3));
    set3(8, 3, 2, E(2), E(3), E(4)); set3(8, 1, 1, E(2), E(3), E(4));
    set2(8, 2, 1, E(3), E(4));       set2(8, 0, 0, E(3), E(4));
    set3(8, 3, 1, E(3), E(4), E(5)); set3(8, 1, 0, E(3), E(4), E(5));
    set3(8, 2, 0, E(4), E(5), E(6));
    set3(8, 3, 0, E(5), E(6), E(7));
    // B_HU (prediction.rs:480-499)
    set2(9, 0, 0, eL(0), eL(1));
    set3(9, 1, 0, eL(0), eL(1), eL(2));
    set2(9, 2, 0, eL(1), eL(2));        set2(9, 0, 1, eL(1), eL(2));
    set3(9, 3, 0, eL(1), eL(2), eL(3)); set3(9, 1, 1, eL(1), eL(2), eL(3));
    set2(9, 2, 1, eL(2), eL(3));        set2(9, 0, 2, eL(2), eL(3));
    set3(9, 3, 1, eL(2), eL(3), eL(3)); set3(9, 1, 2, eL(2), eL(3), eL(3));
    set1(9, 2, 2, eL(3)); set1(9, 3, 2, eL(3));
    set1(9, 0, 3, eL(3)); set1(9, 1, 3, eL(3)); set1(9, 2, 3, eL(3)); set1(9, 3, 3, eL(3));
    return t;
}

__constant__ BpredTable c_bpred;

}  // namespace

void bpred_table_eval(int mode, const uint8_t edge[13], uint8_t out[16]) {
    static const BpredTable t = make_bpred_table();
    uint8_t ws[6 * kWsStride];
    std::memset(ws, 0, sizeof(ws));
    const int org = 1 * kWsStride + 4;
    ws[org + eP().off] = edge[0];
    for (int i = 0; i < 8; i++) ws[org + eA(i).off] = edge[1 + i];
    for (int j = 0; j < 4; j++) ws[org + eL(j).off] = edge[9 + j];
    for (int p = 0; p < 16; p++) {
        const uint32_t w = t.w[mode * 16 + p];
        const int a = ws[org + (int8_t)(w & 0xff)], b = ws[org + (int8_t)((w >> 8) & 0xff)], c = ws[org + (int8_t)((w >> 16) & 0xff)];
        int v = (w >> 24) ? a + b - c : (a + 2 * b + c + 2) >> 2;
        out[p] = (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
    }
}

cudaError_t upload_tables_once() {
    static const BpredTable t = make_bpred_table();
    return cudaMemcpyToSymbol(c_bpred, &t, sizeof(t));
}

namespace {

// ---------------------------------------------------------------------------------------------
// Inverse transforms (one lane = one 4x4 block, values in registers)
// ---------------------------------------------------------------------------------------------
// transform.rs:1-4, 20-25: MUL2(x) = (x*35468)>>16 and MUL1(x) = x + ((x*20091)>>16) with 64-bit
// products.  As high halves of 32x32 products: (x*20091)>>16 = mulhi(x, 20091<<16), and because
// 35468<<16 does not fit a signed word, (x*35468)>>16 = mulhi(x, (35468<<16) - 2^32) + x.  Exact
// for every int32 x (floor((x*K)/2^32) with K a multiple of 2^16).
__device__ __forceinline__ int mul_c2(int x) { return __mulhi(x, (int)0x8A8C0000u) + x; }
__device__ __forceinline__ int mul_c1(int x) { return __mulhi(x, 0x4E7B0000) + x; }

// transform.rs:7-51.  Vertical pass first (no rounding), then rows with (x + 4) >> 3.  Sums stay
// far inside int32 for every level the bitstream can carry (|level| <= 2114, quantisers <= 440);
// only the products need 64 bits.
__device__ __forceinline__ void idct4x4(int* c) {
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int a1 = c[i] + c[8 + i], b1 = c[i] - c[8 + i];
        const int c1 = mul_c2(c[4 + i]) - mul_c1(c[12 + i]);
        const int d1 = mul_c1(c[4 + i]) + mul_c2(c[12 + i]);
        c[i] = a1 + d1;
        c[4 + i] = b1 + c1;
        c[12 + i] = a1 - d1;
        c[8 + i] = b1 - c1;
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int a1 = c[4 * i] + c[4 * i + 2], b1 = c[4 * i] - c[4 * i + 2];
        const int c1 = mul_c2(c[4 * i + 1]) - mul_c1(c[4 * i + 3]);
        const int d1 = mul_c1(c[4 * i + 1]) + mul_c2(c[4 * i + 3]);
        c[4 * i] = (a1 + d1 + 4) >> 3;
        c[4 * i + 3] = (a1 - d1 + 4) >> 3;
        c[4 * i + 1] = (b1 + c1 + 4) >> 3;
        c[4 * i + 2] = (b1 - c1 + 4) >> 3;
    }
}

// transform.rs:54-86
__device__ __forceinline__ void iwht4x4(int* c) {
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int a1 = c[i] + c[12 + i], b1 = c[4 + i] + c[8 + i];
        const int c1 = c[4 + i] - c[8 + i], d1 = c[i] - c[12 + i];
        c[i] = a1 + b1;
        c[4 + i] = c1 + d1;
        c[8 + i] = a1 - b1;
        c[12 + i] = d1 - c1;
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int a1 = c[4 * i] + c[4 * i + 3], b1 = c[4 * i + 1] + c[4 * i + 2];
        const int c1 = c[4 * i + 1] - c[4 * i + 2], d1 = c[4 * i] - c[4 * i + 3];
        c[4 * i] = (a1 + b1 + 3) >> 3;
        c[4 * i + 1] = (c1 + d1 + 3) >> 3;
        c[4 * i + 2] = (a1 - b1 + 3) >> 3;
        c[4 * i + 3] = (d1 - c1 + 3) >> 3;
    }
}

__device__ __forceinline__ void unpack8(const uint4& v, int* c) {
    c[0] = (int)(short)(v.x & 0xffff); c[1] = (int)v.x >> 16;
    c[2] = (int)(short)(v.y & 0xffff); c[3] = (int)v.y >> 16;
    c[4] = (int)(short)(v.z & 0xffff); c[5] = (int)v.z >> 16;
    c[6] = (int)(short)(v.w & 0xffff); c[7] = (int)v.w >> 16;
}
// residuals only ever meet a pixel through clamp(p + r, 0, 255) (prediction.rs:133-147), so
// saturating r first (to anything of magnitude >= 255) changes nothing and makes it fit 16 bits
// (any saturation bound >= 255 does: one cvt.pack.sat saturates both to int16 and packs them)
__device__ __forceinline__ uint32_t pack_res(int a, int b) {
    uint32_t d;
    asm("cvt.pack.sat.s16.s32 %0, %1, %2;" : "=r"(d) : "r"(b), "r"(a));
    return d;
}

// ---------------------------------------------------------------------------------------------
// k_predict
// ---------------------------------------------------------------------------------------------
// Per-warp workspace in shared memory.  Luma: 17 rows x 32 bytes; pixel (x, y) of the macroblock
// sits at row y+1, column x+4; column 3 is the left border, row 0 the row above (columns 0..3 the
// word holding the corner at column 3, 4..19 the 16 pixels above, 20..23 the 4 above-right);
// rows 4, 8, 12 carry the above-right pixels again in columns 20..23 for the last sub-block
// column (prediction.rs:47-51).  Chroma: 9 rows x 16 bytes, pixel (x, y) at row y+1, column x+4.
struct __align__(16) PredictSmem {
    uint8_t yw[17 * 32];
    uint8_t uw[9 * 16];
    uint8_t vw[9 * 16];
    int16_t res[2][24 * 16];  // residuals of the 24 blocks of the even / odd macroblock of a pair, raster inside a block
    int32_t wht[2][16];
    uint32_t quant[16];                  // this frame's quantisers: [segment * 4 + class] = dc | ac << 16
    int16_t lv[2][32 * 16];              // the next pair's levels (25 x 16 used each), prefetched with cp.async
    uint4 desc[2][2];                    // descriptors: [pair parity][macroblock of the pair], fetched two pairs ahead
    uint8_t map[64];                     // transform slot -> (macroblock of the pair) << 5 | block
};

struct PredictArgs {
    const webpb200_frame_desc* frames;
    const webpb200_mb_desc* mbs;
    const int16_t* levels;
    uint8_t* planes;
    unsigned int* ticket;
    unsigned int* abort_word;
    uint2* edges;             // edge records, 8 units per slot (units 0-3 bottom luma row, 4-5 U, 6-7 V)
    const uint32_t* recbase;  // row id -> first slot of that row
    const uint32_t* rowmap;
    const uint32_t* order;
    uint8_t* mbinfo;
    uint32_t total_rows;
};

// Two builds of the same kernel.  kMinBlocks = 3 (80 registers, 24 warps per SM) is the fast one when the batch
// keeps every resident warp busy (many images: issue-bound).  kMinBlocks = 2 (128 registers, 16 warps per SM)
// has fewer instructions on each macroblock's dependency chain and is the fast one when the rows of a few
// large frames wait on each other (latency-bound): cfg3 2.56 -> 1.85 ms, cfg2 0.93 -> 1.02 ms.
#ifndef PREDICT_BLOCKS_T
#define PREDICT_BLOCKS_T 3
#endif
#ifndef PREDICT_BLOCKS_L
#define PREDICT_BLOCKS_L 2
#endif
constexpr int kPredictBlocksThroughput = PREDICT_BLOCKS_T, kPredictBlocksLatency = PREDICT_BLOCKS_L;

// Level records.  Dense (the boundary format of include/webpb200.h, what the host entropy decoder and
// webpb200_reconstruct_device callers supply): block b of a macroblock at b * 32 bytes of its 800-byte record.
// Packed (what k_entropy writes, the section 8(f) format): only the blocks that carry tokens, 32 bytes each, in
// decode order (Y2 first, then 0..23) from the start of the record; the descriptor's nflags says which.
template <bool kPacked>
__device__ __forceinline__ int level_slot(uint32_t nflags, int blk) {
    if (!kPacked) return blk;
    return blk == 24 ? 0 : (int)((nflags >> 24) & 1u) + __popc(nflags & ((1u << blk) - 1u));
}

// levels of macroblocks `mbi` (n0 slots) and `mbi + 1` (n1 slots, 0 = no such macroblock) -> the warp's
// staging buffers, asynchronously; descriptors of the pair two further on (if `dmbi` != 0 count) into `dslot`
__device__ __forceinline__ void prefetch_pair(PredictSmem& S, const PredictArgs& A, size_t mbi, int n0, int n1, int dslot, size_t dmbi, int dcount,
                                              int lane) {
    if (lane < 25) {
        const char* src = reinterpret_cast<const char*>(A.levels + mbi * WEBPB200_LEVELS_PER_MB) + lane * 32;
        char* dst = reinterpret_cast<char*>(S.lv[0]) + lane * 32;
        if (lane < n0) {
            cp_async16(dst, src);
            cp_async16(dst + 16, src + 16);
        }
        if (lane < n1) {
            cp_async16(dst + sizeof(S.lv[0]), src + 2 * WEBPB200_LEVELS_PER_MB);
            cp_async16(dst + sizeof(S.lv[0]) + 16, src + 2 * WEBPB200_LEVELS_PER_MB + 16);
        }
    } else if (lane - 25 < dcount) {  // lanes 25, 26
        cp_async16(&S.desc[dslot][lane - 25], A.mbs + dmbi + (lane - 25));
    }
    cp_async_commit();
}

template <int kMinBlocks, bool kPacked>
__global__ void __launch_bounds__(kThreads, kMinBlocks) k_predict(PredictArgs A) {
    __shared__ PredictSmem smem[kWarpsPerBlock];
    __shared__ uint32_t s_bpred[160];
    for (int i = threadIdx.x; i < 160; i += kThreads) s_bpred[i] = c_bpred.w[i];
    __syncthreads();

    const int lane = threadIdx.x & 31;
    PredictSmem& S = smem[threadIdx.x >> 5];
    uint32_t* yw32 = reinterpret_cast<uint32_t*>(S.yw);
    uint32_t* uw32 = reinterpret_cast<uint32_t*>(S.uw);
    uint32_t* vw32 = reinterpret_cast<uint32_t*>(S.vw);
    // border roles, fixed per lane.  Lanes 0..7 each hold one 4-pixel unit of the edge record of the
    // macroblock above: 0..3 = its bottom luma row, 4..5 = U, 6..7 = V; `top_dst` is where that unit goes
    // in the workspace's top row.  Lanes 3, 5 and 7 hold the unit whose last byte is the corner of the
    // NEXT macroblock (prediction.rs:30-33, 96-99).  Every lane moves one byte of the right column to
    // the left border.
    uint32_t* const top_dst = lane < 4 ? yw32 + 1 + lane : lane < 6 ? uw32 + (lane - 3) : vw32 + (lane - 5);
    uint32_t* const corner_dst = lane == 3 ? yw32 : lane == 5 ? uw32 : vw32;
    const bool corner_lane = lane == 3 || lane == 5 || lane == 7;
    // ... and of the macroblock just finished: its bottom row, for the row below
    const uint32_t* const bottom_src = lane < 4 ? yw32 + 16 * 8 + 1 + lane : lane < 6 ? uw32 + 8 * 4 + (lane - 3) : vw32 + 8 * 4 + (lane - 5);
    uint8_t* const lb_dst = lane < 16 ? S.yw + (1 + lane) * 32 + 3 : lane < 24 ? S.uw + (lane - 15) * 16 + 3 : S.vw + (lane - 23) * 16 + 3;
    const int lb_gap = lane < 16 ? 16 : 8;  // right column = left border + gap
    // B_PRED per-lane constants: half-warp, pixel inside a sub-block, the edge pixel summed for B_DC; `_h` = as seen
    // from the half-warp's own sub-block of a two-sub-block step (the lower one sits 4 rows down, 2 sub-blocks left)
    const int bp_half = lane >> 4, bp_p = lane & 15;
    const int bp_pixoff = (bp_p >> 2) * 32 + (bp_p & 3);
    const int bp_dcoff = (bp_p & 4) ? (bp_p & 3) * 32 - 1 : (bp_p & 3) - 32;
    const uint8_t* const bp_yw_h = S.yw + bp_half * 120;
    const uint8_t* const bp_dc = S.yw + bp_dcoff;
    const uint8_t* const bp_dc_h = bp_yw_h + bp_dcoff;
    uint8_t* const bp_pix = S.yw + bp_pixoff;
    uint8_t* const bp_pix_h = S.yw + bp_half * 120 + bp_pixoff;
    const int16_t* const bp_res = &S.res[0][0] + bp_p;
    const int16_t* const bp_res_h = bp_res + bp_half * 32;
    const uint32_t* const bp_tab = s_bpred + bp_p;

    for (;;) {
        unsigned int t = 0;
        if (lane == 0) t = atomicAdd(A.ticket, 1u);
        t = __shfl_sync(kFull, t, 0);
        if (t >= A.total_rows) break;
        const uint32_t rid = __ldg(A.order + t);  // row id = image's row_base + macroblock row
        const uint32_t img = __ldg(A.rowmap + rid);
        const webpb200_frame_desc* fd = A.frames + img;
        const int mbw = fd->mbw;
        const int mby = (int)(rid - fd->row_base);
        const size_t ystride = (size_t)mbw * 16, cstride = (size_t)mbw * 8;
        uint8_t* Y = A.planes + fd->y_off;
        uint8_t* U = A.planes + fd->u_off;
        uint8_t* V = A.planes + fd->v_off;
        const size_t mb_row0 = (size_t)fd->mb_base + (size_t)mby * mbw;
        // edge records: mine (written for the row below, if there is one) and the row above's
        const uint32_t tag_above = (uint32_t)mby, tag_mine = (uint32_t)mby + 1u;
        const bool hand_down = mby + 1 < (int)fd->mbh;
        uint2* const edge_mine = A.edges + (size_t)__ldg(A.recbase + rid) * 8 + (lane & 7);
        const uint2* const edge_above = A.edges + (size_t)__ldg(A.recbase + (mby > 0 ? rid - 1 : rid)) * 8 + (lane & 7);
        // cur = my unit of the record above macroblock mbx, nxt = of the one above mbx + 1 (as loaded: its tag
        // is checked when it is needed), prev = of the one above mbx - 1
        uint2 nxt = make_uint2(0u, 0u);
        uint32_t cur = 0, prev = 0;
        if (mby > 0) {
            uint2 first = make_uint2(0u, 0u);
            if (lane < 8) first = ld_edge(edge_above);
            cur = wait_edge(edge_above, first, tag_above, lane < 8, A.abort_word);
            if (lane < 8 && mbw > 1) nxt = ld_edge(edge_above + 8);
        }
        uint8_t* const yout = Y + (size_t)(mby * 16 + (lane >> 1)) * ystride + (lane & 1) * 8;
        uint8_t* const cout = (lane < 16 ? U : V) + (size_t)(mby * 8 + ((lane & 15) >> 1)) * cstride + (lane & 1) * 4;

        // first pair: whole records (the packed slots in use are not known yet), with its descriptors and the next pair's
        prefetch_pair(S, A, mb_row0, 25, mbw > 1 ? 25 : 0, 0, mb_row0, min(mbw, 2), lane);
        if (mbw > 2) prefetch_pair(S, A, mb_row0, 0, 0, 1, mb_row0 + 2, min(mbw - 2, 2), lane);
        // the frame's quantisers, once per row (class 0 = Y, 1 = U/V, 2 = Y2): descriptor loads would
        // miss L1 on every macroblock because the acquire fences keep invalidating it
        if (lane < 16) {
            const webpb200_segment* sg = fd->seg + (lane >> 2);
            const int cls = lane & 3;
            const int dc = cls == 0 ? sg->ydc : cls == 1 ? sg->uvdc : sg->y2dc;
            const int ac = cls == 0 ? sg->yac : cls == 1 ? sg->uvac : sg->y2ac;
            S.quant[lane] = (uint32_t)(dc & 0xffff) | ((uint32_t)ac << 16);
        }
        __syncwarp();
        const int qcls = lane < 16 ? 0 : lane < 24 ? 1 : 2;
        uint4 desc_odd = make_uint4(0, 0, 0, 0);  // descriptor of the pair's second macroblock, kept for its turn
        for (int mbx = 0; mbx < mbw; mbx++) {
            WF_STAMP(0, mbx);
            const size_t mbi = mb_row0 + mbx;
            uint4 mbw4 = desc_odd;
            if ((mbx & 1) == 0) {
                // ---- dequantise + inverse transforms for the pair (mbx, mbx + 1) ----
                // read_coefficients mod.rs:865-866 (the multiply), read_residual_data 888-952.  On average
                // 13 of a macroblock's 24 blocks carry coefficients, so the blocks that need a transform
                // of BOTH macroblocks usually fit the warp's 32 lanes: one pass instead of two.
                const bool two = mbx + 1 < mbw;
                cp_async_wait_all();
                __syncwarp();
                const int dsl = (mbx >> 1) & 1;
                mbw4 = S.desc[dsl][0];
                desc_odd = two ? S.desc[dsl][1] : make_uint4(0x01000000u, 0, 0, 0);  // a missing partner counts as skipped
                const uint32_t d0 = mbw4.x, d1 = desc_odd.x;
                const bool use0 = !((d0 >> 24) & 1), use1 = !((d1 >> 24) & 1);       // not skipped
                const bool y2_0 = use0 && (d0 & 0xff) != 4, y2_1 = use1 && (d1 & 0xff) != 4;
                // Y2 first (iwht, transform.rs:54-86): lane 24 for the even macroblock, lane 25 for the odd one
                if ((lane == 24 && y2_0) || (lane == 25 && y2_1)) {
                    const int wh = lane - 24;
                    // (packed: the Y2 block has a slot, the first, only if it carried tokens)
                    const uint4* lp = reinterpret_cast<const uint4*>(S.lv[wh] + level_slot<kPacked>(wh ? desc_odd.y : mbw4.y, 24) * 16);
                    int c[16];
                    if (!kPacked || (((wh ? desc_odd.y : mbw4.y) >> 24) & 1u)) {
                        unpack8(lp[0], c);
                        unpack8(lp[1], c + 8);
                    } else {
#pragma unroll
                        for (int i = 0; i < 16; i++) c[i] = 0;
                    }
                    const uint32_t qw = S.quant[(((wh ? d1 : d0) >> 16) & 3) * 4 + 2];
                    const int dcq = qw & 0xffff, acq = qw >> 16;
                    c[0] *= dcq;
#pragma unroll
                    for (int i = 1; i < 16; i++) c[i] *= acq;
                    iwht4x4(c);
#pragma unroll
                    for (int i = 0; i < 16; i++) S.wht[wh][i] = c[i];
                }
                __syncwarp();
                // Which blocks get the IDCT (mod.rs:920, 949: DC != 0 or the "has coefficients" flag).  A block
                // without the flag has no coefficients of its own; if its DC (from the Y2 block) is not zero the
                // transform of a lone DC is the constant (dc + 4) >> 3, written here directly -- those blocks
                // (most luma blocks of a 16x16-predicted macroblock) then do not take a lane of the IDCT pass.
                bool nz0 = false, nz1 = false, need0 = false, need1 = false;
                if (lane < 24) {
                    // the DC of a block WITHOUT the flag (the only case it is used below): from the Y2 block; else, packed,
                    // 0 (a block has levels only if it had tokens), dense, whatever the caller's record holds
                    const int dc0 = (y2_0 && lane < 16) ? S.wht[0][lane] : kPacked ? 0 : (int)S.lv[0][lane * 16];
                    const int dc1 = (y2_1 && lane < 16) ? S.wht[1][lane] : kPacked ? 0 : (int)S.lv[1][lane * 16];
                    need0 = use0 && ((mbw4.y >> lane) & 1);
                    need1 = use1 && ((desc_odd.y >> lane) & 1);
                    nz0 = need0 || (use0 && dc0 != 0);
                    nz1 = need1 || (use1 && dc1 != 0);
                    uint4* r0 = reinterpret_cast<uint4*>(S.res[0] + lane * 16);
                    uint4* r1 = reinterpret_cast<uint4*>(S.res[1] + lane * 16);
                    if (!need0) {
                        const int r = nz0 ? (dc0 + 4) >> 3 : 0;
                        const uint32_t w = pack_res(r, r);
                        r0[0] = make_uint4(w, w, w, w); r0[1] = make_uint4(w, w, w, w);
                    }
                    if (!need1) {
                        const int r = nz1 ? (dc1 + 4) >> 3 : 0;
                        const uint32_t w = pack_res(r, r);
                        r1[0] = make_uint4(w, w, w, w); r1[1] = make_uint4(w, w, w, w);
                    }
                }
                const unsigned m0 = __ballot_sync(kFull, need0), m1 = __ballot_sync(kFull, need1);
                const bool any0 = __any_sync(kFull, nz0), any1 = __any_sync(kFull, nz1);
                const int n0 = __popc(m0), total = n0 + __popc(m1);
                const unsigned below = (1u << lane) - 1u;
                if (need0) S.map[__popc(m0 & below)] = (uint8_t)lane;
                if (need1) S.map[n0 + __popc(m1 & below)] = (uint8_t)(32 + lane);
                if (lane == 0) {
                    A.mbinfo[mbi] = any0 ? 1 : 0;
                    if (two) A.mbinfo[mbi + 1] = any1 ? 1 : 0;
                }
                __syncwarp();
                for (int base = 0; base < total; base += 32) {
                    const int slot = base + lane;
                    if (slot < total) {
                        const int mm = S.map[slot], wh = mm >> 5, blk = mm & 31;
                        const uint32_t dw = wh ? d1 : d0;
                        const uint4* lp = reinterpret_cast<const uint4*>(S.lv[wh] + level_slot<kPacked>(wh ? desc_odd.y : mbw4.y, blk) * 16);
                        int c[16];
                        unpack8(lp[0], c);
                        unpack8(lp[1], c + 8);
                        const uint32_t qw = S.quant[((dw >> 16) & 3) * 4 + (blk < 16 ? 0 : 1)];
                        const int dcq = qw & 0xffff, acq = qw >> 16;  // quantisers are positive and below 2^15
                        c[0] *= dcq;
#pragma unroll
                        for (int i = 1; i < 16; i++) c[i] *= acq;
                        if (blk < 16 && (wh ? y2_1 : y2_0)) c[0] = S.wht[wh][blk];
                        idct4x4(c);
                        uint4 o0, o1;
                        o0.x = pack_res(c[0], c[1]);  o0.y = pack_res(c[2], c[3]);
                        o0.z = pack_res(c[4], c[5]);  o0.w = pack_res(c[6], c[7]);
                        o1.x = pack_res(c[8], c[9]);  o1.y = pack_res(c[10], c[11]);
                        o1.z = pack_res(c[12], c[13]); o1.w = pack_res(c[14], c[15]);
                        uint4* rp = reinterpret_cast<uint4*>(S.res[wh] + blk * 16);
                        rp[0] = o0;
                        rp[1] = o1;
                    }
                }
                __syncwarp();
                if (mbx + 2 < mbw) {
                    // the next pair's levels (its descriptors arrived with this pair's levels) and the descriptors
                    // of the pair after it, into the descriptor slot this pair has just left
                    const uint32_t nf0 = S.desc[dsl ^ 1][0].y, nf1 = mbx + 3 < mbw ? S.desc[dsl ^ 1][1].y : 0u;
                    const int n0 = kPacked ? __popc(nf0 & 0x1ffffffu) : 25;
                    const int n1 = mbx + 3 < mbw ? (kPacked ? __popc(nf1 & 0x1ffffffu) : 25) : 0;
                    prefetch_pair(S, A, mbi + 2, n0, n1, dsl, mbi + 4, max(0, min(mbw - mbx - 4, 2)), lane);
                }
            }
            const int16_t* const res = S.res[mbx & 1];
            const int ymode = mbw4.x & 0xff, uvmode = (mbw4.x >> 8) & 0xff;
            const uint32_t bpred_lo = mbw4.z, bpred_hi = mbw4.w;

            // ---- neighbours: the edge record of the above-right macroblock must be there, then build the borders ----
            // create_border_luma / create_border_chroma, prediction.rs:13-123
            if (mby == 0) {
                // first macroblock row: 127 everywhere, corner and above-right included; constant along the row
                if (mbx == 0 && lane < 12) {
                    uint32_t* d = lane < 6 ? yw32 + lane : lane < 9 ? uw32 + (lane - 6) : vw32 + (lane - 9);
                    *d = 0x7f7f7f7fu;
                    if (lane == 5) { yw32[4 * 8 + 5] = 0x7f7f7f7fu; yw32[8 * 8 + 5] = 0x7f7f7f7fu; yw32[12 * 8 + 5] = 0x7f7f7f7fu; }
                }
            } else {
                const bool last = mbx == mbw - 1;
                const uint32_t nx = wait_edge(edge_above + (size_t)(mbx + 1) * 8, nxt, tag_above, lane < 8 && !last, A.abort_word);
                if (lane < 8) {
                    *top_dst = cur;
                    // corner: no macroblock to the left -> 129, else the last pixel above it
                    if (corner_lane) *corner_dst = mbx == 0 ? 0x81818181u : prev;
                    // above-right (luma only): the first 4 pixels above the next macroblock; for the last
                    // macroblock its own last pixel above, four times (prediction.rs:35-38)
                    if (lane == (last ? 3 : 0)) {
                        const uint32_t w = last ? (cur >> 24) * 0x01010101u : nx;
                        yw32[5] = w; yw32[4 * 8 + 5] = w; yw32[8 * 8 + 5] = w; yw32[12 * 8 + 5] = w;
                    }
                    prev = cur;
                    cur = nx;
                    // the unit after that, requested now and looked at in the next round
                    if (mbx + 2 < mbw) nxt = ld_edge(edge_above + (size_t)(mbx + 2) * 8);
                }
            }
            WF_STAMP(1, mbx);
            if (mbx == 0) *lb_dst = 129;  // left border 129 (prediction.rs:54-57, 104-107)
            __syncwarp();

            // ---- luma ----
            if (ymode == 4) {
                // predict_4x4 (prediction.rs:159-183) as a 10-step sx + 2*sy wavefront over the
                // 16 sub-blocks; half-warp h works on sub-block row sy_lo + h, one lane per pixel.
                // Everything that depends on the lane only (which half, which pixel) is folded into
                // base pointers once (bp_*), so that a step's addresses are constants off those.
                const unsigned long long bp = ((unsigned long long)bpred_hi << 32) | bpred_lo;
                const unsigned long long bph = bp >> (8 * bp_half);  // the lower sub-block of a two-sub-block step is 2 further on
#pragma unroll kBpredUnroll
                for (int s = 0; s < 10; s++) {
                    // steps with a single sub-block (0, 1, 8, 9): both half-warps compute it and store
                    // the same bytes, which keeps the step free of divergence
                    const int sy_lo = s < 2 ? 0 : (s - 2) >> 1, sy_hi = (s >> 1) < 3 ? (s >> 1) : 3;
                    const bool two = sy_hi > sy_lo;
                    const int sx = s - 2 * sy_lo, sb = sy_lo * 4 + sx;  // the upper (or only) sub-block; the lower one: sb + 2
                    const int org = (1 + 4 * sy_lo) * 32 + 4 + 4 * sx;  // ... its origin; the lower one: + 4 * 32 - 8
                    const uint8_t* const ew = two ? bp_yw_h : S.yw;     // workspace seen from my sub-block
                    const int mode = (int)((two ? bph : bp) >> (4 * sb)) & 15;
                    // B_DC, predict_bdcpred prediction.rs:271-288: lanes p and p^8 hold the same edge pixel
                    // (computed in every step: skipping it unless a sub-block of the step is B_DC -- a vote and a branch --
                    // was slower, 0.943 against 0.920 ms on cfg2)
                    int dc = (two ? bp_dc_h : bp_dc)[org];
                    dc += __shfl_xor_sync(kFull, dc, 1);
                    dc += __shfl_xor_sync(kFull, dc, 2);
                    dc += __shfl_xor_sync(kFull, dc, 4);
                    const uint32_t w = bp_tab[mode * 16];
                    const int a = ew[org + (int)(int8_t)(w & 0xff)];
                    const int b = ew[org + (int)(int8_t)((w >> 8) & 0xff)];
                    const int cc = ew[org + (int)(int8_t)((w >> 16) & 0xff)];
                    int pred = (w >> 24) ? clamp255(a + b - cc) : (a + 2 * b + cc + 2) >> 2;
                    if (mode == 0) pred = (dc + 4) >> 3;
                    const int r = (two ? bp_res_h : bp_res)[(mbx & 1) * (24 * 16) + sb * 16];
                    (two ? bp_pix_h : bp_pix)[org] = (uint8_t)clamp255(pred + r);  // add_residue
                    __syncwarp();
                }
            } else {
                // 16x16 predictors (prediction.rs:185-269): lane = (row, 8-pixel half)
                const int row = lane >> 1, x0 = (lane & 1) * 8;
                int pr[8];
                if (ymode == 0) {  // DC
                    int v = 0;
                    if (lane < 16) v = (mby > 0 ? (int)S.yw[4 + lane] : 0) + (mbx > 0 ? (int)S.yw[(1 + lane) * 32 + 3] : 0);
                    const int sum = __reduce_add_sync(kFull, (unsigned)v);
                    const int shf = 3 + (mby > 0) + (mbx > 0);
                    const int dc = (mby == 0 && mbx == 0) ? 128 : (sum + (1 << (shf - 1))) >> shf;
#pragma unroll
                    for (int i = 0; i < 8; i++) pr[i] = dc;
                } else if (ymode == 1) {  // V
#pragma unroll
                    for (int i = 0; i < 8; i++) pr[i] = S.yw[4 + x0 + i];
                } else if (ymode == 2) {  // H
                    const int l = S.yw[(1 + row) * 32 + 3];
#pragma unroll
                    for (int i = 0; i < 8; i++) pr[i] = l;
                } else {  // TM
                    const int lmp = (int)S.yw[(1 + row) * 32 + 3] - (int)S.yw[3];
#pragma unroll
                    for (int i = 0; i < 8; i++) pr[i] = clamp255(lmp + (int)S.yw[4 + x0 + i]);
                }
                __syncwarp();  // all reads of the borders done before the block is written
                const int blk = (row >> 2) * 4 + (x0 >> 2);
                const uint2 r0 = *reinterpret_cast<const uint2*>(res + blk * 16 + (row & 3) * 4);
                const uint2 r1 = *reinterpret_cast<const uint2*>(res + (blk + 1) * 16 + (row & 3) * 4);
                const int rr[8] = {(int)(short)(r0.x & 0xffff), (int)r0.x >> 16, (int)(short)(r0.y & 0xffff), (int)r0.y >> 16,
                                   (int)(short)(r1.x & 0xffff), (int)r1.x >> 16, (int)(short)(r1.y & 0xffff), (int)r1.y >> 16};
                const uint32_t o0 = clamp_pack4(pr[0] + rr[0], pr[1] + rr[1], pr[2] + rr[2], pr[3] + rr[3]);
                const uint32_t o1 = clamp_pack4(pr[4] + rr[4], pr[5] + rr[5], pr[6] + rr[6], pr[7] + rr[7]);
                yw32[(1 + row) * 8 + 1 + (x0 >> 2)] = o0;
                yw32[(1 + row) * 8 + 2 + (x0 >> 2)] = o1;
                __syncwarp();
            }
            {  // macroblock -> unfiltered luma plane (mod.rs:713-720)
                const int row = lane >> 1, x0 = (lane & 1) * 8;
                uint2 o;
                o.x = yw32[(1 + row) * 8 + 1 + (x0 >> 2)];
                o.y = yw32[(1 + row) * 8 + 2 + (x0 >> 2)];
                *reinterpret_cast<uint2*>(yout + (size_t)mbx * 16) = o;
            }

            // ---- chroma (intra_predict_chroma, mod.rs:723-783): lane = (plane, row, 4-pixel half)
            {
                const int plane = lane >> 4, j = lane & 15, row = j >> 1, x0 = (j & 1) * 4;
                uint8_t* cw = plane ? S.vw : S.uw;
                uint32_t* cw32 = plane ? vw32 : uw32;
                const unsigned gmask = plane ? 0xffff0000u : 0x0000ffffu;
                int pr[4];
                if (uvmode == 0) {
                    int v;
                    if (j < 8) v = mby > 0 ? (int)cw[4 + j] : 0;
                    else v = mbx > 0 ? (int)cw[(1 + j - 8) * 16 + 3] : 0;
                    const int sum = __reduce_add_sync(gmask, (unsigned)v);
                    const int shf = 2 + (mby > 0) + (mbx > 0);
                    const int dc = (mby == 0 && mbx == 0) ? 128 : (sum + (1 << (shf - 1))) >> shf;
                    pr[0] = pr[1] = pr[2] = pr[3] = dc;
                } else if (uvmode == 1) {
#pragma unroll
                    for (int i = 0; i < 4; i++) pr[i] = cw[4 + x0 + i];
                } else if (uvmode == 2) {
                    const int l = cw[(1 + row) * 16 + 3];
                    pr[0] = pr[1] = pr[2] = pr[3] = l;
                } else {
                    const int lmp = (int)cw[(1 + row) * 16 + 3] - (int)cw[3];
#pragma unroll
                    for (int i = 0; i < 4; i++) pr[i] = clamp255(lmp + (int)cw[4 + x0 + i]);
                }
                __syncwarp();
                const int blk = 16 + plane * 4 + (row >> 2) * 2 + (x0 >> 2);
                const uint2 r = *reinterpret_cast<const uint2*>(res + blk * 16 + (row & 3) * 4);
                const int rr[4] = {(int)(short)(r.x & 0xffff), (int)r.x >> 16, (int)(short)(r.y & 0xffff), (int)r.y >> 16};
                const uint32_t o = clamp_pack4(pr[0] + rr[0], pr[1] + rr[1], pr[2] + rr[2], pr[3] + rr[3]);
                cw32[(1 + row) * 4 + 1 + (x0 >> 2)] = o;
                *reinterpret_cast<uint32_t*>(cout + (size_t)mbx * 8) = o;
            }
            __syncwarp();
            // right column becomes the next macroblock's left border (mod.rs:700-704, 1309-1314)
            *lb_dst = lb_dst[lb_gap];
            // bottom row -> edge record for the row below (unfiltered pixels, as intra_predict needs them)
            if (hand_down && lane < 8) st_edge(edge_mine + (size_t)mbx * 8, *bottom_src, tag_mine);
            __syncwarp();
            WF_STAMP(2, mbx);
#ifdef WF_TRACE
            if (mbx == 0 && rid >= (uint32_t)(WF_TRACE) && rid < (uint32_t)(WF_TRACE) + 4 && lane == 0) { unsigned sm; asm volatile("mov.u32 %0, %smid;" : "=r"(sm)); g_wf_trace[rid - (WF_TRACE)][3][0] = sm; }
#endif
        }
    }
}

// ---------------------------------------------------------------------------------------------
// k_loopfilter
// ---------------------------------------------------------------------------------------------
// loop_filter.rs:4-23
// (the clamps as single saturating conversions: c(v) = sat_s8(v), s2u(v) = c(v) + 128 = sat_u8(v + 128))
__device__ __forceinline__ int lf_c(int v) {
    int r;
    asm("cvt.sat.s8.s32 %0, %1;" : "=r"(r) : "r"(v));
    return r;
}
__device__ __forceinline__ int u2s(int v) { return v - 128; }
__device__ __forceinline__ int s2u(int v) {
    int r;
    asm("cvt.sat.u8.s32 %0, %1;" : "=r"(r) : "r"(v + 128));
    return r;
}

// One 8-pixel line across an edge; q0 is at p[0], neighbours at multiples of `st` bytes.
// kind 0: simple_segment (loop_filter.rs:144-162), 1: subblock_filter (167-206),
// 2: macroblock_filter (212-295); thresholds per 79-139; common_adjust 28-76.
__device__ __forceinline__ void lf_line(uint8_t* p, int st, int kind, int edge_limit, int interior, int hev_thr) {
    const int P1 = p[-2 * st], P0 = p[-st], Q0 = p[0], Q1 = p[st];
    {   // |p0-q0| * 2 + |p1-q1| / 2 (loop_filter.rs:79-83) as sum-of-absolute-difference instructions
        const unsigned d0 = __usad(P0, Q0, 0u);
        if (__usad(P0, Q0, d0 + (__usad(P1, Q1, 0u) >> 1)) > (unsigned)edge_limit) return;
    }
    int P3 = 0, P2 = 0, Q2 = 0, Q3 = 0;
    bool hv = true;  // the simple filter always uses the outer taps
    if (kind != 0) {
        P3 = p[-4 * st]; P2 = p[-3 * st]; Q2 = p[2 * st]; Q3 = p[3 * st];
        const unsigned dp = __usad(P1, P0, 0u), dq = __usad(Q1, Q0, 0u), lim = (unsigned)interior;
        if (__usad(P3, P2, 0u) > lim || __usad(P2, P1, 0u) > lim || dp > lim || __usad(Q3, Q2, 0u) > lim || __usad(Q2, Q1, 0u) > lim || dq > lim)
            return;
        hv = dp > (unsigned)hev_thr || dq > (unsigned)hev_thr;
    }
    const int p1 = u2s(P1), p0 = u2s(P0), q0 = u2s(Q0), q1 = u2s(Q1);
    if (kind == 2 && !hv) {
        const int p2 = u2s(P2), q2 = u2s(Q2);
        const int w = lf_c(lf_c(p1 - q1) + 3 * (q0 - p0));
        int a = lf_c((27 * w + 63) >> 7);
        p[0] = (uint8_t)s2u(q0 - a);
        p[-st] = (uint8_t)s2u(p0 + a);
        a = lf_c((18 * w + 63) >> 7);
        p[st] = (uint8_t)s2u(q1 - a);
        p[-2 * st] = (uint8_t)s2u(p1 + a);
        a = lf_c((9 * w + 63) >> 7);
        p[2 * st] = (uint8_t)s2u(q2 - a);
        p[-3 * st] = (uint8_t)s2u(p2 + a);
        return;
    }
    // common_adjust(use_outer_taps = hv)
    const int outer = hv ? lf_c(p1 - q1) : 0;
    int a = lf_c(outer + 3 * (q0 - p0));
    const int b = lf_c(a + 3) >> 3;
    a = lf_c(a + 4) >> 3;
    p[0] = (uint8_t)s2u(q0 - a);
    p[-st] = (uint8_t)s2u(p0 + b);
    if (kind == 1 && !hv) {
        a = (a + 1) >> 1;
        p[st] = (uint8_t)s2u(q1 - a);
        p[-2 * st] = (uint8_t)s2u(p1 + a);
    }
}

// Per-warp tile: luma rows -4..15 of the macroblock row, two 16-column panels side by side
// (32-byte rows): the previous macroblock of the row on the left, the current one on the right.
// Chroma: 12 rows, two 8-column panels (16-byte rows), per plane.  The previous panel is written
// back (one 16-byte store per row) only after the current macroblock's left edge has been filtered,
// then the current panel slides left.  The 4 top rows belong to the row above: they arrive through
// that row's edge record (32 units: 16 luma, 8 U, 8 V; one per lane), are filtered here (top
// macroblock edge) and written to the plane by this row; the bottom 4 rows of this row go the same
// way to the row below instead of to the plane (the last row of a frame writes its own).
struct __align__(16) FilterSmem {
    uint8_t y[20 * 32];
    uint8_t c[2][12 * 16];
    uint4 ynext[20];     // the next macroblock's luma columns, prefetched with cp.async (L2 path) while this one is filtered
};

struct FilterArgs {
    const webpb200_frame_desc* frames;
    const webpb200_mb_desc* mbs;
    uint8_t* planes;
    unsigned int* ticket;
    unsigned int* abort_word;
    uint2* edges;             // edge records, 32 units per slot
    const uint32_t* recbase;  // row id -> first slot of that row
    const uint32_t* rowmap;
    const uint32_t* order;
    const uint8_t* mbinfo;
    uint32_t total_rows;
};

// Two register budgets, like k_predict: four blocks per SM (56 registers) when every resident warp has work -- more
// warps to hide latency behind: cfg2 0.504 -> 0.467 ms, cfg3 256 x 4K 8.85 -> 8.27 ms -- and three (72 registers, fewer
// instructions on each macroblock's dependency chain) when the rows of a few large frames wait on each other (32 x 4K:
// 1.76 ms against 1.93).  Five blocks (46 registers) and six (40, spills) are slower on both.
#ifndef FILTER_BLOCKS_T
#define FILTER_BLOCKS_T 4
#endif
#ifndef FILTER_BLOCKS_L
#define FILTER_BLOCKS_L 3
#endif
constexpr int kFilterBlocksThroughput = FILTER_BLOCKS_T, kFilterBlocksLatency = FILTER_BLOCKS_L;

template <int kMinBlocks>
__global__ void __launch_bounds__(kThreads, kMinBlocks) k_loopfilter(FilterArgs A) {
    __shared__ FilterSmem smem[kWarpsPerBlock];
    const int lane = threadIdx.x & 31;
    FilterSmem& S = smem[threadIdx.x >> 5];

    for (;;) {
        unsigned int t = 0;
        if (lane == 0) t = atomicAdd(A.ticket, 1u);
        t = __shfl_sync(kFull, t, 0);
        if (t >= A.total_rows) break;
        const uint32_t rid = __ldg(A.order + t);
        const uint32_t img = __ldg(A.rowmap + rid);
        const webpb200_frame_desc* fd = A.frames + img;
        const int frame_level = fd->filter_level;
        if (frame_level == 0) continue;  // mod.rs:1195-1198: nothing is filtered in this frame (no row of it runs, none waits)
        const int mbw = fd->mbw;
        const int mby = (int)(rid - fd->row_base);
        const size_t ystride = (size_t)mbw * 16, cstride = (size_t)mbw * 8;
        const size_t mb_row0 = (size_t)fd->mb_base + (size_t)mby * mbw;
        const bool simple = fd->filter_type != 0;
        const int sharp = fd->sharpness;
        const int row_lo = mby > 0 ? 0 : 4;  // first tile row that exists
        const bool hand_down = mby + 1 < (int)fd->mbh;  // my bottom 4 rows go to the row below, not to the plane
        // my row of the tile in the planes: lanes 0..19 luma rows, lanes 0..23 chroma (plane, row)
        uint8_t* const yrow = A.planes + fd->y_off + (size_t)(mby * 16 - 4 + lane) * ystride;
        const int cpl = lane >= 12, crow = lane - 12 * cpl;
        uint8_t* const crow_ptr = A.planes + (cpl ? fd->v_off : fd->u_off) + (size_t)(mby * 8 - 4 + crow) * cstride;
        const bool yown = lane >= 4 && lane < 20;                 // lanes that fetch this row's own pixels
        const bool cown = !simple && lane < 24 && crow >= 4;
        const bool clane = !simple && lane < 24;
        uint4* const ytile = reinterpret_cast<uint4*>(S.y + lane * 32);            // [0] previous, [1] current panel
        uint2* const ctile = reinterpret_cast<uint2*>(S.c[cpl] + crow * 16);
        // edge records.  Consumer side: lane u takes unit u of the record above the current macroblock
        // into the current panel (units 0..15: luma tile row u/4, word u%4; 16..23: U row, word; 24..31: V).
        const uint32_t tag_above = (uint32_t)mby, tag_mine = (uint32_t)mby + 1u;
        const bool up_lane = mby > 0 && (lane < 16 || !simple);
        uint32_t* const up_dst = lane < 16 ? reinterpret_cast<uint32_t*>(S.y + (lane >> 2) * 32 + 16) + (lane & 3)
                                           : reinterpret_cast<uint32_t*>(S.c[(lane - 16) >> 3] + (((lane - 16) & 7) >> 1) * 16 + 8) + (lane & 1);
        const uint2* const edge_above = A.edges + (size_t)__ldg(A.recbase + (mby > 0 ? rid - 1 : rid)) * 32 + lane;
        // Producer side: the lane that owns a bottom tile row writes that row's units itself (it holds them
        // in registers for the write-back anyway): luma tile row 16 + r -> units 4r..4r+3, chroma tile row
        // 8 + r of plane p -> units 16 + 8p + 2r, + 1.
        uint2* const edge_mine = A.edges + (size_t)__ldg(A.recbase + rid) * 32;
        uint2* const yedge = edge_mine + (lane - 16) * 4;
        uint2* const cedge = edge_mine + 16 + cpl * 8 + (crow - 8) * 2;
        uint2 upv = make_uint2(0u, 0u);
        if (up_lane) upv = ld_edge(edge_above);

        // calculate_filter_parameters, mod.rs:1191-1217, once per row: the level of a macroblock
        // depends only on its segment and on whether it is B_PRED -> 2 x 4 bytes
        uint32_t lvl_plain = 0, lvl_bpred = 0;
        for (int sgi = 0; sgi < 4; sgi++) {
            int level = frame_level;
            if (fd->segments_enabled) {
                const int sl = fd->seg[sgi].lf_level;
                level = fd->seg[sgi].delta_values ? level + sl : sl;
            }
            level = min(max(level, 0), 63);
            int lb = level;
            if (fd->lf_adj_enabled) {
                level += fd->ref_delta0;
                lb = level + fd->mode_delta0;
            }
            lvl_plain |= (uint32_t)min(max(level, 0), 63) << (8 * sgi);
            lvl_bpred |= (uint32_t)min(max(lb, 0), 63) << (8 * sgi);
        }

        // This row's own pixels were written by k_predict (an earlier launch): macroblock x + 1 is always
        // prefetched while x is filtered.  Chroma in registers (an 8-byte cp.async would go through L1).
        uint2 cnext = make_uint2(0, 0);
        uint32_t mbword = __ldg(reinterpret_cast<const uint32_t*>(A.mbs + mb_row0));
        uint32_t mbnz = A.mbinfo[mb_row0];

        for (int mbx = 0; mbx < mbw; mbx++) {
            // ---- the 16 new columns into the current panel: rows 0..15 from the plane ... ----
            if (mbx > 0) {
                cp_async_wait_all();
                if (yown) ytile[1] = S.ynext[lane];
                if (cown) ctile[1] = cnext;
            } else {
                if (yown) ytile[1] = ld_cg_u128(yrow);
                if (cown) ctile[1] = ld_cg_u64(crow_ptr);
            }
            if (mbx + 1 < mbw) {
                if (yown) cp_async16(&S.ynext[lane], yrow + (size_t)(mbx + 1) * 16);
                if (cown) cnext = ld_cg_u64(crow_ptr + (size_t)(mbx + 1) * 8);
                cp_async_commit();
            }
            // ---- ... rows -4..-1 from the edge record of the macroblock above (there once the row above
            // has filtered the left edge of ITS next macroblock); the record after it is requested now ----
#ifndef LF_WAIT_UNIFORM
#define LF_WAIT_UNIFORM 0   // 1: the normal filter waits by vote too (A/B)
#endif
            if (mby > 0) {
                if (simple || LF_WAIT_UNIFORM) {  // half of the lanes have nothing to wait for: wait as one warp
                    const uint32_t strip = wait_edge(edge_above + (size_t)mbx * 32, upv, tag_above, up_lane, A.abort_word);
                    if (up_lane) {
                        *up_dst = strip;
                        if (mbx + 1 < mbw) upv = ld_edge(edge_above + (size_t)(mbx + 1) * 32);
                    }
                } else {                          // every lane has its unit
                    *up_dst = wait_edge_each(edge_above + (size_t)mbx * 32, upv, tag_above, A.abort_word);
                    if (mbx + 1 < mbw) upv = ld_edge(edge_above + (size_t)(mbx + 1) * 32);
                }
            }

            const int ymode = mbword & 0xff, segment = (mbword >> 16) & 3;
            const bool skipped = (mbword >> 24) & 1;
            const bool do_sub = ymode == 4 || (!skipped && mbnz != 0);  // mod.rs:977-978
            const int level = ((ymode == 4 ? lvl_bpred : lvl_plain) >> (8 * segment)) & 0xff;
            if (mbx + 1 < mbw) {  // next macroblock's descriptor word and non_zero_dct flag, early
                mbword = __ldg(reinterpret_cast<const uint32_t*>(A.mbs + mb_row0 + mbx + 1));
                mbnz = A.mbinfo[mb_row0 + mbx + 1];
            }
            // interior limit and thresholds, mod.rs:1219-1243
            int interior = level;
            if (sharp > 0) {
                interior >>= sharp > 4 ? 2 : 1;
                interior = min(interior, 9 - sharp);
            }
            if (interior == 0) interior = 1;
            const int hev_thr = level >= 40 ? 2 : level >= 15 ? 1 : 0;
            const int mbedge = (level + 2) * 2 + interior, subedge = level * 2 + interior;
            __syncwarp();

            // The previous macroblock's bottom 4 rows are final as soon as this macroblock's LEFT edge has been
            // filtered (nothing after that touches the previous panel): they go into the edge record right then,
            // not at the end of the step, which takes most of a step off the lag between consecutive rows.
            auto hand_prev = [&]() {
                if (mbx == 0 || !hand_down) return;
                if (lane >= 16 && lane < 20) {
                    const uint4 prv = ytile[0];
                    uint2* r = yedge + (size_t)(mbx - 1) * 32;
                    st_edge(r, prv.x, tag_mine); st_edge(r + 1, prv.y, tag_mine); st_edge(r + 2, prv.z, tag_mine); st_edge(r + 3, prv.w, tag_mine);
                }
                if (clane && crow >= 8) {
                    const uint2 prv = ctile[0];
                    uint2* r = cedge + (size_t)(mbx - 1) * 32;
                    st_edge(r, prv.x, tag_mine); st_edge(r + 1, prv.y, tag_mine);
                }
            };
            // ---- loop_filter, mod.rs:966-1188: four edge groups in the reference's order ----
            if (level == 0) hand_prev();
            if (level > 0) {
                uint8_t* const y0 = S.y + 4 * 32 + 16;  // pixel (0, 0) of the current macroblock
                if (simple) {
                    // the simple filter reads 4 and writes 2 pixels per line, so the parallel
                    // edges of one direction do not touch each other's pixels: do them together
                    const int r = lane & 15, e0 = (lane >> 4) * 2;  // my line, my two edges (e0, e0 + 1)
                    if (e0 == 0 ? mbx > 0 : do_sub) lf_line(y0 + r * 32 + 4 * e0, 1, 0, e0 == 0 ? mbedge : subedge, 0, 0);
                    if (do_sub) lf_line(y0 + r * 32 + 4 * e0 + 4, 1, 0, subedge, 0, 0);
                    __syncwarp();
                    hand_prev();
                    if (e0 == 0 ? mby > 0 : do_sub) lf_line(y0 + (4 * e0) * 32 + r, 32, 0, e0 == 0 ? mbedge : subedge, 0, 0);
                    if (do_sub) lf_line(y0 + (4 * e0 + 4) * 32 + r, 32, 0, subedge, 0, 0);
                    __syncwarp();
                } else {
                    // lanes 0..15: luma lines; 16..23: U lines; 24..31: V lines
                    uint8_t* base;
                    int st_row, idx;
                    if (lane < 16) { base = y0; st_row = 32; idx = lane; }
                    else { base = S.c[(lane - 16) >> 3] + 4 * 16 + 8; st_row = 16; idx = (lane - 16) & 7; }
                    if (mbx > 0) lf_line(base + idx * st_row, 1, 2, mbedge, interior, hev_thr);
                    __syncwarp();
                    hand_prev();
                    if (do_sub) {
                        lf_line(base + idx * st_row + 4, 1, 1, subedge, interior, hev_thr);
                        __syncwarp();
                        if (lane < 16) lf_line(base + idx * st_row + 8, 1, 1, subedge, interior, hev_thr);
                        __syncwarp();
                        if (lane < 16) lf_line(base + idx * st_row + 12, 1, 1, subedge, interior, hev_thr);
                        __syncwarp();
                    }
                    if (mby > 0) lf_line(base + idx, st_row, 2, mbedge, interior, hev_thr);
                    __syncwarp();
                    if (do_sub) {
                        lf_line(base + idx + 4 * st_row, st_row, 1, subedge, interior, hev_thr);
                        __syncwarp();
                        if (lane < 16) lf_line(base + idx + 8 * 32, 32, 1, subedge, interior, hev_thr);
                        __syncwarp();
                        if (lane < 16) lf_line(base + idx + 12 * 32, 32, 1, subedge, interior, hev_thr);
                        __syncwarp();
                    }
                }
            }

            // ---- the previous macroblock is final now: its rows go to the plane, or (bottom 4 rows, if a
            // row below exists) into the edge record; then the current panel slides left ----
            const bool last = mbx == mbw - 1;
            if (lane < 20) {
                const uint4 cur = ytile[1];
                if (lane >= row_lo) {
                    const uint4 prv = ytile[0];
                    if (lane < 16 || !hand_down) {
                        if (mbx > 0) *reinterpret_cast<uint4*>(yrow + (size_t)mbx * 16 - 16) = prv;
                        if (last) *reinterpret_cast<uint4*>(yrow + (size_t)mbx * 16) = cur;
                    } else {  // (the previous panel's went out right after the left edge, see hand_prev)
                        if (last) {
                            uint2* r = yedge + (size_t)mbx * 32;
                            st_edge(r, cur.x, tag_mine); st_edge(r + 1, cur.y, tag_mine); st_edge(r + 2, cur.z, tag_mine); st_edge(r + 3, cur.w, tag_mine);
                        }
                    }
                }
                ytile[0] = cur;
            }
            if (clane) {
                const uint2 cur = ctile[1];
                if (crow >= row_lo) {
                    const uint2 prv = ctile[0];
                    if (crow < 8 || !hand_down) {
                        if (mbx > 0) *reinterpret_cast<uint2*>(crow_ptr + (size_t)mbx * 8 - 8) = prv;
                        if (last) *reinterpret_cast<uint2*>(crow_ptr + (size_t)mbx * 8) = cur;
                    } else {
                        if (last) {
                            uint2* r = cedge + (size_t)mbx * 32;
                            st_edge(r, cur.x, tag_mine); st_edge(r + 1, cur.y, tag_mine);
                        }
                    }
                }
                ctile[0] = cur;
            }
            __syncwarp();
        }
    }
}

// ---------------------------------------------------------------------------------------------
// k_yuv2rgb: one thread = 8 horizontally adjacent pixels of the two rows (2q-1, 2q) that share
// their pair of chroma rows (q = 0: row 0 alone)
// ---------------------------------------------------------------------------------------------
// yuv.rs:30-71: r = clip((mulhi(y, 19077) + mulhi(v, 26149) - 14234) >> 6) etc. with mulhi(v, c) = (v * c) >> 8.
// The kernel is bound by the ALU pipe (shifts, adds, logic), not by the multiplier, so the floors are arranged to
// need one shift per channel: floor(A / 256) + floor(B / 256) == floor(((A & ~255) + B) / 256) because the masked
// term is a multiple of 256; the constants are multiples of 256 and go into the masked term; the final >> 6 merges
// with the >> 8 (nested floors); a SUBTRACTED term uses -floor(B / 256) == floor((255 - B) / 256), i.e. the running
// sum is OR-ed with 255 instead of masked.  (Tried and slower: the high half of (v << 24) * c as the mulhi, 0.48 ms
// against 0.37 on cfg2 -- IMAD.HI issues at a quarter of the IMAD rate; the three luma terms from a 256-entry
// shared-memory table, 0.45 ms -- sixteen 16-byte gathers per thread with random bank conflicts.)
__device__ __forceinline__ uint32_t yuv_to_rgb(int y, int u, int v) {  // r | g << 8 | b << 16
    const int yy = y * 19077;
    const int r = (v * 26149 + ((yy - 14234 * 256) & ~255)) >> 14;
    const int g = ((((yy + 8708 * 256) | 255) - u * 6419) | 255) - v * 13320 >> 14;
    const int b = (u * 33050 + ((yy - 17685 * 256) & ~255)) >> 14;
    // clip(x) = clamp(x, 0, 255) (yuv.rs:66-71): the clamps and the packing are two saturating cvt.pack
    uint32_t hi, px;
    asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(hi) : "r"(0), "r"(b), "r"(0));
    asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(px) : "r"(g), "r"(r), "r"(hi));
    return px;
}
__device__ __forceinline__ uint32_t mad_lo(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

struct RgbArgs {
    const webpb200_frame_desc* frames;
    const uint8_t* planes;
    uint8_t* out;
    int q_first, q_end;  // row pairs [q_first, q_end) only (pair q = rows 2q-1, 2q)
    // ALPH merge (decoder.rs:718-734): alpha of a frame with an ALPH plane = the un-filtered plane (alpha_planar) or, for
    // filter method none, the residual plane itself (alpha_res); both null = byte 3 stays 255
    const uint8_t* alpha_res;
    const uint8_t* alpha_planar;
};

// Horizontal half of the fancy upsampler for 8 pixels of one chroma row, TWO pixels per register (16-bit fields:
// pixel i in the low half, pixel i + 4 in the high half; every sum stays below 2^16).  c[0..5] are the chroma
// samples at columns 4k-1 .. 4k+4 (clamped to the visible chroma width, which is how yuv.rs:163-170 and 204-219
// treat the first and last pixel).  Pixel 2j has main column c[j+1] and second column c[j], pixel 2j+1 has
// c[j+1] and c[j+2]; h = 3 * main + second, so that (3 * h_main_row + h_second_row + 8) >> 4 ==
// (9*a + 3*b + 3*c + d + 8) >> 4 of yuv.rs:276-282.  `rep` replicates the last visible column (lim < 3 only).
__device__ __forceinline__ void chroma_row8(const uint8_t* row, int k4, int lim, uint32_t rep, uint32_t* h) {
    uint32_t wd = __ldg(reinterpret_cast<const uint32_t*>(row + k4));
    const uint32_t c0 = __ldg(row + max(k4 - 1, 0)), c5 = __ldg(row + k4 + min(4, lim));
    if (lim < 3) wd = __byte_perm(wd, 0, rep);     // the row's last thread: columns past the visible width repeat the last one
    const uint32_t A = __byte_perm(wd, 0, 0x4240);   // c1 | c3 << 16
    const uint32_t B = __byte_perm(wd, 0, 0x4341);   // c2 | c4 << 16
    const uint32_t L = __byte_perm(wd, c0, 0x5154);  // c0 | c2 << 16
    const uint32_t R = __byte_perm(wd, c5, 0x5452);  // c3 | c5 << 16
    h[0] = 3 * A + L;  // pixels 0 | 4
    h[1] = 3 * A + B;  // pixels 1 | 5
    h[2] = 3 * B + A;  // pixels 2 | 6
    h[3] = 3 * B + R;  // pixels 3 | 7
}
// Vertical half: t = 16 * (3 * main + second + 8), i.e. the upsampled sample is byte 1 (pixel i) and byte 3
// (pixel i + 4) of t: one PRMT each to extract.  (Both products as multiply-adds: the multiplier has room, the ALU pipe not.)
__device__ __forceinline__ void chroma_blend(const uint32_t* hm, const uint32_t* hs, uint32_t* t) {
#pragma unroll
    for (int j = 0; j < 4; j++) t[j] = mad_lo(hs[j], 16u, mad_lo(hm[j], 48u, 0x00800080u));
}

// 8 pixels -> tightly packed RGB8 / RGBA8 (yuv.rs:285-289; byte 3 of RGBA belongs to the ALPH merge
// that always follows, decoder.rs:704-734: 255 until then)
// (al = the 8 alpha bytes of the row segment, RGBA only)
__device__ __forceinline__ void store_row8(uint8_t* out, const uint32_t* px, int n, int bpp, uint2 al) {
    if (bpp == 4) {
        uint32_t q[8];
#pragma unroll
        for (int i = 0; i < 8; i++) q[i] = __byte_perm(px[i], i < 4 ? al.x : al.y, 0x4210 | ((i & 3) << 12));  // r g b | alpha byte i
        if (n == 8 && (reinterpret_cast<uintptr_t>(out) & 15) == 0) {
            uint4* o = reinterpret_cast<uint4*>(out);
            o[0] = make_uint4(q[0], q[1], q[2], q[3]);
            o[1] = make_uint4(q[4], q[5], q[6], q[7]);
        } else {
#pragma unroll
            for (int i = 0; i < 8; i++)
                if (i < n) reinterpret_cast<uint32_t*>(out)[i] = q[i];
        }
        return;
    }
    if (n == 8 && (reinterpret_cast<uintptr_t>(out) & 3) == 0) {
        uint32_t wv[6];
#pragma unroll
        for (int j = 0; j < 2; j++) {
            const uint32_t a = px[4 * j], b = px[4 * j + 1], c = px[4 * j + 2], d = px[4 * j + 3];
            wv[3 * j] = __byte_perm(a, b, 0x4210);      // a0 a1 a2 b0
            wv[3 * j + 1] = __byte_perm(b, c, 0x5421);  // b1 b2 c0 c1
            wv[3 * j + 2] = __byte_perm(c, d, 0x6542);  // c2 d0 d1 d2
        }
        if ((reinterpret_cast<uintptr_t>(out) & 7) == 0) {
            uint2* o = reinterpret_cast<uint2*>(out);
            o[0] = make_uint2(wv[0], wv[1]); o[1] = make_uint2(wv[2], wv[3]); o[2] = make_uint2(wv[4], wv[5]);
        } else {
            uint32_t* o = reinterpret_cast<uint32_t*>(out);
#pragma unroll
            for (int j = 0; j < 6; j++) o[j] = wv[j];
        }
    } else {
        for (int i = 0; i < n; i++) { out[3 * i] = (uint8_t)px[i]; out[3 * i + 1] = (uint8_t)(px[i] >> 8); out[3 * i + 2] = (uint8_t)(px[i] >> 16); }
    }
}

// the 8 alpha bytes at p (n valid), by one 8-byte load when the address allows
__device__ __forceinline__ uint2 load_alpha8(const uint8_t* p, int n) {
    if (n == 8 && (reinterpret_cast<uintptr_t>(p) & 7) == 0) return __ldg(reinterpret_cast<const uint2*>(p));
    uint32_t lo = 0, hi = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        if (i < n) lo |= (uint32_t)__ldg(p + i) << (8 * i);
        if (i + 4 < n) hi |= (uint32_t)__ldg(p + i + 4) << (8 * i);
    }
    return make_uint2(lo, hi);
}

#ifndef RGB_PAIRS_PER_BLOCK
#define RGB_PAIRS_PER_BLOCK 2
#endif
constexpr int kRgbPairsPerBlock = RGB_PAIRS_PER_BLOCK;  // block = 64 x 2 threads = 512 pixels x 4 rows
__global__ void __launch_bounds__(64 * kRgbPairsPerBlock) k_yuv2rgb(RgbArgs A) {
    const webpb200_frame_desc* fd = A.frames + blockIdx.z;
    const int w = fd->width, h = fd->height;
    const int q = A.q_first + blockIdx.y * kRgbPairsPerBlock + threadIdx.y;
    const int x0 = (blockIdx.x * 64 + threadIdx.x) * 8;
    const int ya = q == 0 ? 0 : 2 * q - 1, yb = 2 * q;
    if (x0 >= w || ya >= h || q >= A.q_end) return;
    const bool has_b = q > 0 && yb < h;
    const int ystride = fd->mbw * 16, cstride = fd->mbw * 8;
    const int bpp = fd->out_bpp;
    const int n = min(8, w - x0);
    const int k4 = x0 >> 1;  // first chroma column of this thread
    const uint8_t* Y = A.planes + fd->y_off + x0;
    const uint2 ywa = __ldg(reinterpret_cast<const uint2*>(Y + (size_t)ya * ystride));
    uint2 ywb = make_uint2(0, 0);
    if (has_b) ywb = __ldg(reinterpret_cast<const uint2*>(Y + (size_t)yb * ystride));
    uint32_t ua[8], va[8], ub[8], vb[8];
    if (fd->upsampling == WEBPB200_UPSAMPLING_SIMPLE) {  // fill_rgb_buffer_simple, yuv.rs:293-382
        const uint32_t u0 = __ldg(reinterpret_cast<const uint32_t*>(A.planes + fd->u_off + (size_t)(ya >> 1) * cstride + k4));
        const uint32_t v0 = __ldg(reinterpret_cast<const uint32_t*>(A.planes + fd->v_off + (size_t)(ya >> 1) * cstride + k4));
        const int rb = min(yb >> 1, ((h + 1) >> 1) - 1);  // row 2q may not exist: stay inside the plane
        const uint32_t u1 = __ldg(reinterpret_cast<const uint32_t*>(A.planes + fd->u_off + (size_t)rb * cstride + k4));
        const uint32_t v1 = __ldg(reinterpret_cast<const uint32_t*>(A.planes + fd->v_off + (size_t)rb * cstride + k4));
#pragma unroll
        for (int i = 0; i < 8; i++) {
            ua[i] = __byte_perm(u0, 0, 0x4440 | (i >> 1)); va[i] = __byte_perm(v0, 0, 0x4440 | (i >> 1));
            ub[i] = __byte_perm(u1, 0, 0x4440 | (i >> 1)); vb[i] = __byte_perm(v1, 0, 0x4440 | (i >> 1));
        }
    } else {
        // fill_rgb_buffer_fancy, yuv.rs:76-282.  Row 2q-1 takes chroma row q-1 as main and q as second,
        // row 2q the other way round; row 0 and an even-height image's last row use one chroma row
        // for both (yuv.rs:94, 137-151) -- which the clamp of the second row's index gives.
        const int cw = (w + 1) >> 1, chh = (h + 1) >> 1;
        const int ca = q == 0 ? 0 : q - 1, cb = min(q, chh - 1);
        const int lim = cw - 1 - k4;  // last visible chroma column relative to k4
        const uint32_t rep = lim == 0 ? 0x0000u : lim == 1 ? 0x1110u : 0x2210u;
        uint32_t ha[4], hb[4], ta[4], tb[4];
        const uint8_t* U = A.planes + fd->u_off;
        chroma_row8(U + (size_t)ca * cstride, k4, lim, rep, ha);
        chroma_row8(U + (size_t)cb * cstride, k4, lim, rep, hb);
        chroma_blend(ha, hb, ta);
        chroma_blend(hb, ha, tb);
#pragma unroll
        for (int i = 0; i < 4; i++) {
            ua[i] = __byte_perm(ta[i], 0, 0x4441); ua[i + 4] = ta[i] >> 24;
            ub[i] = __byte_perm(tb[i], 0, 0x4441); ub[i + 4] = tb[i] >> 24;
        }
        const uint8_t* V = A.planes + fd->v_off;
        chroma_row8(V + (size_t)ca * cstride, k4, lim, rep, ha);
        chroma_row8(V + (size_t)cb * cstride, k4, lim, rep, hb);
        chroma_blend(ha, hb, ta);
        chroma_blend(hb, ha, tb);
#pragma unroll
        for (int i = 0; i < 4; i++) {
            va[i] = __byte_perm(ta[i], 0, 0x4441); va[i + 4] = ta[i] >> 24;
            vb[i] = __byte_perm(tb[i], 0, 0x4441); vb[i + 4] = tb[i] >> 24;
        }
    }
    uint32_t px[8];
#pragma unroll
    for (int i = 0; i < 8; i++) px[i] = yuv_to_rgb((int)__byte_perm(i < 4 ? ywa.x : ywa.y, 0, 0x4440 | (i & 3)), (int)ua[i], (int)va[i]);
    // ALPH merge: the plane's bytes for this row segment, or 255
    const uint8_t* asrc = nullptr;
    if (bpp == 4 && fd->alpha_mode) asrc = fd->alpha_filter == 0 ? A.alpha_res : A.alpha_planar;
    if (asrc) asrc += fd->alpha_off + x0;
    const uint2 opaque = make_uint2(0xffffffffu, 0xffffffffu);
    store_row8(A.out + fd->out_off + ((size_t)ya * w + x0) * bpp, px, n, bpp, asrc ? load_alpha8(asrc + (size_t)ya * w, n) : opaque);
    if (has_b) {
#pragma unroll
        for (int i = 0; i < 8; i++) px[i] = yuv_to_rgb((int)__byte_perm(i < 4 ? ywb.x : ywb.y, 0, 0x4440 | (i & 3)), (int)ub[i], (int)vb[i]);
        store_row8(A.out + fd->out_off + ((size_t)yb * w + x0) * bpp, px, n, bpp, asrc ? load_alpha8(asrc + (size_t)yb * w, n) : opaque);
    }
}

// ---------------------------------------------------------------------------------------------
// ALPH un-filter + merge (decoder.rs:718-734, get_alpha_predictor extended.rs:131-191).
// alpha = predictor + residual (mod 256), predictor from already-merged neighbours.
//   none:       alpha = residual
//   horizontal: row prefix sums, each row started from the pixel above its first pixel
//   vertical:   column prefix sums, row 0 is a horizontal prefix sum
//   gradient:   clamp(L + T - TL) -- a true 2-D recurrence, done as a row-skewed wavefront
// ---------------------------------------------------------------------------------------------
// The un-filtered alpha goes to a PLANAR arena (same offsets as the residual arena); k_yuv2rgb, which runs after
// these kernels, puts it into byte 3 of the RGBA words it writes anyway.  (Writing byte 3 of finished pixels from
// here cost a 4-sector store per 32 useful bytes and a second pass over the output.)  Filter "none" has no kernel:
// k_yuv2rgb reads the residual plane itself.
struct AlphaArgs {
    const webpb200_frame_desc* frames;
    const uint8_t* alpha;
    uint8_t* out;     // planar alpha arena
    const int* list;  // frame indices handled by this launch
};

// first column (horizontal filter) / first row (vertical filter): one warp per image does the
// serial prefix sum in 32-element steps
__global__ void k_alpha_seed(AlphaArgs A, int vertical) {
    const webpb200_frame_desc* fd = A.frames + A.list[blockIdx.x];
    const int w = fd->width, h = fd->height;
    const uint8_t* a = A.alpha + fd->alpha_off;
    uint8_t* o = A.out + fd->alpha_off;
    const int lane = threadIdx.x;
    const int n = vertical ? w : h;
    const size_t stride = vertical ? 1 : (size_t)w;
    unsigned carry = 0;
    for (int base = 0; base < n; base += 32) {
        const int i = base + lane;
        unsigned v = i < n ? a[(size_t)i * stride] : 0;
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned u = __shfl_up_sync(kFull, v, d);
            if (lane >= d) v += u;
        }
        v += carry;
        if (i < n) o[(size_t)i * stride] = (uint8_t)v;
        carry = __shfl_sync(kFull, v, 31) & 0xff;
    }
}

// horizontal: one warp per row, continuing from the seeded first pixel
__global__ void k_alpha_rows(AlphaArgs A) {
    const webpb200_frame_desc* fd = A.frames + A.list[blockIdx.y];
    const int w = fd->width, h = fd->height;
    const int y = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (y >= h) return;
    const int lane = threadIdx.x & 31;
    const uint8_t* a = A.alpha + fd->alpha_off + (size_t)y * w;
    uint8_t* o = A.out + fd->alpha_off + (size_t)y * w;
    unsigned carry = o[0];
    for (int base = 1; base < w; base += 32) {
        const int x = base + lane;
        unsigned v = x < w ? a[x] : 0;
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned u = __shfl_up_sync(kFull, v, d);
            if (lane >= d) v += u;
        }
        v += carry;
        if (x < w) o[x] = (uint8_t)v;
        carry = __shfl_sync(kFull, v, 31) & 0xff;
    }
}

// vertical: one thread per column, walking down from the seeded first row
__global__ void k_alpha_cols(AlphaArgs A) {
    const webpb200_frame_desc* fd = A.frames + A.list[blockIdx.y];
    const int w = fd->width, h = fd->height;
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= w) return;
    const uint8_t* a = A.alpha + fd->alpha_off;
    uint8_t* o = A.out + fd->alpha_off;
    unsigned v = o[x];
    for (int y0 = 1; y0 < h; y0 += 8) {  // 8 independent loads in flight, then the serial sums
        unsigned r[8];
#pragma unroll
        for (int i = 0; i < 8; i++) r[i] = y0 + i < h ? a[(size_t)(y0 + i) * w + x] : 0;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (y0 + i < h) {
                v = (v + r[i]) & 0xff;
                o[(size_t)(y0 + i) * w + x] = (uint8_t)v;
            }
        }
    }
}

// gradient: clamp(L + T - TL) is a recurrence in both directions.  A warp owns a band of 32 rows,
// lane j = row j running j columns behind lane j-1, so T and TL arrive by shuffle (anti-diagonal
// order).  Bands are pipelined down the image: lane 31 hands each finished pixel of the band's
// last row to the band below through the hand-off array, as a 16-bit word that carries its own
// valid bit (alpha | 0x100, array cleared before the launch) -- one relaxed store, one relaxed
// load, no fence, and the consumer trails the producer by about two tiles.  Bands are handed out
// by a ticket counter, band-major across the images of the launch, so a band only ever waits for
// a ticket drawn earlier.  Residuals and results are staged in 32x32-byte tiles: global traffic
// is one coalesced row segment per instruction, not 32 scattered bytes.
constexpr int kGradWarps = 4;

struct GradArgs {
    AlphaArgs a;
    int n_img, max_bands;
    uint16_t* handoff;
    unsigned int* ticket;
    unsigned int* abort_word;
};

struct __align__(16) GradSmem {
    uint8_t tile[2][32 * 32];  // [tile & 1][row][column]: residual in, alpha out (in place)
    uint8_t top[2][32];        // [tile & 1]: row above the band, columns 32c .. 32c+31
};

__device__ __forceinline__ unsigned ld_relaxed_u16(const uint16_t* p) {
    unsigned short v;
    asm volatile("ld.relaxed.gpu.global.u16 %0, [%1];" : "=h"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u16(uint16_t* p, unsigned v) {
    asm volatile("st.relaxed.gpu.global.u16 [%0], %1;" ::"l"(p), "h"((unsigned short)v) : "memory");
}

__global__ void __launch_bounds__(kGradWarps * 32, 4) k_alpha_gradient(GradArgs G) {
    __shared__ GradSmem sm[kGradWarps];
    const int lane = threadIdx.x & 31;
    GradSmem& S = sm[threadIdx.x >> 5];
    const unsigned total = (unsigned)G.n_img * (unsigned)G.max_bands;

    for (;;) {
        unsigned t = 0;
        if (lane == 0) t = atomicAdd(G.ticket, 1u);
        t = __shfl_sync(kFull, t, 0);
        if (t >= total) return;
        const int g = (int)(t / (unsigned)G.n_img);
        const webpb200_frame_desc* fd = G.a.frames + G.a.list[t % (unsigned)G.n_img];
        const int w = fd->width, h = fd->height;
        const int nbands = (h + 31) >> 5, ntiles = (w + 31) >> 5;
        if (g >= nbands) continue;
        const uint8_t* a = G.a.alpha + fd->alpha_off;
        uint8_t* o = G.a.out + fd->alpha_off;
        uint16_t* hand = G.handoff + (fd->alpha_off >> 4);  // row g of it = last row of band g
        const uint16_t* above = hand + (size_t)(g - 1) * w;
        uint16_t* below = g + 1 < nbands ? hand + (size_t)g * w : nullptr;  // written by lane 31 only

        const int y0 = g * 32, nrows = min(32, h - y0);
        const unsigned wl = lane < nrows ? (unsigned)w : 0u;  // columns this lane works on (0 = idle row)
        const bool hands = lane == 31 && below != nullptr;
        // Zero start values make the edge rules of extended.rs:163-170 fall out of the general one:
        // x == 0 -> L = TL = 0 -> pred = T (0 on the first row); y == 0 -> T = TL = 0 -> pred = L.
        int cur = 0;       // my alpha at column x-1 (L)
        int TL = 0;        // T of the previous step
        int flushed = 0;   // tiles of this band already written out
        unsigned nxt = 0;  // hand-off word of the next tile's column, fetched early
        uint8_t pre[32];   // residual tile fetched one tile ahead (lane = column, index = row)

        auto fetch = [&](int c) {
            const int x = 32 * c + lane;
#pragma unroll
            for (int r = 0; r < 32; r++) pre[r] = (r < nrows && x < w) ? a[(size_t)(y0 + r) * w + x] : 0;
        };
        auto flush = [&](int c) {  // finished tile c -> the planar alpha arena
            const int x = 32 * c + lane;
            const uint8_t* tl = S.tile[c & 1];
            if (x < w) {
#pragma unroll 8
                for (int r = 0; r < nrows; r++) o[(size_t)(y0 + r) * w + x] = tl[r * 32 + lane];
            }
        };

        fetch(0);
        if (g > 0 && lane < w) nxt = ld_relaxed_u16(above + lane);
        if (g == 0) S.top[0][lane] = 0, S.top[1][lane] = 0;
        const int last_tile = (w + 30) >> 5;  // lane 31 reaches column w-1 at step w+30
        for (int c = 0; c <= last_tile; c++) {
            // ---- lane 0 enters tile c
            __syncwarp();
            if (c >= 2) {
                flush(c - 2);
                flushed = c - 1;
            }
            if (c < ntiles) {
                uint8_t* tl = S.tile[c & 1];
#pragma unroll
                for (int r = 0; r < 32; r++) tl[r * 32 + lane] = pre[r];
                if (c + 1 < ntiles) fetch(c + 1);
                if (g > 0) {  // the row above this tile, once the band above has produced it
                    const int x = 32 * c + lane;
                    unsigned v = x < w ? nxt : 0x100u;
                    unsigned spins = 0;
                    while (!__all_sync(kFull, v & 0x100u)) {
                        __nanosleep(40);
                        if (x < w && !(v & 0x100u)) v = ld_relaxed_u16(above + x);
                        if ((++spins & 0x3ff) == 0) {
                            if (spins > (1u << 24)) atomicExch(G.abort_word, 1u);
                            if (*reinterpret_cast<volatile unsigned int*>(G.abort_word)) return;
                        }
                    }
                    S.top[c & 1][lane] = (uint8_t)v;
                    if (x + 32 < w) nxt = ld_relaxed_u16(above + x + 32);
                }
            }
            __syncwarp();
            // ---- 32 anti-diagonal steps
            const uint8_t* top = S.top[c & 1];
            const unsigned row = (unsigned)__cvta_generic_to_shared(S.tile[0] + lane * 32);
            int x = 32 * c - lane;
#pragma unroll 4
            for (int i = 0; i < 32; i++, x++) {
                const int up = __shfl_up_sync(kFull, cur, 1);
                const int T = lane == 0 ? (int)top[i] : up;
                const unsigned cell = row + (x & 31) + ((x & 32) << 5);
                unsigned res;
                asm volatile("ld.shared.u8 %0, [%1];" : "=r"(res) : "r"(cell));
                const int v = (min(max(cur + T - TL, 0), 255) + (int)res) & 0xff;
                TL = T;
                const unsigned act = (unsigned)x < wl, hand = act & (unsigned)hands;
                cur = act ? v : cur;
                // predicated stores: a branch here would sit on the one-warp critical path
                asm volatile("{ .reg .pred p, q; setp.ne.u32 p, %2, 0; setp.ne.u32 q, %3, 0;\n"
                             "  @p st.shared.u8 [%0], %1;\n"
                             "  @q st.relaxed.gpu.global.u16 [%4], %5; }"
                             :: "r"(cell), "r"(v), "r"(act), "r"(hand), "l"(below + x), "h"((unsigned short)(v | 0x100)) : "memory");
            }
        }
        __syncwarp();
        for (int c = flushed; c < ntiles; c++) flush(c);
    }
}

// ---------------------------------------------------------------------------------------------
// k_composite: extended.rs:30-129 + alpha_blending.rs:28-75, one thread per canvas pixel
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t div255(uint32_t v) { return (((v + 0x80) >> 8) + v + 0x80) >> 8; }

__global__ void k_composite(CompositeArgs a) {
    const uint32_t x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= a.canvas_w || y >= a.canvas_h) return;
    const size_t ci = ((size_t)y * a.canvas_w + x) * 4;
    const bool full = a.fx == 0 && a.fy == 0 && a.fw == a.canvas_w && a.fh == a.canvas_h;
    uint32_t c = *reinterpret_cast<const uint32_t*>(a.canvas + ci);
    const bool in_frame = x >= a.fx && x < a.fx + a.fw && y >= a.fy && y < a.fy + a.fh;
    uint32_t src = 0;
    if (in_frame) {
        const size_t fi = (size_t)(y - a.fy) * a.fw + (x - a.fx);
        if (a.frame_has_alpha) src = *reinterpret_cast<const uint32_t*>(a.frame + fi * 4);
        else src = a.frame[fi * 3] | (a.frame[fi * 3 + 1] << 8) | (a.frame[fi * 3 + 2] << 16) | 0xff000000u;
    }
    if (full && !a.use_blend) {
        c = src;
    } else {
        if (a.has_clear && (full || (x >= a.px && x < a.px + a.pw && y >= a.py && y < a.py + a.ph)))
            c = a.clear[0] | (a.clear[1] << 8) | (a.clear[2] << 16) | ((uint32_t)a.clear[3] << 24);
        if (in_frame) {
            if (a.frame_has_alpha && a.use_blend) {
                const uint32_t sa = src >> 24;
                if (sa != 0) {
                    const uint32_t da = c >> 24;
                    const uint32_t dfa = div255(da * (255 - sa));
                    const uint32_t ba = sa + dfa;
                    const uint32_t scale = (1u << 24) / ba;
                    const uint32_t dfa8 = dfa & 0xff;
                    uint32_t o = ba << 24;
                    for (int k = 0; k < 3; k++) {
                        const uint32_t sc = (src >> (8 * k)) & 0xff, dc = (c >> (8 * k)) & 0xff;
                        o |= (((sc * sa + dc * dfa8) * scale) >> 24 & 0xff) << (8 * k);
                    }
                    c = o;
                }
            } else {
                c = src;
            }
        }
    }
    *reinterpret_cast<uint32_t*>(a.canvas + ci) = c;
    if (a.out) {
        if (a.out_bpp == 4) {
            *reinterpret_cast<uint32_t*>(a.out + ci) = c;
        } else {
            uint8_t* o = a.out + ((size_t)y * a.canvas_w + x) * 3;
            o[0] = (uint8_t)c; o[1] = (uint8_t)(c >> 8); o[2] = (uint8_t)(c >> 16);
        }
    }
}

}  // namespace

void wavefront_blocks_per_sm(int* predict_blocks, int* predict_blocks_latency, int* filter_blocks) {
    int p = 0, l = 0, f = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&p, reinterpret_cast<const void*>(k_predict<kPredictBlocksThroughput, false>), kThreads, 0);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&l, reinterpret_cast<const void*>(k_predict<kPredictBlocksLatency, false>), kThreads, 0);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&f, reinterpret_cast<const void*>(k_loopfilter<kFilterBlocksThroughput>), kThreads, 0);
    *predict_blocks = p < 1 ? 1 : p;
    *predict_blocks_latency = l < 1 ? 1 : l;
    *filter_blocks = f < 1 ? 1 : f;
}

// grid of a persistent wavefront kernel: every block resident, no more blocks than rows to hand out
static int wavefront_grid(const BatchDev& b, int blocks_per_sm) {
    int grid = (blocks_per_sm > 0 ? blocks_per_sm : 1) * (b.sm_count > 0 ? b.sm_count : 1);
    const int need = (int)((b.total_rows + kWarpsPerBlock - 1) / kWarpsPerBlock);
    if (grid > need) grid = need;
    return grid < 1 ? 1 : grid;
}
#ifdef WF_GRID_ENV  // experiments only: grid sizes from the environment
#include <cstdlib>
static int env_grid(const char* name, int dflt) { const char* e = getenv(name); return e && atoi(e) > 0 ? atoi(e) : dflt; }
#define WF_GRID(name, g) env_grid(name, g)
#else
#define WF_GRID(name, g) (g)
#endif

cudaError_t launch_predict(const BatchDev& b, cudaStream_t s) {
    PredictArgs a;
    a.frames = b.frames; a.mbs = b.mbs; a.levels = b.levels; a.planes = b.planes;
    a.ticket = reinterpret_cast<unsigned int*>(b.scratch + b.lay.tickets_off);
    a.abort_word = reinterpret_cast<unsigned int*>(b.scratch + b.lay.tickets_off) + 2;
    a.edges = reinterpret_cast<uint2*>(b.scratch + b.lay.edge_pred_off);
    a.recbase = reinterpret_cast<const uint32_t*>(b.scratch + b.lay.recbase_off);
    a.rowmap = reinterpret_cast<const uint32_t*>(b.scratch + b.lay.rowmap_off);
    a.order = reinterpret_cast<const uint32_t*>(b.scratch + b.lay.order_off) + b.order_first;
    a.mbinfo = b.scratch + b.lay.mbinfo_off;
    a.total_rows = b.total_rows;
    // few rows for the machine (they wait on each other): the build with the shorter dependency chain
    const int lat_blocks = b.predict_blocks_latency > 0 ? b.predict_blocks_latency : 1;
    const bool latency_bound = b.predict_blocks_latency > 0 && (uint64_t)b.total_rows <= 2ull * lat_blocks * (b.sm_count > 0 ? b.sm_count : 1) * kWarpsPerBlock;
    if (latency_bound) {
        if (b.levels_packed) k_predict<kPredictBlocksLatency, true><<<WF_GRID("WF_PGRID", wavefront_grid(b, lat_blocks)), kThreads, 0, s>>>(a);
        else k_predict<kPredictBlocksLatency, false><<<WF_GRID("WF_PGRID", wavefront_grid(b, lat_blocks)), kThreads, 0, s>>>(a);
    } else {
        if (b.levels_packed) k_predict<kPredictBlocksThroughput, true><<<wavefront_grid(b, b.predict_blocks), kThreads, 0, s>>>(a);
        else k_predict<kPredictBlocksThroughput, false><<<wavefront_grid(b, b.predict_blocks), kThreads, 0, s>>>(a);
    }
    return cudaGetLastError();
}

cudaError_t launch_loopfilter(const BatchDev& b, cudaStream_t s) {
    FilterArgs a;
    a.frames = b.frames; a.mbs = b.mbs; a.planes = b.planes;
    a.ticket = reinterpret_cast<unsigned int*>(b.scratch + b.lay.tickets_off) + 1;
    a.abort_word = reinterpret_cast<unsigned int*>(b.scratch + b.lay.tickets_off) + 2;
    a.edges = reinterpret_cast<uint2*>(b.scratch + b.lay.edge_lf_off);
    a.recbase = reinterpret_cast<const uint32_t*>(b.scratch + b.lay.recbase_off);
    a.rowmap = reinterpret_cast<const uint32_t*>(b.scratch + b.lay.rowmap_off);
    a.order = reinterpret_cast<const uint32_t*>(b.scratch + b.lay.order_off) + b.order_first;
    a.mbinfo = b.scratch + b.lay.mbinfo_off;
    a.total_rows = b.total_rows;
#ifndef LF_LATENCY_ROWS
#define LF_LATENCY_ROWS 2   // the latency build up to this many times the rows it keeps resident
#endif
    const int sms = b.sm_count > 0 ? b.sm_count : 1;
    const bool latency_bound = (uint64_t)b.total_rows <= (uint64_t)LF_LATENCY_ROWS * kFilterBlocksLatency * sms * kWarpsPerBlock;
    if (latency_bound) k_loopfilter<kFilterBlocksLatency><<<WF_GRID("WF_FGRID", wavefront_grid(b, std::min(b.filter_blocks, kFilterBlocksLatency))), kThreads, 0, s>>>(a);
    else k_loopfilter<kFilterBlocksThroughput><<<WF_GRID("WF_FGRID", wavefront_grid(b, b.filter_blocks)), kThreads, 0, s>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_yuv2rgb(const BatchDev& b, cudaStream_t s) {
    RgbArgs a;
    a.frames = b.frames + b.frame_first; a.planes = b.planes; a.out = b.out;
    a.alpha_res = b.merge_alpha ? b.alpha : nullptr;
    a.alpha_planar = b.merge_alpha ? b.alpha_planar : nullptr;
    // thread = 8 pixels x 2 rows; block = 64 x 2 threads = 512 pixels x 4 rows; row pairs: h/2 + 1
    const int q_all = (int)b.max_h / 2 + 1;
    a.q_first = b.q_end > b.q_first ? b.q_first : 0;
    a.q_end = b.q_end > b.q_first ? (b.q_end < q_all ? b.q_end : q_all) : q_all;
    if (a.q_end <= a.q_first) return cudaSuccess;
    dim3 grid((b.max_w + 511) / 512, (a.q_end - a.q_first + kRgbPairsPerBlock - 1) / kRgbPairsPerBlock, b.n_frames);
    k_yuv2rgb<<<grid, dim3(64, kRgbPairsPerBlock), 0, s>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_alpha(const BatchDev& b, cudaStream_t s, int* launches) {
    // frames grouped by ALPH filter method; the index lists were uploaded with the batch
    const int* const counts = b.alpha_counts;
    if (counts[1] + counts[2] + counts[3] == 0) return cudaSuccess;  // (filter none: k_yuv2rgb reads the residual plane)
    if (!b.alpha_planar) return cudaErrorInvalidValue;
    AlphaArgs a;
    a.frames = b.frames; a.alpha = b.alpha; a.out = b.alpha_planar;
    if (counts[1]) {
        a.list = b.alpha_lists + b.n_frames;
        k_alpha_seed<<<counts[1], 32, 0, s>>>(a, 0);
        k_alpha_rows<<<dim3((b.alpha_maxh[1] + 7) / 8, counts[1]), 256, 0, s>>>(a);
        *launches += 2;
    }
    if (counts[2]) {
        a.list = b.alpha_lists + 2 * b.n_frames;
        k_alpha_seed<<<counts[2], 32, 0, s>>>(a, 1);
        k_alpha_cols<<<dim3((b.alpha_maxw[2] + 127) / 128, counts[2]), 128, 0, s>>>(a);
        *launches += 2;
    }
    if (counts[3]) {
        a.list = b.alpha_lists + 3 * b.n_frames;
        if (!b.alpha_handoff) return cudaErrorInvalidValue;
        GradArgs g;
        g.a = a;
        g.n_img = counts[3];
        g.max_bands = (int)((b.alpha_maxh[3] + 31) / 32);
        g.handoff = b.alpha_handoff;
        g.ticket = reinterpret_cast<unsigned int*>(b.scratch + b.lay.tickets_off) + 3;
        g.abort_word = reinterpret_cast<unsigned int*>(b.scratch + b.lay.tickets_off) + 2;
        cudaError_t e = cudaMemsetAsync(g.handoff, 0, b.alpha_handoff_bytes, s);
        if (e == cudaSuccess) e = cudaMemsetAsync(g.ticket, 0, 4, s);
        if (e != cudaSuccess) return e;
        const unsigned want = ((unsigned)g.n_img * (unsigned)g.max_bands + kGradWarps - 1) / kGradWarps;
        k_alpha_gradient<<<min(want, (unsigned)b.sm_count * 4u), kGradWarps * 32, 0, s>>>(g);
        ++*launches;
    }
    return cudaGetLastError();
}

#ifdef WF_TRACE
extern "C" int webpb200_debug_wf_trace(long long* out) {
    return (int)cudaMemcpyFromSymbol(out, g_wf_trace, sizeof(g_wf_trace));
}
#endif

cudaError_t launch_composite(const CompositeArgs& a, cudaStream_t s) {
    dim3 grid((a.canvas_w + 255) / 256, a.canvas_h);
    k_composite<<<grid, 256, 0, s>>>(a);
    return cudaGetLastError();
}

}  // namespace webpb200
