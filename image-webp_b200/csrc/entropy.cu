// entropy.cu -- device entropy stage: macroblock headers and coefficient tokens of VP8 keyframes
// decoded on the GPU (sm_100a).
//
// What is computed is exactly host_vp8.cc's macroblock loop (reference lossy/mod.rs:615-671
// read_macroblock_header, 785-872 read_coefficients, 874-963 read_residual_data with its
// neighbour contexts, 1255-1281 the partition-per-row loop and the skip rule; boolean decoder
// lossy/arithmetic_decoder.rs) with the same outputs: webpb200_mb_desc and the raw levels.
//
// A boolean decoder is a serial chain (every symbol needs the range and value the previous one
// left), so nothing inside one partition runs in parallel, here or anywhere.  What the GPU has is
// room: one block per frame, warp 0 walks the first partition (modes), warps 1..P walk the P token
// partitions, each with ONE working lane -- a divergent 32-lane decode would serialise anyway.  A
// batch of a thousand frames keeps a thousand such chains in flight, against one per host core,
// and the 800 bytes per macroblock of levels never cross PCIe.
//
//   modes warp  -> mb_desc in global memory, `modes_done` (shared) counts finished macroblocks
//   token warp p-> rows p, p+P, ... ; waits for modes_done > its macroblock; with P > 1 row y also
//                  waits for row y-1 to have passed the same column (the "top" non-zero context,
//                  mod.rs:124-129), handed over in a shared-memory ring of P+1 context rows
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>

#include "entropy.h"
#include "kernels.cuh"
#include "vp8_constants.h"

// Build-time variants for A/B timing (tools/ab_build.sh); the defaults are the measured best.
// Tried and dropped because they were slower on B200 (every added instruction lengthens the one
// lane's dependency chain): speculative normalisation of both outcomes beside the comparison
// (+7 %), probability rows held in registers via LDS.128 (+8 %), the first four probabilities of
// a row in one 32-bit load (+4 %), requesting bitstream bytes one refill ahead (+3 %), the refill as
// an out-of-line call to shrink the code (+12 %: the call sequence costs more than the fetches it
// saves), a packed per-block table for the context bit positions instead of arithmetic (+2 %),
// folding the "window empty" test into the one-arm through a saturating shift (+8 %), requesting the
// EOB / zero / one probabilities of the next rows (16-byte entries, no extraction) ahead of use (+7 %).
// (one inlined copy of the block decoder in a 25-block loop instead of three copies: less code to fetch, -2 %)
#ifndef ENT_WIDE
#define ENT_WIDE 2       // token probabilities stored as {p << 24, (256 - p) << 24}: split = hi32(range * a + b), one IMAD.WIDE; 2 = 16-byte entries
#endif
#ifndef ENT_LUT
#define ENT_LUT 1        // band / zigzag of coefficient n from a shared-memory table instead of immediates
#endif

namespace webpb200 {
namespace {

__constant__ uint8_t c_bmode_probs[900];  // [top][left][node 9], common.rs KEYFRAME_BPRED_MODE_PROBS

// trees as in host_vp8.cc (common.rs:20-42, 147-172): index > 0 = node, <= 0 = -leaf
__constant__ int8_t c_segment_tree[6] = {2, 4, -0, -1, -2, -3};
__constant__ int8_t c_ymode_tree[8] = {-4, 2, 4, 6, -0, -1, -2, -3};
__constant__ int8_t c_bmode_tree[18] = {-0, 2, -1, 4, -2, 6, 8, 12, -3, 10, -5, -6, -4, 14, -7, 16, -8, -9};
__constant__ int8_t c_uvmode_tree[6] = {-0, 2, -1, 4, -2, -3};
__constant__ uint8_t c_ymode_probs[4] = {145, 156, 163, 128};
__constant__ uint8_t c_uvmode_probs[3] = {142, 114, 183};
__constant__ uint8_t c_cat_probs[4][12] = {{173, 148, 140, 0}, {176, 155, 140, 135, 0}, {180, 157, 141, 134, 130, 0},
                                           {254, 254, 243, 230, 196, 177, 153, 140, 133, 130, 129, 0}};
__constant__ int c_cat_base[4] = {11, 19, 35, 67};

// Boolean decoder, 32-bit window: `bits` + 8 bits of `value` are valid; refilled with up to three
// bytes when bits < 0.  Symbol for symbol the host reader of host_vp8.cc.
//
// One lane runs this alone, so what counts is the number of instructions per symbol: measured on
// B200 (tools/micro/latbench.cu) a dependent ALU op costs 4-5 cycles, FLO 17, LDS 30, a taken
// branch about 15, and ncu shows the warp issuing one instruction every 5.5 cycles.
struct DevBool {
    const uint8_t* base;
    uint32_t pos, len;
    uint32_t value, range;
    int bits;
    bool phantom, eof;

    __device__ __forceinline__ void init(const uint8_t* data, uint32_t n) {
        base = data; pos = 0; len = n; value = 0; range = 255; bits = -8; phantom = false; eof = false;
    }
    __device__ __forceinline__ void load(const uint8_t* data, uint32_t n, const BoolState& s) {
        base = data; len = n; pos = s.pos; value = s.value; range = s.range; bits = s.bits;
        phantom = s.phantom != 0; eof = s.eof != 0;
    }
    // Fast path straight-line: three bytes are always fetched (the bitstream arena is padded, so
    // reading past the end of a partition is safe) and the rare overshoot at the end of a partition
    // is undone and replayed byte by byte with the end-of-data rules (arithmetic_decoder.rs:249-276).
    __device__ __forceinline__ void refill() {
        const uint32_t b0 = __ldg(base + pos), b1 = __ldg(base + pos + 1), b2 = __ldg(base + pos + 2);
        value = (value << 24) | (b0 << 16) | (b1 << 8) | b2;
        pos += 3;
        bits += 24;
        if (pos > len) refill_tail();
    }
    __device__ __forceinline__ void refill_tail() {
        const uint32_t over = pos - len;  // 1..3 bytes that do not exist
        value >>= 8 * over;
        bits -= 8 * (int)over;
        pos = len;
        if (bits >= 0) return;  // the real bytes were enough
        if (!phantom && !eof) {
            phantom = true;
            value <<= 8;
            bits += 8;
        } else {
            eof = true;
        }
    }
    // The decision and the renormalisation, given split = 1 + (((range - 1) * prob) >> 8).
    // Two arms, each with its own renormalisation: every caller branches on the result at once, so
    // the comparison is made once and nothing has to be selected (-4 % against the select form).
    // (Folding the bits < 0 test into the one-arm through a saturating shift was slower: +8 %.)
    __device__ __forceinline__ bool decide(uint32_t split) {
        const uint32_t big = split << bits;
        if (value >= big) {
            value -= big;
            const uint32_t r = range - split;
            const int sh = __clz(r) - 24;
            range = r << sh;
            bits -= sh;
            return true;
        }
        const int sh = __clz(split) - 24;
        range = split << sh;
        bits -= sh;
        return false;
    }
    __device__ __forceinline__ bool get(uint32_t prob) {
        if (bits < 0) {
            refill();
            if (bits < 0) return false;
        }
        return decide((range * prob + (256u - prob)) >> 8);  // = 1 + (((range - 1) * prob) >> 8)
    }
    // prob given as e = {p << 24, (256 - p) << 24} (p = 0 stored as 1: both give split = 1 for any
    // range <= 255): hi32(range * e.x + e.y) = (range * p + 256 - p) >> 8 in one wide multiply-add
    __device__ __forceinline__ bool get(uint2 e) {
        if (bits < 0) {
            refill();
            if (bits < 0) return false;
        }
        return decide((uint32_t)(((unsigned long long)range * e.x + e.y) >> 32));
    }
    __device__ __forceinline__ bool get(uint4 e) {
        if (bits < 0) {
            refill();
            if (bits < 0) return false;
        }
        const unsigned long long addend = (unsigned long long)e.x | ((unsigned long long)e.y << 32);
        return decide((uint32_t)(((unsigned long long)range * e.z + addend) >> 32));
    }
    __device__ __forceinline__ int tree(const int8_t* t, const uint8_t* probs) {
        int i = 0;
        do {
            i = t[i + (int)get((uint32_t)probs[i >> 1])];
        } while (i > 0);
        return -i;
    }
};

// COEFF_BANDS and ZIGZAG (common.rs) as nibble tables in immediates: no table load on the chain
__device__ __forceinline__ uint32_t band_of(int n) { return ((n < 8 ? 0x65463210u : 0x76666666u) >> ((n & 7) * 4)) & 7u; }
__device__ __forceinline__ uint32_t zigzag_of(int n) { return ((n < 8 ? 0x63258410u : 0xFEB7ADC9u) >> ((n & 7) * 4)) & 15u; }

#if ENT_WIDE == 2
// 16-byte entries {(256 - p) << 24, 0, p << 24, -}: one LDS.128 puts the 64-bit addend of the wide
// multiply-add into an aligned register pair and the multiplier next to it, no moves needed
typedef uint4 TokenProb;
__device__ __forceinline__ TokenProb make_token_prob(uint32_t p) {
    if (p == 0) p = 1;
    return make_uint4((256u - p) << 24, 0u, p << 24, 0u);
}
#elif ENT_WIDE
typedef uint2 TokenProb;   // {p << 24, (256 - p) << 24}
__device__ __forceinline__ TokenProb make_token_prob(uint32_t p) {
    if (p == 0) p = 1;
    return make_uint2(p << 24, (256u - p) << 24);
}
#else
typedef uint8_t TokenProb;
__device__ __forceinline__ TokenProb make_token_prob(uint32_t p) { return (uint8_t)p; }
#endif
constexpr int kTokenProbs = 4 * 8 * 3 * 11;  // [plane][band][ctx][node], common.rs COEFF_PROBS layout
constexpr int kProbBytes = ((int)sizeof(TokenProb) * kTokenProbs + 15) & ~15;

// Tokens of one 4x4 block (host_vp8.cc read_block; reference mod.rs:785-872).  probs = the plane's
// [band][ctx][node] table in shared memory.  Returns "has coefficients".
__device__ __forceinline__ int read_block(DevBool& br, const TokenProb* probs, const uint32_t* lut, int ctx, int first, int16_t* out) {
#if ENT_LUT
    // band(n) and zigzag(n) come from one shared-memory word fetched when n advances: the load is
    // issued a whole symbol before its result is needed, and replaces about eight ALU instructions
    const char* base = reinterpret_cast<const char*>(probs);
    int n = first;
    uint32_t w = lut[n];
    const TokenProb* p = reinterpret_cast<const TokenProb*>(base + (w & 0xffffu)) + ctx * 11;
    if (!br.get(p[0])) return 0;  // DCT_EOB straight away
    for (;;) {
        while (!br.get(p[1])) {  // DCT_0: no EOB test before the next token
            if (++n == 16) return 1;
            w = lut[n];
            p = reinterpret_cast<const TokenProb*>(base + (w & 0xffffu));
        }
        const uint32_t wn = lut[n + 1];
        int v;
        if (!br.get(p[2])) {
            v = 1;
            ctx = 1;
        } else {
            if (!br.get(p[3])) {
                v = !br.get(p[4]) ? 2 : 3 + (int)br.get(p[5]);
            } else if (!br.get(p[6])) {
                if (!br.get(p[7])) {
                    v = 5 + (int)br.get(159u);
                } else {
                    v = 7 + 2 * (int)br.get(165u);
                    v += (int)br.get(145u);
                }
            } else {
                const int b1 = br.get(p[8]);
                const int b0 = br.get(p[9 + b1]);
                const int cat = 2 * b1 + b0;
                v = 0;
                for (const uint8_t* t = c_cat_probs[cat]; *t; ++t) v = v + v + (int)br.get((uint32_t)*t);
                v += c_cat_base[cat];
            }
            ctx = 2;
        }
        if (br.get(128u)) v = -v;
        *reinterpret_cast<int16_t*>(reinterpret_cast<char*>(out) + (w >> 16)) = (int16_t)v;
        if (++n == 16) return 1;
        w = wn;
        p = reinterpret_cast<const TokenProb*>(base + (w & 0xffffu)) + ctx * 11;
        if (!br.get(p[0])) return 1;  // DCT_EOB
    }
#else
    (void)lut;
    int n = first;
    const TokenProb* p = probs + (band_of(n) * 3 + ctx) * 11;
    if (!br.get(p[0])) return 0;  // DCT_EOB straight away
    for (;;) {
        while (!br.get(p[1])) {  // DCT_0: no EOB test before the next token
            if (++n == 16) return 1;
            p = probs + band_of(n) * 33;
        }
        int v;
        if (!br.get(p[2])) {
            v = 1;
            ctx = 1;
        } else {
            if (!br.get(p[3])) {
                v = !br.get(p[4]) ? 2 : 3 + (int)br.get(p[5]);
            } else if (!br.get(p[6])) {
                if (!br.get(p[7])) {
                    v = 5 + (int)br.get(159u);
                } else {
                    v = 7 + 2 * (int)br.get(165u);
                    v += (int)br.get(145u);
                }
            } else {
                const int b1 = br.get(p[8]);
                const int b0 = br.get(p[9 + b1]);
                const int cat = 2 * b1 + b0;
                v = 0;
                for (const uint8_t* t = c_cat_probs[cat]; *t; ++t) v = v + v + (int)br.get((uint32_t)*t);
                v += c_cat_base[cat];
            }
            ctx = 2;
        }
        if (br.get(128u)) v = -v;
        out[zigzag_of(n)] = (int16_t)v;
        if (++n == 16) return 1;
        p = probs + (band_of(n) * 3 + ctx) * 11;
        if (!br.get(p[0])) return 1;  // DCT_EOB
    }
#endif
}

struct EntropyArgs {
    const webpb200_frame_desc* frames;
    const Vp8EntropyHdr* hdrs;
    const uint8_t* bits;
    webpb200_mb_desc* mbs;
    int16_t* levels;
    int* status;  // per frame; written only on failure
    int frame_first;
    int max_parts;
    int max_mbw;
    int n_frames;
    int frames_per_block;
    int shared_probs;  // every frame of the launch carries the same token probabilities: one table per block instead of one per frame
    unsigned int* rows_done;  // host-mapped, [frame * 8 + partition] = macroblock rows that partition has finished; or null
    int wave;                 // split layout only: blocks per wave (the SM count)
    int faults;               // test hook (webpb200_debug_entropy_faults): bit 0 = the split layout's mode blocks do nothing
    unsigned int* sync;       // split layout only, zero at launch: [frame] = macroblock rows whose descriptors are in global memory (~0u: the
                              // frame failed), [frames + frame] = somebody has taken the frame's mode partition
};

struct EntropyCtl {
    int modes_done;   // macroblocks whose descriptor is in global memory
    int abort;        // some warp hit the end of its partition: the frame is a BitStreamError
    int row_prog[9];  // ring slot -> (row << 16) | macroblocks finished, monotone per slot
    uint32_t lut[17]; // coefficient n -> byte offset of band(n)'s context-0 row | zigzag(n) * 2 << 16  (ENT_LUT)
};
constexpr int kCtlBytes = 128;
// per frame: [TokenProb[1056] unless the block shares one table] | EntropyCtl | top_bpred u32[max_mbw] | top_nz u16[(P+1)][max_mbw]
__host__ __device__ inline size_t entropy_smem_per_frame(int max_parts, int max_mbw, bool own_probs) {
    return ((own_probs ? kProbBytes : 0) + kCtlBytes + 4 * (size_t)max_mbw + 2 * (size_t)(max_parts + 1) * (size_t)max_mbw + 15) & ~(size_t)15;
}

// spin on a shared-memory word until it reaches `need` or the frame is aborted
// (a wait that never ends would be a scheduling bug, not a data property: after a few seconds the
// frame is failed with WEBPB200_ERR_CUDA instead of hanging the device)
__device__ __forceinline__ bool wait_shared(volatile int* word, int need, volatile int* abort, int* status) {
    unsigned spins = 0;
    while (*word < need) {
        if (*abort) return false;
        __nanosleep(64);
        if (++spins > (1u << 26)) {
            *status = WEBPB200_ERR_CUDA;
            *abort = 1;
            return false;
        }
    }
    __threadfence_block();
    return true;
}

// the same on a global word another block writes (split layout)
__device__ __forceinline__ bool wait_global(const volatile unsigned int* word, unsigned int need, int* status) {
    unsigned spins = 0;
    while (*word < need) {
        __nanosleep(200);
        if (++spins > (1u << 24)) {
            *status = WEBPB200_ERR_CUDA;
            return false;
        }
    }
    __threadfence();
    return true;
}

// one frame's working set in shared memory
struct FrameSmem {
    TokenProb* probs;
    EntropyCtl* ctl;
    uint32_t* top_bpred;
    uint16_t* top_nz;
};

// ---------------- macroblock headers of frame f (read_macroblock_header, mod.rs:615-671) ----------------
// kSplit = false: the token warps of the frame sit in the same block and follow `modes_done` in shared memory,
// macroblock by macroblock.  kSplit = true: they sit in another block and follow A.sync[1 + f] in global memory,
// macroblock row by macroblock row (one device-scope fence per row).
template <bool kSplit>
__device__ __forceinline__ void decode_modes(const EntropyArgs& A, int f, const FrameSmem& M) {
    const Vp8EntropyHdr& H = A.hdrs[f];
    const webpb200_frame_desc& fd = A.frames[f];
    const int mbw = fd.mbw, mbh = fd.mbh;
    webpb200_mb_desc* mbs = A.mbs + fd.mb_base;
    volatile int* v_abort = &M.ctl->abort;
    volatile int* v_modes = &M.ctl->modes_done;
    volatile unsigned int* g_rows = kSplit ? A.sync + f : nullptr;
    if (kSplit && atomicCAS(A.sync + A.frame_first + A.n_frames + f, 0u, 1u) != 0u) return;  // the frame's token warp got tired of waiting and does it
    uint32_t* s_top_bpred = M.top_bpred;
    DevBool b;
    b.load(A.bits + H.data_off + H.first_off, H.first_len, H.first);
    const bool update_map = H.update_map, has_skip = H.has_prob_skip;
    const int prob_skip = H.prob_skip;
    const uint8_t seg_probs[3] = {H.seg_probs[0], H.seg_probs[1], H.seg_probs[2]};
    int done = 0;
    for (int mby = 0; mby < mbh; mby++) {
        uint32_t left = 0;  // four B modes, one per byte (row y in byte y)
        for (int mbx = 0; mbx < mbw; mbx++) {
            uint32_t w0 = 0, bp0 = 0, bp1 = 0;  // the descriptor's 16 bytes: {w0, nflags = 0, bp0, bp1}
            if (update_map) {
                int i = 0;
                do {
                    i = c_segment_tree[i + (int)b.get(seg_probs[i >> 1])];
                } while (i > 0);
                w0 |= (uint32_t)(-i) << 16;
            }
            const int skipped = has_skip ? (int)b.get(prob_skip) : 0;
            w0 |= (uint32_t)skipped << 24;
            const int ymode = b.tree(c_ymode_tree, c_ymode_probs);
            w0 |= (uint32_t)ymode;
            uint32_t top = s_top_bpred[mbx];  // column x in byte x
            if (ymode == 4) {
                uint32_t nt = 0;
#pragma unroll 1
                for (int y = 0; y < 4; y++) {
                    int l = (left >> (8 * y)) & 0xff;
                    uint32_t row = 0;
#pragma unroll 1
                    for (int x = 0; x < 4; x++) {
                        const int t = (top >> (8 * x)) & 0xff;
                        const int bm = b.tree(c_bmode_tree, c_bmode_probs + (t * 10 + l) * 9);
                        row |= (uint32_t)bm << (8 * x);
                        l = bm;
                    }
                    top = row;
                    if (y == 3) nt = row;
                    left = (left & ~(0xffu << (8 * y))) | ((uint32_t)l << (8 * y));
                    // nibbles of sub-blocks 4y..4y+3
                    const uint32_t nib = (row & 0xf) | ((row >> 4) & 0xf0) | ((row >> 8) & 0xf00) | ((row >> 12) & 0xf000);
                    if (y < 2) bp0 |= nib << (16 * y);
                    else bp1 |= nib << (16 * (y - 2));
                }
                s_top_bpred[mbx] = nt;
            } else {
                const uint32_t m = (0x01030200u >> (8 * ymode)) & 0xff;  // DC->B_DC V->B_VE H->B_HE TM->B_TM
                left = m * 0x01010101u;
                s_top_bpred[mbx] = left;
            }
            w0 |= (uint32_t)b.tree(c_uvmode_tree, c_uvmode_probs) << 8;
            if (b.eof) {
                A.status[f] = WEBPB200_ERR_BITSTREAM;
                if (kSplit) {
                    __threadfence();
                    *g_rows = ~0u;
                } else {
                    *v_abort = 1;
                }
                return;
            }
            *reinterpret_cast<uint4*>(mbs + (size_t)mby * mbw + mbx) = make_uint4(w0, 0u, bp0, bp1);
            if (!kSplit) {
                __threadfence_block();
                *v_modes = ++done;
            }
        }
        if (kSplit) {
            __threadfence();
            *g_rows = (unsigned int)mby + 1u;
        }
    }
}

// ---------------- coefficient tokens of partition p of frame f (read_residual_data, mod.rs:874-963) ----------------
template <bool kSplit>
__device__ __forceinline__ void decode_tokens(const EntropyArgs& A, int f, int p, int max_mbw, const FrameSmem& M) {
    const Vp8EntropyHdr& H = A.hdrs[f];
    const int P = H.num_parts;
    const webpb200_frame_desc& fd = A.frames[f];
    const int mbw = fd.mbw, mbh = fd.mbh;
    webpb200_mb_desc* mbs = A.mbs + fd.mb_base;
    EntropyCtl* ctl = M.ctl;
    volatile int* v_abort = &ctl->abort;
    volatile int* v_modes = &ctl->modes_done;
    const TokenProb* s_probs = M.probs;
    uint16_t* s_top_nz = M.top_nz;
    DevBool br;
    br.init(A.bits + H.data_off + H.part_off[p], H.part_len[p]);
    int16_t* levels = A.levels + (size_t)fd.mb_base * WEBPB200_LEVELS_PER_MB;
    const int ring = P + 1;
    int seen_modes = 0;
    unsigned int seen_rows = 0;
    for (int mby = p; mby < mbh; mby += P) {
        uint16_t* cur = s_top_nz + (size_t)(mby % ring) * max_mbw;
        const uint16_t* prev = s_top_nz + (size_t)((mby + ring - 1) % ring) * max_mbw;
        volatile int* my_prog = &ctl->row_prog[mby % ring];
        volatile int* up_prog = &ctl->row_prog[(mby + ring - 1) % ring];
        if (kSplit && seen_rows <= (unsigned int)mby) {  // this row's descriptors (written by the frame's mode warp in another block)
            const volatile unsigned int* g_rows = A.sync + f;
            if (!wait_global(g_rows, (unsigned int)mby + 1u, A.status + f)) return;
            seen_rows = *g_rows;
            if (seen_rows == ~0u) return;  // the mode partition ran out of bits
        }
        uint32_t l = 0;  // left context: bit 0 Y2, 1-4 Y rows, 5-6 U rows, 7-8 V rows
        for (int mbx = 0; mbx < mbw; mbx++) {
            const int mbi = mby * mbw + mbx;
            if (!kSplit && seen_modes <= mbi) {
                if (!wait_shared(v_modes, mbi + 1, v_abort, A.status + f)) return;
                seen_modes = *v_modes;
            }
            if (P > 1 && mby > 0 && !wait_shared(up_prog, ((mby - 1) << 16) | (mbx + 1), v_abort, A.status + f)) return;
            uint32_t t = mby > 0 ? prev[mbx] : 0;  // top context: bit 0 Y2, 1-4 Y columns, 5-6 U, 7-8 V
            const uint32_t w0 = *reinterpret_cast<const volatile uint32_t*>(mbs + mbi);
            const int ymode = w0 & 0xff;
            const bool skipped = (w0 >> 24) & 1;
            if (skipped) {  // mod.rs:1263-1272: everything but a B macroblock's Y2 context is cleared
                t = ymode != 4 ? 0 : (t & 1);
                l = ymode != 4 ? 0 : (l & 1);
            } else {
                int16_t* lv = levels + (size_t)mbi * WEBPB200_LEVELS_PER_MB;
                uint32_t nflags = 0;
                // 25 blocks in bitstream order: Y2 (if any), 16 Y, 4 U, 4 V.  Context bit of block k:
                // top  = 1 + column (Y), 5 + column (U), 7 + column (V), 0 (Y2); left likewise with rows.
                // Levels go out PACKED (the section 8(f) wire format): a block takes a 32-byte slot of its
                // macroblock's 800-byte record only if it carries tokens, slots in this decode order; nflags
                // says which blocks those are.  A slot is cleared right before it is decoded into, so nothing
                // has to be cleared ahead of the kernel and the blocks without tokens are never written or read.
                const bool has_y2 = ymode != 4;
                int16_t* slot = lv;
                bool clean = false;
#pragma unroll 1
                for (int k = has_y2 ? -1 : 0; k < 24; k++) {
                    int tb, lb, plane, first = 0, blk = k;
                    if (k < 0) { tb = 0; lb = 0; plane = 1; blk = 24; }
                    else if (k < 16) { tb = 1 + (k & 3); lb = 1 + (k >> 2); plane = has_y2 ? 0 : 3; first = has_y2; }
                    else { const int c = k - 16; tb = 5 + 2 * (c >> 2) + (c & 1); lb = 5 + 2 * (c >> 2) + ((c >> 1) & 1); plane = 2; }
                    if (!clean) {
                        reinterpret_cast<uint4*>(slot)[0] = make_uint4(0u, 0u, 0u, 0u);
                        reinterpret_cast<uint4*>(slot)[1] = make_uint4(0u, 0u, 0u, 0u);
                        clean = true;
                    }
                    const int n = read_block(br, s_probs + plane * 264, ctl->lut, ((t >> tb) & 1) + ((l >> lb) & 1), first, slot);
                    nflags |= (uint32_t)n << blk;
                    t = (t & ~(1u << tb)) | ((uint32_t)n << tb);
                    l = (l & ~(1u << lb)) | ((uint32_t)n << lb);
                    if (n) { slot += 16; clean = false; }
                }
                mbs[mbi].nflags = nflags;
                if (br.eof) {
                    A.status[f] = WEBPB200_ERR_BITSTREAM;
                    if (!kSplit) *v_abort = 1;
                    return;
                }
            }
            cur[mbx] = (uint16_t)t;
            if (P > 1) {
                __threadfence_block();
                *my_prog = (mby << 16) | (mbx + 1);
            }
        }
        if (A.rows_done) {
            // This row's levels and descriptors are complete: tell the host, which starts reconstructing
            // finished bands of rows while the rest is still being decoded.  The observer is the host, so
            // the fence that orders this row's global writes before the flag is system-scope; the flag
            // itself is a plain posted store into this partition's own slot -- system-scope ATOMICS on
            // host memory cut concurrent D2H copies from 57 to 6 GB/s (tools/micro/copybench.cu), a
            // fence once per macroblock row does not.
            __threadfence_system();
            *reinterpret_cast<volatile unsigned int*>(A.rows_done + (size_t)f * 8 + p) = (unsigned int)(mby / P + 1);
        }
    }
}

// shared tables of the block's frames [frame0, frame0 + F), filled by the whole block (ends with a barrier)
__device__ __forceinline__ void fill_frame_tables(const EntropyArgs& A, uint8_t* smem_all, int frame0, int F, bool own, size_t per_frame) {
    const size_t frames_at = own ? 0 : kProbBytes;
    const size_t ctl_at = own ? kProbBytes : 0;  // offset of EntropyCtl inside a frame's region
    // which of the block's frames the shared table is filled from: the first one the host did not drop
    // (a dropped frame's header is all zero)
    int table_from = 0;
    if (!own)
        for (int sl = F - 1; sl >= 0; sl--)
            if (frame0 + sl < A.n_frames && A.hdrs[A.frame_first + frame0 + sl].num_parts) table_from = sl;
    for (int sl = 0; sl < F; sl++) {
        const int fi = frame0 + sl;
        if (fi >= A.n_frames) break;
        uint8_t* sm = smem_all + frames_at + sl * per_frame;
        const Vp8EntropyHdr& Hs = A.hdrs[A.frame_first + fi];
        TokenProb* pr = reinterpret_cast<TokenProb*>(own ? sm : smem_all);
        EntropyCtl* cs = reinterpret_cast<EntropyCtl*>(sm + ctl_at);
        uint32_t* tb = reinterpret_cast<uint32_t*>(sm + ctl_at + kCtlBytes);
        if (own || sl == table_from)
            for (int i = threadIdx.x; i < kTokenProbs; i += blockDim.x) pr[i] = make_token_prob(Hs.probs[i]);
        for (int i = threadIdx.x; i < (int)(sizeof(EntropyCtl) / 4); i += blockDim.x) reinterpret_cast<int*>(cs)[i] = 0;
        for (int i = threadIdx.x; i < A.max_mbw; i += blockDim.x) tb[i] = 0;
    }
    __syncthreads();
    for (int sl = 0; sl < F; sl++) {
        EntropyCtl* cs = reinterpret_cast<EntropyCtl*>(smem_all + frames_at + sl * per_frame + ctl_at);
        if (threadIdx.x < 17) cs->lut[threadIdx.x] = band_of(threadIdx.x) * 33u * (uint32_t)sizeof(TokenProb) | (zigzag_of(threadIdx.x & 15) * 2u) << 16;
    }
    __syncthreads();
}
__device__ __forceinline__ FrameSmem frame_smem(const EntropyArgs& A, uint8_t* smem_all, int slot, bool own, size_t per_frame) {
    uint8_t* smem = smem_all + (own ? 0 : kProbBytes) + slot * per_frame;
    const size_t ctl_at = own ? kProbBytes : 0;
    FrameSmem M;
    M.probs = reinterpret_cast<TokenProb*>(own ? smem : smem_all);
    M.ctl = reinterpret_cast<EntropyCtl*>(smem + ctl_at);
    M.top_bpred = reinterpret_cast<uint32_t*>(smem + ctl_at + kCtlBytes);
    M.top_nz = reinterpret_cast<uint16_t*>(smem + ctl_at + kCtlBytes + 4 * (size_t)A.max_mbw);
    return M;
}

// Block layouts of k_entropy.  frames_per_block = 1 (frames with several token partitions): one frame, warp 0 =
// modes, warps 1..P = token partitions.  frames_per_block = 4 (batches whose frames all have one token partition):
// warps 0-3 = the token partitions of four frames, warps 4-7 = their mode partitions.  With two warps per block the
// token warps all sat on odd warp slots, i.e. on two of the SM's four schedulers, while the other two idled once the
// short mode partitions were done; four consecutive token warps always cover all four schedulers (ncu: "not
// selected" stalls 0.36 -> 0.1 per instruction, 35.6 -> 33 ms on cfg2).
// Batches whose frames all carry the same token probabilities (no probability updates: the usual case)
// keep ONE table per block (17 KB) instead of one per frame.
__global__ void __launch_bounds__(288) k_entropy(EntropyArgs A) {
    extern __shared__ __align__(16) uint8_t smem_all[];
    const int F = A.frames_per_block;
    const int warp_all = threadIdx.x >> 5;
    const int slot = F == 1 ? 0 : (warp_all & 3);                 // which of the block's frames this warp serves
    const int warp = F == 1 ? warp_all : (warp_all < 4 ? 1 : 0);  // role: 0 = modes, 1.. = token partition + 1
    const bool own = !A.shared_probs;
    const size_t per_frame = entropy_smem_per_frame(A.max_parts, A.max_mbw, own);
    fill_frame_tables(A, smem_all, blockIdx.x * F, F, own, per_frame);
    const int fi = blockIdx.x * F + slot;
    if ((threadIdx.x & 31) != 0 || fi >= A.n_frames) return;  // one working lane per partition
    const FrameSmem M = frame_smem(A, smem_all, slot, own, per_frame);
    const int f = A.frame_first + fi;
    const int P = A.hdrs[f].num_parts;
    if (P == 0 || warp > P) return;  // failed on the host / no such partition
    if (warp == 0) decode_modes<false>(A, f, M);
    else decode_tokens<false>(A, f, warp - 1, A.max_mbw, M);
}

// Batches whose frames all have ONE token partition (split layout): blocks of four warps, four frames per block,
// and the mode partitions of those four frames in ANOTHER block of four warps.  A warp's registers are allocated for
// all 32 lanes although one lane works, so the number of resident chains is set by registers per block: with the
// mode warps (a fifth of the work, finished early) in the same block as the token warps, half of every block's
// registers sat idle for most of the kernel and an SM held 16 token chains; now it holds 36, which is what lets the
// chains of several batches in flight share the issue slots (webpb200_decode_batch_begin).
//
// A chain slows down with every other chain on its SM (28.5 ms with three others, 33 ms among sixteen), so the
// token blocks must be spread evenly.  The hardware places the blocks of a grid breadth-first in blockIdx order, one
// per SM and round again (tools/micro/placement.cu) -- as long as every block stays -- so the roles go by blockIdx in
// waves of one block per SM: modes of groups 0..S-1, tokens of groups 0..S-1, modes of the next S groups, ...
// (Roles by atomic ticket, i.e. in arrival order, put 256 token blocks on 94 of 148 SMs: 35 instead of 32 ms.)
// Token warps follow their frame's mode warp row by row through A.sync (global memory, one device-scope fence per
// macroblock row on the producer's side).  Nothing guarantees that a lower blockIdx starts first, so whoever gets to a
// frame's mode partition first claims it: a token warp that sees no progress decodes the modes itself.
constexpr int kSplitFrames = 4;
__global__ void __launch_bounds__(32 * kSplitFrames) k_entropy_split(EntropyArgs A) {
    extern __shared__ __align__(16) uint8_t smem_all[];
    const int groups = (A.n_frames + kSplitFrames - 1) / kSplitFrames;
    // waves of A.wave blocks: an even wave holds mode blocks, the odd wave after it the token blocks of the same groups;
    // the last pair of waves is short (no block without work)
    const int per_pair = 2 * A.wave;
    const int pair = (int)blockIdx.x / per_pair, in_pair = (int)blockIdx.x % per_pair;
    const int in_wave = min(A.wave, groups - pair * A.wave);  // groups in this pair of waves
    const bool modes = in_pair < in_wave;
    const int frame0 = (pair * A.wave + (modes ? in_pair : in_pair - in_wave)) * kSplitFrames;
    const bool own = !A.shared_probs;
    const size_t per_frame = entropy_smem_per_frame(1, A.max_mbw, own);
    fill_frame_tables(A, smem_all, frame0, kSplitFrames, own, per_frame);
    const int slot = threadIdx.x >> 5;
    const int fi = frame0 + slot;
    if ((threadIdx.x & 31) != 0 || fi >= A.n_frames) return;  // one working lane per chain
    const FrameSmem M = frame_smem(A, smem_all, slot, own, per_frame);
    const int f = A.frame_first + fi;
    if (A.hdrs[f].num_parts == 0) return;  // failed on the host
    if (modes && (A.faults & 1)) return;   // test hook: as if this block had never been scheduled
#ifdef ENT_DEBUG
    unsigned long long t0, t1; unsigned sm, wid;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    asm volatile("mov.u32 %0, %%smid;" : "=r"(sm));
    asm volatile("mov.u32 %0, %%warpid;" : "=r"(wid));
#endif
    if (!modes) {
        // nobody may have started on the frame's modes (its mode block is not resident yet, and nothing says it will be
        // before this block ends): after a while, do them here (a no-op if the mode warp has claimed the frame meanwhile)
        const volatile unsigned int* g_rows = A.sync + f;
        unsigned spins = 0;
        while (*g_rows == 0u && ++spins < 2000u) __nanosleep(200);
    }
    if (modes || *(const volatile unsigned int*)(A.sync + f) == 0u) decode_modes<true>(A, f, M);
    if (!modes) decode_tokens<true>(A, f, 0, A.max_mbw, M);
#ifdef ENT_DEBUG
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    printf("ENT %d %d %u %u %llu %llu\n", modes ? 1 : 0, fi, sm, wid, t0, t1);
#endif
}

}  // namespace

static int g_entropy_faults = 0;
void set_entropy_faults(int flags) { g_entropy_faults = flags; }

void upload_entropy_tables_once() {
    cudaMemcpyToSymbol(c_bmode_probs, kBModeProbs, sizeof(kBModeProbs));
}

cudaError_t launch_entropy(const BatchDev& b, const Vp8EntropyHdr* d_hdrs, const uint8_t* d_bits, webpb200_mb_desc* d_mbs,
                           int16_t* d_levels, int* d_status, int max_parts, bool same_probs, unsigned int* rows_done, unsigned int* d_sync,
                           cudaStream_t s) {
    if (b.n_frames <= 0) return cudaSuccess;
    EntropyArgs a;
    a.frames = b.frames; a.hdrs = d_hdrs; a.bits = d_bits; a.mbs = d_mbs; a.levels = d_levels; a.status = d_status;
    a.frame_first = b.frame_first;
    a.max_parts = max_parts;
    a.max_mbw = (int)((b.max_w + 15) / 16);
    a.n_frames = b.n_frames;
    // single-partition batches: four frames per block; split = their mode warps in blocks of their own (see k_entropy_split)
    const bool four = max_parts == 1 && b.n_frames >= 8;
    const bool split = four && d_sync != nullptr;
    a.frames_per_block = four ? kSplitFrames : 1;
    a.rows_done = rows_done;
    a.shared_probs = same_probs ? 1 : 0;
    a.sync = d_sync;
    a.wave = 1;
    a.faults = g_entropy_faults;
    static_assert(sizeof(EntropyCtl) <= kCtlBytes, "EntropyCtl outgrew its slot");
    const int F = a.frames_per_block;
    const size_t smem = entropy_smem_per_frame(split ? 1 : max_parts, a.max_mbw, !same_probs) * (size_t)F + (same_probs ? kProbBytes : 0);
    if (split) {
        if (smem > 48 * 1024) {  // per device and function: cheap enough to repeat
            cudaError_t e = cudaFuncSetAttribute(k_entropy_split, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) return e;
        }
        const int groups = (b.n_frames + F - 1) / F;
        a.wave = b.sm_count > 0 ? b.sm_count : 1;
        k_entropy_split<<<2 * groups, 32 * kSplitFrames, smem, s>>>(a);
    } else {
        if (smem > 48 * 1024) {
            cudaError_t e = cudaFuncSetAttribute(k_entropy, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) return e;
        }
        k_entropy<<<(b.n_frames + F - 1) / F, F == 1 ? 32 * (1 + max_parts) : 256, smem, s>>>(a);
    }
    return cudaGetLastError();
}

}  // namespace webpb200
