// vp8l.cu -- sm_100a kernels that undo the VP8L transforms of compressed ALPH planes (see vp8l.cuh).
// Arithmetic follows reference lossless/decoder/reverse_transform.rs; pixels are ARGB words (the reference keeps
// RGBA bytes: its pixel[0..3] = bits 16-23, 8-15, 0-7, 24-31 here).
#include "vp8l.cuh"

namespace webpb200 {
namespace {

constexpr unsigned kFull = 0xffffffffu;
constexpr int kWarps = 4;

// per-channel wrapping add (reverse_transform.rs: `wrapping_add` on every byte)
__device__ __forceinline__ uint32_t add4(uint32_t a, uint32_t b) {
    return (((a & 0xff00ff00u) + (b & 0xff00ff00u)) & 0xff00ff00u) | (((a & 0x00ff00ffu) + (b & 0x00ff00ffu)) & 0x00ff00ffu);
}
// average2 (reverse_transform.rs:580-590): floor((a + b) / 2) per channel
__device__ __forceinline__ uint32_t avg4(uint32_t a, uint32_t b) { return (((a ^ b) & 0xfefefefeu) >> 1) + (a & b); }
__device__ __forceinline__ int chan(uint32_t p, int c) { return (int)((p >> (8 * c)) & 0xffu); }
// clamp_add_subtract_full (593-597)
__device__ __forceinline__ uint32_t clamp_full(uint32_t l, uint32_t t, uint32_t tl) {
    uint32_t r = 0;
#pragma unroll
    for (int c = 0; c < 4; c++) r |= (uint32_t)min(max(chan(l, c) + chan(t, c) - chan(tl, c), 0), 255) << (8 * c);
    return r;
}
// clamp_add_subtract_half (600-604) of (l + t) / 2 and tl; the i16 division truncates towards zero
__device__ __forceinline__ uint32_t clamp_half(uint32_t l, uint32_t t, uint32_t tl) {
    uint32_t r = 0;
#pragma unroll
    for (int c = 0; c < 4; c++) {
        const int a = (chan(l, c) + chan(t, c)) >> 1;
        const int d = a - chan(tl, c);
        r |= (uint32_t)min(max(a + (d < 0 ? -((-d) >> 1) : d >> 1), 0), 255) << (8 * c);
    }
    return r;
}
// predictor 11 (235-290): left if sum |t - tl| < sum |l - tl| over the four channels, else top
__device__ __forceinline__ uint32_t select4(uint32_t l, uint32_t t, uint32_t tl) {
    return __vsadu4(t, tl) < __vsadu4(l, tl) ? l : t;
}

// The fourteen predictors (reverse_transform.rs:79-351).  The lanes of a warp sit in different blocks of the
// predictor image (their columns are skewed), so a `switch` on the mode would run every case that occurs, one
// after the other (measured: 2400 cycles per step on random modes); instead the candidates are computed
// straight-line (they share their averages) and the mode picks one.
__device__ __forceinline__ uint32_t predict_mixed(int mode, uint32_t L, uint32_t T, uint32_t TL, uint32_t TR) {
    const uint32_t a_l_tl = avg4(L, TL), a_t_tr = avg4(T, TR);
    uint32_t p = 0xff000000u;                       // 0
    p = mode == 1 ? L : p;
    p = mode == 2 ? T : p;
    p = mode == 3 ? TR : p;
    p = mode == 4 ? TL : p;
    p = mode == 5 ? avg4(avg4(L, TR), T) : p;
    p = mode == 6 ? a_l_tl : p;
    p = mode == 7 ? avg4(L, T) : p;
    p = mode == 8 ? avg4(TL, T) : p;
    p = mode == 9 ? a_t_tr : p;
    p = mode == 10 ? avg4(a_l_tl, a_t_tr) : p;
    p = mode == 11 ? select4(L, T, TL) : p;
    if (mode >= 12) p = mode == 12 ? clamp_full(L, T, TL) : clamp_half(L, T, TL);  // the two per-channel ones: rarer, kept behind a branch
    return p;
}

// the same for a mode that every lane of the step shares (a plane with one predictor, or lanes inside one block):
// only that predictor is computed
__device__ __forceinline__ uint32_t predict_one(int mode, uint32_t L, uint32_t T, uint32_t TL, uint32_t TR) {
    switch (mode) {
        case 0: return 0xff000000u;
        case 1: return L;
        case 2: return T;
        case 3: return TR;
        case 4: return TL;
        case 5: return avg4(avg4(L, TR), T);
        case 6: return avg4(L, TL);
        case 7: return avg4(L, T);
        case 8: return avg4(TL, T);
        case 9: return avg4(T, TR);
        case 10: return avg4(avg4(L, TL), avg4(T, TR));
        case 11: return select4(L, T, TL);
        case 12: return clamp_full(L, T, TL);
        case 13: return clamp_half(L, T, TL);
        default: return 0u;  // the reference leaves such pixels as they are
    }
}

__device__ __forceinline__ uint2 ld_unit(const uint2* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p));
    return make_uint2((uint32_t)v, (uint32_t)(v >> 32));
}
__device__ __forceinline__ void st_unit(uint2* p, uint32_t px, uint32_t tag) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"((unsigned long long)px | ((unsigned long long)tag << 32)) : "memory");
}

struct StageArgs {
    Vp8lBatch b;
    int stage;
};

// which of the plane's buffers holds the image before stage s (vp8l.cuh): the upload, then [1] and [2] in turn
__device__ __forceinline__ int buffer_before(int s) { return s == 0 ? 0 : 1 + ((s - 1) & 1); }

// One stage of every plane.  Work unit = a band of 32 rows of one plane, handed out band-major across the planes
// by a ticket counter, so that a predictor band only ever waits for a band with a lower ticket.
//   predictor: lane j owns row 32g + j and runs two columns behind lane j-1, so that top-right, top and top-left
//   arrive by one shuffle per step; lane 0 takes them from the hand-off records of the band above (8-byte units
//   {pixel, tag}, the same protocol as the macroblock wavefront's edge records), the band's last row writes them.
//   colour / subtract-green / palette: the warp sweeps its 32 rows, lanes along x.
__global__ void __launch_bounds__(kWarps * 32) k_vp8l_stage(StageArgs A) {
    // predictor stage: per lane (= row) a ring of 16 residuals, filled 8 at a time one block of 8 ahead of use
    // (17-word rows: lanes fall into different banks)
    __shared__ uint32_t s_res[kWarps][32][17];
    const int lane = threadIdx.x & 31;
    uint32_t* const ring = s_res[threadIdx.x >> 5][lane];
    const Vp8lBatch& B = A.b;
    const int s = A.stage;
    const unsigned max_bands = (B.max_height + 31) >> 5;
    const unsigned total = max_bands * (unsigned)B.n_planes;
    for (;;) {
        unsigned t = 0;
        if (lane == 0) t = atomicAdd(B.ticket, 1u);
        t = __shfl_sync(kFull, t, 0);
        if (t >= total) return;
        const int g = (int)(t / (unsigned)B.n_planes);
        const Vp8lPlane& P = B.planes[t % (unsigned)B.n_planes];
        if (s >= (int)P.n_stages || g * 32 >= (int)P.height) continue;
        const Vp8lStage& S = P.st[s];
        const int h = (int)P.height, w = (int)S.w_in;
        const uint32_t* im = B.pixels + P.buf_off[buffer_before(s)];
        uint32_t* om = B.pixels + P.buf_off[1 + (s & 1)];
        const uint32_t* data = B.aux + S.data_off;
        const int y0 = g * 32, rows = min(32, h - y0);

        if (S.kind == VP8L_SUBTRACT_GREEN) {  // reverse_transform.rs:393-398
            for (int r = 0; r < rows; r++) {
                const uint32_t* row = im + (size_t)(y0 + r) * w;
                uint32_t* orow = om + (size_t)(y0 + r) * w;
                for (int x0 = lane; x0 < w; x0 += 128) {
                    uint32_t p[4];
#pragma unroll
                    for (int k = 0; k < 4; k++) p[k] = x0 + 32 * k < w ? row[x0 + 32 * k] : 0u;
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        const uint32_t gg = (p[k] >> 8) & 0xffu;
                        if (x0 + 32 * k < w) orow[x0 + 32 * k] = (p[k] & 0xff00ff00u) | (((p[k] & 0x00ff00ffu) + (gg << 16 | gg)) & 0x00ff00ffu);
                    }
                }
            }
        } else if (S.kind == VP8L_COLOR) {  // apply_color_transform, 353-391
            const int bx = (w + (1 << S.bits) - 1) >> S.bits;
            for (int r = 0; r < rows; r++) {
                const int y = y0 + r;
                const uint32_t* row = im + (size_t)y * w;
                uint32_t* orow = om + (size_t)y * w;
                const uint32_t* mrow = data + (size_t)(y >> S.bits) * bx;
                for (int x0 = lane; x0 < w; x0 += 128) {
                  uint32_t pp[4];
#pragma unroll
                  for (int k = 0; k < 4; k++) pp[k] = x0 + 32 * k < w ? row[x0 + 32 * k] : 0u;
#pragma unroll
                  for (int k = 0; k < 4; k++) {
                    const int x = x0 + 32 * k;
                    if (x >= w) break;
                    const uint32_t m = mrow[x >> S.bits], p = pp[k];
                    const int g2r = (int)(int8_t)(m & 0xff), g2b = (int)(int8_t)((m >> 8) & 0xff), r2b = (int)(int8_t)((m >> 16) & 0xff);
                    const int gg = (int)(int8_t)((p >> 8) & 0xff);
                    int rr = (int)((p >> 16) & 0xff), bb = (int)(p & 0xff);
                    rr = (rr + ((g2r * gg) >> 5)) & 0xff;
                    bb = (bb + ((g2b * gg) >> 5)) & 0xff;
                    bb = (bb + ((r2b * (int)(int8_t)rr) >> 5)) & 0xff;
                    orow[x] = (p & 0xff00ff00u) | ((uint32_t)rr << 16) | (uint32_t)bb;
                  }
                }
            }
        } else if (S.kind == VP8L_PALETTE) {  // apply_color_indexing_transform, 400-578: into the other buffer
            uint32_t* out = om;
            const int W = (int)P.width, n = (int)S.table_size;
            const int pack = (int)S.bits, bpe = 8 >> pack, mask = (1 << bpe) - 1;
            for (int r = 0; r < rows; r++) {
                const uint32_t* row = im + (size_t)(y0 + r) * w;
                uint32_t* orow = out + (size_t)(y0 + r) * W;
                for (int x0 = lane; x0 < W; x0 += 128) {
                    uint32_t q[4];
#pragma unroll
                    for (int k = 0; k < 4; k++) q[k] = x0 + 32 * k < W ? row[(x0 + 32 * k) >> pack] : 0u;
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const int x = x0 + 32 * j;
                        if (x >= W) break;
                        const int k = ((int)((q[j] >> 8) & 0xffu) >> ((x & ((1 << pack) - 1)) * bpe)) & mask;
                        orow[x] = k < n ? data[k] : 0u;
                    }
                }
            }
        } else {  // predictor transform, apply_predictor_transform 24-78 + the fourteen predictors 79-351
            const int bx = (w + (1 << S.bits) - 1) >> S.bits;
            const int y = y0 + lane;
            const bool row_ok = lane < rows;
            const uint32_t* row = im + (size_t)(row_ok ? y : y0) * w;
            uint32_t* orow = om + (size_t)(row_ok ? y : y0) * w;
            const uint32_t* mrow = data + (size_t)((row_ok ? y : y0) >> S.bits) * bx;
            const uint32_t tag_in = ((uint32_t)(s + 1) << 16) | (uint32_t)g, tag_out = ((uint32_t)(s + 1) << 16) | (uint32_t)(g + 1);
            const uint2* above = B.handoff + P.hand_off + (size_t)((g + 1) & 1) * w;  // written by band g - 1
            uint2* below = B.handoff + P.hand_off + (size_t)(g & 1) * w;
            const bool hands = lane == rows - 1 && y0 + rows < h;                      // the band's last row, if a band follows
            uint32_t L = 0, T = 0, TL = 0, mine = 0, first = 0;
            uint2 nxt = make_uint2(0u, 0u);
            unsigned spins_total = 0;
            // the whole warp waits with lane 0 (a loop that one lane runs alone leaves the warp on the compiler's
            // "not converged" code path afterwards, several times slower): the exit is a warp-uniform vote
            auto wait_unit = [&](const uint2* u, uint2 v, bool want) -> uint32_t {
                unsigned spins = 0;
                bool ok = !want || v.y == tag_in;
                while (!__all_sync(kFull, ok)) {
                    __nanosleep(40);
                    if (!ok) {
                        v = ld_unit(u);
                        ok = v.y == tag_in;
                    }
                    if ((++spins & 0x3ff) == 0) {
                        spins_total += 0x400;
                        if (spins_total > (1u << 24)) atomicExch(B.abort_word, 1u);
                        if (*reinterpret_cast<volatile unsigned int*>(B.abort_word)) break;
                    }
                }
                return v.x;
            };
            if (g > 0) {
                // lane 0's "lane above" is the last row of the band above: its pixel 0 goes into the pipeline as
                // `top` now, the others arrive one per step as `top-right`
                uint2 first_unit = make_uint2(0u, 0u);
                if (lane == 0) first_unit = ld_unit(above);
                const uint32_t t0 = wait_unit(above, first_unit, lane == 0);
                if (lane == 0) {
                    T = t0;
                    if (w > 1) nxt = ld_unit(above + 1);
                }
            }
            // residuals: block b (columns 8b..8b+7) of my row sits in the ring while it is used, block b + 1 is in
            // flight in `ahead` and is put into the ring when block b + 1 begins, block b + 2 is requested then
            uint32_t ahead[8];
            auto load8 = [&](int xb) {
#pragma unroll
                for (int k = 0; k < 8; k++) ahead[k] = (row_ok && xb + k < w) ? row[xb + k] : 0u;
            };
            auto store8 = [&](int xb) {
#pragma unroll
                for (int k = 0; k < 8; k++) ring[(xb + k) & 15] = ahead[k];
            };
            load8(0);
            store8(0);
            load8(8);
            // predictor mode of my row's current block and of the next one
            uint32_t mode_w = mrow[0], mode_next = bx > 1 ? mrow[1] : 0u;
            const int bmask = (1 << S.bits) - 1;
            const int steps = w + 2 * 31;
            for (int i = 0; i < steps; i++) {
                // top-right of my column: what the lane above finished in the previous step (its column x + 1)
                uint32_t tr = __shfl_up_sync(kFull, mine, 1);
                const int x = i - 2 * lane;
                if (g > 0) {
                    // (x of lane 0 is i: the condition below is the same for every lane that evaluates it)
                    const bool want = lane == 0 && i + 1 < w;
                    const uint32_t unit = wait_unit(above + (lane == 0 ? i + 1 : 0), nxt, want);  // requested one step earlier
                    if (lane == 0) {
                        tr = want ? unit : 0u;
                        if (i + 2 < w) nxt = ld_unit(above + i + 2);
                    }
                } else if (lane == 0) {
                    tr = 0;
                }
                const uint32_t TRv = tr;
                const bool in = row_ok && x >= 0 && x < w;
                if (in && (x & bmask) == 0 && x > 0) {
                    mode_w = mode_next;
                    const int nb = (x >> S.bits) + 1;
                    mode_next = nb < bx ? mrow[nb] : 0u;
                }
                // do all the lanes that predict in this step use the same predictor?  (a vote: the warp is converged here)
                const int mode = (int)((mode_w >> 8) & 0xffu);
                const bool general = in && y > 0 && x > 0;
                const int mode0 = (int)__reduce_max_sync(kFull, general ? (unsigned)mode : 0u);
                const bool uniform = __all_sync(kFull, !general || mode == mode0);
                if (in) {
                    if ((x & 7) == 0 && x > 0) {
                        store8(x);
                        load8(x + 8);
                    }
                    const uint32_t res = ring[x & 15];
                    uint32_t pred;
                    if (y == 0) pred = x == 0 ? 0xff000000u : L;          // top row: left; first pixel: opaque black
                    else if (x == 0) pred = T;                              // left column: top
                    else {
                        // the last column's top-right is the first pixel of the CURRENT row (memory order, 3 | 108-116)
                        const uint32_t trp = x == w - 1 ? first : TRv;
                        if (uniform) {
                            pred = predict_one(mode0, L, T, TL, trp);
                        } else {
                            pred = predict_mixed(mode, L, T, TL, trp);
                            if (mode >= 14) pred = 0u;  // the reference leaves such pixels as they are
                        }
                    }
                    mine = add4(res, pred);
                    orow[x] = mine;
                    if (x == 0) first = mine;
                    L = mine;
                    if (hands) st_unit(below + x, mine, tag_out);
                }
                // shift the pipeline: next step's top-left = this step's top, top = this step's top-right
                TL = T;
                T = TRv;
            }
        }
    }
}

__global__ void k_vp8l_green(Vp8lBatch B) {
    const Vp8lPlane& P = B.planes[blockIdx.y];
    const size_t n = (size_t)P.width * P.height;
    const uint32_t* im = B.pixels + P.buf_off[buffer_before((int)P.n_stages)];
    uint8_t* a = B.alpha + P.alpha_off;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) a[i] = (uint8_t)(im[i] >> 8);
}

}  // namespace

cudaError_t launch_vp8l_inverse(const Vp8lBatch& b, cudaStream_t s, int* launches) {
    if (b.n_planes <= 0) return cudaSuccess;
    const unsigned bands = ((b.max_height + 31) >> 5) * (unsigned)b.n_planes;
    const unsigned want = (bands + kWarps - 1) / kWarps;
    const unsigned grid = want < (unsigned)b.sm_count * 8u ? want : (unsigned)b.sm_count * 8u;
    for (int st = 0; st < b.max_stages; st++) {
        cudaError_t e = cudaMemsetAsync(b.ticket, 0, 4, s);
        if (e != cudaSuccess) return e;
        StageArgs a;
        a.b = b;
        a.stage = st;
        k_vp8l_stage<<<grid ? grid : 1, kWarps * 32, 0, s>>>(a);
        ++*launches;
    }
    k_vp8l_green<<<dim3(b.sm_count * 2, b.n_planes), 256, 0, s>>>(b);
    ++*launches;
    return cudaGetLastError();
}

}  // namespace webpb200
