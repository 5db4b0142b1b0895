// host_vp8.cc -- host entropy stage of the lossy path (see host_vp8.h).
//
// Behaviour follows the reference decoder (image-webp v0.2.4): frame header
// lossy/mod.rs:510-613, segment / loop-filter / quantiser fields 397-508, partitions 369-395,
// macroblock header 615-671, coefficient tokens 785-872 with the neighbour contexts of
// 874-963 and the skip handling of 1261-1275, boolean decoder lossy/arithmetic_decoder.rs
// (one zero byte past the end of a partition is tolerated, any further read is a
// BitStreamError, 249-276).  What differs is the output: raw levels, modes and per-block
// "has coefficients" flags for the GPU instead of reconstructed pixels.
#include "host_vp8.h"

#include <cstring>
#include <vector>

#include "vp8_constants.h"

namespace webpb200 {
namespace {

// Boolean decoder with a 64-bit window.  `bits` + 8 bits of `value` are valid.
struct BoolReader {
    const uint8_t* p = nullptr;
    const uint8_t* end = nullptr;
    uint64_t value = 0;
    uint32_t range = 255;
    int bits = -8;
    bool zero_byte_used = false;
    bool eof = true;  // an unset reader is past its end (arithmetic_decoder.rs:60-76)

    void init(const uint8_t* data, size_t len) {
        p = data;
        end = data + len;
        value = 0;
        range = 255;
        bits = -8;
        zero_byte_used = false;
        eof = false;
    }

    bool refill() {
        if (end - p >= 8) {  // 7 bytes at once (an 8-byte load, the eighth byte is not consumed)
            uint64_t v;
            std::memcpy(&v, p, 8);
            v = __builtin_bswap64(v) >> 8;
            p += 7;
            value = (value << 56) | v;
            bits += 56;
            return true;
        }
        if (p < end) {
            value = (value << 8) | *p++;
            bits += 8;
            return true;
        }
        if (!zero_byte_used && !eof) {
            zero_byte_used = true;
            value <<= 8;
            bits += 8;
            return true;
        }
        eof = true;
        return false;
    }

    // One boolean.  The chain multiply -> shift -> compare -> normalise is serial by nature
    // (about 12 cycles per symbol); a branch-free form measured 6 % slower on the B200 host.
    inline int get(int prob) {
        if (__builtin_expect(bits < 0, 0) && !refill()) return 0;
        const uint32_t split = 1 + (((range - 1) * (uint32_t)prob) >> 8);
        const uint64_t big = (uint64_t)split << bits;
        int bit;
        if (value >= big) {
            range -= split;
            value -= big;
            bit = 1;
        } else {
            range = split;
            bit = 0;
        }
        const int shift = __builtin_clz(range) - 24;
        range <<= shift;
        bits -= shift;
        return bit;
    }
    inline int flag() { return get(128); }
    int literal(int n) {
        int v = 0;
        while (n-- > 0) v = (v << 1) | flag();
        return v;
    }
    int opt_signed(int n) {  // flag, magnitude, sign (1 = negative)
        if (!flag()) return 0;
        int m = literal(n);
        return flag() ? -m : m;
    }
    int tree(const int8_t* t, const uint8_t* probs) {
        int i = 0;
        do {
            i = t[i + get(probs[i >> 1])];
        } while (i > 0);
        return -i;
    }
};

// The same decoder with a 32-bit window refilled one byte at a time: slow, but its state between
// two symbols is exactly what the device reader (entropy.cu) continues from.  Header parse only.
struct ByteBoolReader {
    const uint8_t* base = nullptr;
    uint32_t pos = 0, len = 0;
    uint32_t value = 0, range = 255;
    int bits = -8;
    bool phantom = false;
    bool eof = true;

    void init(const uint8_t* data, size_t n) {
        base = data;
        pos = 0;
        len = (uint32_t)n;
        value = 0;
        range = 255;
        bits = -8;
        phantom = false;
        eof = false;
    }
    bool refill() {
        if (pos < len) {
            value = (value << 8) | base[pos++];
            bits += 8;
            return true;
        }
        if (!phantom && !eof) {
            phantom = true;
            value <<= 8;
            bits += 8;
            return true;
        }
        eof = true;
        return false;
    }
    int get(int prob) {
        if (bits < 0 && !refill()) return 0;
        const uint32_t split = 1 + (((range - 1) * (uint32_t)prob) >> 8);
        const uint32_t big = split << bits;
        int bit;
        if (value >= big) {
            range -= split;
            value -= big;
            bit = 1;
        } else {
            range = split;
            bit = 0;
        }
        const int shift = __builtin_clz(range) - 24;
        range <<= shift;
        bits -= shift;
        return bit;
    }
    int flag() { return get(128); }
    int literal(int n) {
        int v = 0;
        while (n-- > 0) v = (v << 1) | flag();
        return v;
    }
    int opt_signed(int n) {
        if (!flag()) return 0;
        int m = literal(n);
        return flag() ? -m : m;
    }
    int tree(const int8_t* t, const uint8_t* probs) {
        int i = 0;
        do {
            i = t[i + get(probs[i >> 1])];
        } while (i > 0);
        return -i;
    }
    void resume(const uint8_t* data, size_t n, const BoolState& s) {  // what the device reader does with the header's state
        base = data; len = (uint32_t)n; pos = s.pos; value = s.value; range = s.range; bits = s.bits;
        phantom = s.phantom != 0; eof = s.eof != 0;
    }
};

const int8_t kSegmentTree[6] = {2, 4, -0, -1, -2, -3};
const int8_t kYModeTree[8] = {-4, 2, 4, 6, -0, -1, -2, -3};
const int8_t kBModeTree[18] = {-0, 2, -1, 4, -2, 6, 8, 12, -3, 10, -5, -6, -4, 14, -7, 16, -8, -9};
const int8_t kUVModeTree[6] = {-0, 2, -1, 4, -2, -3};

// extra-bit probabilities of DCT_CAT3..6 (RFC 6386 13.2), zero-terminated
const uint8_t kCat3[] = {173, 148, 140, 0};
const uint8_t kCat4[] = {176, 155, 140, 135, 0};
const uint8_t kCat5[] = {180, 157, 141, 134, 130, 0};
const uint8_t kCat6[] = {254, 254, 243, 230, 196, 177, 153, 140, 133, 130, 129, 0};
const uint8_t* const kCat3456[4] = {kCat3, kCat4, kCat5, kCat6};
const int kCatBase3456[4] = {11, 19, 35, 67};

typedef uint8_t TokenProbs[8][3][11];  // [band][ctx][node] of one plane

// Tokens of one 4x4 block.  Same token tree as the reference (common.rs:659-662) written
// out as nested branches.  Returns "has coefficients" (mod.rs:809, 823-828, 868).
template <class Reader>
inline int read_block(Reader& br, const TokenProbs& probs, int ctx, int first, int16_t* out) {
    int n = first;
    const uint8_t* p = probs[kCoeffBands[n]][ctx];
    if (!br.get(p[0])) return 0;  // DCT_EOB straight away
    for (;;) {
        while (!br.get(p[1])) {  // DCT_0: no EOB test before the next token
            if (++n == 16) return 1;
            p = probs[kCoeffBands[n]][0];
        }
        int v;
        if (!br.get(p[2])) {
            v = 1;
            ctx = 1;
        } else {
            if (!br.get(p[3])) {
                v = !br.get(p[4]) ? 2 : 3 + br.get(p[5]);
            } else if (!br.get(p[6])) {
                if (!br.get(p[7])) {
                    v = 5 + br.get(159);
                } else {
                    v = 7 + 2 * br.get(165);
                    v += br.get(145);
                }
            } else {
                int b1 = br.get(p[8]);
                int b0 = br.get(p[9 + b1]);
                int cat = 2 * b1 + b0;
                v = 0;
                for (const uint8_t* t = kCat3456[cat]; *t; ++t) v = v + v + br.get(*t);
                v += kCatBase3456[cat];
            }
            ctx = 2;
        }
        if (br.flag()) v = -v;
        out[kZigzag[n]] = (int16_t)v;
        if (++n == 16) return 1;
        p = probs[kCoeffBands[n]][ctx];
        if (!br.get(p[0])) return 1;  // DCT_EOB
    }
}

inline int clamp127(int v) { return v < 0 ? 0 : v > 127 ? 127 : v; }

struct NeighbourCtx {
    uint8_t bpred[4];
    uint8_t nz[9];  // y2, y0..y3, u0,u1, v0,v1 (PreviousMacroBlock::complexity, mod.rs:124-129)
};

}  // namespace

int vp8_probe(const uint8_t* d, size_t len, Vp8Dims* dims) {
    if (len < 3) return WEBPB200_ERR_IO;
    uint32_t tag = d[0] | (d[1] << 8) | ((uint32_t)d[2] << 16);
    if (tag & 1) return WEBPB200_ERR_UNSUPPORTED;  // non-keyframe
    if (len < 6) return WEBPB200_ERR_IO;
    if (d[3] != 0x9d || d[4] != 0x01 || d[5] != 0x2a) return WEBPB200_ERR_VP8_MAGIC;
    if (len < 10) return WEBPB200_ERR_IO;
    dims->width = (d[6] | (d[7] << 8)) & 0x3FFF;
    dims->height = (d[8] | (d[9] << 8)) & 0x3FFF;
    dims->mbw = (dims->width + 15) >> 4;
    dims->mbh = (dims->height + 15) >> 4;
    return WEBPB200_OK;
}

namespace {

struct HeaderFields {  // what the macroblock loop needs from the frame header
    bool segments_enabled = false, update_map = false;
    uint8_t seg_probs[3] = {255, 255, 255};
    int num_parts = 1;
    size_t first_off = 0, first_len = 0;
    size_t part_off[8] = {}, part_len[8] = {};
    uint8_t probs[1056];
    int prob_skip = -1;
};

// Frame tag to mb_no_coeff_skip (mod.rs:510-613), with reader `b` left at the first macroblock.
template <class Reader>
int parse_frame_header(const uint8_t* d, size_t len, Reader& b, webpb200_frame_desc* fd, Vp8Dims& dims, HeaderFields& h) {
    int st = vp8_probe(d, len, &dims);
    if (st) return st;
    const uint32_t tag = d[0] | (d[1] << 8) | ((uint32_t)d[2] << 16);
    const size_t first_size = tag >> 5;
    size_t pos = 10;
    if (len - pos < first_size) return WEBPB200_ERR_IO;

    b.init(d + pos, first_size);
    h.first_off = pos;
    h.first_len = first_size;
    pos += first_size;

    fd->width = (uint16_t)dims.width;
    fd->height = (uint16_t)dims.height;
    fd->mbw = (uint16_t)dims.mbw;
    fd->mbh = (uint16_t)dims.mbh;

    // ---- frame header (mod.rs:569-611) ----
    const int color_space = b.literal(1);
    (void)b.literal(1);  // pixel_type: parsed, never used in reconstruction
    if (color_space != 0) return WEBPB200_ERR_COLOR_SPACE;

    struct SegHdr {
        bool delta = false;
        int quant = 0, lf = 0;
    } seg[4];
    const bool segments_enabled = b.flag();
    if (segments_enabled) {  // read_segment_updates, 469-508
        h.update_map = b.flag();
        if (b.flag()) {
            const bool absolute = b.flag();
            for (auto& s : seg) s.delta = !absolute;
            for (auto& s : seg) s.quant = (int8_t)b.opt_signed(7);
            for (auto& s : seg) s.lf = (int8_t)b.opt_signed(6);
        }
        if (h.update_map)
            for (int i = 0; i < 3; i++) h.seg_probs[i] = b.flag() ? (uint8_t)b.literal(8) : 255;
        if (b.eof) return WEBPB200_ERR_BITSTREAM;
    }
    h.segments_enabled = segments_enabled;
    fd->segments_enabled = segments_enabled;
    fd->filter_type = (uint8_t)b.flag();
    fd->filter_level = (uint8_t)b.literal(6);
    fd->sharpness = (uint8_t)b.literal(3);
    int ref_delta[4] = {0, 0, 0, 0}, mode_delta[4] = {0, 0, 0, 0};
    const bool lf_adj = b.flag();
    if (lf_adj) {  // read_loop_filter_adjustments, 453-467
        if (b.flag()) {
            for (int& v : ref_delta) v = b.opt_signed(6);
            for (int& v : mode_delta) v = b.opt_signed(6);
        }
        if (b.eof) return WEBPB200_ERR_BITSTREAM;
    }
    fd->lf_adj_enabled = lf_adj;
    fd->ref_delta0 = (int8_t)ref_delta[0];
    fd->mode_delta0 = (int8_t)mode_delta[0];
    const int num_parts = 1 << b.literal(2);
    if (b.eof) return WEBPB200_ERR_BITSTREAM;
    h.num_parts = num_parts;

    // ---- token partitions (init_partitions, 369-395) ----
    if (num_parts > 1) {
        const size_t table = 3 * (size_t)(num_parts - 1);
        if (len - pos < table) return WEBPB200_ERR_IO;
        const uint8_t* sz = d + pos;
        pos += table;
        for (int i = 0; i < num_parts - 1; i++) {
            size_t s = sz[3 * i] | (sz[3 * i + 1] << 8) | ((size_t)sz[3 * i + 2] << 16);
            if (len - pos < s) return WEBPB200_ERR_IO;
            h.part_off[i] = pos;
            h.part_len[i] = s;
            pos += s;
        }
    }
    h.part_off[num_parts - 1] = pos;
    h.part_len[num_parts - 1] = len - pos;

    // ---- quantiser indices (read_quantization_indices, 397-451) ----
    {
        const int yac_abs = b.literal(7);
        const int ydc_d = b.opt_signed(4), y2dc_d = b.opt_signed(4), y2ac_d = b.opt_signed(4);
        const int uvdc_d = b.opt_signed(4), uvac_d = b.opt_signed(4);
        std::memset(fd->seg, 0, sizeof(fd->seg));
        const int nseg = segments_enabled ? 4 : 1;
        for (int i = 0; i < nseg; i++) {
            int base = yac_abs;
            if (segments_enabled) base = seg[i].delta ? seg[i].quant + yac_abs : seg[i].quant;
            webpb200_segment& s = fd->seg[i];
            s.ydc = kDcQuant[clamp127(base + ydc_d)];
            s.yac = kAcQuant[clamp127(base)];
            s.y2dc = (int16_t)(kDcQuant[clamp127(base + y2dc_d)] * 2);
            s.y2ac = (int16_t)(kAcQuant[clamp127(base + y2ac_d)] * 155 / 100);
            if (s.y2ac < 8) s.y2ac = 8;
            s.uvdc = kDcQuant[clamp127(base + uvdc_d)];
            if (s.uvdc > 132) s.uvdc = 132;
            s.uvac = kAcQuant[clamp127(base + uvac_d)];
        }
        for (int i = 0; i < 4; i++) {
            fd->seg[i].lf_level = (int8_t)seg[i].lf;
            fd->seg[i].delta_values = seg[i].delta;
        }
        if (b.eof) return WEBPB200_ERR_BITSTREAM;
    }
    (void)b.literal(1);  // refresh-entropy bit, ignored (599-600)

    // ---- token probabilities (update_token_probabilities, 352-367) ----
    std::memcpy(h.probs, kCoeffProbsDefault, sizeof(h.probs));
    for (int i = 0; i < 1056; i++)
        if (b.get(kCoeffUpdateProbs[i])) h.probs[i] = (uint8_t)b.literal(8);
    if (b.eof) return WEBPB200_ERR_BITSTREAM;

    if (b.literal(1) == 1) h.prob_skip = b.literal(8);
    if (b.eof) return WEBPB200_ERR_BITSTREAM;
    return WEBPB200_OK;
}

}  // namespace

int vp8_parse_header(const uint8_t* d, size_t len, webpb200_frame_desc* fd, Vp8EntropyHdr* eh) {
    Vp8Dims dims;
    HeaderFields h;
    ByteBoolReader b;
    std::memset(eh, 0, sizeof(*eh));
    const int st = parse_frame_header(d, len, b, fd, dims, h);
    if (st) return st;
    eh->first_off = (uint32_t)h.first_off;
    eh->first_len = (uint32_t)h.first_len;
    for (int i = 0; i < 8; i++) {
        eh->part_off[i] = (uint32_t)h.part_off[i];
        eh->part_len[i] = (uint32_t)h.part_len[i];
    }
    eh->first.pos = b.pos;
    eh->first.value = b.value;
    eh->first.range = b.range;
    eh->first.bits = b.bits;
    eh->first.phantom = b.phantom;
    eh->first.eof = b.eof;
    eh->num_parts = (uint8_t)h.num_parts;
    eh->update_map = h.segments_enabled && h.update_map;
    eh->has_prob_skip = h.prob_skip >= 0;
    eh->prob_skip = (uint8_t)(h.prob_skip >= 0 ? h.prob_skip : 0);
    std::memcpy(eh->seg_probs, h.seg_probs, 3);
    std::memcpy(eh->probs, h.probs, sizeof(eh->probs));
    return WEBPB200_OK;
}

namespace {
// The macroblock loop (decode_frame_, mod.rs:1255-1281): `b` is the first-partition reader standing at
// the first macroblock, `parts` the token-partition readers.
template <class Reader>
int decode_macroblocks(Reader& b, Reader* parts, const HeaderFields& h, const Vp8Dims& dims, webpb200_mb_desc* mbs, int16_t* levels) {
    const bool segments_enabled = h.segments_enabled, update_map = h.update_map;
    const uint8_t* seg_probs = h.seg_probs;
    const int num_parts = h.num_parts, prob_skip = h.prob_skip;
    const TokenProbs* probs = reinterpret_cast<const TokenProbs*>(h.probs);  // [plane]

    // ---- macroblocks (decode_frame_, 1255-1281) ----
    const int mbw = dims.mbw, mbh = dims.mbh;
    std::vector<NeighbourCtx> top((size_t)mbw);
    std::memset(top.data(), 0, top.size() * sizeof(NeighbourCtx));
    static const uint8_t kImpliedBMode[4] = {0, 2, 3, 1};  // DC->B_DC V->B_VE H->B_HE TM->B_TM (common.rs:75-83)

    for (int mby = 0; mby < mbh; mby++) {
        Reader& tp = parts[mby % num_parts];
        NeighbourCtx left;
        std::memset(&left, 0, sizeof(left));
        for (int mbx = 0; mbx < mbw; mbx++) {
            const size_t mbi = (size_t)mby * mbw + mbx;
            webpb200_mb_desc& mb = mbs[mbi];
            int16_t* lv = levels + mbi * WEBPB200_LEVELS_PER_MB;
            NeighbourCtx& tc = top[mbx];
            std::memset(&mb, 0, sizeof(mb));
            std::memset(lv, 0, WEBPB200_LEVELS_PER_MB * sizeof(int16_t));

            // read_macroblock_header, 615-671
            if (segments_enabled && update_map) mb.segment = (uint8_t)b.tree(kSegmentTree, seg_probs);
            const bool skipped = prob_skip >= 0 ? b.get(prob_skip) : false;
            mb.flags = skipped ? 1 : 0;
            const int ymode = b.tree(kYModeTree, kYModeProbs);
            mb.ymode = (uint8_t)ymode;
            if (ymode == 4) {
                for (int y = 0; y < 4; y++)
                    for (int x = 0; x < 4; x++) {
                        const int bm = b.tree(kBModeTree, kBModeProbs + (tc.bpred[x] * 10 + left.bpred[y]) * 9);
                        const int k = y * 4 + x;
                        mb.bpred[k >> 1] |= (uint8_t)(bm << (4 * (k & 1)));
                        tc.bpred[x] = (uint8_t)bm;
                        left.bpred[y] = (uint8_t)bm;
                    }
            } else {
                const uint8_t m = kImpliedBMode[ymode];
                for (int i = 0; i < 4; i++) tc.bpred[i] = left.bpred[i] = m;
            }
            mb.uvmode = (uint8_t)b.tree(kUVModeTree, kUVModeProbs);
            if (b.eof) return WEBPB200_ERR_BITSTREAM;

            if (skipped) {  // 1263-1272
                if (ymode != 4) left.nz[0] = tc.nz[0] = 0;
                for (int i = 1; i < 9; i++) left.nz[i] = tc.nz[i] = 0;
                continue;
            }

            // read_residual_data, 874-963 (dequantisation and transforms run on the GPU)
            uint32_t nflags = 0;
            int plane = 3;  // Plane::YCoeff0
            if (ymode != 4) {
                const int n = read_block(tp, probs[1], tc.nz[0] + left.nz[0], 0, lv + 24 * 16);
                tc.nz[0] = left.nz[0] = (uint8_t)n;
                nflags |= (uint32_t)n << 24;
                plane = 0;  // Plane::YCoeff1
            }
            const int first = plane == 0 ? 1 : 0;
            for (int y = 0; y < 4; y++) {
                int l = left.nz[y + 1];
                for (int x = 0; x < 4; x++) {
                    const int k = y * 4 + x;
                    const int n = read_block(tp, probs[plane], tc.nz[x + 1] + l, first, lv + k * 16);
                    nflags |= (uint32_t)n << k;
                    l = n;
                    tc.nz[x + 1] = (uint8_t)n;
                }
                left.nz[y + 1] = (uint8_t)l;
            }
            for (int j = 5; j <= 7; j += 2) {
                const int blk0 = j == 5 ? 16 : 20;
                for (int y = 0; y < 2; y++) {
                    int l = left.nz[y + j];
                    for (int x = 0; x < 2; x++) {
                        const int k = blk0 + y * 2 + x;
                        const int n = read_block(tp, probs[2], tc.nz[x + j] + l, 0, lv + k * 16);
                        nflags |= (uint32_t)n << k;
                        l = n;
                        tc.nz[x + j] = (uint8_t)n;
                    }
                    left.nz[y + j] = (uint8_t)l;
                }
            }
            mb.nflags = nflags;
            if (tp.eof) return WEBPB200_ERR_BITSTREAM;
        }
    }
    return WEBPB200_OK;
}
}  // namespace

int vp8_entropy_decode(const uint8_t* d, size_t len, webpb200_frame_desc* fd, webpb200_mb_desc* mbs,
                       int16_t* levels) {
    Vp8Dims dims;
    HeaderFields h;
    BoolReader b;
    const int st = parse_frame_header(d, len, b, fd, dims, h);
    if (st) return st;
    BoolReader parts[8];
    for (int i = 0; i < h.num_parts; i++) parts[i].init(d + h.part_off[i], h.part_len[i]);
    return decode_macroblocks(b, parts, h, dims, mbs, levels);
}

int vp8_entropy_decode_resumed(const uint8_t* d, size_t len, webpb200_frame_desc* fd, webpb200_mb_desc* mbs, int16_t* levels) {
    Vp8EntropyHdr eh;
    int st = vp8_parse_header(d, len, fd, &eh);
    if (st) return st;
    // everything the loop needs is rebuilt from the exported header alone, as on the device
    Vp8Dims dims{fd->width, fd->height, fd->mbw, fd->mbh};
    HeaderFields h;
    h.segments_enabled = eh.update_map != 0;  // only ever used together with update_map
    h.update_map = eh.update_map != 0;
    std::memcpy(h.seg_probs, eh.seg_probs, 3);
    h.num_parts = eh.num_parts;
    h.prob_skip = eh.has_prob_skip ? eh.prob_skip : -1;
    std::memcpy(h.probs, eh.probs, sizeof(h.probs));
    ByteBoolReader b, parts[8];
    b.resume(d + eh.first_off, eh.first_len, eh.first);
    for (int i = 0; i < eh.num_parts; i++) parts[i].init(d + eh.part_off[i], eh.part_len[i]);
    return decode_macroblocks(b, parts, h, dims, mbs, levels);
}

}  // namespace webpb200
