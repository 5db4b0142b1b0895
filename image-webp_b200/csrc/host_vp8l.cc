// host_vp8l.cc -- host-side VP8L (lossless) decoder, used only to expand VP8L-compressed
// ALPH chunks into the alpha residual plane the GPU merges (reference extended.rs:284-294).
// Entropy-coded and serial, so it stays on the host like the boolean decoder.
//
// Pixels are 32-bit ARGB words here (the reference keeps RGBA bytes); results follow the
// reference's decoder, lossless/decoder/{mod,huffman,reverse_transform}.rs, including the
// places where it departs from libwebp: predictor selector = the whole green byte with
// 14 and 15 doing nothing (reverse_transform.rs:50-70), and a colour cache that is not
// refreshed by cache hits or distance-1 copies (decoder/mod.rs:537-600).
#include "host_vp8l.h"

#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>

#include "../../include/webpb200.h"

namespace webpb200 {
namespace {

struct Bits {  // LSB-first bit reader (decoder/mod.rs:697-774)
    const uint8_t* p;
    const uint8_t* end;
    uint64_t acc = 0;
    int n = 0;
    // BitReader::fill with the reference's exact bookkeeping (decoder/mod.rs:718-738): with 8 or more bytes left a
    // 64-bit lookahead is ORed in, the position moves by (63 - n) / 8 bytes and n |= 56; otherwise byte by byte
    // while n < 56.  So 56..63 bits count as available after a fill -- which decides what the reference REJECTS:
    // a backward reference can need 58 bits (15 + 10 + 15 + 18) and get_copy_distance (619-633) does not refill.
    void fill() {
        if (end - p >= 8) {
            uint64_t look;
            std::memcpy(&look, p, 8);  // little-endian host
            p += (63 - n) / 8;
            acc |= look << n;
            n |= 56;
        } else {
            while (p < end && n < 56) {
                acc |= (uint64_t)*p++ << n;
                n += 8;
            }
        }
    }
    bool drop(int k) {
        if (n < k) return false;
        acc >>= k;
        n -= k;
        return true;
    }
    bool read(int k, uint32_t* v) {
        if (n < k) fill();
        *v = (uint32_t)(acc & ((1ull << k) - 1));
        return drop(k);
    }
};

struct Code {  // prefix code: flat lookup on the next `width` bits
    bool single = true;
    uint16_t only = 0;
    int width = 0;
    std::vector<uint32_t> lut;  // symbol | length << 16

    int build(const std::vector<uint8_t>& len) {  // huffman.rs:46-177
        int hist[16] = {0}, used = 0;
        for (uint8_t l : len) {
            hist[l]++;
            used += l != 0;
        }
        if (used == 0) return WEBPB200_ERR_HUFFMAN;
        if (used == 1) {
            single = true;
            for (size_t s = 0; s < len.size(); s++)
                if (len[s]) only = (uint16_t)s;
            return 0;
        }
        int maxl = 15;
        while (maxl > 1 && !hist[maxl]) maxl--;
        uint32_t space = 0;
        for (int l = 1; l <= maxl; l++) space = (space << 1) + (uint32_t)hist[l];
        if (space != (1u << maxl)) return WEBPB200_ERR_HUFFMAN;
        single = false;
        width = maxl;
        lut.assign((size_t)1 << maxl, 0);
        uint32_t next[17] = {0}, c = 0;
        for (int l = 1; l <= maxl; l++) {
            next[l] = c;
            c = (c + (uint32_t)hist[l]) << 1;
        }
        for (size_t s = 0; s < len.size(); s++) {
            int l = len[s];
            if (!l) continue;
            uint32_t code = next[l]++, rev = 0;
            for (int i = 0; i < l; i++) rev |= ((code >> i) & 1) << (l - 1 - i);
            for (uint32_t k = rev; k < (1u << maxl); k += 1u << l) lut[k] = (uint32_t)s | ((uint32_t)l << 16);
        }
        return 0;
    }
    void pair(uint16_t zero, uint16_t one) {  // huffman.rs:183-189
        single = false;
        width = 1;
        lut = {zero | (1u << 16), one | (1u << 16)};
    }
    bool get(Bits& b, uint32_t* sym) const {
        if (single) {
            *sym = only;
            return true;
        }
        uint32_t e = lut[b.acc & ((1u << width) - 1)];
        *sym = e & 0xffff;
        return b.drop((int)(e >> 16));
    }
};

struct Group {
    Code c[5];  // green+length+cache, red, blue, alpha, distance
};

const uint16_t kAlphabet[5] = {280, 256, 256, 256, 40};
const uint8_t kLengthOrder[19] = {17, 18, 0, 1, 2, 3, 4, 5, 16, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15};
// (dx, dy) of the 120 short distance codes (lossless/mod.rs:15-31)
const int8_t kNear[120][2] = {
    {0, 1},  {1, 0},  {1, 1},  {-1, 1}, {0, 2},  {2, 0},  {1, 2},  {-1, 2}, {2, 1},  {-2, 1}, {2, 2},  {-2, 2},
    {0, 3},  {3, 0},  {1, 3},  {-1, 3}, {3, 1},  {-3, 1}, {2, 3},  {-2, 3}, {3, 2},  {-3, 2}, {0, 4},  {4, 0},
    {1, 4},  {-1, 4}, {4, 1},  {-4, 1}, {3, 3},  {-3, 3}, {2, 4},  {-2, 4}, {4, 2},  {-4, 2}, {0, 5},  {3, 4},
    {-3, 4}, {4, 3},  {-4, 3}, {5, 0},  {1, 5},  {-1, 5}, {5, 1},  {-5, 1}, {2, 5},  {-2, 5}, {5, 2},  {-5, 2},
    {4, 4},  {-4, 4}, {3, 5},  {-3, 5}, {5, 3},  {-5, 3}, {0, 6},  {6, 0},  {1, 6},  {-1, 6}, {6, 1},  {-6, 1},
    {2, 6},  {-2, 6}, {6, 2},  {-6, 2}, {4, 5},  {-4, 5}, {5, 4},  {-5, 4}, {3, 6},  {-3, 6}, {6, 3},  {-6, 3},
    {0, 7},  {7, 0},  {1, 7},  {-1, 7}, {5, 5},  {-5, 5}, {7, 1},  {-7, 1}, {4, 6},  {-4, 6}, {6, 4},  {-6, 4},
    {2, 7},  {-2, 7}, {7, 2},  {-7, 2}, {3, 7},  {-3, 7}, {7, 3},  {-7, 3}, {5, 6},  {-5, 6}, {6, 5},  {-6, 5},
    {8, 0},  {4, 7},  {-4, 7}, {7, 4},  {-7, 4}, {8, 1},  {8, 2},  {6, 6},  {-6, 6}, {8, 3},  {5, 7},  {-5, 7},
    {7, 5},  {-7, 5}, {8, 4},  {6, 7},  {-6, 7}, {7, 6},  {-7, 6}, {8, 5},  {7, 7},  {-7, 7}, {8, 6},  {8, 7}};

inline int ceil_shift(int v, int bits) { return (int)(((uint32_t)v + (1u << bits) - 1) >> bits); }

#define NEED(expr)                                    \
    do {                                              \
        if (!(expr)) return WEBPB200_ERR_BITSTREAM;   \
    } while (0)

struct Decoder {
    Bits br;
    int width = 0, height = 0;

    struct Xform {
        int kind = -1, bits = 0, table_size = 0;
        std::vector<uint32_t> data;
    };
    Xform xf[4];
    bool have[4] = {false, false, false, false};
    int order[4], norder = 0;

    int read_code(int alphabet, Code* out) {  // decoder/mod.rs:342-382
        uint32_t v;
        NEED(br.read(1, &v));
        if (v) {
            uint32_t two, wide, s0, s1;
            NEED(br.read(1, &two));
            NEED(br.read(1, &wide));
            NEED(br.read(wide ? 8 : 1, &s0));
            if ((int)s0 >= alphabet) return WEBPB200_ERR_BITSTREAM;
            if (!two) {
                out->single = true;
                out->only = (uint16_t)s0;
                return 0;
            }
            NEED(br.read(8, &s1));
            if ((int)s1 >= alphabet) return WEBPB200_ERR_BITSTREAM;
            out->pair((uint16_t)s0, (uint16_t)s1);
            return 0;
        }
        std::vector<uint8_t> cl(19, 0);
        uint32_t ncl;
        NEED(br.read(4, &ncl));
        for (uint32_t i = 0; i < 4 + ncl; i++) {
            NEED(br.read(3, &v));
            cl[kLengthOrder[i]] = (uint8_t)v;
        }
        Code meta;
        int st = meta.build(cl);
        if (st) return st;
        // read_huffman_code_lengths, decoder/mod.rs:385-452
        int budget = alphabet;
        NEED(br.read(1, &v));
        if (v) {
            uint32_t lb, m;
            NEED(br.read(3, &lb));
            NEED(br.read(2 + 2 * (int)lb, &m));
            if ((int)m > alphabet - 2) return WEBPB200_ERR_BITSTREAM;
            budget = 2 + (int)m;
        }
        std::vector<uint8_t> len((size_t)alphabet, 0);
        int prev = 8, s = 0;
        while (s < alphabet && budget-- > 0) {
            br.fill();
            uint32_t c;
            NEED(meta.get(br, &c));
            if (c < 16) {
                len[s++] = (uint8_t)c;
                if (c) prev = (int)c;
                continue;
            }
            static const int extra[3] = {2, 3, 7}, base[3] = {3, 3, 11};
            uint32_t r;
            NEED(br.read(extra[c - 16], &r));
            int rep = (int)r + base[c - 16];
            if (s + rep > alphabet) return WEBPB200_ERR_BITSTREAM;
            const uint8_t fill = c == 16 ? (uint8_t)prev : 0;
            while (rep-- > 0) len[s++] = fill;
        }
        return out->build(len);
    }

    // One entropy-coded image (decode_image_stream, decoder/mod.rs:159-174 + 272-340 + 455-601)
    int read_pixels(int xs, int ys, bool top_level, uint32_t* out) {
        uint32_t v;
        int cache_bits = 0;
        NEED(br.read(1, &v));
        if (v) {
            NEED(br.read(4, &v));
            if (v < 1 || v > 11) return WEBPB200_ERR_COLOR_CACHE;
            cache_bits = (int)v;
        }
        int mbits = 0, mxs = 1;
        std::vector<uint16_t> meta;
        uint32_t ngroups = 1;
        if (top_level) {
            NEED(br.read(1, &v));
            if (v) {
                NEED(br.read(3, &v));
                mbits = (int)v + 2;
                mxs = ceil_shift(xs, mbits);
                const int mys = ceil_shift(ys, mbits);
                std::vector<uint32_t> img((size_t)mxs * mys);
                int st = read_pixels(mxs, mys, false, img.data());
                if (st) return st;
                meta.resize(img.size());
                for (size_t i = 0; i < img.size(); i++) {
                    meta[i] = (uint16_t)((img[i] >> 8) & 0xffff);
                    if (meta[i] >= ngroups) ngroups = meta[i] + 1u;
                }
            }
        }
        std::vector<Group> groups(ngroups);
        for (auto& g : groups)
            for (int j = 0; j < 5; j++) {
                int st = read_code(kAlphabet[j] + (j == 0 && cache_bits ? (1 << cache_bits) : 0), &g.c[j]);
                if (st) return st;
            }
        std::vector<uint32_t> cache;
        if (cache_bits) cache.assign((size_t)1 << cache_bits, 0);
        auto remember = [&](uint32_t argb) { cache[(0x1e35a7bdu * argb) >> (32 - cache_bits)] = argb; };

        const size_t total = (size_t)xs * ys;
        const uint32_t mask = mbits ? (1u << mbits) - 1 : 0xffffu;
        const Group* g = &groups[0];
        size_t i = 0, block_end = 0;
        while (i < total) {
            br.fill();
            if (i >= block_end) {
                const size_t x = i % (size_t)xs, y = i / (size_t)xs;
                size_t xe = x | mask;
                if (xe > (size_t)xs - 1) xe = (size_t)xs - 1;
                block_end = xe + y * (size_t)xs + 1;
                g = &groups[mbits ? meta[(y >> mbits) * (size_t)mxs + (x >> mbits)] : 0];
                if (g->c[0].single && g->c[1].single && g->c[2].single && g->c[3].single && g->c[0].only < 256) {
                    const uint32_t px = ((uint32_t)g->c[3].only << 24) | ((uint32_t)(g->c[1].only & 0xff) << 16) |
                                        ((uint32_t)g->c[0].only << 8) | (g->c[2].only & 0xff);
                    const size_t run = mbits ? block_end - i : total;
                    for (size_t k = 0; k < run; k++) out[i + k] = px;
                    if (cache_bits) remember(px);
                    i += run;
                    continue;
                }
            }
            uint32_t s;
            NEED(g->c[0].get(br, &s));
            if (s < 256) {
                uint32_t px;
                if (g->c[1].single && g->c[2].single && g->c[3].single) {
                    // only green is coded (the usual ALPH plane): the other channels cost no bits, and nothing the
                    // reference does between them (decoder/mod.rs:505-520) can fail
                    px = ((uint32_t)(g->c[3].only & 0xff) << 24) | ((uint32_t)(g->c[1].only & 0xff) << 16) | (s << 8) | (g->c[2].only & 0xff);
                } else {
                    uint32_t r, b, a;
                    NEED(g->c[1].get(br, &r));
                    NEED(g->c[2].get(br, &b));
                    if (br.n < 15) br.fill();
                    NEED(g->c[3].get(br, &a));
                    px = ((a & 0xff) << 24) | ((r & 0xff) << 16) | (s << 8) | (b & 0xff);
                }
                out[i++] = px;
                if (cache_bits) remember(px);
            } else if (s < 280) {
                size_t length, dcode;
                if (!copy_value(s - 256, &length)) return WEBPB200_ERR_BITSTREAM;
                uint32_t ds;
                // (no refill here: the reference has none either, decoder/mod.rs:536-541, so a reference that needs
                // more bits than one fill left is a BitStreamError there and here)
                NEED(g->c[4].get(br, &ds));
                if (!copy_value(ds, &dcode)) return WEBPB200_ERR_BITSTREAM;
                size_t dist;
                if (dcode > 120) {
                    dist = dcode - 120;
                } else {
                    const int dd = kNear[dcode - 1][0] + kNear[dcode - 1][1] * xs;
                    dist = dd < 1 ? 1 : (size_t)dd;
                }
                if (i < dist || total - i < length) return WEBPB200_ERR_BITSTREAM;
                for (size_t k = 0; k < length; k++) out[i + k] = out[i + k - dist];
                if (cache_bits && dist != 1)
                    for (size_t k = 0; k < length; k++) remember(out[i + k]);
                i += length;
            } else {
                if (!cache_bits || s - 280 >= cache.size()) return WEBPB200_ERR_BITSTREAM;
                out[i++] = cache[s - 280];
            }
        }
        return 0;
    }

    bool copy_value(uint32_t prefix, size_t* out) {  // get_copy_distance
        if (prefix < 4) {
            *out = prefix + 1;
            return true;
        }
        const int extra = (int)(prefix - 2) >> 1;
        const size_t offset = (size_t)(2 + (prefix & 1)) << extra;
        const size_t bits = (size_t)(br.acc & ((1ull << extra) - 1));
        if (!br.drop(extra)) return false;
        *out = offset + bits + 1;
        return true;
    }

    int read_transforms(int* xs_out) {  // decoder/mod.rs:177-261
        int xs = width;
        for (;;) {
            uint32_t v;
            NEED(br.read(1, &v));
            if (!v) break;
            NEED(br.read(2, &v));
            if (have[v]) return WEBPB200_ERR_TRANSFORM;
            have[v] = true;
            order[norder++] = (int)v;
            Xform& t = xf[v];
            t.kind = (int)v;
            if (v == 0 || v == 1) {
                uint32_t sb;
                NEED(br.read(3, &sb));
                t.bits = (int)sb + 2;
                const int bx = ceil_shift(xs, t.bits), by = ceil_shift(height, t.bits);
                t.data.assign((size_t)bx * by, 0);
                int st = read_pixels(bx, by, false, t.data.data());
                if (st) return st;
            } else if (v == 3) {
                uint32_t n;
                NEED(br.read(8, &n));
                t.table_size = (int)n + 1;
                t.data.assign((size_t)t.table_size, 0);
                int st = read_pixels(t.table_size, 1, false, t.data.data());
                if (st) return st;
                const int pack = t.table_size <= 2 ? 3 : t.table_size <= 4 ? 2 : t.table_size <= 16 ? 1 : 0;
                xs = ceil_shift(xs, pack);
                for (int i = 1; i < t.table_size; i++) t.data[i] = add(t.data[i], t.data[i - 1]);
            }
        }
        *xs_out = xs;
        return 0;
    }

    static uint32_t add(uint32_t a, uint32_t b) {  // per-channel wrapping add
        return (((a & 0xff00ff00u) + (b & 0xff00ff00u)) & 0xff00ff00u) | (((a & 0x00ff00ffu) + (b & 0x00ff00ffu)) & 0x00ff00ffu);
    }
    static uint32_t avg(uint32_t a, uint32_t b) { return (((a ^ b) & 0xfefefefeu) >> 1) + (a & b); }
    static int ch(uint32_t p, int c) { return (int)((p >> (8 * c)) & 0xff); }
    static uint32_t clampsum_full(uint32_t l, uint32_t t, uint32_t tl) {
        uint32_t r = 0;
        for (int c = 0; c < 4; c++) {
            int v = ch(l, c) + ch(t, c) - ch(tl, c);
            v = v < 0 ? 0 : v > 255 ? 255 : v;
            r |= (uint32_t)v << (8 * c);
        }
        return r;
    }
    static uint32_t clampsum_half(uint32_t l, uint32_t t, uint32_t tl) {
        uint32_t r = 0;
        for (int c = 0; c < 4; c++) {
            const int a = (ch(l, c) + ch(t, c)) / 2;
            int v = a + (a - ch(tl, c)) / 2;
            v = v < 0 ? 0 : v > 255 ? 255 : v;
            r |= (uint32_t)v << (8 * c);
        }
        return r;
    }
    static uint32_t select(uint32_t l, uint32_t t, uint32_t tl) {
        int dl = 0, dt = 0;
        for (int c = 0; c < 4; c++) {
            dl += std::abs(ch(t, c) - ch(tl, c));  // |predict - L|
            dt += std::abs(ch(l, c) - ch(tl, c));  // |predict - T|
        }
        return dl < dt ? l : t;
    }

    void undo_predictor(uint32_t* im, int w, const Xform& t) {  // reverse_transform.rs:24-351
        const int bx = ceil_shift(w, t.bits);
        im[0] = add(im[0], 0xff000000u);
        for (int x = 1; x < w; x++) im[x] = add(im[x], im[x - 1]);
        for (int y = 1; y < height; y++) {
            uint32_t* row = im + (size_t)y * w;
            const uint32_t* up = row - w;
            row[0] = add(row[0], up[0]);
            for (int x = 1; x < w; x++) {
                const int mode = (int)((t.data[(size_t)(y >> t.bits) * bx + (x >> t.bits)] >> 8) & 0xff);
                const uint32_t L = row[x - 1], T = up[x], TL = up[x - 1], TR = up[x + 1];
                uint32_t pred;
                switch (mode) {
                    case 0: pred = 0xff000000u; break;
                    case 1: pred = L; break;
                    case 2: pred = T; break;
                    case 3: pred = TR; break;
                    case 4: pred = TL; break;
                    case 5: pred = avg(avg(L, TR), T); break;
                    case 6: pred = avg(L, TL); break;
                    case 7: pred = avg(L, T); break;
                    case 8: pred = avg(TL, T); break;
                    case 9: pred = avg(T, TR); break;
                    case 10: pred = avg(avg(L, TL), avg(T, TR)); break;
                    case 11: pred = select(L, T, TL); break;
                    case 12: pred = clampsum_full(L, T, TL); break;
                    case 13: pred = clampsum_half(L, T, TL); break;
                    default: continue;  // the reference leaves such pixels as they are
                }
                row[x] = add(row[x], pred);
            }
        }
    }

    void undo_color(uint32_t* im, int w, size_t rows, const Xform& t) {  // reverse_transform.rs:353-391
        const int bx = ceil_shift(w, t.bits);
        for (size_t y = 0; y < rows; y++)
            for (int x = 0; x < w; x++) {
                const uint32_t m = t.data[(y >> t.bits) * (size_t)bx + (x >> t.bits)];
                const int8_t g2r = (int8_t)(m & 0xff), g2b = (int8_t)((m >> 8) & 0xff), r2b = (int8_t)((m >> 16) & 0xff);
                uint32_t& p = im[y * (size_t)w + x];
                const int8_t g = (int8_t)((p >> 8) & 0xff);
                int r = (int)((p >> 16) & 0xff), b = (int)(p & 0xff);
                r = (r + ((g2r * g) >> 5)) & 0xff;
                b = (b + ((g2b * g) >> 5)) & 0xff;
                b = (b + ((r2b * (int8_t)r) >> 5)) & 0xff;
                p = (p & 0xff00ff00u) | ((uint32_t)r << 16) | (uint32_t)b;
            }
    }

    void undo_palette(uint32_t* im, const Xform& t) {  // reverse_transform.rs:400-578
        const int n = t.table_size;
        auto look = [&](int k) { return k < n ? t.data[k] : 0u; };
        if (n > 16) {
            for (size_t i = 0; i < (size_t)width * height; i++) im[i] = look((int)((im[i] >> 8) & 0xff));
            return;
        }
        const int pack = n <= 2 ? 3 : n <= 4 ? 2 : 1;
        const int per = 1 << pack, bpe = 8 >> pack, m = (1 << bpe) - 1;
        const int pw = (width + per - 1) / per;
        std::vector<uint8_t> idx((size_t)pw);
        for (int y = height - 1; y >= 0; y--) {
            for (int k = 0; k < pw; k++) idx[k] = (uint8_t)((im[(size_t)y * pw + k] >> 8) & 0xff);
            for (int x = 0; x < width; x++) im[(size_t)y * width + x] = look((idx[x / per] >> ((x % per) * bpe)) & m);
        }
    }

    // the entropy-coded part only: `im` receives the image in the transformed domain (xs x height), `plan` the
    // transforms in the order they have to be undone
    int run_entropy(uint32_t* im, Vp8lPlan* plan) {
        int xs;
        int st = read_transforms(&xs);
        if (st) return st;
        st = read_pixels(xs, height, true, im);
        if (st) return st;
        plan->n_stages = 0;
        int w = xs;
        for (int k = norder - 1; k >= 0; k--) {
            Xform& t = xf[order[k]];
            Vp8lPlan::Stage& ps = plan->st[plan->n_stages++];
            ps.kind = t.kind;
            ps.bits = t.bits;
            ps.table_size = t.table_size;
            ps.w_in = w;
            ps.data.swap(t.data);
            if (t.kind == 3) {
                ps.bits = t.table_size <= 2 ? 3 : t.table_size <= 4 ? 2 : t.table_size <= 16 ? 1 : 0;
                w = width;
            }
        }
        return 0;
    }

    int run(uint32_t* im) {
        int xs;
        int st = read_transforms(&xs);
        if (st) return st;
        st = read_pixels(xs, height, true, im);
        if (st) return st;
        int w = xs;
        for (int k = norder - 1; k >= 0; k--) {
            const Xform& t = xf[order[k]];
            switch (t.kind) {
                case 0: undo_predictor(im, w, t); break;
                case 1: undo_color(im, w, (size_t)height, t); break;
                case 2:
                    for (size_t i = 0; i < (size_t)w * height; i++) {
                        const uint32_t g = (im[i] >> 8) & 0xff;
                        im[i] = (im[i] & 0xff00ff00u) | ((((im[i] & 0x00ff00ffu) + (g << 16 | g))) & 0x00ff00ffu);
                    }
                    break;
                default:
                    undo_palette(im, t);
                    w = width;
                    break;
            }
        }
        return 0;
    }
};

}  // namespace

int vp8l_decode_green(const uint8_t* data, size_t len, int width, int height, uint8_t* green) {
    if (width <= 0 || height <= 0) return WEBPB200_ERR_INVALID_PARAM;
    Decoder d;
    d.br.p = data;
    d.br.end = data + len;
    d.width = width;
    d.height = height;
    // +1 word: the top-right read of a row's last pixel looks one past it (the reference
    // reads the next row's first pixel there; for the very last pixel that word is unused)
    std::vector<uint32_t> im((size_t)width * height + 1, 0);
    int st = d.run(im.data());
    if (st) return st;
    for (size_t i = 0; i < (size_t)width * height; i++) green[i] = (uint8_t)((im[i] >> 8) & 0xff);
    return 0;
}

int vp8l_decode_entropy(const uint8_t* data, size_t len, int width, int height, uint32_t* argb, Vp8lPlan* plan) {
    if (width <= 0 || height <= 0) return WEBPB200_ERR_INVALID_PARAM;
    Decoder d;
    d.br.p = data;
    d.br.end = data + len;
    d.width = width;
    d.height = height;
    return d.run_entropy(argb, plan);
}

}  // namespace webpb200
