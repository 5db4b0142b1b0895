/* oracle/ora_vp8l.c -- TEST INFRASTRUCTURE, not product code.
 *
 * Restatement in C of the reference's VP8L (lossless) decoder, which the lossy path needs
 * for compressed ALPH chunks (extended.rs:284-294) and which the oracle also uses for the
 * lossless golden files:
 *   src/lossless/mod.rs, src/lossless/decoder/mod.rs, huffman.rs, reverse_transform.rs.
 * Pixels are kept as RGBA bytes exactly as the reference does (decoder/mod.rs:17-21).
 * Reference quirks kept on purpose (they define the oracle): predictor modes 14/15 are
 * no-ops (reverse_transform.rs:69), predictor mode = whole green byte (50), colour-cache
 * is not refreshed by cache hits or distance-1 copies (decoder/mod.rs:548-574, 580-600).
 */
#include "ora.h"
#include <stdlib.h>
#include <string.h>

/* ---- BitReader, decoder/mod.rs:697-774 ------------------------------------------- */
typedef struct {
    const uint8_t* p; size_t len, pos;
    uint64_t buffer; int nbits;
} bitrd;

/* fill, 718-738, with the reference's exact bookkeeping: with 8 or more bytes left it ORs a 64-bit lookahead in,
 * advances by (63 - nbits) / 8 bytes and sets nbits |= 56 (so 56..63 bits count as available, although more are
 * in the buffer); otherwise byte by byte while nbits < 56.  It matters for what the reference REJECTS: a backward
 * reference may need 58 bits (15 + 10 + 15 + 18) between two fills and get_copy_distance (619-633) does not refill. */
static void br_fill(bitrd* b) {
    if (b->len - b->pos >= 8) {
        uint64_t look = 0;
        for (int i = 7; i >= 0; i--) look = (look << 8) | b->p[b->pos + i];
        b->pos += (size_t)((63 - b->nbits) / 8);
        b->buffer |= look << b->nbits;
        b->nbits |= 56;
    } else {
        while (b->pos < b->len && b->nbits < 56) {
            b->buffer |= (uint64_t)b->p[b->pos++] << b->nbits;
            b->nbits += 8;
        }
    }
}
static uint64_t br_peek(const bitrd* b, int n) { return b->buffer & (((uint64_t)1 << n) - 1); }
static int br_consume(bitrd* b, int n) { /* 751-759 */
    if (b->nbits < n) return ORA_ERR_BITSTREAM;
    b->buffer >>= n;
    b->nbits -= n;
    return 0;
}
static int br_read(bitrd* b, int n, uint32_t* out) { /* 762-773 */
    if (b->nbits < n) br_fill(b);
    *out = (uint32_t)br_peek(b, n);
    return br_consume(b, n);
}
#define RD(n, var) do { int e_ = br_read(br, (n), &(var)); if (e_) return e_; } while (0)

/* ---- HuffmanTree, huffman.rs ---------------------------------------------------------
 * The reference builds a 10-bit primary table plus secondary tables; decoded symbols are
 * those of the canonical prefix code read LSB-first, which is all this restatement keeps:
 * a flat table indexed by the next max_length (<= 15) bits. */
typedef struct {
    int single;          /* HuffmanTreeInner::Single */
    uint16_t symbol;     /* the single symbol */
    int max_len;
    uint32_t* tab;       /* entries sym | len << 16, indexed by the next max_len bits */
} huff;

static void huff_free(huff* h) { free(h->tab); h->tab = NULL; }

static uint32_t bitrev(uint32_t v, int n) {
    uint32_t r = 0;
    for (int i = 0; i < n; i++) { r = (r << 1) | (v & 1); v >>= 1; }
    return r;
}

/* build_implicit, huffman.rs:46-177 */
static int huff_build(huff* h, const uint16_t* lengths, int n) {
    memset(h, 0, sizeof(*h));
    int hist[16] = {0};
    int num = 0;
    for (int i = 0; i < n; i++) { hist[lengths[i]]++; if (lengths[i]) num++; }
    if (num == 0) return ORA_ERR_HUFFMAN;
    if (num == 1) {
        for (int i = 0; i < n; i++) if (lengths[i]) { h->single = 1; h->symbol = (uint16_t)i; break; }
        return 0;
    }
    int max_len = 15;
    while (max_len > 1 && hist[max_len] == 0) max_len--;
    uint32_t used = 0;
    for (int i = 1; i < max_len; i++) used = (used << 1) + (uint32_t)hist[i];
    used = (used << 1) + (uint32_t)hist[max_len];
    if (used != (1u << max_len)) return ORA_ERR_HUFFMAN;
    h->max_len = max_len;
    h->tab = (uint32_t*)malloc(sizeof(uint32_t) << max_len);
    if (!h->tab) return ORA_ERR_NOMEM;
    uint32_t code = 0;
    for (int len = 1; len <= max_len; len++) {
        for (int s = 0; s < n; s++) {
            if (lengths[s] != len) continue;
            uint32_t rev = bitrev(code, len);
            for (uint32_t k = rev; k < (1u << max_len); k += 1u << len)
                h->tab[k] = (uint32_t)s | ((uint32_t)len << 16);
            code++;
        }
        code <<= 1;
    }
    return 0;
}
static void huff_single(huff* h, uint16_t sym) { memset(h, 0, sizeof(*h)); h->single = 1; h->symbol = sym; }
/* build_two_node, huffman.rs:183-189 */
static int huff_two(huff* h, uint16_t zero, uint16_t one) {
    memset(h, 0, sizeof(*h));
    h->max_len = 1;
    h->tab = (uint32_t*)malloc(2 * sizeof(uint32_t));
    if (!h->tab) return ORA_ERR_NOMEM;
    h->tab[0] = zero | (1u << 16);
    h->tab[1] = one | (1u << 16);
    return 0;
}
/* read_symbol, huffman.rs:211-233 (caller has filled the bit reader) */
static int huff_read(const huff* h, bitrd* b, uint16_t* sym) {
    if (h->single) { *sym = h->symbol; return 0; }
    uint32_t e = h->tab[b->buffer & ((1u << h->max_len) - 1)];
    *sym = (uint16_t)(e & 0xffff);
    return br_consume(b, (int)(e >> 16));
}

/* ---- decoder -------------------------------------------------------------------------- */
enum { GREEN = 0, RED = 1, BLUE = 2, ALPHA = 3, DIST = 4 };
static const uint16_t kAlphabet[5] = {256 + 24, 256, 256, 256, 40}; /* decoder/mod.rs:27 */
static const uint8_t kCodeLengthOrder[19] = {17, 18, 0, 1, 2, 3, 4, 5, 16, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15}; /* lossless/mod.rs:10-12 */
/* lossless/mod.rs:15-31 */
static const int8_t kDistMap[120][2] = {
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

typedef struct { int kind; int size_bits; uint8_t* data; int table_size; } transform_t;

typedef struct {
    bitrd br;
    transform_t tr[4]; int have[4];
    uint8_t order[4]; int norder;
    int width, height;
} vp8l;

typedef struct {
    int xsize; int bits; uint32_t mask;
    int cache_bits; uint8_t (*cache)[4];
    uint16_t* image; int ngroups;
    huff (*groups)[5];
} huff_info;

static void info_free(huff_info* hi) {
    if (hi->groups) for (int i = 0; i < hi->ngroups; i++) for (int j = 0; j < 5; j++) huff_free(&hi->groups[i][j]);
    free(hi->groups); free(hi->image); free(hi->cache);
    memset(hi, 0, sizeof(*hi));
}

static int subsample_size(int size, int bits) { return (int)(((uint32_t)size + (1u << bits) - 1) >> bits); } /* 30-34 */

static int decode_image_stream(vp8l* d, int xsize, int ysize, int is_argb, uint8_t* data);

/* read_huffman_code_lengths, decoder/mod.rs:385-452 */
static int read_code_lengths(vp8l* d, const uint16_t* cl_lengths, int num_symbols, uint16_t* out) {
    bitrd* br = &d->br;
    huff table;
    int e = huff_build(&table, cl_lengths, 19);
    if (e) return e;
    uint32_t v;
    int max_symbol;
    e = br_read(br, 1, &v); if (e) goto done;
    if (v == 1) {
        uint32_t lb; e = br_read(br, 3, &lb); if (e) goto done;
        int length_nbits = 2 + 2 * (int)lb;
        uint32_t mm2; e = br_read(br, length_nbits, &mm2); if (e) goto done;
        if ((int)mm2 > num_symbols - 2) { e = ORA_ERR_BITSTREAM; goto done; }
        max_symbol = 2 + (int)mm2;
    } else {
        max_symbol = num_symbols;
    }
    memset(out, 0, sizeof(uint16_t) * (size_t)num_symbols);
    int prev = 8, symbol = 0;
    while (symbol < num_symbols) {
        if (max_symbol == 0) break;
        max_symbol--;
        br_fill(br);
        uint16_t code_len;
        e = huff_read(&table, br, &code_len); if (e) goto done;
        if (code_len < 16) {
            out[symbol++] = code_len;
            if (code_len != 0) prev = code_len;
        } else {
            int use_prev = code_len == 16;
            int slot = code_len - 16;
            int extra_bits, repeat_offset;
            if (slot == 0) { extra_bits = 2; repeat_offset = 3; }
            else if (slot == 1) { extra_bits = 3; repeat_offset = 3; }
            else if (slot == 2) { extra_bits = 7; repeat_offset = 11; }
            else { e = ORA_ERR_BITSTREAM; goto done; }
            uint32_t rb; e = br_read(br, extra_bits, &rb); if (e) goto done;
            int repeat = (int)rb + repeat_offset;
            if (symbol + repeat > num_symbols) { e = ORA_ERR_BITSTREAM; goto done; }
            int length = use_prev ? prev : 0;
            while (repeat-- > 0) out[symbol++] = (uint16_t)length;
        }
    }
    e = 0;
done:
    huff_free(&table);
    return e;
}

/* read_huffman_code, decoder/mod.rs:342-382 */
static int read_huffman_code(vp8l* d, int alphabet_size, huff* h) {
    bitrd* br = &d->br;
    uint32_t simple;
    RD(1, simple);
    if (simple == 1) {
        uint32_t ns, f8, zero;
        RD(1, ns);
        int num_symbols = (int)ns + 1;
        RD(1, f8);
        RD(1 + 7 * (int)f8, zero);
        if ((int)zero >= alphabet_size) return ORA_ERR_BITSTREAM;
        if (num_symbols == 1) { huff_single(h, (uint16_t)zero); return 0; }
        uint32_t one;
        RD(8, one);
        if ((int)one >= alphabet_size) return ORA_ERR_BITSTREAM;
        return huff_two(h, (uint16_t)zero, (uint16_t)one);
    }
    uint16_t cl[19] = {0};
    uint32_t ncl;
    RD(4, ncl);
    int num = 4 + (int)ncl;
    for (int i = 0; i < num; i++) { uint32_t v; RD(3, v); cl[kCodeLengthOrder[i]] = (uint16_t)v; }
    uint16_t* lengths = (uint16_t*)malloc(sizeof(uint16_t) * (size_t)alphabet_size);
    if (!lengths) return ORA_ERR_NOMEM;
    int e = read_code_lengths(d, cl, alphabet_size, lengths);
    if (!e) e = huff_build(h, lengths, alphabet_size);
    free(lengths);
    return e;
}

/* read_huffman_codes, decoder/mod.rs:272-340 */
static int read_huffman_codes(vp8l* d, int read_meta, int xsize, int ysize, int cache_bits, huff_info* hi) {
    bitrd* br = &d->br;
    memset(hi, 0, sizeof(*hi));
    uint32_t ngroups = 1;
    int hbits = 0, hx = 1;
    uint32_t v = 0;
    if (read_meta) RD(1, v);
    if (read_meta && v == 1) {
        uint32_t hb; RD(3, hb);
        hbits = (int)hb + 2;
        hx = subsample_size(xsize, hbits);
        int hy = subsample_size(ysize, hbits);
        size_t n = (size_t)hx * hy;
        uint8_t* data = (uint8_t*)calloc(n ? n * 4 : 1, 1);
        if (!data) return ORA_ERR_NOMEM;
        int e = decode_image_stream(d, hx, hy, 0, data);
        if (e) { free(data); return e; }
        hi->image = (uint16_t*)malloc(sizeof(uint16_t) * (n ? n : 1));
        if (!hi->image) { free(data); return ORA_ERR_NOMEM; }
        for (size_t i = 0; i < n; i++) {
            uint16_t code = (uint16_t)((data[4 * i] << 8) | data[4 * i + 1]);
            if ((uint32_t)code >= ngroups) ngroups = (uint32_t)code + 1;
            hi->image[i] = code;
        }
        free(data);
    }
    hi->groups = (huff(*)[5])calloc(ngroups, sizeof(huff[5]));
    if (!hi->groups) return ORA_ERR_NOMEM;
    hi->ngroups = (int)ngroups;
    for (uint32_t g = 0; g < ngroups; g++)
        for (int j = 0; j < 5; j++) {
            int alphabet = kAlphabet[j];
            if (j == 0 && cache_bits) alphabet += 1 << cache_bits;
            int e = read_huffman_code(d, alphabet, &hi->groups[g][j]);
            if (e) return e;
        }
    hi->xsize = hx;
    hi->bits = hbits;
    hi->mask = hbits == 0 ? 0xffffu : (1u << hbits) - 1; /* u16 !0 */
    hi->cache_bits = cache_bits;
    if (cache_bits) {
        hi->cache = (uint8_t(*)[4])calloc((size_t)1 << cache_bits, 4);
        if (!hi->cache) return ORA_ERR_NOMEM;
    }
    return 0;
}

static int huff_index(const huff_info* hi, int x, int y) { /* 619-628 */
    if (hi->bits == 0) return 0;
    return hi->image[(size_t)(y >> hi->bits) * hi->xsize + (x >> hi->bits)];
}
static void cache_insert(huff_info* hi, const uint8_t c[4]) { /* 638-645 */
    uint32_t col = ((uint32_t)c[0] << 16) | ((uint32_t)c[1] << 8) | c[2] | ((uint32_t)c[3] << 24);
    uint32_t idx = (0x1e35a7bdu * col) >> (32 - hi->cache_bits);
    memcpy(hi->cache[idx], c, 4);
}

/* get_copy_distance, decoder/mod.rs:618-632 (line numbers of the fn in the file: 576-590) */
static int get_copy_distance(bitrd* br, int prefix, size_t* out) {
    if (prefix < 4) { *out = (size_t)prefix + 1; return 0; }
    int extra_bits = (prefix - 2) >> 1;
    size_t offset = (size_t)(2 + (prefix & 1)) << extra_bits;
    size_t bits = (size_t)br_peek(br, extra_bits);
    int e = br_consume(br, extra_bits);
    if (e) return e;
    *out = offset + bits + 1;
    return 0;
}
static size_t plane_code_to_distance(int xsize, size_t plane_code) {
    if (plane_code > 120) return plane_code - 120;
    int xo = kDistMap[plane_code - 1][0], yo = kDistMap[plane_code - 1][1];
    int dist = xo + yo * xsize;
    return dist < 1 ? 1 : (size_t)dist;
}

/* decode_image_data, decoder/mod.rs:455-601 */
static int decode_image_data(vp8l* d, int width, int height, huff_info* hi, uint8_t* data) {
    bitrd* br = &d->br;
    size_t num_values = (size_t)width * height;
    huff* tree = hi->groups[huff_index(hi, 0, 0)];
    size_t index = 0, next_block_start = 0;
    int e;
    while (index < num_values) {
        br_fill(br);
        if (index >= next_block_start) {
            size_t x = index % (size_t)width, y = index / (size_t)width;
            size_t xm = x | hi->mask;
            if (xm > (size_t)(width - 1)) xm = (size_t)(width - 1);
            next_block_start = xm + y * (size_t)width + 1;
            tree = hi->groups[huff_index(hi, (int)x, (int)y)];
            if (tree[0].single && tree[1].single && tree[2].single && tree[3].single) {
                uint16_t code = tree[GREEN].symbol;
                if (code < 256) {
                    size_t n = hi->bits == 0 ? num_values : next_block_start - index;
                    uint8_t value[4] = {(uint8_t)tree[RED].symbol, (uint8_t)code, (uint8_t)tree[BLUE].symbol, (uint8_t)tree[ALPHA].symbol};
                    for (size_t i = 0; i < n; i++) memcpy(data + (index + i) * 4, value, 4);
                    if (hi->cache) cache_insert(hi, value);
                    index += n;
                    continue;
                }
            }
        }
        uint16_t code;
        e = huff_read(&tree[GREEN], br, &code); if (e) return e;
        if (code < 256) {
            uint16_t red, blue, alpha;
            e = huff_read(&tree[RED], br, &red); if (e) return e;
            e = huff_read(&tree[BLUE], br, &blue); if (e) return e;
            if (br->nbits < 15) br_fill(br);
            e = huff_read(&tree[ALPHA], br, &alpha); if (e) return e;
            uint8_t px[4] = {(uint8_t)red, (uint8_t)code, (uint8_t)blue, (uint8_t)alpha};
            memcpy(data + index * 4, px, 4);
            if (hi->cache) cache_insert(hi, px);
            index++;
        } else if (code < 256 + 24) {
            size_t length, dist_code;
            e = get_copy_distance(br, code - 256, &length); if (e) return e;
            uint16_t dist_symbol;
            e = huff_read(&tree[DIST], br, &dist_symbol); if (e) return e;
            e = get_copy_distance(br, dist_symbol, &dist_code); if (e) return e;
            size_t dist = plane_code_to_distance(width, dist_code);
            if (index < dist || num_values - index < length) return ORA_ERR_BITSTREAM;
            /* the reference's three copy strategies (537-565) all equal a forward byte copy */
            for (size_t i = 0; i < length * 4; i++) data[index * 4 + i] = data[index * 4 + i - dist * 4];
            if (dist != 1 && hi->cache)
                for (size_t i = 0; i < length; i++) cache_insert(hi, data + (index + i) * 4);
            index += length;
        } else {
            if (!hi->cache) return ORA_ERR_BITSTREAM;
            memcpy(data + index * 4, hi->cache[code - 280], 4);
            index++;
            /* 586-597: when the next code is another cache hit the reference takes it from
             * the primary table at once; the symbols are those of going round this loop. */
        }
    }
    return 0;
}

/* decode_image_stream, decoder/mod.rs:159-174; read_color_cache 604-616 */
static int decode_image_stream(vp8l* d, int xsize, int ysize, int is_argb, uint8_t* data) {
    bitrd* br = &d->br;
    uint32_t v;
    int cache_bits = 0;
    RD(1, v);
    if (v == 1) {
        uint32_t cb; RD(4, cb);
        if (cb < 1 || cb > 11) return ORA_ERR_COLOR_CACHE;
        cache_bits = (int)cb;
    }
    huff_info hi;
    int e = read_huffman_codes(d, is_argb, xsize, ysize, cache_bits, &hi);
    if (!e) e = decode_image_data(d, xsize, ysize, &hi, data);
    info_free(&hi);
    return e;
}

/* read_transforms, decoder/mod.rs:177-261 */
static int read_transforms(vp8l* d, int* out_xsize) {
    bitrd* br = &d->br;
    int xsize = d->width;
    for (;;) {
        uint32_t more; RD(1, more);
        if (more != 1) break;
        uint32_t tt; RD(2, tt);
        if (d->have[tt]) return ORA_ERR_TRANSFORM;
        d->order[d->norder++] = (uint8_t)tt;
        transform_t* t = &d->tr[tt];
        t->kind = (int)tt;
        if (tt == 0 || tt == 1) {
            uint32_t sb; RD(3, sb);
            t->size_bits = (int)sb + 2;
            int bx = subsample_size(xsize, t->size_bits), by = subsample_size(d->height, t->size_bits);
            size_t n = (size_t)bx * by;
            t->data = (uint8_t*)calloc(n ? n * 4 : 1, 1);
            if (!t->data) return ORA_ERR_NOMEM;
            d->have[tt] = 1;
            int e = decode_image_stream(d, bx, by, 0, t->data);
            if (e) return e;
        } else if (tt == 2) {
            d->have[tt] = 1;
        } else {
            uint32_t cs; RD(8, cs);
            int table_size = (int)cs + 1;
            t->table_size = table_size;
            t->data = (uint8_t*)calloc((size_t)table_size * 4, 1);
            if (!t->data) return ORA_ERR_NOMEM;
            d->have[tt] = 1;
            int e = decode_image_stream(d, table_size, 1, 0, t->data);
            if (e) return e;
            int bits = table_size <= 2 ? 3 : table_size <= 4 ? 2 : table_size <= 16 ? 1 : 0;
            xsize = subsample_size(xsize, bits);
            for (int i = 4; i < table_size * 4; i++) t->data[i] = (uint8_t)(t->data[i] + t->data[i - 4]); /* adjust_color_map 264-268 */
        }
    }
    *out_xsize = xsize;
    return 0;
}

/* ---- reverse_transform.rs ---------------------------------------------------------------- */
static uint8_t average2(uint8_t a, uint8_t b) { return (uint8_t)(((unsigned)a + b) / 2); }
static uint8_t clamp_add_sub_full(int a, int b, int c) { int v = a + b - c; return (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v); }
static uint8_t clamp_add_sub_half(int a, int b) { int v = a + (a - b) / 2; return (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v); }

/* apply_predictor_transform_N over bytes [start, end), reverse_transform.rs:75-351.
 * One byte at a time in memory order reproduces every variant, including those the
 * reference writes pixel-wise, because each output byte only reads bytes before it. */
static void predict_range(uint8_t* im, size_t start, size_t end, size_t width, int mode) {
    size_t w4 = width * 4;
    switch (mode) {
    case 0: for (size_t i = start + 3; i < end; i += 4) im[i] = (uint8_t)(im[i] + 0xff); break;
    case 1: for (size_t i = start; i < end; i++) im[i] = (uint8_t)(im[i] + im[i - 4]); break;
    case 2: for (size_t i = start; i < end; i++) im[i] = (uint8_t)(im[i] + im[i - w4]); break;
    case 3: for (size_t i = start; i < end; i++) im[i] = (uint8_t)(im[i] + im[i - w4 + 4]); break;
    case 4: for (size_t i = start; i < end; i++) im[i] = (uint8_t)(im[i] + im[i - w4 - 4]); break;
    case 5: for (size_t i = start; i < end; i++) im[i] = (uint8_t)(im[i] + average2(average2(im[i - 4], im[i - w4 + 4]), im[i - w4])); break;
    case 6: for (size_t i = start; i < end; i++) im[i] = (uint8_t)(im[i] + average2(im[i - 4], im[i - w4 - 4])); break;
    case 7: for (size_t i = start; i < end; i++) im[i] = (uint8_t)(im[i] + average2(im[i - 4], im[i - w4])); break;
    case 8: for (size_t i = start; i < end; i++) im[i] = (uint8_t)(im[i] + average2(im[i - w4 - 4], im[i - w4])); break;
    case 9: for (size_t i = start; i < end; i++) im[i] = (uint8_t)(im[i] + average2(im[i - w4], im[i - w4 + 4])); break;
    case 10: for (size_t i = start; i < end; i++) im[i] = (uint8_t)(im[i] + average2(average2(im[i - 4], im[i - w4 - 4]), average2(im[i - w4], im[i - w4 + 4]))); break;
    case 11:
        for (size_t i = start; i < end; i += 4) {
            int pl = 0, pt = 0;
            for (int c = 0; c < 4; c++) {
                int l = im[i + c - 4], t = im[i + c - w4], tl = im[i + c - w4 - 4];
                int predict = l + t - tl;
                pl += abs(predict - l);
                pt += abs(predict - t);
            }
            for (int c = 0; c < 4; c++)
                im[i + c] = (uint8_t)(im[i + c] + (pl < pt ? im[i + c - 4] : im[i + c - w4]));
        }
        break;
    case 12: for (size_t i = start; i < end; i++) im[i] = (uint8_t)(im[i] + clamp_add_sub_full(im[i - 4], im[i - w4], im[i - w4 - 4])); break;
    case 13: for (size_t i = start; i < end; i++) im[i] = (uint8_t)(im[i] + clamp_add_sub_half((im[i - 4] + im[i - w4]) / 2, im[i - w4 - 4])); break;
    default: break;
    }
}

/* apply_predictor_transform, reverse_transform.rs:24-74 */
static void apply_predictor(uint8_t* im, int width, int height, int size_bits, const uint8_t* pdata) {
    size_t bx = (size_t)subsample_size(width, size_bits);
    size_t w = (size_t)width, h = (size_t)height;
    im[3] = (uint8_t)(im[3] + 255);
    predict_range(im, 4, w * 4, w, 1);
    for (size_t y = 1; y < h; y++)
        for (int i = 0; i < 4; i++) im[y * w * 4 + i] = (uint8_t)(im[y * w * 4 + i] + im[(y - 1) * w * 4 + i]);
    for (size_t y = 1; y < h; y++)
        for (size_t b = 0; b < bx; b++) {
            size_t bi = (y >> size_bits) * bx + b;
            int predictor = pdata[bi * 4 + 1];
            size_t xs = b << size_bits; if (xs < 1) xs = 1;
            size_t xe = (b + 1) << size_bits; if (xe > w) xe = w;
            predict_range(im, (y * w + xs) * 4, (y * w + xe) * 4, w, predictor);
        }
}

/* apply_color_transform, reverse_transform.rs:353-391 */
static uint32_t ct_delta(int8_t t, int8_t c) { return (uint32_t)((int32_t)t * (int32_t)c) >> 5; }
static void apply_color(uint8_t* im, size_t nbytes, int width, int size_bits, const uint8_t* tdata) {
    size_t bx = (size_t)subsample_size(width, size_bits);
    size_t w = (size_t)width;
    size_t rows = nbytes / (w * 4);
    for (size_t y = 0; y < rows; y++)
        for (size_t x = 0; x < w; x++) {
            const uint8_t* t = tdata + ((y >> size_bits) * bx + (x >> size_bits)) * 4;
            uint8_t* px = im + (y * w + x) * 4;
            uint8_t r2b = t[0], g2b = t[1], g2r = t[2];
            uint32_t green = px[1], tr = px[0], tb = px[2];
            tr += ct_delta((int8_t)g2r, (int8_t)green);
            tb += ct_delta((int8_t)g2b, (int8_t)green);
            tb += ct_delta((int8_t)r2b, (int8_t)tr);
            px[0] = (uint8_t)(tr & 0xff);
            px[2] = (uint8_t)(tb & 0xff);
        }
}

/* apply_color_indexing_transform, reverse_transform.rs:400-578 */
static void apply_color_indexing(uint8_t* im, int width, int height, int table_size, const uint8_t* table) {
    size_t w = (size_t)width, h = (size_t)height;
    if (table_size > 16) {
        for (size_t i = 0; i < w * h; i++) {
            int k = im[4 * i + 1];
            if (k < table_size) memcpy(im + 4 * i, table + 4 * k, 4); else memset(im + 4 * i, 0, 4);
        }
        return;
    }
    int wbits = table_size <= 2 ? 3 : table_size <= 4 ? 2 : 1;
    int ppb = 1 << wbits, bpe = 8 / ppb, mask = (1 << bpe) - 1;
    size_t pw = (w + (size_t)ppb - 1) / (size_t)ppb;
    uint8_t* packed = (uint8_t*)malloc(pw ? pw : 1);
    if (!packed || w == 0 || h == 0) { free(packed); return; }
    for (size_t yr = 0; yr < h; yr++) {
        size_t y = h - 1 - yr;
        for (size_t i = 0; i < pw; i++) packed[i] = im[(y * pw + i) * 4 + 1];
        for (size_t x = 0; x < w; x++) {
            int k = (packed[x / (size_t)ppb] >> ((x % (size_t)ppb) * (size_t)bpe)) & mask;
            uint8_t* px = im + (y * w + x) * 4;
            if (k < table_size) memcpy(px, table + 4 * k, 4); else memset(px, 0, 4);
        }
    }
    free(packed);
}

/* LosslessDecoder::decode_frame, decoder/mod.rs:67-152 */
int ora_vp8l_decode(const uint8_t* data, size_t len, uint32_t width, uint32_t height, int implicit, uint8_t* buf) {
    vp8l d;
    memset(&d, 0, sizeof(d));
    d.br.p = data; d.br.len = len;
    bitrd* br = &d.br;
    int e = 0;
    if (implicit) {
        d.width = (int)(uint16_t)width;
        d.height = (int)(uint16_t)height;
    } else {
        uint32_t sig, w, h, au, ver;
        e = br_read(br, 8, &sig); if (e) return e;
        if (sig != 0x2f) return ORA_ERR_LOSSLESS_SIG;
        e = br_read(br, 14, &w); if (e) return e;
        e = br_read(br, 14, &h); if (e) return e;
        d.width = (int)w + 1; d.height = (int)h + 1;
        if ((uint32_t)d.width != width || (uint32_t)d.height != height) return ORA_ERR_INCONSISTENT_SIZE;
        e = br_read(br, 1, &au); if (e) return e;
        e = br_read(br, 3, &ver); if (e) return e;
        if (ver != 0) return ORA_ERR_VERSION;
    }
    int tw = d.width;
    e = read_transforms(&d, &tw);
    if (e) goto done;
    size_t tsize = (size_t)tw * d.height * 4;
    e = decode_image_stream(&d, tw, d.height, 1, buf);
    if (e) goto done;
    {
        size_t image_size = tsize;
        int w = tw;
        for (int k = d.norder - 1; k >= 0; k--) {
            transform_t* t = &d.tr[d.order[k]];
            switch (t->kind) {
            case 0: apply_predictor(buf, w, d.height, t->size_bits, t->data); break;
            case 1: apply_color(buf, image_size, w, t->size_bits, t->data); break;
            case 2:
                for (size_t i = 0; i + 3 < image_size; i += 4) {
                    buf[i] = (uint8_t)(buf[i] + buf[i + 1]);
                    buf[i + 2] = (uint8_t)(buf[i + 2] + buf[i + 1]);
                }
                break;
            default:
                w = d.width;
                image_size = (size_t)w * d.height * 4;
                apply_color_indexing(buf, w, d.height, t->table_size, t->data);
                break;
            }
        }
    }
done:
    for (int i = 0; i < 4; i++) free(d.tr[i].data);
    return e;
}
