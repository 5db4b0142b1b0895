/* harness/vp8synth.c -- synthetic VP8 keyframe writer (test and bench input generator).
 *
 * Writes valid VP8 keyframe payloads from random modes and coefficients with a boolean
 * entropy ENCODER, so that the decoders under test (the CUDA product, the CPU oracle and
 * libwebp) all get the same bitstreams.  It is the mirror image of what the reference's
 * decoder reads: frame header lossy/mod.rs:510-613, segment/LF/quant fields 397-508,
 * partition table 369-395, macroblock header 615-671, tokens 785-963, contexts 1255-1275.
 * The bool encoder follows RFC 6386 section 7 (same algorithm as the reference's
 * lossy/arithmetic_encoder.rs:39-161; its known-answer bytes are checked in
 * tests/test_synth.py).  Not part of the decode path.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    int32_t width, height;
    uint64_t seed;
    int32_t filter_type;   /* 1 = simple, 0 = normal (mod.rs:164) */
    int32_t filter_level;  /* 0..63 */
    int32_t sharpness;     /* 0..7 */
    int32_t segments;      /* 0 off | 1 abs data + map | 2 delta data + map | 3 abs data, no map | 4 map, no data */
    int32_t seg_quant[4];  /* -127..127 */
    int32_t seg_lf[4];     /* -63..63 */
    int32_t seg_probs[3];  /* -1 = keep 255 */
    int32_t lf_adj;        /* 0 off | 1 on, no update | 2 on with deltas */
    int32_t ref_delta[4], mode_delta[4]; /* -63..63 */
    int32_t log2_parts;    /* 0..3 */
    int32_t yac_abs;       /* 0..127 */
    int32_t q_delta[5];    /* ydc, y2dc, y2ac, uvdc, uvac: -15..15 */
    int32_t prob_skip;     /* -1: mb_no_skip_coeff = 0, else the probability byte */
    int32_t p_skip_1024;   /* P(macroblock skipped) * 1024 (only with prob_skip >= 0) */
    int32_t profile;       /* 0 realistic | 1 uniform | 2 extreme levels | 3 all TM | 4 B_PRED LD/VL | 5 zero-run tails */
    int32_t update_probs;  /* 1: send some random token probability updates */
    int32_t version;       /* 0..3, ignored by the decoder (mod.rs:521) */
    int32_t density_1024;  /* scales coefficient density; 1024 = profile default */
} vp8synth_params;

/* ---- rng ---------------------------------------------------------------------------- */
typedef struct { uint64_t s; } rng_t;
static uint64_t rnd64(rng_t* r) { /* splitmix64 */
    uint64_t z = (r->s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static uint32_t rnd(rng_t* r, uint32_t n) { return (uint32_t)((rnd64(r) >> 32) * (uint64_t)n >> 32); }
static int chance(rng_t* r, uint32_t p1024) { return rnd(r, 1024) < p1024; }

/* ---- bool encoder (RFC 6386 7.3) ------------------------------------------------------- */
typedef struct {
    uint8_t* buf; size_t cap, n;
    uint32_t bottom, range; int bit_num; int overflow;
} benc;

static void benc_init(benc* e, uint8_t* buf, size_t cap) {
    e->buf = buf; e->cap = cap; e->n = 0; e->bottom = 0; e->range = 255; e->bit_num = 24; e->overflow = 0;
}
static void benc_carry(benc* e) {
    size_t i = e->n;
    while (i > 0) {
        i--;
        if (e->buf[i] < 255) { e->buf[i]++; return; }
        e->buf[i] = 0;
    }
}
static void benc_push(benc* e, uint8_t b) {
    if (e->n < e->cap) e->buf[e->n++] = b; else e->overflow = 1;
}
static void put(benc* e, int bit, int prob) {
    uint32_t split = 1 + (((e->range - 1) * (uint32_t)prob) >> 8);
    if (bit) { e->bottom += split; e->range -= split; } else e->range = split;
    while (e->range < 128) {
        e->range <<= 1;
        if (e->bottom & 0x80000000u) benc_carry(e);
        e->bottom <<= 1;
        if (--e->bit_num == 0) {
            benc_push(e, (uint8_t)(e->bottom >> 24));
            e->bottom &= 0xffffff;
            e->bit_num = 8;
        }
    }
}
static void put_flag(benc* e, int bit) { put(e, bit, 128); }
static void put_literal(benc* e, int nbits, uint32_t v) {
    for (int b = nbits - 1; b >= 0; b--) put(e, (v >> b) & 1, 128);
}
/* flag, magnitude, sign with 1 = negative (what the decoder expects,
 * arithmetic_decoder.rs:350-368); value 0 is sent as "absent" unless force is set */
static void put_opt_signed(benc* e, int nbits, int v, int force) {
    if (v == 0 && !force) { put_flag(e, 0); return; }
    put_flag(e, 1);
    put_literal(e, nbits, (uint32_t)(v < 0 ? -v : v));
    put_flag(e, v < 0);
}
static size_t benc_flush(benc* e) {
    int c = e->bit_num;
    uint32_t v = e->bottom;
    if (v & (1u << (32 - c))) benc_carry(e);
    v <<= c & 7;
    c = (c >> 3) - 1;
    while (c-- >= 0) v <<= 8;
    for (c = 0; c < 4; c++) { benc_push(e, (uint8_t)(v >> 24)); v <<= 8; }
    return e->n;
}
/* walk a tree (RFC layout: positive = next index, <= 0 = -leaf) from `start` */
static int tree_path(const int8_t* tree, int i, int value, int* bits, int* nodes, int depth) {
    for (int b = 0; b < 2; b++) {
        int t = tree[i + b];
        nodes[depth] = i; bits[depth] = b;
        if (t <= 0) { if (-t == value) return depth + 1; }
        else { int d = tree_path(tree, t, value, bits, nodes, depth + 1); if (d) return d; }
    }
    return 0;
}
static void put_tree(benc* e, const int8_t* tree, const uint8_t* probs, int value, int start) {
    int bits[16], nodes[16];
    int d = tree_path(tree, start, value, bits, nodes, 0);
    for (int k = 0; k < d; k++) put(e, bits[k], probs[nodes[k] >> 1]);
}

/* exported for the known-answer tests of the encoder (arithmetic_encoder.rs:177-240) */
typedef struct { benc e; uint8_t buf[256]; } kat_enc;
void* vp8synth_kat_new(void) { kat_enc* k = (kat_enc*)calloc(1, sizeof(*k)); benc_init(&k->e, k->buf, sizeof(k->buf)); return k; }
void vp8synth_kat_bool(void* k, int bit, int prob) { put(&((kat_enc*)k)->e, bit, prob); }
void vp8synth_kat_literal(void* k, int n, int v) { put_literal(&((kat_enc*)k)->e, n, (uint32_t)v); }
void vp8synth_kat_opt_signed(void* k, int n, int v, int present) { if (!present) put_flag(&((kat_enc*)k)->e, 0); else put_opt_signed(&((kat_enc*)k)->e, n, v, 1); }
void vp8synth_kat_ymode(void* k, int mode);
int vp8synth_kat_finish(void* k, uint8_t* out) { kat_enc* q = (kat_enc*)k; int n = (int)benc_flush(&q->e); memcpy(out, q->buf, (size_t)n); free(q); return n; }

/* ---- VP8 constants (RFC 6386; same numbers as reference lossy/common.rs) ---------------- */
#include "vp8synth_tables.h"

static const int8_t kSegmentTree[6] = {2, 4, -0, -1, -2, -3};
static const int8_t kYModeTree[8] = {-4, 2, 4, 6, -0, -1, -2, -3};
static const int8_t kBModeTree[18] = {-0, 2, -1, 4, -2, 6, 8, 12, -3, 10, -5, -6, -4, 14, -7, 16, -8, -9};
static const int8_t kUVModeTree[6] = {-0, 2, -1, 4, -2, -3};
static const int8_t kTokenTree[22] = {-11, 2, -0, 4, -1, 6, 8, 12, -2, 10, -3, -4, 14, 16, -5, -6, 18, 20, -7, -8, -9, -10};
static const uint8_t kCatProb[6][12] = {
    {159, 0}, {165, 145, 0}, {173, 148, 140, 0}, {176, 155, 140, 135, 0}, {180, 157, 141, 134, 130, 0},
    {254, 254, 243, 230, 196, 177, 153, 140, 133, 130, 129, 0}};
static const int kCatBase[6] = {5, 7, 11, 19, 35, 67};
static const int kCatBits[6] = {1, 2, 3, 4, 5, 11};

void vp8synth_kat_ymode(void* k, int mode) { put_tree(&((kat_enc*)k)->e, kYModeTree, kYModeProb, mode, 0); }

/* ---- frame writer ------------------------------------------------------------------------ */
typedef struct { uint8_t bpred[4]; uint8_t cplx[9]; } ctx_t;

typedef struct {
    const vp8synth_params* p;
    rng_t rng;
    uint8_t token_probs[4][8][3][11];
    benc hdr, part[8];
    int nparts;
} writer;

static int pick_luma_mode(writer* w) {
    switch (w->p->profile) {
    case 1: return (int)rnd(&w->rng, 5);
    case 3: return 3;
    case 4: return 4;
    default: {
        uint32_t r = rnd(&w->rng, 100);
        return r < 70 ? 4 : r < 88 ? 0 : r < 91 ? 1 : r < 96 ? 2 : 3;
    }
    }
}
static int pick_bmode(writer* w) {
    switch (w->p->profile) {
    case 1: return (int)rnd(&w->rng, 10);
    case 3: return 1;
    case 4: { uint32_t r = rnd(&w->rng, 8); return r < 3 ? 4 : r < 6 ? 7 : (int)rnd(&w->rng, 10); }
    default: return chance(&w->rng, 410) ? 0 : 1 + (int)rnd(&w->rng, 9);
    }
}
static int pick_chroma_mode(writer* w) {
    switch (w->p->profile) {
    case 1: return (int)rnd(&w->rng, 4);
    case 3: return 3;
    default: { uint32_t r = rnd(&w->rng, 100); return r < 65 ? 0 : r < 77 ? 1 : r < 95 ? 2 : 3; }
    }
}

/* magnitude of one non-zero level */
static int pick_level(writer* w) {
    int prof = w->p->profile;
    if (prof == 2) {
        uint32_t r = rnd(&w->rng, 8);
        if (r == 0) return 2114;                        /* DCT_CAT6 maximum, common.rs:664-673 */
        if (r == 1) return 67 + (int)rnd(&w->rng, 2048);
        if (r == 2) return 1 + (int)rnd(&w->rng, 66);
    }
    int v = 1;
    while (v < 64 && chance(&w->rng, 455)) v++; /* 1 + geometric, mean about 1.8 */
    return v;
}

/* fill one block's scan-order levels; returns nothing, zero blocks allowed */
static void pick_block(writer* w, int first, int16_t* lv, int* tail_zero_run) {
    memset(lv, 0, 16 * sizeof(int16_t));
    *tail_zero_run = 0;
    uint32_t dens = w->p->density_1024 > 0 ? (uint32_t)w->p->density_1024 : 1024;
    uint32_t pzero = w->p->profile == 2 ? 200 : 460;
    if (chance(&w->rng, pzero)) {
        if (w->p->profile == 5 && chance(&w->rng, 300)) *tail_zero_run = 1; /* all DCT_0, no EOB */
        return;
    }
    for (int i = first; i < 16; i++) {
        uint32_t pnz = i < 4 ? 460 : (i < 13 ? 460 - (uint32_t)(i - 3) * 36 : 100);
        pnz = pnz * dens >> 10;
        if (pnz > 1000) pnz = 1000;
        if (chance(&w->rng, pnz)) {
            int m = pick_level(w);
            lv[i] = (int16_t)(chance(&w->rng, 512) ? -m : m);
        }
    }
    if (w->p->profile == 5 && chance(&w->rng, 300)) *tail_zero_run = 1;
}

/* tokens of one block, mirror of read_coefficients (mod.rs:785-872); returns has_coefficients */
static int put_block(writer* w, benc* e, int plane, int complexity, const int16_t* lv, int tail_zero_run) {
    int first = plane == 0 ? 1 : 0;
    int last = -1;
    for (int i = first; i < 16; i++) if (lv[i]) last = i;
    int end = tail_zero_run ? 15 : last; /* tokens are written for scan positions first..end */
    int skip = 0, has = 0;
    for (int i = first; i < 16; i++) {
        const uint8_t* probs = w->token_probs[plane][kBands[i]][complexity];
        if (i > end) { put_tree(e, kTokenTree, probs, 11, 0); return has; } /* EOB; never follows a DCT_0 */
        int v = lv[i], a = v < 0 ? -v : v;
        has = 1;
        if (a == 0) {
            put_tree(e, kTokenTree, probs, 0, skip ? 2 : 0);
            skip = 1; complexity = 0;
            continue;
        }
        int token, cat = -1;
        if (a <= 4) token = a;
        else {
            cat = a < 7 ? 0 : a < 11 ? 1 : a < 19 ? 2 : a < 35 ? 3 : a < 67 ? 4 : 5;
            token = 5 + cat;
        }
        put_tree(e, kTokenTree, probs, token, skip ? 2 : 0);
        if (cat >= 0) {
            int extra = a - kCatBase[cat];
            for (int b = kCatBits[cat] - 1, t = 0; b >= 0; b--, t++) put(e, (extra >> b) & 1, kCatProb[cat][t]);
        }
        put_flag(e, v < 0);
        skip = 0;
        complexity = a == 1 ? 1 : 2;
    }
    return has;
}

static void put_u24(uint8_t* p, uint32_t v) { p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)(v >> 16); }

/* Upper bound for the payload of a w x h frame (worst case about 1.3 KiB per macroblock
 * for extreme profiles). */
size_t vp8synth_bound(int width, int height) {
    size_t mbs = (size_t)((width + 15) / 16) * (size_t)((height + 15) / 16);
    return 4096 + mbs * 1400;
}

/* Returns the payload size, or 0 when `cap` is too small / the parameters are invalid. */
size_t vp8synth_frame(const vp8synth_params* p, uint8_t* out, size_t cap) {
    if (p->width < 1 || p->width > 16383 || p->height < 1 || p->height > 16383) return 0;
    int mbw = (p->width + 15) / 16, mbh = (p->height + 15) / 16;
    size_t mbs = (size_t)mbw * mbh;
    writer w;
    memset(&w, 0, sizeof(w));
    w.p = p;
    w.rng.s = p->seed * 0x2545F4914F6CDD1Dull + 0xB200;
    memcpy(w.token_probs, kCoeffProb0, sizeof(w.token_probs));
    w.nparts = 1 << (p->log2_parts & 3);

    size_t hdr_cap = 2048 + mbs * 40, part_cap = 64 + (mbs * 1400) / (size_t)w.nparts + (size_t)mbw * 1400;
    uint8_t* hbuf = (uint8_t*)malloc(hdr_cap);
    uint8_t* pbuf[8] = {0};
    ctx_t* top = (ctx_t*)calloc((size_t)mbw, sizeof(ctx_t));
    int ok = hbuf && top;
    for (int i = 0; i < w.nparts; i++) { pbuf[i] = (uint8_t*)malloc(part_cap); ok = ok && pbuf[i]; }
    size_t total = 0;
    if (!ok) goto done;
    benc_init(&w.hdr, hbuf, hdr_cap);
    for (int i = 0; i < w.nparts; i++) benc_init(&w.part[i], pbuf[i], part_cap);
    benc* h = &w.hdr;

    /* frame header, mod.rs:569-611 */
    put_literal(h, 1, 0);                          /* color space */
    put_literal(h, 1, (uint32_t)(p->seed & 1));    /* pixel_type / clamping type: ignored */
    int seg_on = p->segments != 0;
    int seg_map = p->segments == 1 || p->segments == 2 || p->segments == 4;
    int seg_data = p->segments == 1 || p->segments == 2 || p->segments == 3;
    uint8_t seg_probs[3] = {255, 255, 255};
    put_flag(h, seg_on);
    if (seg_on) { /* read_segment_updates, mod.rs:469-508 */
        put_flag(h, seg_map);
        put_flag(h, seg_data);
        if (seg_data) {
            put_flag(h, p->segments != 2); /* 1 = absolute values */
            for (int i = 0; i < 4; i++) put_opt_signed(h, 7, p->seg_quant[i], 0);
            for (int i = 0; i < 4; i++) put_opt_signed(h, 6, p->seg_lf[i], 0);
        }
        if (seg_map)
            for (int i = 0; i < 3; i++) {
                int upd = p->seg_probs[i] >= 0;
                put_flag(h, upd);
                if (upd) { put_literal(h, 8, (uint32_t)p->seg_probs[i]); seg_probs[i] = (uint8_t)p->seg_probs[i]; }
            }
    }
    put_flag(h, p->filter_type);
    put_literal(h, 6, (uint32_t)p->filter_level);
    put_literal(h, 3, (uint32_t)p->sharpness);
    put_flag(h, p->lf_adj != 0);
    if (p->lf_adj) { /* read_loop_filter_adjustments, mod.rs:453-467 */
        put_flag(h, p->lf_adj == 2);
        if (p->lf_adj == 2) {
            for (int i = 0; i < 4; i++) put_opt_signed(h, 6, p->ref_delta[i], 0);
            for (int i = 0; i < 4; i++) put_opt_signed(h, 6, p->mode_delta[i], 0);
        }
    }
    put_literal(h, 2, (uint32_t)(p->log2_parts & 3));
    put_literal(h, 7, (uint32_t)p->yac_abs); /* read_quantization_indices, mod.rs:397-451 */
    for (int i = 0; i < 5; i++) put_opt_signed(h, 4, p->q_delta[i], 0);
    put_literal(h, 1, 0); /* refresh entropy probs (ignored) */
    for (int i = 0; i < 4; i++) /* update_token_probabilities, mod.rs:352-367 */
        for (int j = 0; j < 8; j++)
            for (int k = 0; k < 3; k++)
                for (int t = 0; t < 11; t++) {
                    int upd = p->update_probs && rnd(&w.rng, 40) == 0;
                    put(h, upd, kCoeffUpdateProb[i][j][k][t]);
                    if (upd) {
                        uint32_t np = 1 + rnd(&w.rng, 255);
                        put_literal(h, 8, np);
                        w.token_probs[i][j][k][t] = (uint8_t)np;
                    }
                }
    put_literal(h, 1, p->prob_skip >= 0);
    if (p->prob_skip >= 0) put_literal(h, 8, (uint32_t)p->prob_skip);

    /* macroblocks, raster order: header into the first partition, tokens into partition
     * mby % nparts (mod.rs:1255-1281) */
    for (int mby = 0; mby < mbh; mby++) {
        benc* tp = &w.part[mby % w.nparts];
        ctx_t left;
        memset(&left, 0, sizeof(left));
        for (int mbx = 0; mbx < mbw; mbx++) {
            ctx_t* tc = &top[mbx];
            /* read_macroblock_header, mod.rs:615-671 */
            if (seg_on && seg_map) put_tree(h, kSegmentTree, seg_probs, (int)rnd(&w.rng, 4), 0);
            int skipped = 0;
            if (p->prob_skip >= 0) {
                skipped = chance(&w.rng, (uint32_t)p->p_skip_1024);
                put(h, skipped, p->prob_skip);
            }
            int ymode = pick_luma_mode(&w);
            put_tree(h, kYModeTree, kYModeProb, ymode, 0);
            if (ymode == 4) {
                for (int y = 0; y < 4; y++)
                    for (int x = 0; x < 4; x++) {
                        int bm = pick_bmode(&w);
                        put_tree(h, kBModeTree, kBModeProb[tc->bpred[x]][left.bpred[y]], bm, 0);
                        tc->bpred[x] = (uint8_t)bm;
                        left.bpred[y] = (uint8_t)bm;
                    }
            } else {
                static const uint8_t implied[4] = {0, 2, 3, 1}; /* DC->B_DC V->B_VE H->B_HE TM->B_TM */
                for (int i = 0; i < 4; i++) { tc->bpred[i] = implied[ymode]; left.bpred[i] = implied[ymode]; }
            }
            put_tree(h, kUVModeTree, kUVModeProb, pick_chroma_mode(&w), 0);

            if (skipped) { /* mod.rs:1264-1272 */
                if (ymode != 4) { left.cplx[0] = 0; tc->cplx[0] = 0; }
                for (int i = 1; i < 9; i++) { left.cplx[i] = 0; tc->cplx[i] = 0; }
                continue;
            }
            /* read_residual_data, mod.rs:874-963 */
            int16_t lv[16]; int tz;
            int plane = ymode == 4 ? 3 : 1;
            if (plane == 1) {
                pick_block(&w, 0, lv, &tz);
                int n = put_block(&w, tp, 1, tc->cplx[0] + left.cplx[0], lv, tz);
                left.cplx[0] = (uint8_t)n; tc->cplx[0] = (uint8_t)n;
                plane = 0;
            }
            for (int y = 0; y < 4; y++) {
                int l = left.cplx[y + 1];
                for (int x = 0; x < 4; x++) {
                    pick_block(&w, plane == 0 ? 1 : 0, lv, &tz);
                    int n = put_block(&w, tp, plane, tc->cplx[x + 1] + l, lv, tz);
                    l = n; tc->cplx[x + 1] = (uint8_t)n;
                }
                left.cplx[y + 1] = (uint8_t)l;
            }
            for (int j = 5; j <= 7; j += 2)
                for (int y = 0; y < 2; y++) {
                    int l = left.cplx[y + j];
                    for (int x = 0; x < 2; x++) {
                        pick_block(&w, 0, lv, &tz);
                        int n = put_block(&w, tp, 2, tc->cplx[x + j] + l, lv, tz);
                        l = n; tc->cplx[x + j] = (uint8_t)n;
                    }
                    left.cplx[y + j] = (uint8_t)l;
                }
        }
    }
    {
        size_t hsz = benc_flush(h), psz[8], need = 10 + hsz + 3 * (size_t)(w.nparts - 1);
        int bad = h->overflow || hsz >= (1u << 19);
        for (int i = 0; i < w.nparts; i++) { psz[i] = benc_flush(&w.part[i]); need += psz[i]; bad = bad || w.part[i].overflow || psz[i] >= (1u << 24); }
        if (bad || need > cap) goto done;
        uint8_t* o = out;
        put_u24(o, ((uint32_t)hsz << 5) | (1u << 4) | ((uint32_t)(p->version & 7) << 1)); /* frame tag, mod.rs:511-524 */
        o[3] = 0x9d; o[4] = 0x01; o[5] = 0x2a;
        o[6] = (uint8_t)p->width; o[7] = (uint8_t)(p->width >> 8);
        o[8] = (uint8_t)p->height; o[9] = (uint8_t)(p->height >> 8);
        o += 10;
        memcpy(o, hbuf, hsz); o += hsz;
        for (int i = 0; i + 1 < w.nparts; i++) { put_u24(o, (uint32_t)psz[i]); o += 3; } /* mod.rs:369-385 */
        for (int i = 0; i < w.nparts; i++) { memcpy(o, pbuf[i], psz[i]); o += psz[i]; }
        total = (size_t)(o - out);
    }
done:
    free(hbuf); free(top);
    for (int i = 0; i < 8; i++) free(pbuf[i]);
    return total;
}
