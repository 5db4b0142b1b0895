/* oracle/ora_vp8.c -- TEST INFRASTRUCTURE, not product code.
 *
 * Restatement in C of the reference's lossy VP8 keyframe decoder, in the reference's own
 * order of operations (entropy decode and reconstruction interleaved per macroblock with
 * border arrays, then a second raster pass for the loop filter, then YUV->RGB):
 *   src/lossy/arithmetic_decoder.rs, src/lossy/mod.rs, src/lossy/transform.rs,
 *   src/lossy/prediction.rs, src/lossy/loop_filter.rs, src/lossy/yuv.rs.
 * Parity pinned on the reference's golden images (see ora.h).
 */
#define _POSIX_C_SOURCE 200809L
#include "ora.h"
#include "vp8_tables.h"
#include <stdlib.h>
#include <string.h>
#include <time.h>

/* ------------------------------------------------------------------------------------
 * Boolean decoder.  arithmetic_decoder.rs:39-320.  The reference loads 4 bytes at a time
 * while whole chunks remain and single bytes at the tail; decoded bits depend only on the
 * byte string, so this restatement loads byte-wise.  End-of-data rule (249-276): one zero
 * byte past the end is tolerated, the next load marks the decoder past-EOF for good and
 * every later read returns false/0 (BitResult::err, 29-36).  A decoder that was never
 * initialised is past-EOF from the start (60-76, test at 675-682).
 * ---------------------------------------------------------------------------------- */
struct ora_bool {
    const uint8_t* p;
    size_t len, pos;       /* pos may reach len+1 (the tolerated zero byte) */
    uint64_t value;
    uint32_t range;
    int bit_count;
    int past_eof;
};

static void bool_init(struct ora_bool* b, const uint8_t* p, size_t len) {
    b->p = p; b->len = len; b->pos = 0;
    b->value = 0; b->range = 255; b->bit_count = -8; b->past_eof = 0;
}
static void bool_uninit(struct ora_bool* b) {
    bool_init(b, NULL, 0);
    b->past_eof = 1; /* FINAL_BYTES_REMAINING_EOF, arithmetic_decoder.rs:72 */
}

/* cold_read_bit, arithmetic_decoder.rs:278-320 */
static int bool_read(struct ora_bool* b, int prob) {
    if (b->past_eof) return 0;
    if (b->bit_count < 0) {
        if (b->pos < b->len) {
            b->value = (b->value << 8) | b->p[b->pos++];
            b->bit_count += 8;
        } else if (b->pos == b->len) {
            b->pos++;
            b->value <<= 8;
            b->bit_count += 8;
        } else {
            b->past_eof = 1;
            return 0;
        }
    }
    uint32_t split = 1 + (((b->range - 1) * (uint32_t)prob) >> 8);
    uint64_t bigsplit = (uint64_t)split << b->bit_count;
    int ret;
    if (b->value >= bigsplit) {
        b->range -= split;
        b->value -= bigsplit;
        ret = 1;
    } else {
        b->range = split;
        ret = 0;
    }
    /* shift = leading_zeros(range).saturating_sub(24) */
    int shift = 0;
    while ((b->range << shift) < 128) shift++;
    b->range <<= shift;
    b->bit_count -= shift;
    return ret;
}
static int bool_flag(struct ora_bool* b) { return bool_read(b, 128); }
/* cold_read_literal 336-348 */
static int bool_literal(struct ora_bool* b, int n) {
    int v = 0;
    for (int i = 0; i < n; i++) v = ((v << 1) + bool_flag(b)) & 0xff;
    return v;
}
/* cold_read_optional_signed_value 350-368: flag, magnitude, sign (1 = negative) */
static int bool_opt_signed(struct ora_bool* b, int n) {
    if (!bool_flag(b)) return 0;
    int mag = bool_literal(b, n);
    int sign = bool_flag(b);
    return sign ? -mag : mag;
}
/* cold_read_with_tree 370-389, tree in RFC 6386 layout (common.rs:151-165); `start` is the
 * array index of the first node (node k of the reference = array index 2k). */
static int bool_tree(struct ora_bool* b, const int8_t* tree, const uint8_t* probs, int start) {
    int i = start;
    do {
        i = tree[i + bool_read(b, probs[i >> 1])];
    } while (i > 0);
    return -i;
}

ora_bool* ora_bool_new(const uint8_t* data, size_t len) {
    ora_bool* b = (ora_bool*)malloc(sizeof(*b));
    if (!b) return NULL;
    if (data) bool_init(b, data, len); else bool_uninit(b);
    return b;
}
void ora_bool_free(ora_bool* b) { free(b); }
int ora_bool_read_bool(ora_bool* b, int prob) { return bool_read(b, prob); }
int ora_bool_read_flag(ora_bool* b) { return bool_flag(b); }
int ora_bool_read_literal(ora_bool* b, int n) { return bool_literal(b, n); }
int ora_bool_past_eof(ora_bool* b) { return b->past_eof; }

/* trees, common.rs:151-165, 292, 659-662 */
static const int8_t kSegmentTree[6] = {2, 4, -0, -1, -2, -3};
static const int8_t kYModeTree[8] = {-4, 2, 4, 6, -0, -1, -2, -3};
static const int8_t kBModeTree[18] = {-0, 2, -1, 4, -2, 6, 8, 12, -3, 10, -5, -6, -4, 14, -7, 16, -8, -9};
static const int8_t kUVModeTree[6] = {-0, 2, -1, 4, -2, -3};
static const int8_t kTokenTree[22] = {-11, 2, -0, 4, -1, 6, 8, 12, -2, 10, -3, -4,
                                      14, 16, -5, -6, 18, 20, -7, -8, -9, -10};
/* common.rs:664-673 */
static const uint8_t kCatProb[6][12] = {
    {159, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0},
    {165, 145, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0},
    {173, 148, 140, 0, 0, 0, 0, 0, 0, 0, 0, 0},
    {176, 155, 140, 135, 0, 0, 0, 0, 0, 0, 0, 0},
    {180, 157, 141, 134, 130, 0, 0, 0, 0, 0, 0, 0},
    {254, 254, 243, 230, 196, 177, 153, 140, 133, 130, 129, 0}};
static const uint8_t kCatBase[6] = {5, 7, 11, 19, 35, 67};

enum { PLANE_YCOEFF1 = 0, PLANE_Y2 = 1, PLANE_CHROMA = 2, PLANE_YCOEFF0 = 3 }; /* common.rs:9-18 */
enum { DCT_EOB = 11, DCT_0 = 0, DCT_CAT1 = 5 };
enum { LUMA_B = 4 };

/* ------------------------------------------------------------------------------------
 * transform.rs
 * ---------------------------------------------------------------------------------- */
void ora_idct4x4(int32_t* block) { /* transform.rs:7-51 */
    const int64_t C1 = 20091, C2 = 35468;
    for (int i = 0; i < 4; i++) {
        int64_t a1 = (int64_t)block[i] + block[8 + i];
        int64_t b1 = (int64_t)block[i] - block[8 + i];
        int64_t t1 = ((int64_t)block[4 + i] * C2) >> 16;
        int64_t t2 = (int64_t)block[12 + i] + (((int64_t)block[12 + i] * C1) >> 16);
        int64_t c1 = t1 - t2;
        t1 = (int64_t)block[4 + i] + (((int64_t)block[4 + i] * C1) >> 16);
        t2 = ((int64_t)block[12 + i] * C2) >> 16;
        int64_t d1 = t1 + t2;
        block[i] = (int32_t)(a1 + d1);
        block[4 + i] = (int32_t)(b1 + c1);
        block[12 + i] = (int32_t)(a1 - d1);
        block[8 + i] = (int32_t)(b1 - c1);
    }
    for (int i = 0; i < 4; i++) {
        int64_t a1 = (int64_t)block[4 * i] + block[4 * i + 2];
        int64_t b1 = (int64_t)block[4 * i] - block[4 * i + 2];
        int64_t t1 = ((int64_t)block[4 * i + 1] * C2) >> 16;
        int64_t t2 = (int64_t)block[4 * i + 3] + (((int64_t)block[4 * i + 3] * C1) >> 16);
        int64_t c1 = t1 - t2;
        t1 = (int64_t)block[4 * i + 1] + (((int64_t)block[4 * i + 1] * C1) >> 16);
        t2 = ((int64_t)block[4 * i + 3] * C2) >> 16;
        int64_t d1 = t1 + t2;
        block[4 * i] = (int32_t)((a1 + d1 + 4) >> 3);
        block[4 * i + 3] = (int32_t)((a1 - d1 + 4) >> 3);
        block[4 * i + 1] = (int32_t)((b1 + c1 + 4) >> 3);
        block[4 * i + 2] = (int32_t)((b1 - c1 + 4) >> 3);
    }
}

/* i32 arithmetic wraps in the reference's release build; go through uint32_t so the C
 * is defined for the same inputs. */
static int32_t wadd(int32_t a, int32_t b) { return (int32_t)((uint32_t)a + (uint32_t)b); }
static int32_t wsub(int32_t a, int32_t b) { return (int32_t)((uint32_t)a - (uint32_t)b); }

void ora_iwht4x4(int32_t* block) { /* transform.rs:54-86 */
    for (int i = 0; i < 4; i++) {
        int32_t a1 = wadd(block[i], block[12 + i]);
        int32_t b1 = wadd(block[4 + i], block[8 + i]);
        int32_t c1 = wsub(block[4 + i], block[8 + i]);
        int32_t d1 = wsub(block[i], block[12 + i]);
        block[i] = wadd(a1, b1);
        block[4 + i] = wadd(c1, d1);
        block[8 + i] = wsub(a1, b1);
        block[12 + i] = wsub(d1, c1);
    }
    for (int i = 0; i < 4; i++) {
        int32_t* r = block + 4 * i;
        int32_t a1 = wadd(r[0], r[3]);
        int32_t b1 = wadd(r[1], r[2]);
        int32_t c1 = wsub(r[1], r[2]);
        int32_t d1 = wsub(r[0], r[3]);
        int32_t a2 = wadd(a1, b1), b2 = wadd(c1, d1), c2 = wsub(a1, b1), d2 = wsub(d1, c1);
        r[0] = wadd(a2, 3) >> 3;
        r[1] = wadd(b2, 3) >> 3;
        r[2] = wadd(c2, 3) >> 3;
        r[3] = wadd(d2, 3) >> 3;
    }
}

/* ------------------------------------------------------------------------------------
 * prediction.rs
 * ---------------------------------------------------------------------------------- */
#define LUMA_STRIDE 21  /* prediction.rs:10 */
#define LUMA_WS (21 * 17)
#define CHROMA_STRIDE 9 /* prediction.rs:77 */
#define CHROMA_WS 81

static uint8_t clamp255(int v) { return (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v); }

/* create_border_luma, prediction.rs:13-74 */
static void create_border_luma(uint8_t* ws, int mbx, int mby, int mbw, const uint8_t* top,
                               const uint8_t* left) {
    const int stride = LUMA_STRIDE;
    memset(ws, 0, LUMA_WS);
    uint8_t* above = ws + 1;
    if (mby == 0) {
        for (int i = 0; i < stride - 1; i++) above[i] = 127;
    } else {
        for (int i = 0; i < 16; i++) above[i] = top[mbx * 16 + i];
        if (mbx == mbw - 1) {
            for (int i = 16; i < 20; i++) above[i] = top[mbx * 16 + 15];
        } else {
            for (int i = 16; i < 20; i++) above[i] = top[mbx * 16 + i];
        }
    }
    for (int i = 17; i < stride; i++) {
        ws[4 * stride + i] = ws[i];
        ws[8 * stride + i] = ws[i];
        ws[12 * stride + i] = ws[i];
    }
    if (mbx == 0) {
        for (int i = 0; i < 16; i++) ws[(i + 1) * stride] = 129;
    } else {
        for (int i = 0; i < 16; i++) ws[(i + 1) * stride] = left[1 + i];
    }
    ws[0] = mby == 0 ? 127 : mbx == 0 ? 129 : left[0];
}

/* create_border_chroma, prediction.rs:80-123 */
static void create_border_chroma(uint8_t* ws, int mbx, int mby, const uint8_t* top,
                                 const uint8_t* left) {
    const int stride = CHROMA_STRIDE;
    memset(ws, 0, CHROMA_WS);
    if (mby == 0) {
        for (int i = 1; i < stride; i++) ws[i] = 127;
    } else {
        for (int i = 0; i < 8; i++) ws[1 + i] = top[mbx * 8 + i];
    }
    if (mbx == 0) {
        for (int y = 0; y < 8; y++) ws[(y + 1) * stride] = 129;
    } else {
        for (int y = 0; y < 8; y++) ws[(y + 1) * stride] = left[1 + y];
    }
    ws[0] = mby == 0 ? 127 : mbx == 0 ? 129 : left[0];
}

/* add_residue, prediction.rs:133-147 */
void ora_add_residue(uint8_t* ws, const int32_t* rb, int y0, int x0, int stride) {
    int pos = y0 * stride + x0;
    for (int r = 0; r < 4; r++) {
        for (int c = 0; c < 4; c++) {
            int64_t s = (int64_t)rb[4 * r + c] + ws[pos + c];
            ws[pos + c] = (uint8_t)(s < 0 ? 0 : s > 255 ? 255 : s);
        }
        pos += stride;
    }
}

static uint8_t avg3(uint8_t l, uint8_t t, uint8_t r) { return (uint8_t)((l + 2 * t + r + 2) >> 2); }
static uint8_t avg2(uint8_t t, uint8_t r) { return (uint8_t)((t + r + 1) >> 1); }

/* predict_vpred, prediction.rs:185-195 (the chunk tail beyond `size` columns is scratch:
 * the reference writes above_slice over the whole chunk[1..]; for the 16x16 / 8x8 calls
 * x0 = 1 so that is exactly the block plus, for luma, the 4 top-right columns copied down
 * into cols 17..20 of every row.  Rows 4, 8, 12 hold the same values there already and the
 * other rows' cols 17..20 are never read, so the effect on results is nil -- reproduced
 * anyway.) */
static void predict_vpred(uint8_t* a, int size, int x0, int y0, int stride) {
    for (int y = 0; y < size; y++) {
        uint8_t* row = a + (y0 + y) * stride;
        for (int i = 0; 1 + i < stride; i++) row[1 + i] = a[(y0 - 1) * stride + x0 + i];
    }
}
/* predict_hpred, prediction.rs:197-203 */
static void predict_hpred(uint8_t* a, int size, int x0, int y0, int stride) {
    for (int y = 0; y < size; y++) {
        uint8_t* row = a + (y0 + y) * stride;
        uint8_t left = row[x0 - 1];
        for (int i = x0; i < stride; i++) row[i] = left;
    }
}
/* predict_dcpred, prediction.rs:205-234 */
static void predict_dcpred(uint8_t* a, int size, int stride, int above, int left) {
    uint32_t sum = 0;
    int shf = size == 8 ? 2 : 3;
    if (left) {
        for (int y = 0; y < size; y++) sum += a[(y + 1) * stride];
        shf++;
    }
    if (above) {
        for (int x = 1; x <= size; x++) sum += a[x];
        shf++;
    }
    uint32_t dc = (!left && !above) ? 128 : (sum + (1u << (shf - 1))) >> shf;
    for (int y = 0; y < size; y++)
        for (int x = 0; x < size; x++) a[1 + stride * (y + 1) + x] = (uint8_t)dc;
}
/* predict_tmpred, prediction.rs:238-269 */
static void predict_tmpred(uint8_t* a, int size, int x0, int y0, int stride) {
    int p = a[(y0 - 1) * stride + x0 - 1];
    for (int y = 0; y < size; y++) {
        int lmp = a[(y0 + y) * stride + x0 - 1] - p;
        for (int x = 0; x < size; x++)
            a[(y0 + y) * stride + x0 + x] = clamp255(lmp + a[(y0 - 1) * stride + x0 + x]);
    }
}
/* predict_bdcpred, prediction.rs:271-288 */
static void predict_bdcpred(uint8_t* a, int x0, int y0, int stride) {
    uint32_t v = 4;
    for (int i = 0; i < 4; i++) v += a[(y0 - 1) * stride + x0 + i];
    for (int i = 0; i < 4; i++) v += a[(y0 + i) * stride + x0 - 1];
    v >>= 3;
    for (int y = 0; y < 4; y++)
        for (int x = 0; x < 4; x++) a[(y0 + y) * stride + x0 + x] = (uint8_t)v;
}

/* predict_b*pred, prediction.rs:339-499; pixel helpers 290-337 */
void ora_predict_bmode(uint8_t* a, int stride, int x0, int y0, int mode) {
#define PX(x, y) a[(y0 + (y)) * stride + x0 + (x)]
    const uint8_t* top = a + (y0 - 1) * stride + x0; /* A0.. */
    uint8_t P = top[-1];
    uint8_t L0 = PX(-1, 0), L1 = PX(-1, 1), L2 = PX(-1, 2), L3 = PX(-1, 3);
    switch (mode) {
    case 0: predict_bdcpred(a, x0, y0, stride); break;
    case 1: predict_tmpred(a, 4, x0, y0, stride); break;
    case 2: { /* VE 339-354 */
        uint8_t v[4] = {avg3(P, top[0], top[1]), avg3(top[0], top[1], top[2]),
                        avg3(top[1], top[2], top[3]), avg3(top[2], top[3], top[4])};
        for (int y = 0; y < 4; y++) for (int x = 0; x < 4; x++) PX(x, y) = v[x];
    } break;
    case 3: { /* HE 356-374 */
        uint8_t v[4] = {avg3(P, L0, L1), avg3(L0, L1, L2), avg3(L1, L2, L3), avg3(L2, L3, L3)};
        for (int y = 0; y < 4; y++) for (int x = 0; x < 4; x++) PX(x, y) = v[y];
    } break;
    case 4: { /* LD 376-395 */
        uint8_t v[7] = {avg3(top[0], top[1], top[2]), avg3(top[1], top[2], top[3]),
                        avg3(top[2], top[3], top[4]), avg3(top[3], top[4], top[5]),
                        avg3(top[4], top[5], top[6]), avg3(top[5], top[6], top[7]),
                        avg3(top[6], top[7], top[7])};
        for (int y = 0; y < 4; y++) for (int x = 0; x < 4; x++) PX(x, y) = v[x + y];
    } break;
    case 5: { /* RD 397-415; edge e0..e8 = L3,L2,L1,L0,P,A0..A3 */
        uint8_t e[9] = {L3, L2, L1, L0, P, top[0], top[1], top[2], top[3]};
        uint8_t v[7];
        for (int i = 0; i < 7; i++) v[i] = avg3(e[i], e[i + 1], e[i + 2]);
        for (int y = 0; y < 4; y++) for (int x = 0; x < 4; x++) PX(x, y) = v[3 - y + x];
    } break;
    case 6: { /* VR 417-436 */
        uint8_t e1 = L2, e2 = L1, e3 = L0, e4 = P, e5 = top[0], e6 = top[1], e7 = top[2], e8 = top[3];
        PX(0, 3) = avg3(e1, e2, e3);
        PX(0, 2) = avg3(e2, e3, e4);
        PX(1, 3) = PX(0, 1) = avg3(e3, e4, e5);
        PX(1, 2) = PX(0, 0) = avg2(e4, e5);
        PX(2, 3) = PX(1, 1) = avg3(e4, e5, e6);
        PX(2, 2) = PX(1, 0) = avg2(e5, e6);
        PX(3, 3) = PX(2, 1) = avg3(e5, e6, e7);
        PX(3, 2) = PX(2, 0) = avg2(e6, e7);
        PX(3, 1) = avg3(e6, e7, e8);
        PX(3, 0) = avg2(e7, e8);
    } break;
    case 7: { /* VL 438-457 */
        const uint8_t* t = top;
        PX(0, 0) = avg2(t[0], t[1]);
        PX(0, 1) = avg3(t[0], t[1], t[2]);
        PX(0, 2) = PX(1, 0) = avg2(t[1], t[2]);
        PX(1, 1) = PX(0, 3) = avg3(t[1], t[2], t[3]);
        PX(1, 2) = PX(2, 0) = avg2(t[2], t[3]);
        PX(1, 3) = PX(2, 1) = avg3(t[2], t[3], t[4]);
        PX(2, 2) = PX(3, 0) = avg2(t[3], t[4]);
        PX(2, 3) = PX(3, 1) = avg3(t[3], t[4], t[5]);
        PX(3, 2) = avg3(t[4], t[5], t[6]);
        PX(3, 3) = avg3(t[5], t[6], t[7]);
    } break;
    case 8: { /* HD 459-478 */
        uint8_t e0 = L3, e1 = L2, e2 = L1, e3 = L0, e4 = P, e5 = top[0], e6 = top[1], e7 = top[2];
        PX(0, 3) = avg2(e0, e1);
        PX(1, 3) = avg3(e0, e1, e2);
        PX(0, 2) = PX(2, 3) = avg2(e1, e2);
        PX(1, 2) = PX(3, 3) = avg3(e1, e2, e3);
        PX(2, 2) = PX(0, 1) = avg2(e2, e3);
        PX(3, 2) = PX(1, 1) = avg3(e2, e3, e4);
        PX(2, 1) = PX(0, 0) = avg2(e3, e4);
        PX(3, 1) = PX(1, 0) = avg3(e3, e4, e5);
        PX(2, 0) = avg3(e4, e5, e6);
        PX(3, 0) = avg3(e5, e6, e7);
    } break;
    case 9: { /* HU 480-499 */
        PX(0, 0) = avg2(L0, L1);
        PX(1, 0) = avg3(L0, L1, L2);
        PX(2, 0) = PX(0, 1) = avg2(L1, L2);
        PX(3, 0) = PX(1, 1) = avg3(L1, L2, L3);
        PX(2, 1) = PX(0, 2) = avg2(L2, L3);
        PX(3, 1) = PX(1, 2) = avg3(L2, L3, L3);
        PX(2, 2) = PX(3, 2) = PX(0, 3) = PX(1, 3) = PX(2, 3) = PX(3, 3) = L3;
    } break;
    }
#undef PX
}

/* predict_4x4, prediction.rs:159-183 */
static void predict_4x4(uint8_t* ws, int stride, const uint8_t* modes, const int32_t* resdata) {
    for (int sby = 0; sby < 4; sby++)
        for (int sbx = 0; sbx < 4; sbx++) {
            int i = sbx + sby * 4, y0 = sby * 4 + 1, x0 = sbx * 4 + 1;
            ora_predict_bmode(ws, stride, x0, y0, modes[i]);
            ora_add_residue(ws, resdata + i * 16, y0, x0, stride);
        }
}

/* ------------------------------------------------------------------------------------
 * loop_filter.rs
 * ---------------------------------------------------------------------------------- */
static int lf_c(int v) { return v < -128 ? -128 : v > 127 ? 127 : v; }
static int u2s(uint8_t v) { return (int)v - 128; }
static uint8_t s2u(int v) { return (uint8_t)(lf_c(v) + 128); }
static int adiff(uint8_t a, uint8_t b) { return a > b ? a - b : b - a; }

/* Both orientations share one body: px[k*step] for k = -4..3 around the edge. The
 * reference has *_vertical (stride) and *_horizontal (contiguous) twins, 28-295. */
static int common_adjust(int use_outer_taps, uint8_t* q0p, ptrdiff_t s) { /* 28-76 */
    int p1 = u2s(q0p[-2 * s]), p0 = u2s(q0p[-s]), q0 = u2s(q0p[0]), q1 = u2s(q0p[s]);
    int outer = use_outer_taps ? lf_c(p1 - q1) : 0;
    int a = lf_c(outer + 3 * (q0 - p0));
    int b = lf_c(a + 3) >> 3;
    a = lf_c(a + 4) >> 3;
    q0p[0] = s2u(q0 - a);
    q0p[-s] = s2u(p0 + b);
    return a;
}
static int simple_threshold(int limit, const uint8_t* q0p, ptrdiff_t s) { /* 79-95 */
    return adiff(q0p[-s], q0p[0]) * 2 + adiff(q0p[-2 * s], q0p[s]) / 2 <= limit;
}
static int should_filter(int interior, int edge, const uint8_t* q0p, ptrdiff_t s) { /* 97-128 */
    return simple_threshold(edge, q0p, s) && adiff(q0p[-4 * s], q0p[-3 * s]) <= interior &&
           adiff(q0p[-3 * s], q0p[-2 * s]) <= interior && adiff(q0p[-2 * s], q0p[-s]) <= interior &&
           adiff(q0p[3 * s], q0p[2 * s]) <= interior && adiff(q0p[2 * s], q0p[s]) <= interior &&
           adiff(q0p[s], q0p[0]) <= interior;
}
static int hev(int thr, const uint8_t* q0p, ptrdiff_t s) { /* 130-139 */
    return adiff(q0p[-2 * s], q0p[-s]) > thr || adiff(q0p[s], q0p[0]) > thr;
}
static void simple_segment(int edge, uint8_t* q0p, ptrdiff_t s) { /* 144-162 */
    if (simple_threshold(edge, q0p, s)) common_adjust(1, q0p, s);
}
static void subblock_filter(int thr, int interior, int edge, uint8_t* q0p, ptrdiff_t s) { /* 167-206 */
    if (should_filter(interior, edge, q0p, s)) {
        int hv = hev(thr, q0p, s);
        int a = (common_adjust(hv, q0p, s) + 1) >> 1;
        if (!hv) {
            q0p[s] = s2u(u2s(q0p[s]) - a);
            q0p[-2 * s] = s2u(u2s(q0p[-2 * s]) + a);
        }
    }
}
static void macroblock_filter(int thr, int interior, int edge, uint8_t* q0p, ptrdiff_t s) { /* 212-295 */
    if (should_filter(interior, edge, q0p, s)) {
        if (!hev(thr, q0p, s)) {
            int p2 = u2s(q0p[-3 * s]), p1 = u2s(q0p[-2 * s]), p0 = u2s(q0p[-s]);
            int q0 = u2s(q0p[0]), q1 = u2s(q0p[s]), q2 = u2s(q0p[2 * s]);
            int w = lf_c(lf_c(p1 - q1) + 3 * (q0 - p0));
            int a = lf_c((27 * w + 63) >> 7);
            q0p[0] = s2u(q0 - a);
            q0p[-s] = s2u(p0 + a);
            a = lf_c((18 * w + 63) >> 7);
            q0p[s] = s2u(q1 - a);
            q0p[-2 * s] = s2u(p1 + a);
            a = lf_c((9 * w + 63) >> 7);
            q0p[2 * s] = s2u(q2 - a);
            q0p[-3 * s] = s2u(p2 + a);
        } else {
            common_adjust(1, q0p, s);
        }
    }
}
void ora_lf_simple_h(int e, uint8_t* px) { simple_segment(e, px + 4, 1); }
void ora_lf_simple_v(int e, uint8_t* px, size_t point, size_t stride) { simple_segment(e, px + point, (ptrdiff_t)stride); }
void ora_lf_subblock_h(int h, int i, int e, uint8_t* px) { subblock_filter(h, i, e, px + 4, 1); }
void ora_lf_subblock_v(int h, int i, int e, uint8_t* px, size_t point, size_t stride) { subblock_filter(h, i, e, px + point, (ptrdiff_t)stride); }
void ora_lf_macroblock_h(int h, int i, int e, uint8_t* px) { macroblock_filter(h, i, e, px + 4, 1); }
void ora_lf_macroblock_v(int h, int i, int e, uint8_t* px, size_t point, size_t stride) { macroblock_filter(h, i, e, px + point, (ptrdiff_t)stride); }

/* ------------------------------------------------------------------------------------
 * Vp8Decoder, lossy/mod.rs:252-1297
 * ---------------------------------------------------------------------------------- */
typedef struct { uint8_t bpred[4]; uint8_t complexity[9]; } prev_mb; /* mod.rs:124-129 */

typedef struct {
    const uint8_t* r; size_t rlen, rpos;   /* the reader */
    struct ora_bool b;
    struct ora_bool partitions[8];
    int num_partitions;
    ora_frame* f;
    uint8_t segment_tree_probs[3];
    uint8_t token_probs[4][8][3][11];
    prev_mb* top;
    prev_mb left;
    uint8_t *top_border_y, *top_border_u, *top_border_v;
    uint8_t left_border_y[17], left_border_u[9], left_border_v[9];
    ora_mb* macroblocks;
    int keep;
} vp8dec;

static int rd_exact(vp8dec* d, const uint8_t** p, size_t n) {
    if (d->rlen - d->rpos < n) return 0;
    *p = d->r + d->rpos;
    d->rpos += n;
    return 1;
}
#define CHECK(dec) do { if ((dec)->past_eof) return ORA_ERR_BITSTREAM; } while (0)

/* read_segment_updates, mod.rs:469-508 */
static int read_segment_updates(vp8dec* d) {
    ora_frame* f = d->f;
    f->segments_update_map = bool_flag(&d->b);
    int update_data = bool_flag(&d->b);
    if (update_data) {
        int mode = bool_flag(&d->b);
        for (int i = 0; i < 4; i++) f->segment[i].delta_values = !mode;
        for (int i = 0; i < 4; i++) f->segment[i].quantizer_level = (int8_t)bool_opt_signed(&d->b, 7);
        for (int i = 0; i < 4; i++) f->segment[i].loopfilter_level = (int8_t)bool_opt_signed(&d->b, 6);
    }
    if (f->segments_update_map) {
        for (int i = 0; i < 3; i++) {
            int update = bool_flag(&d->b);
            d->segment_tree_probs[i] = update ? (uint8_t)bool_literal(&d->b, 8) : 255;
        }
    }
    CHECK(&d->b);
    return 0;
}

/* read_loop_filter_adjustments, mod.rs:453-467 */
static int read_lf_adjustments(vp8dec* d) {
    if (bool_flag(&d->b)) {
        for (int i = 0; i < 4; i++) d->f->ref_delta[i] = bool_opt_signed(&d->b, 6);
        for (int i = 0; i < 4; i++) d->f->mode_delta[i] = bool_opt_signed(&d->b, 6);
    }
    CHECK(&d->b);
    return 0;
}

/* init_partitions, mod.rs:369-395 */
static int init_partitions(vp8dec* d, int n) {
    if (n > 1) {
        const uint8_t* sizes;
        if (!rd_exact(d, &sizes, (size_t)(3 * n - 3))) return ORA_ERR_IO;
        for (int i = 0; i < n - 1; i++) {
            size_t size = sizes[3 * i] | (sizes[3 * i + 1] << 8) | ((size_t)sizes[3 * i + 2] << 16);
            const uint8_t* p;
            if (!rd_exact(d, &p, size)) return ORA_ERR_IO;
            bool_init(&d->partitions[i], p, size);
        }
    }
    bool_init(&d->partitions[n - 1], d->r + d->rpos, d->rlen - d->rpos); /* read_to_end */
    d->rpos = d->rlen;
    return 0;
}

static int16_t dc_quant(int idx) { return kDcQ[idx < 0 ? 0 : idx > 127 ? 127 : idx]; }
static int16_t ac_quant(int idx) { return kAcQ[idx < 0 ? 0 : idx > 127 ? 127 : idx]; }

/* read_quantization_indices, mod.rs:397-451 */
static int read_quant_indices(vp8dec* d) {
    ora_frame* f = d->f;
    int yac_abs = bool_literal(&d->b, 7);
    int ydc_delta = bool_opt_signed(&d->b, 4);
    int y2dc_delta = bool_opt_signed(&d->b, 4);
    int y2ac_delta = bool_opt_signed(&d->b, 4);
    int uvdc_delta = bool_opt_signed(&d->b, 4);
    int uvac_delta = bool_opt_signed(&d->b, 4);
    int n = f->segments_enabled ? 4 : 1;
    for (int i = 0; i < n; i++) {
        ora_segment* s = &f->segment[i];
        int base;
        if (f->segments_enabled)
            base = s->delta_values ? (int)s->quantizer_level + yac_abs : (int)s->quantizer_level;
        else
            base = yac_abs;
        s->ydc = dc_quant(base + ydc_delta);
        s->yac = ac_quant(base);
        s->y2dc = (int16_t)(dc_quant(base + y2dc_delta) * 2);
        s->y2ac = (int16_t)((int32_t)ac_quant(base + y2ac_delta) * 155 / 100);
        s->uvdc = dc_quant(base + uvdc_delta);
        s->uvac = ac_quant(base + uvac_delta);
        if (s->y2ac < 8) s->y2ac = 8;
        if (s->uvdc > 132) s->uvdc = 132;
    }
    CHECK(&d->b);
    return 0;
}

/* update_token_probabilities, mod.rs:352-367 */
static int update_token_probs(vp8dec* d) {
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 8; j++)
            for (int k = 0; k < 3; k++)
                for (int t = 0; t < 11; t++)
                    if (bool_read(&d->b, kCoeffUpdateProb[i][j][k][t]))
                        d->token_probs[i][j][k][t] = (uint8_t)bool_literal(&d->b, 8);
    CHECK(&d->b);
    return 0;
}

/* read_frame_header, mod.rs:510-613 */
static int read_frame_header(vp8dec* d) {
    ora_frame* f = d->f;
    const uint8_t* p;
    if (!rd_exact(d, &p, 3)) return ORA_ERR_IO;
    uint32_t tag = p[0] | (p[1] << 8) | ((uint32_t)p[2] << 16);
    if (tag & 1) return ORA_ERR_UNSUPPORTED;
    f->version = (tag >> 1) & 7;
    f->for_display = (tag >> 4) & 1;
    size_t first_partition_size = tag >> 5;
    if (!rd_exact(d, &p, 3)) return ORA_ERR_IO;
    if (p[0] != 0x9d || p[1] != 0x01 || p[2] != 0x2a) return ORA_ERR_VP8_MAGIC;
    if (!rd_exact(d, &p, 4)) return ORA_ERR_IO;
    int w = p[0] | (p[1] << 8), h = p[2] | (p[3] << 8);
    f->width = w & 0x3FFF;
    f->height = h & 0x3FFF;
    f->mbw = (f->width + 15) / 16;
    f->mbh = (f->height + 15) / 16;
    int mbw = f->mbw, mbh = f->mbh;

    d->top = (prev_mb*)calloc((size_t)(mbw ? mbw : 1), sizeof(prev_mb));
    memset(&d->left, 0, sizeof(d->left));
    size_t ysz = (size_t)mbw * 16 * mbh * 16, csz = (size_t)mbw * 8 * mbh * 8;
    f->ybuf = (uint8_t*)calloc(ysz ? ysz : 1, 1);
    f->ubuf = (uint8_t*)calloc(csz ? csz : 1, 1);
    f->vbuf = (uint8_t*)calloc(csz ? csz : 1, 1);
    d->top_border_y = (uint8_t*)malloc((size_t)f->width + 4 + 16);
    d->top_border_u = (uint8_t*)malloc((size_t)8 * mbw + 1);
    d->top_border_v = (uint8_t*)malloc((size_t)8 * mbw + 1);
    if (!d->top || !f->ybuf || !f->ubuf || !f->vbuf || !d->top_border_y || !d->top_border_u || !d->top_border_v)
        return ORA_ERR_NOMEM;
    memset(d->top_border_y, 127, (size_t)f->width + 4 + 16);
    memset(d->top_border_u, 127, (size_t)8 * mbw + 1);
    memset(d->top_border_v, 127, (size_t)8 * mbw + 1);
    memset(d->left_border_y, 129, 17);
    memset(d->left_border_u, 129, 9);
    memset(d->left_border_v, 129, 9);

    if (!rd_exact(d, &p, first_partition_size)) return ORA_ERR_IO;
    bool_init(&d->b, p, first_partition_size);

    int color_space = bool_literal(&d->b, 1);
    f->pixel_type = bool_literal(&d->b, 1);
    if (color_space != 0) return ORA_ERR_COLOR_SPACE;

    f->segments_enabled = bool_flag(&d->b);
    if (f->segments_enabled) {
        int e = read_segment_updates(d);
        if (e) return e;
    }
    f->filter_type = bool_flag(&d->b);
    f->filter_level = bool_literal(&d->b, 6);
    f->sharpness_level = bool_literal(&d->b, 3);
    f->lf_adj_enabled = bool_flag(&d->b);
    if (f->lf_adj_enabled) {
        int e = read_lf_adjustments(d);
        if (e) return e;
    }
    int num_partitions = 1 << bool_literal(&d->b, 2);
    CHECK(&d->b);
    d->num_partitions = f->num_partitions = num_partitions;
    int e = init_partitions(d, num_partitions);
    if (e) return e;
    e = read_quant_indices(d);
    if (e) return e;
    (void)bool_literal(&d->b, 1); /* refresh entropy probs, ignored (mod.rs:599-600) */
    e = update_token_probs(d);
    if (e) return e;
    int mb_no_skip_coeff = bool_literal(&d->b, 1);
    f->prob_skip_false = mb_no_skip_coeff == 1 ? bool_literal(&d->b, 8) : -1;
    CHECK(&d->b);
    return 0;
}

static int into_intra(int luma_mode) { /* common.rs:75-83 */
    switch (luma_mode) { case 0: return 0; case 1: return 2; case 2: return 3; default: return 1; }
}

/* read_macroblock_header, mod.rs:615-671 */
static int read_mb_header(vp8dec* d, int mbx, ora_mb* mb) {
    ora_frame* f = d->f;
    memset(mb, 0, sizeof(*mb));
    if (f->segments_enabled && f->segments_update_map)
        mb->segmentid = (uint8_t)bool_tree(&d->b, kSegmentTree, d->segment_tree_probs, 0);
    mb->coeffs_skipped = f->prob_skip_false >= 0 ? (uint8_t)bool_read(&d->b, f->prob_skip_false) : 0;
    mb->luma_mode = (uint8_t)bool_tree(&d->b, kYModeTree, kYModeProb, 0);
    if (mb->luma_mode == LUMA_B) {
        for (int y = 0; y < 4; y++)
            for (int x = 0; x < 4; x++) {
                int top = d->top[mbx].bpred[x], left = d->left.bpred[y];
                int bmode = bool_tree(&d->b, kBModeTree, kBModeProb[top][left], 0);
                mb->bpred[x + y * 4] = (uint8_t)bmode;
                d->top[mbx].bpred[x] = (uint8_t)bmode;
                d->left.bpred[y] = (uint8_t)bmode;
            }
    } else {
        int mode = into_intra(mb->luma_mode);
        for (int i = 0; i < 4; i++) {
            mb->bpred[12 + i] = (uint8_t)mode;
            d->left.bpred[i] = (uint8_t)mode;
        }
    }
    mb->chroma_mode = (uint8_t)bool_tree(&d->b, kUVModeTree, kUVModeProb, 0);
    memcpy(d->top[mbx].bpred, mb->bpred + 12, 4);
    CHECK(&d->b);
    return 0;
}

/* read_coefficients, mod.rs:785-872.  `raw` (may be NULL) receives the token values
 * before the multiply at line 866, in the same de-zigzagged slots. */
static int read_coefficients(vp8dec* d, int32_t* block, int16_t* raw, int p, int plane,
                             int complexity, int dcq, int acq, int* has) {
    int first = plane == PLANE_YCOEFF1 ? 1 : 0;
    struct ora_bool* dec = &d->partitions[p];
    int has_coefficients = 0, skip = 0;
    for (int i = first; i < 16; i++) {
        int band = kBands[i];
        const uint8_t* probs = d->token_probs[plane][band][complexity];
        int token = bool_tree(dec, kTokenTree, probs, skip ? 2 : 0);
        int abs_value;
        if (token == DCT_EOB) break;
        if (token == DCT_0) {
            skip = 1;
            has_coefficients = 1;
            complexity = 0;
            continue;
        }
        if (token <= 4) {
            abs_value = token;
        } else {
            const uint8_t* cp = kCatProb[token - DCT_CAT1];
            int extra = 0;
            for (int t = 0; t < 12 && cp[t]; t++) extra = extra + extra + bool_read(dec, cp[t]);
            abs_value = kCatBase[token - DCT_CAT1] + extra;
        }
        skip = 0;
        complexity = abs_value == 0 ? 0 : abs_value == 1 ? 1 : 2;
        if (bool_flag(dec)) abs_value = -abs_value; /* read_sign */
        int zz = kZigzag[i];
        block[zz] = abs_value * (zz > 0 ? acq : dcq);
        if (raw) raw[zz] = (int16_t)abs_value;
        has_coefficients = 1;
    }
    CHECK(dec);
    *has = has_coefficients;
    return 0;
}

/* read_residual_data, mod.rs:874-963 */
static int read_residual_data(vp8dec* d, ora_mb* mb, int mbx, int p, int32_t* blocks, int16_t* raw) {
    ora_frame* f = d->f;
    const ora_segment* seg = &f->segment[mb->segmentid];
    memset(blocks, 0, 384 * sizeof(int32_t));
    int plane = mb->luma_mode == LUMA_B ? PLANE_YCOEFF0 : PLANE_Y2;
    int n, e;
    if (plane == PLANE_Y2) {
        int complexity = d->top[mbx].complexity[0] + d->left.complexity[0];
        int32_t block[16] = {0};
        e = read_coefficients(d, block, raw ? raw + 24 * 16 : NULL, p, plane, complexity, seg->y2dc, seg->y2ac, &n);
        if (e) return e;
        d->left.complexity[0] = (uint8_t)n;
        d->top[mbx].complexity[0] = (uint8_t)n;
        if (n) mb->nflags |= 1u << 24;
        ora_iwht4x4(block);
        for (int k = 0; k < 16; k++) blocks[16 * k] = block[k];
        plane = PLANE_YCOEFF1;
    }
    for (int y = 0; y < 4; y++) {
        int left = d->left.complexity[y + 1];
        for (int x = 0; x < 4; x++) {
            int i = x + y * 4;
            int32_t* block = blocks + i * 16;
            int complexity = d->top[mbx].complexity[x + 1] + left;
            e = read_coefficients(d, block, raw ? raw + i * 16 : NULL, p, plane, complexity, seg->ydc, seg->yac, &n);
            if (e) return e;
            if (n) mb->nflags |= 1u << i;
            if (block[0] != 0 || n) {
                mb->non_zero_dct = 1;
                ora_idct4x4(block);
            }
            left = n;
            d->top[mbx].complexity[x + 1] = (uint8_t)n;
        }
        d->left.complexity[y + 1] = (uint8_t)left;
    }
    plane = PLANE_CHROMA;
    for (int jj = 0; jj < 2; jj++) {
        int j = jj == 0 ? 5 : 7;
        for (int y = 0; y < 2; y++) {
            int left = d->left.complexity[y + j];
            for (int x = 0; x < 2; x++) {
                int i = x + y * 2 + (j == 5 ? 16 : 20);
                int32_t* block = blocks + i * 16;
                int complexity = d->top[mbx].complexity[x + j] + left;
                e = read_coefficients(d, block, raw ? raw + i * 16 : NULL, p, plane, complexity, seg->uvdc, seg->uvac, &n);
                if (e) return e;
                if (n) mb->nflags |= 1u << i;
                if (block[0] != 0 || n) {
                    mb->non_zero_dct = 1;
                    ora_idct4x4(block);
                }
                left = n;
                d->top[mbx].complexity[x + j] = (uint8_t)n;
            }
            d->left.complexity[y + j] = (uint8_t)left;
        }
    }
    return 0;
}

/* intra_predict_luma, mod.rs:673-721 */
static void intra_predict_luma(vp8dec* d, int mbx, int mby, const ora_mb* mb, const int32_t* resdata) {
    const int stride = LUMA_STRIDE;
    int mw = d->f->mbw;
    uint8_t ws[LUMA_WS];
    create_border_luma(ws, mbx, mby, mw, d->top_border_y, d->left_border_y);
    switch (mb->luma_mode) {
    case 1: predict_vpred(ws, 16, 1, 1, stride); break;
    case 2: predict_hpred(ws, 16, 1, 1, stride); break;
    case 3: predict_tmpred(ws, 16, 1, 1, stride); break;
    case 0: predict_dcpred(ws, 16, stride, mby != 0, mbx != 0); break;
    default: predict_4x4(ws, stride, mb->bpred, resdata); break;
    }
    if (mb->luma_mode != LUMA_B)
        for (int y = 0; y < 4; y++)
            for (int x = 0; x < 4; x++)
                ora_add_residue(ws, resdata + (x + y * 4) * 16, 1 + y * 4, 1 + x * 4, stride);
    d->left_border_y[0] = ws[16];
    for (int i = 0; i < 16; i++) d->left_border_y[1 + i] = ws[(i + 1) * stride + 16];
    for (int i = 0; i < 16; i++) d->top_border_y[mbx * 16 + i] = ws[16 * stride + 1 + i];
    for (int y = 0; y < 16; y++)
        memcpy(d->f->ybuf + ((size_t)(mby * 16 + y) * mw * 16 + mbx * 16), ws + (1 + y) * stride + 1, 16);
}

/* set_chroma_border, mod.rs:1301-1322 */
static void set_chroma_border(uint8_t* left, uint8_t* top, const uint8_t* cb, int mbx) {
    const int stride = CHROMA_STRIDE;
    left[0] = cb[8];
    for (int i = 0; i < 8; i++) left[1 + i] = cb[(i + 1) * stride + 8];
    for (int i = 0; i < 8; i++) top[mbx * 8 + i] = cb[8 * stride + 1 + i];
}

/* intra_predict_chroma, mod.rs:723-783 */
static void intra_predict_chroma(vp8dec* d, int mbx, int mby, const ora_mb* mb, const int32_t* resdata) {
    const int stride = CHROMA_STRIDE;
    int mw = d->f->mbw;
    uint8_t uws[CHROMA_WS], vws[CHROMA_WS];
    create_border_chroma(uws, mbx, mby, d->top_border_u, d->left_border_u);
    create_border_chroma(vws, mbx, mby, d->top_border_v, d->left_border_v);
    switch (mb->chroma_mode) {
    case 0:
        predict_dcpred(uws, 8, stride, mby != 0, mbx != 0);
        predict_dcpred(vws, 8, stride, mby != 0, mbx != 0);
        break;
    case 1: predict_vpred(uws, 8, 1, 1, stride); predict_vpred(vws, 8, 1, 1, stride); break;
    case 2: predict_hpred(uws, 8, 1, 1, stride); predict_hpred(vws, 8, 1, 1, stride); break;
    default: predict_tmpred(uws, 8, 1, 1, stride); predict_tmpred(vws, 8, 1, 1, stride); break;
    }
    for (int y = 0; y < 2; y++)
        for (int x = 0; x < 2; x++) {
            int i = x + y * 2;
            ora_add_residue(uws, resdata + 16 * 16 + i * 16, 1 + y * 4, 1 + x * 4, stride);
            ora_add_residue(vws, resdata + 20 * 16 + i * 16, 1 + y * 4, 1 + x * 4, stride);
        }
    set_chroma_border(d->left_border_u, d->top_border_u, uws, mbx);
    set_chroma_border(d->left_border_v, d->top_border_v, vws, mbx);
    for (int y = 0; y < 8; y++) {
        size_t idx = (size_t)(mby * 8 + y) * mw * 8 + mbx * 8;
        memcpy(d->f->ubuf + idx, uws + (1 + y) * stride + 1, 8);
        memcpy(d->f->vbuf + idx, vws + (1 + y) * stride + 1, 8);
    }
}

/* calculate_filter_parameters, mod.rs:1191-1244 */
static void calc_filter_params(const ora_frame* f, const ora_mb* mb, int* level, int* interior, int* hevt) {
    const ora_segment* seg = &f->segment[mb->segmentid];
    int fl = f->filter_level;
    if (fl == 0) { *level = *interior = *hevt = 0; return; }
    if (f->segments_enabled) {
        if (seg->delta_values) fl += seg->loopfilter_level; else fl = seg->loopfilter_level;
    }
    fl = fl < 0 ? 0 : fl > 63 ? 63 : fl;
    if (f->lf_adj_enabled) {
        fl += f->ref_delta[0];
        if (mb->luma_mode == LUMA_B) fl += f->mode_delta[0];
    }
    fl = fl < 0 ? 0 : fl > 63 ? 63 : fl;
    int il = fl;
    if (f->sharpness_level > 0) {
        il >>= f->sharpness_level > 4 ? 2 : 1;
        if (il > 9 - f->sharpness_level) il = 9 - f->sharpness_level;
    }
    if (il == 0) il = 1;
    *level = fl; *interior = il;
    *hevt = fl >= 40 ? 2 : fl >= 15 ? 1 : 0;
}

/* loop_filter, mod.rs:966-1188 */
static void loop_filter_mb(ora_frame* f, int mbx, int mby, const ora_mb* mb) {
    size_t luma_w = (size_t)f->mbw * 16, chroma_w = (size_t)f->mbw * 8;
    int level, interior, hevt;
    calc_filter_params(f, mb, &level, &interior, &hevt);
    if (level <= 0) return;
    int mbedge = (level + 2) * 2 + interior;
    int subedge = level * 2 + interior;
    int do_sub = mb->luma_mode == LUMA_B || (!mb->coeffs_skipped && mb->non_zero_dct);
    uint8_t *Y = f->ybuf, *U = f->ubuf, *V = f->vbuf;
    if (mbx > 0) { /* left MB edge, 981-1024 */
        if (f->filter_type) {
            for (int y = 0; y < 16; y++) simple_segment(mbedge, Y + (mby * 16 + y) * luma_w + mbx * 16, 1);
        } else {
            for (int y = 0; y < 16; y++) macroblock_filter(hevt, interior, mbedge, Y + (mby * 16 + y) * luma_w + mbx * 16, 1);
            for (int y = 0; y < 8; y++) {
                macroblock_filter(hevt, interior, mbedge, U + (mby * 8 + y) * chroma_w + mbx * 8, 1);
                macroblock_filter(hevt, interior, mbedge, V + (mby * 8 + y) * chroma_w + mbx * 8, 1);
            }
        }
    }
    if (do_sub) { /* inner vertical edges, 1027-1074 */
        if (f->filter_type) {
            for (int x = 4; x < 15; x += 4)
                for (int y = 0; y < 16; y++) simple_segment(subedge, Y + (mby * 16 + y) * luma_w + mbx * 16 + x, 1);
        } else {
            for (int x = 4; x < 13; x += 4)
                for (int y = 0; y < 16; y++) subblock_filter(hevt, interior, subedge, Y + (mby * 16 + y) * luma_w + mbx * 16 + x, 1);
            for (int y = 0; y < 8; y++) {
                subblock_filter(hevt, interior, subedge, U + (mby * 8 + y) * chroma_w + mbx * 8 + 4, 1);
                subblock_filter(hevt, interior, subedge, V + (mby * 8 + y) * chroma_w + mbx * 8 + 4, 1);
            }
        }
    }
    if (mby > 0) { /* top MB edge, 1077-1128 */
        if (f->filter_type) {
            for (int x = 0; x < 16; x++) simple_segment(mbedge, Y + (mby * 16) * luma_w + mbx * 16 + x, (ptrdiff_t)luma_w);
        } else {
            for (int x = 0; x < 16; x++) macroblock_filter(hevt, interior, mbedge, Y + (mby * 16) * luma_w + mbx * 16 + x, (ptrdiff_t)luma_w);
            for (int x = 0; x < 8; x++) {
                macroblock_filter(hevt, interior, mbedge, U + (mby * 8) * chroma_w + mbx * 8 + x, (ptrdiff_t)chroma_w);
                macroblock_filter(hevt, interior, mbedge, V + (mby * 8) * chroma_w + mbx * 8 + x, (ptrdiff_t)chroma_w);
            }
        }
    }
    if (do_sub) { /* inner horizontal edges, 1131-1186 */
        if (f->filter_type) {
            for (int y = 4; y < 15; y += 4)
                for (int x = 0; x < 16; x++) simple_segment(subedge, Y + (mby * 16 + y) * luma_w + mbx * 16 + x, (ptrdiff_t)luma_w);
        } else {
            for (int y = 4; y < 13; y += 4)
                for (int x = 0; x < 16; x++) subblock_filter(hevt, interior, subedge, Y + (mby * 16 + y) * luma_w + mbx * 16 + x, (ptrdiff_t)luma_w);
            for (int x = 0; x < 8; x++) {
                subblock_filter(hevt, interior, subedge, U + (mby * 8 + 4) * chroma_w + mbx * 8 + x, (ptrdiff_t)chroma_w);
                subblock_filter(hevt, interior, subedge, V + (mby * 8 + 4) * chroma_w + mbx * 8 + x, (ptrdiff_t)chroma_w);
            }
        }
    }
}

static void dec_cleanup(vp8dec* d) {
    free(d->top); free(d->top_border_y); free(d->top_border_u); free(d->top_border_v);
    free(d->macroblocks);
}

void ora_frame_free(ora_frame* f) {
    if (!f) return;
    free(f->ybuf); free(f->ubuf); free(f->vbuf);
    free(f->ybuf0); free(f->ubuf0); free(f->vbuf0);
    free(f->mbs); free(f->levels); free(f->residue);
    memset(f, 0, sizeof(*f));
}

/* decode_frame_, mod.rs:1252-1297 */
/* Optional per-stage wall-clock split of the CPU baseline (bench.py): when ora_stage_timing is set, every decode
 * adds its nanoseconds to ora_stage_ns[0] boolean-entropy decode incl. dequantisation and inverse transforms (the
 * reference does those inside read_residual_data), [1] intra prediction, [2] loop filter, [3] YUV -> RGB(A). */
int ora_stage_timing = 0;
long long ora_stage_ns[4] = {0, 0, 0, 0};
static long long stage_now(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (long long)ts.tv_sec * 1000000000ll + ts.tv_nsec;
}
void ora_stage_add(int k, long long ns) { __atomic_fetch_add(&ora_stage_ns[k], ns, __ATOMIC_RELAXED); }
long long ora_stage_clock(void) { return stage_now(); }

int ora_vp8_decode(const uint8_t* data, size_t len, int keep, ora_frame* out) {
    vp8dec d;
    memset(&d, 0, sizeof(d));
    memset(out, 0, sizeof(*out));
    d.r = data; d.rlen = len; d.f = out; d.keep = keep;
    bool_uninit(&d.b);
    for (int i = 0; i < 8; i++) bool_uninit(&d.partitions[i]);
    d.segment_tree_probs[0] = d.segment_tree_probs[1] = d.segment_tree_probs[2] = 255;
    memcpy(d.token_probs, kCoeffProb0, sizeof(d.token_probs));
    out->prob_skip_false = -1;
    out->num_partitions = 1;

    int e = read_frame_header(&d);
    if (e) goto fail;
    int mbw = out->mbw, mbh = out->mbh;
    size_t nmb = (size_t)mbw * mbh;
    d.macroblocks = (ora_mb*)calloc(nmb ? nmb : 1, sizeof(ora_mb));
    if (!d.macroblocks) { e = ORA_ERR_NOMEM; goto fail; }
    if (keep) {
        out->levels = (int16_t*)calloc(nmb ? nmb * 400 : 1, sizeof(int16_t));
        out->residue = (int32_t*)calloc(nmb ? nmb * 384 : 1, sizeof(int32_t));
        if (!out->levels || !out->residue) { e = ORA_ERR_NOMEM; goto fail; }
    }
    for (int mby = 0; mby < mbh; mby++) {
        int p = mby % d.num_partitions;
        memset(&d.left, 0, sizeof(d.left));
        for (int mbx = 0; mbx < mbw; mbx++) {
            ora_mb* mb = &d.macroblocks[(size_t)mby * mbw + mbx];
            int32_t blocks[384];
            const long long t_a = ora_stage_timing ? stage_now() : 0;
            e = read_mb_header(&d, mbx, mb);
            if (e) goto fail;
            if (!mb->coeffs_skipped) {
                e = read_residual_data(&d, mb, mbx, p, blocks,
                                       keep ? out->levels + ((size_t)mby * mbw + mbx) * 400 : NULL);
                if (e) goto fail;
            } else {
                if (mb->luma_mode != LUMA_B) {
                    d.left.complexity[0] = 0;
                    d.top[mbx].complexity[0] = 0;
                }
                for (int i = 1; i < 9; i++) {
                    d.left.complexity[i] = 0;
                    d.top[mbx].complexity[i] = 0;
                }
                memset(blocks, 0, sizeof(blocks));
            }
            if (keep) memcpy(out->residue + ((size_t)mby * mbw + mbx) * 384, blocks, sizeof(blocks));
            const long long t_b = ora_stage_timing ? stage_now() : 0;
            intra_predict_luma(&d, mbx, mby, mb, blocks);
            intra_predict_chroma(&d, mbx, mby, mb, blocks);
            if (ora_stage_timing) {
                const long long t_c = stage_now();
                ora_stage_add(0, t_b - t_a);
                ora_stage_add(1, t_c - t_b);
            }
        }
        memset(d.left_border_y, 129, 17);
        memset(d.left_border_u, 129, 9);
        memset(d.left_border_v, 129, 9);
    }
    if (keep) {
        size_t ysz = (size_t)mbw * 16 * mbh * 16, csz = (size_t)mbw * 8 * mbh * 8;
        out->ybuf0 = (uint8_t*)malloc(ysz ? ysz : 1);
        out->ubuf0 = (uint8_t*)malloc(csz ? csz : 1);
        out->vbuf0 = (uint8_t*)malloc(csz ? csz : 1);
        if (!out->ybuf0 || !out->ubuf0 || !out->vbuf0) { e = ORA_ERR_NOMEM; goto fail; }
        memcpy(out->ybuf0, out->ybuf, ysz);
        memcpy(out->ubuf0, out->ubuf, csz);
        memcpy(out->vbuf0, out->vbuf, csz);
    }
    const long long t_lf = ora_stage_timing ? stage_now() : 0;
    for (int mby = 0; mby < mbh; mby++)
        for (int mbx = 0; mbx < mbw; mbx++)
            loop_filter_mb(out, mbx, mby, &d.macroblocks[(size_t)mby * mbw + mbx]);
    if (ora_stage_timing) ora_stage_add(2, stage_now() - t_lf);
    if (keep) { out->mbs = d.macroblocks; d.macroblocks = NULL; }
    dec_cleanup(&d);
    return ORA_OK;
fail:
    dec_cleanup(&d);
    ora_frame_free(out);
    return e;
}

/* ------------------------------------------------------------------------------------
 * yuv.rs
 * ---------------------------------------------------------------------------------- */
static int mulhi(uint8_t v, uint32_t coeff) { return (int)(((uint32_t)v * coeff) >> 8); } /* 30-32 */
static uint8_t clip6(int v) { v >>= 6; return (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v); }   /* 53-56 */
void ora_yuv_to_rgb(uint8_t y, uint8_t u, uint8_t v, uint8_t* rgb) { /* 59-71, 285-289 */
    rgb[0] = clip6(mulhi(y, 19077) + mulhi(v, 26149) - 14234);
    rgb[1] = clip6(mulhi(y, 19077) - mulhi(u, 6419) - mulhi(v, 13320) + 8708);
    rgb[2] = clip6(mulhi(y, 19077) + mulhi(u, 33050) - 17685);
}
static uint8_t fancy(uint8_t m, uint8_t s1, uint8_t s2, uint8_t t) { /* 276-282 */
    return (uint8_t)((9 * m + 3 * s1 + 3 * s2 + t + 8) / 16);
}
/* fill_row_fancy_with_2_uv_rows, yuv.rs:155-220 */
static void fill_row_fancy_2(uint8_t* row, int width, int bpp, const uint8_t* y, const uint8_t* u1,
                             const uint8_t* u2, const uint8_t* v1, const uint8_t* v2, int cw) {
    ora_yuv_to_rgb(y[0], fancy(u1[0], u1[0], u2[0], u2[0]), fancy(v1[0], v1[0], v2[0], v2[0]), row);
    int rest = width - 1;          /* pixels after the first */
    int pairs = rest / 2;
    if (pairs > cw - 1) pairs = cw - 1; /* zip with windows(2) */
    for (int k = 0; k < pairs; k++) {
        uint8_t* rgb = row + bpp + (size_t)k * 2 * bpp;
        ora_yuv_to_rgb(y[1 + 2 * k], fancy(u1[k], u1[k + 1], u2[k], u2[k + 1]),
                       fancy(v1[k], v1[k + 1], v2[k], v2[k + 1]), rgb);
        ora_yuv_to_rgb(y[2 + 2 * k], fancy(u1[k + 1], u1[k], u2[k + 1], u2[k]),
                       fancy(v1[k + 1], v1[k], v2[k + 1], v2[k]), rgb + bpp);
    }
    if (rest % 2 == 1) { /* remainder of chunks_exact(2): one final pixel */
        uint8_t* rgb = row + bpp + (size_t)(rest / 2) * 2 * bpp;
        uint8_t fu1 = u1[cw - 1], fu2 = u2[cw - 1], fv1 = v1[cw - 1], fv2 = v2[cw - 1];
        ora_yuv_to_rgb(y[width - 1], fancy(fu1, fu1, fu2, fu2), fancy(fv1, fv1, fv2, fv2), rgb);
    }
}
/* fill_row_fancy_with_1_uv_row, yuv.rs:222-273 */
static void fill_row_fancy_1(uint8_t* row, int width, int bpp, const uint8_t* y, const uint8_t* u,
                             const uint8_t* v, int cw) {
    ora_yuv_to_rgb(y[0], u[0], v[0], row);
    int rest = width - 1;
    int pairs = rest / 2;
    if (pairs > cw - 1) pairs = cw - 1;
    for (int k = 0; k < pairs; k++) {
        uint8_t* rgb = row + bpp + (size_t)k * 2 * bpp;
        ora_yuv_to_rgb(y[1 + 2 * k], fancy(u[k], u[k + 1], u[k], u[k + 1]), fancy(v[k], v[k + 1], v[k], v[k + 1]), rgb);
        ora_yuv_to_rgb(y[2 + 2 * k], fancy(u[k + 1], u[k], u[k + 1], u[k]), fancy(v[k + 1], v[k], v[k + 1], v[k]), rgb + bpp);
    }
    if (rest % 2 == 1) {
        uint8_t* rgb = row + bpp + (size_t)(rest / 2) * 2 * bpp;
        ora_yuv_to_rgb(y[width - 1], u[cw - 1], v[cw - 1], rgb);
    }
}
/* fill_rgb_buffer_fancy, yuv.rs:76-152 */
void ora_fill_rgb_buffer_fancy(uint8_t* buf, const uint8_t* yb, const uint8_t* ub, const uint8_t* vb,
                               int width, int height, int buffer_width, int bpp) {
    int cbw = buffer_width / 2;
    int cw = (width + 1) / 2;
    size_t rowb = (size_t)width * bpp;
    fill_row_fancy_1(buf, width, bpp, yb, ub, vb, cw);
    int npairs = (height - 1) / 2; /* chunks_exact(2 rows) over rows 1..height */
    for (int j = 0; j < npairs; j++) {
        const uint8_t *u1 = ub + (size_t)j * cbw, *u2 = u1 + cbw;
        const uint8_t *v1 = vb + (size_t)j * cbw, *v2 = v1 + cbw;
        fill_row_fancy_2(buf + rowb * (1 + 2 * j), width, bpp, yb + (size_t)buffer_width * (1 + 2 * j), u1, u2, v1, v2, cw);
        fill_row_fancy_2(buf + rowb * (2 + 2 * j), width, bpp, yb + (size_t)buffer_width * (2 + 2 * j), u2, u1, v2, v1, cw);
    }
    if ((height - 1) % 2 == 1) { /* even height: final row with its single chroma row */
        int ch = (height + 1) / 2;
        fill_row_fancy_1(buf + rowb * (height - 1), width, bpp, yb + (size_t)buffer_width * (height - 1),
                         ub + (size_t)(ch - 1) * cbw, vb + (size_t)(ch - 1) * cbw, cw);
    }
}
/* fill_rgb_buffer_simple + fill_rgba_row_simple, yuv.rs:293-382 */
void ora_fill_rgb_buffer_simple(uint8_t* buf, size_t buf_len, const uint8_t* yb, const uint8_t* ub,
                                const uint8_t* vb, int width, int chroma_width, int buffer_width,
                                int nrows_planes, int bpp) {
    size_t rowb = (size_t)width * bpp;
    size_t nrows = rowb ? buf_len / rowb : 0;
    if (nrows > (size_t)nrows_planes) nrows = (size_t)nrows_planes;
    for (size_t r = 0; r < nrows; r++) {
        uint8_t* row = buf + r * rowb;
        const uint8_t* y = yb + r * buffer_width;
        const uint8_t* u = ub + (r / 2) * (size_t)(buffer_width / 2);
        const uint8_t* v = vb + (r / 2) * (size_t)(buffer_width / 2);
        int pairs = width / 2;
        if (pairs > chroma_width) pairs = chroma_width;
        for (int k = 0; k < pairs; k++) {
            ora_yuv_to_rgb(y[2 * k], u[k], v[k], row + (size_t)2 * k * bpp);
            ora_yuv_to_rgb(y[2 * k + 1], u[k], v[k], row + (size_t)(2 * k + 1) * bpp);
        }
        if ((width & 1) && pairs < chroma_width)
            ora_yuv_to_rgb(y[width - 1], u[pairs], v[pairs], row + (size_t)(width - 1) * bpp);
    }
}
/* Frame::fill_rgb / fill_rgba, mod.rs:184-241 */
void ora_frame_fill(const ora_frame* f, uint8_t* buf, int bpp, int fancy_up) {
    int bw = f->mbw * 16; /* buffer_width(), mod.rs:174-181 */
    const long long t0 = ora_stage_timing ? stage_now() : 0;
    if (fancy_up)
        ora_fill_rgb_buffer_fancy(buf, f->ybuf, f->ubuf, f->vbuf, f->width, f->height, bw, bpp);
    else
        ora_fill_rgb_buffer_simple(buf, (size_t)f->width * f->height * bpp, f->ybuf, f->ubuf, f->vbuf,
                                   f->width, (f->width + 1) / 2, bw, f->mbh * 16, bpp);
    if (ora_stage_timing) ora_stage_add(3, stage_now() - t0);
}
