/* oracle/ora_webp.c -- TEST INFRASTRUCTURE, not product code.
 *
 * Restatement in C of the callers around the lossy path:
 *   src/decoder.rs (RIFF walk, read_image, read_frame), src/extended.rs (VP8X header, ALPH
 *   chunk, alpha predictor, composite_frame), src/alpha_blending.rs.
 */
#include "ora.h"
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ---- alpha_blending.rs ------------------------------------------------------------------- */
static uint32_t div_by_255(uint32_t v) { return (((v + 0x80) >> 8) + v + 0x80) >> 8; } /* 73-75 */

void ora_blend_pixel(const uint8_t src[4], const uint8_t dst[4], uint8_t out[4]) { /* 28-64 */
    uint32_t src_a = src[3];
    if (src_a == 0) { memcpy(out, dst, 4); return; }
    uint32_t dst_a = dst[3];
    uint32_t dst_factor_a = div_by_255(dst_a * (255 - src_a));
    uint32_t blend_a = src_a + dst_factor_a;
    uint32_t scale = (1u << 24) / blend_a;
    uint8_t dfa = (uint8_t)dst_factor_a;
    for (int c = 0; c < 3; c++) {
        uint32_t unscaled = (uint32_t)src[c] * src_a + (uint32_t)dst[c] * dfa;
        out[c] = (uint8_t)((unscaled * scale) >> 24);
    }
    out[3] = (uint8_t)blend_a;
}

/* ---- extended.rs ------------------------------------------------------------------------ */
/* composite_frame, extended.rs:30-129 */
void ora_composite_frame(uint8_t* canvas, uint32_t cw, uint32_t ch, int has_clear, const uint8_t clear[4],
                         const uint8_t* frame, uint32_t fx, uint32_t fy, uint32_t fw, uint32_t fh,
                         int frame_has_alpha, int use_blend, uint32_t pw, uint32_t ph, uint32_t px, uint32_t py) {
    int full = fx == 0 && fy == 0 && fw == cw && fh == ch;
    size_t npx = (size_t)cw * ch;
    if (full && !use_blend) {
        if (frame_has_alpha) {
            memcpy(canvas, frame, npx * 4);
        } else {
            for (size_t i = 0; i < npx; i++) {
                memcpy(canvas + 4 * i, frame + 3 * i, 3);
                canvas[4 * i + 3] = 255;
            }
        }
        return;
    }
    if (has_clear) {
        if (full) {
            for (size_t i = 0; i < npx; i++) memcpy(canvas + 4 * i, clear, 4);
        } else {
            for (size_t y = 0; y < ph; y++)
                for (size_t x = 0; x < pw; x++)
                    memcpy(canvas + ((x + px) + (y + py) * (size_t)cw) * 4, clear, 4);
        }
    }
    uint32_t availw = cw > fx ? cw - fx : 0, availh = ch > fy ? ch - fy : 0;
    size_t width = fw < availw ? fw : availw, height = fh < availh ? fh : availh;
    if (frame_has_alpha && use_blend) {
        for (size_t y = 0; y < height; y++)
            for (size_t x = 0; x < width; x++) {
                const uint8_t* in = frame + (x + y * (size_t)fw) * 4;
                uint8_t* out = canvas + ((x + fx) + (y + fy) * (size_t)cw) * 4;
                uint8_t b[4];
                ora_blend_pixel(in, out, b);
                memcpy(out, b, 4);
            }
    } else if (frame_has_alpha) {
        for (size_t y = 0; y < height; y++)
            memcpy(canvas + (fx + (y + fy) * (size_t)cw) * 4, frame + y * (size_t)fw * 4, width * 4);
    } else {
        for (size_t y = 0; y < height; y++) {
            const uint8_t* in = frame + y * (size_t)fw * 3;
            uint8_t* out = canvas + (fx + (y + fy) * (size_t)cw) * 4;
            for (size_t x = 0; x < width; x++) {
                memcpy(out + 4 * x, in + 3 * x, 3);
                out[4 * x + 3] = 255;
            }
        }
    }
}

/* get_alpha_predictor, extended.rs:131-191 */
static uint8_t alpha_predictor(size_t x, size_t y, size_t width, int filter, const uint8_t* im) {
    switch (filter) {
    case 0: return 0;
    case 1:
        if (x == 0 && y == 0) return 0;
        if (x == 0) return im[((y - 1) * width + x) * 4 + 3];
        return im[(y * width + x - 1) * 4 + 3];
    case 2:
        if (x == 0 && y == 0) return 0;
        if (y == 0) return im[(y * width + x - 1) * 4 + 3];
        return im[((y - 1) * width + x) * 4 + 3];
    default: {
        int left, top, tl;
        if (x == 0 && y == 0) { left = top = tl = 0; }
        else if (x == 0) { left = top = tl = im[((y - 1) * width + x) * 4 + 3]; }
        else if (y == 0) { left = top = tl = im[(y * width + x - 1) * 4 + 3]; }
        else {
            left = im[(y * width + x - 1) * 4 + 3];
            top = im[((y - 1) * width + x) * 4 + 3];
            tl = im[((y - 1) * width + x - 1) * 4 + 3];
        }
        int c = left + top - tl;
        return (uint8_t)(c < 0 ? 0 : c > 255 ? 255 : c);
    }
    }
}

/* decoder.rs:718-734 (still image) and 837-854 (animation frame) */
void ora_alpha_merge(uint8_t* rgba, const uint8_t* alpha_data, int width, int height, int filter) {
    for (size_t y = 0; y < (size_t)height; y++)
        for (size_t x = 0; x < (size_t)width; x++) {
            uint8_t pred = alpha_predictor(x, y, (size_t)width, filter, rgba);
            size_t ai = y * (size_t)width + x;
            rgba[ai * 4 + 3] = (uint8_t)(pred + alpha_data[ai]);
        }
}

/* read_alpha_chunk, extended.rs:253-308 */
int ora_read_alpha_chunk(const uint8_t* data, size_t len, int width, int height, uint8_t* out, int* filter) {
    if (len < 1) return ORA_ERR_IO;
    uint8_t info = data[0];
    int preprocessing = (info & 0x30) >> 4;
    int filtering = (info & 0x0c) >> 2;
    int compression = info & 0x03;
    if (preprocessing > 1) return ORA_ERR_ALPHA_PREPROC;
    if (compression > 1) return ORA_ERR_COMPRESSION;
    *filter = filtering;
    size_t n = (size_t)width * height;
    if (compression == 1) {
        uint8_t* rgba = (uint8_t*)calloc(n ? n * 4 : 1, 1);
        if (!rgba) return ORA_ERR_NOMEM;
        int e = ora_vp8l_decode(data + 1, len - 1, (uint32_t)width, (uint32_t)height, 1, rgba);
        if (e) { free(rgba); return e; }
        for (size_t i = 0; i < n; i++) out[i] = rgba[4 * i + 1];
        free(rgba);
    } else {
        if (len - 1 < n) return ORA_ERR_IO;
        memcpy(out, data + 1, n);
    }
    return 0;
}

/* ---- decoder.rs ----------------------------------------------------------------------------- */
enum { CK_RIFF, CK_WEBP, CK_VP8, CK_VP8L, CK_VP8X, CK_ANIM, CK_ANMF, CK_ALPH, CK_ICCP, CK_EXIF, CK_XMP, CK_UNKNOWN, CK_N };
static int fourcc_kind(const uint8_t* p) { /* decoder.rs:189-205 */
    static const char* names[] = {"RIFF", "WEBP", "VP8 ", "VP8L", "VP8X", "ANIM", "ANMF", "ALPH", "ICCP", "EXIF", "XMP "};
    for (int i = 0; i < 11; i++) if (memcmp(p, names[i], 4) == 0) return i;
    return CK_UNKNOWN;
}

typedef struct { uint64_t start, end; int present; } range_t;

struct ora_decoder {
    const uint8_t* r; size_t len;
    uint32_t width, height;
    int kind;                 /* 0 lossy 1 lossless 2 extended */
    int ext_alpha, ext_animation;
    int has_bg; uint8_t bg[4]; uint8_t bg_hint[4];
    int is_lossy, has_alpha;
    uint32_t num_frames, loop_count; uint64_t loop_duration;
    range_t chunks[CK_N];
    int upsampling;           /* 0 bilinear 1 simple */
    size_t memory_limit;      /* decoder.rs:314, usize::MAX by default (354) */
    /* AnimationState, decoder.rs:240-263 */
    uint32_t next_frame; uint64_t next_frame_start; int dispose_next_frame;
    uint32_t prev_w, prev_h, prev_x, prev_y;
    uint8_t* canvas;
};

/* read_chunk_header, decoder.rs:957-964; returns 0 on EOF */
static int chunk_header(const ora_decoder* d, uint64_t pos, int* kind, uint64_t* size, uint64_t* rounded) {
    if (pos > d->len || d->len - pos < 8) return 0;
    const uint8_t* p = d->r + pos;
    *kind = fourcc_kind(p);
    uint32_t sz = p[4] | (p[5] << 8) | (p[6] << 16) | ((uint32_t)p[7] << 24);
    uint32_t odd = sz & 1;
    uint32_t r = sz > 0xffffffffu - odd ? 0xffffffffu : sz + odd; /* saturating_add */
    *size = sz; *rounded = r;
    return 1;
}
static uint32_t rd24(const uint8_t* p) { return p[0] | (p[1] << 8) | ((uint32_t)p[2] << 16); }

static void chunk_insert(ora_decoder* d, int kind, uint64_t s, uint64_t e) { /* entry().or_insert() */
    if (kind == CK_UNKNOWN || d->chunks[kind].present) return;
    d->chunks[kind].start = s; d->chunks[kind].end = e; d->chunks[kind].present = 1;
}

/* read_data, decoder.rs:365-549 */
static int read_data(ora_decoder* d) {
    int kind; uint64_t riff_size, rr;
    if (!chunk_header(d, 0, &kind, &riff_size, &rr)) return ORA_ERR_IO;
    if (kind != CK_RIFF) return ORA_ERR_RIFF_SIGNATURE;
    if (d->len < 12) return ORA_ERR_IO;
    if (fourcc_kind(d->r + 8) != CK_WEBP) return ORA_ERR_WEBP_SIGNATURE;
    uint64_t csize, crounded;
    if (!chunk_header(d, 12, &kind, &csize, &crounded)) return ORA_ERR_IO;
    uint64_t start = 20;
    const uint8_t* p = d->r + start;
    size_t avail = d->len - (size_t)start;
    if (kind == CK_VP8) {
        /* sequential reads, decoder.rs:380-396: each can hit the end of the data on its own */
        if (avail < 3) return ORA_ERR_IO;
        uint32_t tag = rd24(p);
        if (tag & 1) return ORA_ERR_UNSUPPORTED;
        if (avail < 6) return ORA_ERR_IO;
        if (p[3] != 0x9d || p[4] != 0x01 || p[5] != 0x2a) return ORA_ERR_VP8_MAGIC;
        if (avail < 10) return ORA_ERR_IO;
        d->width = (p[6] | (p[7] << 8)) & 0x3FFF;
        d->height = (p[8] | (p[9] << 8)) & 0x3FFF;
        if (d->width == 0 || d->height == 0) return ORA_ERR_INCONSISTENT_SIZE;
        chunk_insert(d, CK_VP8, start, start + csize);
        d->kind = 0; d->is_lossy = 1;
    } else if (kind == CK_VP8L) {
        if (avail < 1) return ORA_ERR_IO;
        if (p[0] != 0x2f) return ORA_ERR_LOSSLESS_SIG;
        if (avail < 5) return ORA_ERR_IO;
        uint32_t header = p[1] | (p[2] << 8) | (p[3] << 16) | ((uint32_t)p[4] << 24);
        if (header >> 29) return ORA_ERR_VERSION;
        d->width = (header & 0x3FFF) + 1;
        d->height = ((header >> 14) & 0x3FFF) + 1;
        chunk_insert(d, CK_VP8L, start, start + csize);
        d->kind = 1;
        d->has_alpha = (header >> 28) & 1;
    } else if (kind == CK_VP8X) {
        /* read_extended_header, extended.rs:193-228 */
        if (avail < 10) return ORA_ERR_IO;
        uint8_t flags = p[0];
        d->ext_alpha = (flags & 0x10) != 0;
        int exif = (flags & 0x08) != 0, xmp = (flags & 0x04) != 0;
        d->ext_animation = (flags & 0x02) != 0;
        uint32_t cw = rd24(p + 4) + 1, chh = rd24(p + 7) + 1;
        if ((uint64_t)cw * chh > 0xffffffffull) return ORA_ERR_IMAGE_TOO_LARGE;
        d->width = cw; d->height = chh;
        uint64_t position = start + crounded;
        uint64_t max_position = position + (riff_size > 12 ? riff_size - 12 : 0);
        while (position < max_position) {
            int ck; uint64_t sz, szr;
            if (!chunk_header(d, position, &ck, &sz, &szr)) break; /* UnexpectedEof => stop */
            uint64_t rs = position + 8, re = position + 8 + sz;
            uint64_t hdr_pos = position;
            position += 8 + szr;
            chunk_insert(d, ck, rs, re);
            if (ck == CK_ANMF) {
                d->num_frames++;
                if (sz < 24) return ORA_ERR_CHUNK_SIZE;
                if (hdr_pos + 8 + 16 > d->len) return ORA_ERR_IO;
                uint32_t duration = rd24(d->r + hdr_pos + 8 + 12); /* read_u32 & 0xffffff */
                d->loop_duration += duration;
                if (!d->is_lossy) {
                    int sk; uint64_t a, b;
                    if (!chunk_header(d, hdr_pos + 8 + 16, &sk, &a, &b)) return ORA_ERR_IO;
                    if (sk == CK_VP8 || sk == CK_ALPH) d->is_lossy = 1;
                }
            }
        }
        d->is_lossy = d->is_lossy || d->chunks[CK_VP8].present;
        if ((d->ext_animation && (!d->chunks[CK_ANIM].present || !d->chunks[CK_ANMF].present)) ||
            (exif && !d->chunks[CK_EXIF].present) || (xmp && !d->chunks[CK_XMP].present) ||
            (!d->ext_animation && d->chunks[CK_VP8].present == d->chunks[CK_VP8L].present))
            return ORA_ERR_CHUNK_MISSING;
        if (d->ext_animation) {
            range_t a = d->chunks[CK_ANIM];
            if (a.end - a.start > 6) return ORA_ERR_CHUNK_SIZE;
            if (a.end > d->len || a.end - a.start < 6) return ORA_ERR_IO;
            memcpy(d->bg_hint, d->r + a.start, 4);
            d->loop_count = d->r[a.start + 4] | (d->r[a.start + 5] << 8);
            d->next_frame_start = d->chunks[CK_ANMF].start - 8;
        }
        if (d->chunks[CK_ANMF].present) {
            range_t a = d->chunks[CK_ANMF];
            uint64_t pos2 = a.start + 16;
            for (int i = 0; i < 2; i++) {
                int sk; uint64_t sz, szr;
                if (!chunk_header(d, pos2, &sk, &sz, &szr)) return ORA_ERR_IO;
                chunk_insert(d, sk, pos2 + 8, pos2 + 8 + sz);
                pos2 += 8 + szr;
                if (pos2 + 8 > a.end) break;
            }
        }
        d->has_alpha = d->ext_alpha;
        d->kind = 2;
    } else {
        return ORA_ERR_CHUNK_HEADER;
    }
    return 0;
}

int ora_open(const uint8_t* data, size_t len, int upsampling, ora_decoder** out) {
    ora_decoder* d = (ora_decoder*)calloc(1, sizeof(*d));
    if (!d) return ORA_ERR_NOMEM;
    d->r = data; d->len = len; d->upsampling = upsampling;
    d->loop_count = 1; d->dispose_next_frame = 1;
    d->memory_limit = (size_t)-1;
    int e = read_data(d);
    if (e) { free(d); *out = NULL; return e; }
    *out = d;
    return 0;
}
void ora_close(ora_decoder* d) { if (d) { free(d->canvas); free(d); } }
void ora_dimensions(const ora_decoder* d, uint32_t* w, uint32_t* h) { *w = d->width; *h = d->height; }
int ora_has_alpha(const ora_decoder* d) { return d->has_alpha; }
int ora_is_animated(const ora_decoder* d) { return d->kind == 2 && d->ext_animation; }
int ora_is_lossy(const ora_decoder* d) { return d->is_lossy; }
uint32_t ora_num_frames(const ora_decoder* d) { return d->num_frames; }
uint32_t ora_loop_count(const ora_decoder* d) { return d->loop_count; }
uint64_t ora_loop_duration(const ora_decoder* d) { return d->loop_duration; }
size_t ora_output_buffer_size(const ora_decoder* d) { return (size_t)d->width * d->height * (d->has_alpha ? 4 : 3); }
int ora_set_background_color(ora_decoder* d, const uint8_t c[4]) {
    if (d->kind != 2) return ORA_ERR_INVALID_PARAM;
    d->has_bg = 1; memcpy(d->bg, c, 4);
    return 0;
}
void ora_set_lossy_upsampling(ora_decoder* d, int u) { d->upsampling = u; }
void ora_set_memory_limit(ora_decoder* d, size_t limit) { d->memory_limit = limit; } /* decoder.rs:554 */
int ora_background_color_hint(const ora_decoder* d, uint8_t rgba[4]) { /* decoder.rs:559-565 */
    if (d->kind != 2) return ORA_ERR_INVALID_PARAM;
    memcpy(rgba, d->bg_hint, 4);
    return 0;
}
/* read_chunk, decoder.rs:620-640, for ICCP / EXIF / XMP (643-655) */
int ora_metadata(ora_decoder* d, int which, const uint8_t** data, size_t* len) {
    const int kind = which == 0 ? CK_ICCP : which == 1 ? CK_EXIF : CK_XMP;
    *data = NULL; *len = 0;
    if (!d->chunks[kind].present) return 0;
    range_t r = d->chunks[kind];
    if (r.end - r.start > (uint64_t)d->memory_limit) return ORA_ERR_MEMORY_LIMIT;
    if (r.end > d->len) return ORA_ERR_IO; /* read_exact past the end */
    *data = d->r + r.start; *len = (size_t)(r.end - r.start);
    return 0;
}

static int range_ok(const ora_decoder* d, range_t r) { return r.start <= r.end && r.end <= d->len; }

/* decode one VP8 chunk to rgb/rgba (+ALPH merge) */
static int lossy_to_pixels(ora_decoder* d, const uint8_t* vp8, size_t vp8len, const uint8_t* alph, size_t alphlen,
                           int check_w, int check_h, int do_check, int bpp, uint8_t* buf) {
    ora_frame f;
    int e = ora_vp8_decode(vp8, vp8len, 0, &f);
    if (e) return e;
    if (do_check && (f.width != check_w || f.height != check_h)) { ora_frame_free(&f); return ORA_ERR_INCONSISTENT_SIZE; }
    ora_frame_fill(&f, buf, bpp, d->upsampling == 0);
    if (bpp == 4 && !alph) { ora_frame_free(&f); return ORA_ERR_CHUNK_MISSING; } /* decoder.rs:707-711 */
    if (bpp == 4) {
        /* NB decoder.rs:814-857: in an animation the ALPH plane is sized by the ANMF frame
         * while the merge loop runs over the VP8 frame's own size. */
        size_t n = (size_t)check_w * check_h;
        uint8_t* a = (uint8_t*)malloc(n ? n : 1);
        int filter = 0;
        if (!a) { ora_frame_free(&f); return ORA_ERR_NOMEM; }
        e = ora_read_alpha_chunk(alph, alphlen, check_w, check_h, a, &filter);
        if (!e) ora_alpha_merge(buf, a, f.width, f.height, filter);
        free(a);
    }
    ora_frame_free(&f);
    return e;
}

int ora_read_frame(ora_decoder* d, uint8_t* buf, size_t len, uint32_t* duration_ms);

/* read_image, decoder.rs:669-741 */
int ora_read_image(ora_decoder* d, uint8_t* buf, size_t len) {
    if (len != ora_output_buffer_size(d)) return ORA_ERR_IMAGE_TOO_LARGE;
    if (ora_is_animated(d)) {
        /* 674-680: run read_frame on a fresh AnimationState, then restore */
        uint32_t nf = d->next_frame; uint64_t nfs = d->next_frame_start; int disp = d->dispose_next_frame;
        uint32_t pw = d->prev_w, ph = d->prev_h, px = d->prev_x, py = d->prev_y;
        uint8_t* canvas = d->canvas;
        d->next_frame = 0; d->dispose_next_frame = 1; d->prev_w = d->prev_h = d->prev_x = d->prev_y = 0; d->canvas = NULL;
        d->next_frame_start = d->chunks[CK_ANMF].start - 8;
        uint32_t dur;
        int e = ora_read_frame(d, buf, len, &dur);
        free(d->canvas);
        d->next_frame = nf; d->next_frame_start = nfs; d->dispose_next_frame = disp;
        d->prev_w = pw; d->prev_h = ph; d->prev_x = px; d->prev_y = py; d->canvas = canvas;
        return e;
    }
    if (d->chunks[CK_VP8L].present) {
        range_t r = d->chunks[CK_VP8L];
        if (!range_ok(d, r)) return ORA_ERR_IO;
        if (d->has_alpha) return ora_vp8l_decode(d->r + r.start, (size_t)(r.end - r.start), d->width, d->height, 0, buf);
        size_t n = (size_t)d->width * d->height;
        uint8_t* tmp = (uint8_t*)calloc(n ? n * 4 : 1, 1);
        if (!tmp) return ORA_ERR_NOMEM;
        int e = ora_vp8l_decode(d->r + r.start, (size_t)(r.end - r.start), d->width, d->height, 0, tmp);
        if (!e) for (size_t i = 0; i < n; i++) memcpy(buf + 3 * i, tmp + 4 * i, 3);
        free(tmp);
        return e;
    }
    if (!d->chunks[CK_VP8].present) return ORA_ERR_CHUNK_MISSING;
    range_t r = d->chunks[CK_VP8];
    if (!range_ok(d, r)) return ORA_ERR_IO;
    if (d->has_alpha) {
        range_t a = d->chunks[CK_ALPH];
        if (a.present && !range_ok(d, a)) return ORA_ERR_IO;
        return lossy_to_pixels(d, d->r + r.start, (size_t)(r.end - r.start), a.present ? d->r + a.start : NULL,
                               a.present ? (size_t)(a.end - a.start) : 0, (int)d->width, (int)d->height, 1, 4, buf);
    }
    return lossy_to_pixels(d, d->r + r.start, (size_t)(r.end - r.start), NULL, 0, (int)d->width, (int)d->height, 1, 3, buf);
}

/* read_frame, decoder.rs:752-922 */
int ora_read_frame(ora_decoder* d, uint8_t* buf, size_t len, uint32_t* duration_ms) {
    if (!ora_is_animated(d)) return ORA_ERR_NOT_ANIMATED;
    if (len != ora_output_buffer_size(d)) return ORA_ERR_IMAGE_TOO_LARGE;
    if (d->next_frame == d->num_frames) return ORA_ERR_NO_MORE_FRAMES;
    int kind; uint64_t anmf_size, rr;
    uint64_t pos = d->next_frame_start;
    if (!chunk_header(d, pos, &kind, &anmf_size, &rr)) return ORA_ERR_IO;
    if (kind != CK_ANMF || anmf_size < 32) return ORA_ERR_CHUNK_HEADER;
    pos += 8;
    if (pos + 16 > d->len) return ORA_ERR_IO;
    const uint8_t* p = d->r + pos;
    uint32_t frame_x = rd24(p) * 2, frame_y = rd24(p + 3) * 2;
    uint32_t frame_w = rd24(p + 6) + 1, frame_h = rd24(p + 9) + 1;
    if (frame_w > 16384 || frame_h > 16384) return ORA_ERR_IMAGE_TOO_LARGE;
    if (frame_x + frame_w > d->width || frame_y + frame_h > d->height) return ORA_ERR_FRAME_OUTSIDE;
    uint32_t duration = rd24(p + 12);
    uint8_t frame_info = p[15];
    int use_blend = (frame_info & 2) == 0;
    int dispose = (frame_info & 1) != 0;
    pos += 16;
    uint64_t csize, crounded;
    if (!chunk_header(d, pos, &kind, &csize, &crounded)) return ORA_ERR_IO;
    if (crounded + 24 > anmf_size) return ORA_ERR_CHUNK_HEADER;
    pos += 8;
    uint8_t* frame = NULL;
    int frame_has_alpha = 0, e = 0;
    size_t npx = (size_t)frame_w * frame_h;
    if (kind == CK_VP8) {
        if (pos + csize > d->len) return ORA_ERR_IO;
        frame = (uint8_t*)calloc(npx * 3, 1);
        if (!frame) return ORA_ERR_NOMEM;
        e = lossy_to_pixels(d, d->r + pos, (size_t)csize, NULL, 0, (int)frame_w, (int)frame_h, 1, 3, frame);
    } else if (kind == CK_VP8L) {
        if (pos + csize > d->len) return ORA_ERR_IO;
        frame = (uint8_t*)calloc(npx * 4, 1);
        if (!frame) return ORA_ERR_NOMEM;
        e = ora_vp8l_decode(d->r + pos, (size_t)csize, frame_w, frame_h, 0, frame);
        frame_has_alpha = 1;
    } else if (kind == CK_ALPH) {
        if (crounded + 32 > anmf_size) return ORA_ERR_CHUNK_HEADER;
        if (pos + csize > d->len) return ORA_ERR_IO;
        uint64_t next = pos + crounded;
        int nk; uint64_t nsize, nr;
        if (!chunk_header(d, next, &nk, &nsize, &nr)) return ORA_ERR_IO;
        if (csize + nsize + 32 > anmf_size) return ORA_ERR_CHUNK_HEADER;
        if (next + 8 + nsize > d->len) return ORA_ERR_IO;
        frame = (uint8_t*)calloc(npx * 4, 1);
        if (!frame) return ORA_ERR_NOMEM;
        /* no size check of the VP8 frame in this branch (decoder.rs:814-857) */
        {
            ora_frame f;
            e = ora_vp8_decode(d->r + next + 8, (size_t)nsize, 0, &f);
            if (!e) {
                if ((size_t)f.width * f.height != npx) {
                    /* the reference would index out of bounds / mis-stride here; refuse */
                    e = ORA_ERR_INCONSISTENT_SIZE;
                } else {
                    size_t n = npx;
                    uint8_t* a = (uint8_t*)malloc(n ? n : 1);
                    int filter = 0;
                    if (!a) e = ORA_ERR_NOMEM;
                    /* reference order: ALPH is parsed before the VP8 chunk */
                    if (!e) e = ora_read_alpha_chunk(d->r + pos, (size_t)csize, (int)frame_w, (int)frame_h, a, &filter);
                    if (!e) {
                        ora_frame_fill(&f, frame, 4, d->upsampling == 0);
                        ora_alpha_merge(frame, a, f.width, f.height, filter);
                    }
                    free(a);
                }
                ora_frame_free(&f);
            }
        }
        frame_has_alpha = 1;
    } else {
        return ORA_ERR_CHUNK_HEADER;
    }
    if (e) { free(frame); return e; }

    int has_clear = 0; uint8_t clear[4] = {0, 0, 0, 0};
    if (d->dispose_next_frame) { /* 861-869 */
        if (d->has_bg) { has_clear = 1; memcpy(clear, d->bg, 4); }
        else if (frame_has_alpha) { has_clear = 1; }
    }
    if (!d->canvas) { /* 872-882 */
        size_t n = (size_t)d->width * d->height;
        d->canvas = (uint8_t*)calloc(n ? n * 4 : 1, 1);
        if (!d->canvas) { free(frame); return ORA_ERR_NOMEM; }
        if (d->has_bg) for (size_t i = 0; i < n; i++) memcpy(d->canvas + 4 * i, d->bg, 4);
    }
    ora_composite_frame(d->canvas, d->width, d->height, has_clear, clear, frame, frame_x, frame_y, frame_w, frame_h,
                        frame_has_alpha, use_blend, d->prev_w, d->prev_h, d->prev_x, d->prev_y);
    free(frame);
    d->prev_w = frame_w; d->prev_h = frame_h; d->prev_x = frame_x; d->prev_y = frame_y;
    d->dispose_next_frame = dispose;
    d->next_frame_start += anmf_size + 8;
    d->next_frame += 1;
    size_t n = (size_t)d->width * d->height;
    if (d->has_alpha) memcpy(buf, d->canvas, n * 4);
    else for (size_t i = 0; i < n; i++) memcpy(buf + 3 * i, d->canvas + 4 * i, 3);
    if (duration_ms) *duration_ms = duration;
    return 0;
}

int ora_reset_animation(ora_decoder* d) { /* decoder.rs:929-935 */
    if (!ora_is_animated(d)) return ORA_ERR_NOT_ANIMATED;
    d->next_frame = 0;
    d->next_frame_start = d->chunks[CK_ANMF].start - 8;
    d->dispose_next_frame = 1;
    return 0;
}

/* ---- CPU baseline: one image per worker thread at a time -------------------------------- */
typedef struct {
    const uint8_t* const* files; const size_t* lens; int n;
    uint8_t* const* outs; const size_t* out_lens; int* status;
    int next; pthread_mutex_t mu;
} batch_t;

static void* batch_worker(void* arg) {
    batch_t* b = (batch_t*)arg;
    for (;;) {
        pthread_mutex_lock(&b->mu);
        int i = b->next++;
        pthread_mutex_unlock(&b->mu);
        if (i >= b->n) break;
        ora_decoder* d = NULL;
        int e = ora_open(b->files[i], b->lens[i], 0, &d);
        if (!e) { e = ora_read_image(d, b->outs[i], b->out_lens[i]); ora_close(d); }
        b->status[i] = e;
    }
    return NULL;
}

int ora_decode_batch(const uint8_t* const* files, const size_t* lens, int n, uint8_t* const* outs,
                     const size_t* out_lens, int* status, int threads) {
    batch_t b = {files, lens, n, outs, out_lens, status, 0, PTHREAD_MUTEX_INITIALIZER};
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    pthread_t th[256];
    int started = 0;
    for (int t = 0; t < threads - 1; t++)
        if (pthread_create(&th[started], NULL, batch_worker, &b) == 0) started++;
    batch_worker(&b);
    for (int t = 0; t < started; t++) pthread_join(th[t], NULL);
    int bad = 0;
    for (int i = 0; i < n; i++) if (status[i]) bad++;
    return bad;
}
