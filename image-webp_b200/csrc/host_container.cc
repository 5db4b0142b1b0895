// host_container.cc -- see host_container.h.  Follows reference decoder.rs:365-549 (read_data),
// 762-833 (ANMF), extended.rs:193-228 (VP8X header) and 253-283 (ALPH header).
#include "host_container.h"

#include <cstring>

#include "../../include/webpb200.h"

namespace webpb200 {
namespace {

int fourcc(const uint8_t* p) {
    static const char* const names[11] = {"RIFF", "WEBP", "VP8 ", "VP8L", "VP8X", "ANIM", "ANMF", "ALPH", "ICCP", "EXIF", "XMP "};
    for (int i = 0; i < 11; i++)
        if (!std::memcmp(p, names[i], 4)) return i;
    return CK_UNKNOWN;
}
uint32_t le24(const uint8_t* p) { return p[0] | (p[1] << 8) | ((uint32_t)p[2] << 16); }
uint32_t le32(const uint8_t* p) { return p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t)p[3] << 24); }

void remember(Container* c, int kind, uint64_t s, uint64_t e) {  // HashMap entry().or_insert()
    if (kind == CK_UNKNOWN || c->chunks[kind].present) return;
    c->chunks[kind].start = s;
    c->chunks[kind].end = e;
    c->chunks[kind].present = true;
}

}  // namespace

bool read_chunk_header(const uint8_t* data, size_t len, uint64_t pos, int* kind, uint64_t* size, uint64_t* rounded) {
    if (pos > len || len - pos < 8) return false;
    *kind = fourcc(data + pos);
    const uint32_t sz = le32(data + pos + 4);
    const uint32_t odd = sz & 1;
    *size = sz;
    *rounded = sz > 0xffffffffu - odd ? 0xffffffffu : sz + odd;  // saturating_add
    return true;
}

int parse_container(const uint8_t* data, size_t len, Container* c) {
    *c = Container();
    c->data = data;
    c->len = len;
    int kind;
    uint64_t riff_size, tmp;
    if (!read_chunk_header(data, len, 0, &kind, &riff_size, &tmp)) return WEBPB200_ERR_IO;
    if (kind != CK_RIFF) return WEBPB200_ERR_RIFF_SIGNATURE;
    if (len < 12) return WEBPB200_ERR_IO;
    if (fourcc(data + 8) != CK_WEBP) return WEBPB200_ERR_WEBP_SIGNATURE;
    uint64_t size, rounded;
    if (!read_chunk_header(data, len, 12, &kind, &size, &rounded)) return WEBPB200_ERR_IO;
    const uint64_t start = 20;
    const uint8_t* p = data + start;
    const size_t avail = len - (size_t)start;
    switch (kind) {
        case CK_VP8: {
            if (avail < 3) return WEBPB200_ERR_IO;
            if (le24(p) & 1) return WEBPB200_ERR_UNSUPPORTED;
            if (avail < 6) return WEBPB200_ERR_IO;
            if (p[3] != 0x9d || p[4] != 0x01 || p[5] != 0x2a) return WEBPB200_ERR_VP8_MAGIC;
            if (avail < 10) return WEBPB200_ERR_IO;
            c->width = (p[6] | (p[7] << 8)) & 0x3FFF;
            c->height = (p[8] | (p[9] << 8)) & 0x3FFF;
            if (!c->width || !c->height) return WEBPB200_ERR_INCONSISTENT_SIZE;
            remember(c, CK_VP8, start, start + size);
            c->kind = 0;
            c->is_lossy = true;
            return 0;
        }
        case CK_VP8L: {
            if (avail < 1) return WEBPB200_ERR_IO;
            if (p[0] != 0x2f) return WEBPB200_ERR_LOSSLESS_SIG;
            if (avail < 5) return WEBPB200_ERR_IO;
            const uint32_t h = le32(p + 1);
            if (h >> 29) return WEBPB200_ERR_VERSION;
            c->width = (h & 0x3FFF) + 1;
            c->height = ((h >> 14) & 0x3FFF) + 1;
            remember(c, CK_VP8L, start, start + size);
            c->kind = 1;
            c->has_alpha = (h >> 28) & 1;
            return 0;
        }
        case CK_VP8X:
            break;
        default:
            return WEBPB200_ERR_CHUNK_HEADER;
    }
    // VP8X: extended.rs:193-228 then the chunk walk decoder.rs:428-544
    if (avail < 10) return WEBPB200_ERR_IO;
    const uint8_t flags = p[0];
    const bool alpha = flags & 0x10, exif = flags & 0x08, xmp = flags & 0x04, anim = flags & 0x02;
    c->width = le24(p + 4) + 1;
    c->height = le24(p + 7) + 1;
    if ((uint64_t)c->width * c->height > 0xffffffffull) return WEBPB200_ERR_IMAGE_TOO_LARGE;
    uint64_t pos = start + rounded;
    const uint64_t limit = pos + (riff_size > 12 ? riff_size - 12 : 0);
    while (pos < limit) {
        int ck;
        uint64_t sz, szr;
        if (!read_chunk_header(data, len, pos, &ck, &sz, &szr)) break;  // UnexpectedEof ends the walk
        const uint64_t hdr = pos;
        remember(c, ck, pos + 8, pos + 8 + sz);
        pos += 8 + szr;
        if (ck != CK_ANMF) continue;
        if (c->num_frames == 0) c->first_anmf = hdr;
        c->num_frames++;
        if (sz < 24) return WEBPB200_ERR_CHUNK_SIZE;
        if (hdr + 8 + 16 > len) return WEBPB200_ERR_IO;
        c->loop_duration += le24(data + hdr + 8 + 12);
        if (!c->is_lossy) {
            int sk;
            uint64_t a, b;
            if (!read_chunk_header(data, len, hdr + 24, &sk, &a, &b)) return WEBPB200_ERR_IO;
            if (sk == CK_VP8 || sk == CK_ALPH) c->is_lossy = true;
        }
    }
    c->is_lossy = c->is_lossy || c->chunks[CK_VP8].present;
    if ((anim && (!c->chunks[CK_ANIM].present || !c->chunks[CK_ANMF].present)) || (exif && !c->chunks[CK_EXIF].present) ||
        (xmp && !c->chunks[CK_XMP].present) || (!anim && c->chunks[CK_VP8].present == c->chunks[CK_VP8L].present))
        return WEBPB200_ERR_CHUNK_MISSING;
    if (anim) {
        const Range a = c->chunks[CK_ANIM];
        if (a.size() > 6) return WEBPB200_ERR_CHUNK_SIZE;
        if (a.end > len || a.size() < 6) return WEBPB200_ERR_IO;
        std::memcpy(c->background_hint, data + a.start, 4);
        c->loop_count = data[a.start + 4] | (data[a.start + 5] << 8);
    }
    if (c->chunks[CK_ANMF].present) {  // index the first frame's sub-chunks, decoder.rs:526-540
        const Range a = c->chunks[CK_ANMF];
        uint64_t q = a.start + 16;
        for (int i = 0; i < 2; i++) {
            int sk;
            uint64_t sz, szr;
            if (!read_chunk_header(data, len, q, &sk, &sz, &szr)) return WEBPB200_ERR_IO;
            remember(c, sk, q + 8, q + 8 + sz);
            q += 8 + szr;
            if (q + 8 > a.end) break;
        }
    }
    c->has_alpha = alpha;
    c->animated = anim;
    c->kind = 2;
    return 0;
}

int parse_anmf(const Container& c, uint64_t pos, AnmfFrame* f) {
    int kind;
    uint64_t anmf_size, r;
    if (!read_chunk_header(c.data, c.len, pos, &kind, &anmf_size, &r)) return WEBPB200_ERR_IO;
    if (kind != CK_ANMF || anmf_size < 32) return WEBPB200_ERR_CHUNK_HEADER;
    pos += 8;
    if (pos + 16 > c.len) return WEBPB200_ERR_IO;
    const uint8_t* p = c.data + pos;
    f->x = le24(p) * 2;
    f->y = le24(p + 3) * 2;
    f->w = le24(p + 6) + 1;
    f->h = le24(p + 9) + 1;
    if (f->w > 16384 || f->h > 16384) return WEBPB200_ERR_IMAGE_TOO_LARGE;
    if (f->x + f->w > c.width || f->y + f->h > c.height) return WEBPB200_ERR_FRAME_OUTSIDE;
    f->duration = le24(p + 12);
    f->blend = (p[15] & 2) == 0;
    f->dispose = (p[15] & 1) != 0;
    f->next = pos - 8 + anmf_size + 8;
    pos += 16;
    uint64_t size, rounded;
    if (!read_chunk_header(c.data, c.len, pos, &kind, &size, &rounded)) return WEBPB200_ERR_IO;
    if (rounded + 24 > anmf_size) return WEBPB200_ERR_CHUNK_HEADER;
    pos += 8;
    f->vp8 = Range();
    f->alph = Range();
    f->lossless = false;
    if (kind == CK_VP8 || kind == CK_VP8L) {
        if (pos + size > c.len) return WEBPB200_ERR_IO;
        f->vp8.start = pos;
        f->vp8.end = pos + size;
        f->vp8.present = true;
        f->lossless = kind == CK_VP8L;
        return 0;
    }
    if (kind != CK_ALPH) return WEBPB200_ERR_CHUNK_HEADER;
    if (rounded + 32 > anmf_size) return WEBPB200_ERR_CHUNK_HEADER;
    if (pos + size > c.len) return WEBPB200_ERR_IO;
    f->alph.start = pos;
    f->alph.end = pos + size;
    f->alph.present = true;
    const uint64_t next = pos + rounded;
    int nk;
    uint64_t nsize, nr;
    if (!read_chunk_header(c.data, c.len, next, &nk, &nsize, &nr)) return WEBPB200_ERR_IO;
    if (size + nsize + 32 > anmf_size) return WEBPB200_ERR_CHUNK_HEADER;
    if (next + 8 + nsize > c.len) return WEBPB200_ERR_IO;
    f->vp8.start = next + 8;  // the reference decodes whatever follows as VP8 (decoder.rs:831-833)
    f->vp8.end = next + 8 + nsize;
    f->vp8.present = true;
    return 0;
}

int parse_alph_header(const uint8_t* data, size_t len, int* filter, bool* compressed) {
    if (len < 1) return WEBPB200_ERR_IO;
    const uint8_t info = data[0];
    if (((info >> 4) & 3) > 1) return WEBPB200_ERR_ALPHA_PREPROC;
    if ((info & 3) > 1) return WEBPB200_ERR_COMPRESSION;
    *filter = (info >> 2) & 3;
    *compressed = (info & 3) == 1;
    return 0;
}

}  // namespace webpb200
