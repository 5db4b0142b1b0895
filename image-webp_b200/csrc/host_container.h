// host_container.h -- RIFF / VP8X / ANIM / ANMF / ALPH walk (host side, no pixel work).
#pragma once
#include <cstddef>
#include <cstdint>

namespace webpb200 {

struct Range {
    uint64_t start = 0, end = 0;
    bool present = false;
    uint64_t size() const { return end - start; }
};

enum ChunkKind { CK_RIFF, CK_WEBP, CK_VP8, CK_VP8L, CK_VP8X, CK_ANIM, CK_ANMF, CK_ALPH, CK_ICCP, CK_EXIF, CK_XMP, CK_UNKNOWN, CK_COUNT };

// What WebPDecoder::new learns about a file (reference decoder.rs:365-549).
struct Container {
    const uint8_t* data = nullptr;
    size_t len = 0;
    uint32_t width = 0, height = 0;
    int kind = 0;  // 0 simple lossy, 1 simple lossless, 2 extended
    bool animated = false, has_alpha = false, is_lossy = false;
    uint32_t num_frames = 0, loop_count = 1;
    uint64_t loop_duration = 0;
    uint8_t background_hint[4] = {0, 0, 0, 0};
    Range chunks[CK_COUNT];
    uint64_t first_anmf = 0;  // file offset of the first ANMF chunk header
};

int parse_container(const uint8_t* data, size_t len, Container* out);

// chunk header at `pos` (reference decoder.rs:957-964); false at end of data
bool read_chunk_header(const uint8_t* data, size_t len, uint64_t pos, int* kind, uint64_t* size, uint64_t* rounded);

struct AnmfFrame {
    uint32_t x, y, w, h, duration;
    bool blend, dispose;
    Range vp8, alph;      // sub-chunk payloads
    bool lossless = false;  // frame is a VP8L chunk (outside this library's path)
    uint64_t next = 0;    // offset of the following ANMF header
};
// ANMF header + sub-chunk ranges (reference decoder.rs:762-833)
int parse_anmf(const Container& c, uint64_t pos, AnmfFrame* out);

// ALPH chunk header byte (reference extended.rs:253-283); *compressed = VP8L payload follows
int parse_alph_header(const uint8_t* data, size_t len, int* filter, bool* compressed);

}  // namespace webpb200
