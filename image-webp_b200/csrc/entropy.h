// entropy.h -- hand-over between the host header parse and the device entropy stage.
//
// The device entropy stage (entropy.cu) runs the serial boolean decode of one frame's
// macroblock headers and coefficient tokens on the GPU: one warp per bitstream partition with a
// single working lane, thousands of frames side by side, instead of one frame per host core.
// The host still reads the frame header (reference lossy/mod.rs:510-613: a few hundred booleans)
// and passes on what the macroblock loop (mod.rs:1255-1281) needs: where the partitions are, the
// token probabilities, and the boolean decoder of the first partition exactly as the header left it.
#pragma once
#include <cstdint>

namespace webpb200 {

// State of a boolean decoder between two symbols, in the form the device reader uses:
// `value` holds bits + 8 valid bits, refilled one to three bytes at a time when bits < 0
// (arithmetic_decoder.rs semantics: one zero byte past the end is tolerated, 249-276).
struct BoolState {
    uint32_t pos;      // bytes of the partition consumed
    uint32_t value;
    uint32_t range;
    int32_t bits;
    uint8_t phantom;   // the tolerated zero byte has been used
    uint8_t eof;
    uint8_t pad[2];
};

struct Vp8EntropyHdr {
    uint64_t data_off;         // the VP8 chunk's bytes in the bitstream arena
    uint32_t first_off;        // first partition (headers and macroblock modes) inside the chunk
    uint32_t first_len;
    uint32_t part_off[8];      // token partitions inside the chunk (init_partitions, mod.rs:369-395)
    uint32_t part_len[8];
    BoolState first;           // first-partition reader at the first macroblock
    uint8_t num_parts;         // 1, 2, 4 or 8; 0 = frame is not to be decoded (failed on the host)
    uint8_t update_map;        // segment ids are coded per macroblock (mod.rs:624-627)
    uint8_t has_prob_skip;     // mb_no_coeff_skip flag (mod.rs:607-611)
    uint8_t prob_skip;
    uint8_t seg_probs[3];      // segment id tree probabilities
    uint8_t pad;
    uint8_t probs[1056];       // token probabilities [plane 4][band 8][ctx 3][node 11]
    uint8_t pad2[4];
};
static_assert(sizeof(Vp8EntropyHdr) == 1168, "Vp8EntropyHdr layout");

}  // namespace webpb200
