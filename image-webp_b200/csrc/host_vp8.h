// host_vp8.h -- host half of the VP8 keyframe decode: header parse and the serial
// boolean-entropy decode of modes and coefficient levels into the packed boundary arrays
// (include/webpb200.h).  No pixel work happens on the host.
#pragma once
#include <cstddef>
#include <cstdint>

#include "../../include/webpb200.h"
#include "entropy.h"

namespace webpb200 {

struct Vp8Dims {
    int width, height, mbw, mbh;
};

// First 10 bytes of a VP8 chunk: frame tag, start code, dimensions
// (reference lossy/mod.rs:511-540).  Returns a WEBPB200_* status.
int vp8_probe(const uint8_t* data, size_t len, Vp8Dims* dims);

// Everything Vp8Decoder::decode_frame reads from the bitstream (reference
// lossy/mod.rs:510-671, 785-963, 1252-1281), without reconstructing:
//   fd     <- frame-level fields (dims, filter, segments with resolved quantisers, deltas);
//             offsets / out_bpp / upsampling / alpha fields are left untouched
//   mbs    <- mbw*mbh macroblock descriptors, raster order
//   levels <- mbw*mbh*400 raw token values (this function zero-fills them)
int vp8_entropy_decode(const uint8_t* data, size_t len, webpb200_frame_desc* fd, webpb200_mb_desc* mbs,
                       int16_t* levels);

// The header half of vp8_entropy_decode alone, for the device entropy stage: frame-level fields
// into fd as above, and into eh the partition layout, token probabilities and the state of the
// first partition's boolean decoder at the first macroblock (eh->data_off is the caller's).
int vp8_parse_header(const uint8_t* data, size_t len, webpb200_frame_desc* fd, Vp8EntropyHdr* eh);

// vp8_entropy_decode done the way the device entropy stage does it, on the host: the frame header is
// parsed into a Vp8EntropyHdr, and the macroblock loop then runs from that structure alone with the
// byte-wise reader resumed from the exported state.  For tests of the hand-over (no GPU needed).
int vp8_entropy_decode_resumed(const uint8_t* data, size_t len, webpb200_frame_desc* fd, webpb200_mb_desc* mbs, int16_t* levels);

}  // namespace webpb200
