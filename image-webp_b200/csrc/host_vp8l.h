// host_vp8l.h -- VP8L decode of a compressed ALPH payload to its green channel.
#pragma once
#include <cstddef>
#include <cstdint>

namespace webpb200 {

// LosslessDecoder::decode_frame with implicit dimensions, green channel only
// (reference extended.rs:284-294, lossless/decoder/mod.rs:67-152).  green = width*height bytes.
int vp8l_decode_green(const uint8_t* data, size_t len, int width, int height, uint8_t* green);

}  // namespace webpb200
