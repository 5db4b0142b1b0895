// host_vp8l.h -- VP8L decode of a compressed ALPH payload to its green channel.
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

namespace webpb200 {

// LosslessDecoder::decode_frame with implicit dimensions, green channel only
// (reference extended.rs:284-294, lossless/decoder/mod.rs:67-152).  green = width*height bytes.
int vp8l_decode_green(const uint8_t* data, size_t len, int width, int height, uint8_t* green);

// The entropy-coded half alone (prefix codes, backward references, colour cache: decoder/mod.rs:159-601): the
// image in the TRANSFORMED domain into argb (row length plan->st[0].w_in, or width when there is no transform;
// room for width * height + 1 words) and the transforms in the order they are to be undone.  The device undoes
// them (vp8l.cu).
struct Vp8lPlan {
    struct Stage {
        int kind = 0, bits = 0, table_size = 0, w_in = 0;  // kind: 0 predictor 1 colour 2 subtract-green 3 palette (bits = log2 pixels per byte)
        std::vector<uint32_t> data;                        // block image (predictor / colour) or summed colour table
    };
    int n_stages = 0;
    Stage st[4];
};
int vp8l_decode_entropy(const uint8_t* data, size_t len, int width, int height, uint32_t* argb, Vp8lPlan* plan);

}  // namespace webpb200
