// vp8l.cuh -- inverse transforms of VP8L-compressed ALPH planes on the device.
//
// The entropy-coded part of a VP8L stream (prefix codes, LZ77, colour cache) is serial and stays on the host
// (host_vp8l.cc); what it yields is the image in the TRANSFORMED domain plus the transform list.  Undoing the
// transforms (reference lossless/decoder/reverse_transform.rs:24-578) is pixel work: three of them are pointwise,
// the predictor transform is a recurrence over left / top / top-left / top-right neighbours that runs as a
// row-skewed wavefront.  The green channel of the result is the alpha residual plane (extended.rs:284-294).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace webpb200 {

enum { VP8L_PREDICTOR = 0, VP8L_COLOR = 1, VP8L_SUBTRACT_GREEN = 2, VP8L_PALETTE = 3 };

// One transform to undo, in the order the decoder undoes them (the reverse of the order in the stream).
struct Vp8lStage {
    uint32_t kind;        // VP8L_*
    uint32_t bits;        // predictor / colour: block size log2;  palette: pixels-per-byte log2 (0..3)
    uint32_t table_size;  // palette: colours in the table
    uint32_t w_in;        // row length of the image this stage reads (palette: the packed width)
    uint64_t data_off;    // word offset of the stage's data in the aux arena: block image (mode / coefficients in the
                          // ARGB word as the stream coded them) or the colour table (already summed, adjust_color_map)
};

struct Vp8lPlane {
    uint32_t width, height;   // of the finished plane
    uint32_t n_stages;
    uint32_t first_band;      // index of this plane's first 32-row band in the hand-off arena bookkeeping (unused by kernels)
    uint64_t buf_off[3];      // word offsets of the plane's three ARGB buffers in the pixel arena (each width*height + 1 words):
                              // [0] the entropy decoder's output as uploaded (never written by a kernel, so a pass can be
                              // repeated), [1] / [2] written alternately: stage s reads [0] (s = 0) or what stage s - 1 wrote,
                              // and writes [1 + (s & 1)]
    uint64_t hand_off;        // unit offset of the plane's hand-off records (2 * width units of 8 bytes)
    uint64_t alpha_off;       // byte offset of the plane's alpha residuals in the alpha arena (frame_desc.alpha_off)
    Vp8lStage st[4];
};

struct Vp8lBatch {
    const Vp8lPlane* planes;  // device
    int n_planes;
    int max_stages;           // largest n_stages
    uint32_t max_height;
    uint32_t* pixels;         // device pixel arena
    const uint32_t* aux;      // device aux arena
    uint2* handoff;           // device hand-off records, zero before the first launch of a batch
    uint8_t* alpha;           // device alpha arena
    unsigned int* ticket;     // one word, cleared by the launcher before every stage
    unsigned int* abort_word;
    int sm_count;
};

// undoes every plane's transforms (one launch per stage index) and writes the green channel into the alpha arena;
// adds the number of kernels launched to *launches
cudaError_t launch_vp8l_inverse(const Vp8lBatch& b, cudaStream_t s, int* launches);

}  // namespace webpb200
