// kernels.cuh -- launch interface of the sm_100a reconstruction kernels (kernels.cu).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "../../include/webpb200.h"
#include "entropy.h"
#include "vp8l.cuh"

namespace webpb200 {

// Scratch arena of the wavefront kernels.
//   words [0] ticket counter of the prediction wavefront, [1] of the loop filter, [2] watchdog abort
//   word (non-zero after a pass = a wavefront wait timed out), [3] ticket counter of the gradient ALPH
//   filter; then the EDGE RECORDS through which a macroblock row hands its bottom edge to the row
//   below (see kernels.cu, "hand-off"); all of this must be zero before a pass (`zero_bytes`).
// After it: `total_rows` uint32 row id -> image (row id = frame.row_base + macroblock row),
// `total_rows` uint32 ticket -> row id (the order rows are handed out in: any order in which row y of
// an image comes after its row y-1 is valid; across images, row-major is the fast one),
// `total_rows` uint32 row id -> first edge-record slot of that row, then one byte per macroblock
// (non_zero_dct).
//
// Edge-record slots: image i owns 2 * mbw_i consecutive slots starting at 2 * (sum of mbw of the
// images before it); row y of the image uses the mbw slots of parity y & 1 (row y + 2 may overwrite
// what row y wrote: by then row y + 1 has consumed it, because row y + 2 only gets to macroblock x
// after row y + 1 has finished macroblock x + 1).
struct ScratchLayout {
    uint64_t tickets_off, edge_pred_off, edge_lf_off, rowmap_off, order_off, recbase_off, mbinfo_off, total;
    uint64_t zero_bytes;  // prefix that must be cleared before every pass
};
constexpr uint64_t kEdgePredBytes = 64;   // per slot: 8 x {4 pixels, tag}: bottom luma row, bottom U row, bottom V row
constexpr uint64_t kEdgeFilterBytes = 256;  // per slot: 32 x {4 pixels, tag}: bottom 4 luma rows, bottom 4 U rows, bottom 4 V rows
// edge_slots = 2 * sum of mbw over the frames; with_filter = false leaves the loop filter's records out
ScratchLayout scratch_layout(uint64_t total_rows, uint64_t total_mbs, uint64_t edge_slots, bool with_filter = true);
// fills rowmap / order / recbase (each total_rows entries) for frames [0, n); returns edge_slots
uint64_t edge_slots_of(const webpb200_frame_desc* frames, int n);

struct BatchDev {
    const webpb200_frame_desc* frames;  // device
    const webpb200_mb_desc* mbs;
    const int16_t* levels;
    int levels_packed;      // levels are in k_entropy's packed form (kernels.cu, "Level records"), not the dense boundary format
    const uint8_t* alpha;  // alpha residual planes (may be null)
    uint8_t* alpha_planar;  // un-filtered alpha planes, same offsets (written by k_alpha_*, read by k_yuv2rgb; may be null when no frame has ALPH)
    int merge_alpha;        // k_yuv2rgb puts the alpha planes into byte 3 of RGBA pixels (set by the pass that ran the ALPH stage)
    uint8_t* planes;
    uint8_t* out;
    uint8_t* scratch;
    ScratchLayout lay;
    int n_frames;           // frames covered by this launch ...
    int frame_first;        // ... starting at this index of `frames` (a chunk of the batch; 0 = all)
    uint32_t total_rows;    // macroblock rows (tickets) of those frames ...
    uint32_t order_first;   // ... whose tickets start at this index of the order array
    uint32_t max_w, max_h;  // largest frame of the batch (grid sizing)
    int sm_count;
    int predict_blocks, predict_blocks_latency, filter_blocks;  // resident blocks per SM of the wavefront kernels on this device (wavefront_blocks_per_sm)
    int q_first, q_end;     // k_yuv2rgb: row pairs [q_first, q_end) only; 0, 0 = every row
    // frames with an ALPH plane, grouped by filter method m: alpha_lists[m * n_frames + k]
    const int* alpha_lists;  // device
    int alpha_counts[4];
    uint32_t alpha_maxw[4], alpha_maxh[4];
    // gradient ALPH filter: the last row of every 32-row band is handed to the band below through
    // this array (uint16 = alpha | 0x100 once written; cleared before the launch).  The rows of
    // frame f start at entry f.alpha_off >> 4; alpha_handoff_size(end of the alpha arena) bytes.
    uint16_t* alpha_handoff;
    uint64_t alpha_handoff_bytes;
    // VP8L-compressed ALPH planes whose transforms are undone on the device before the un-filter (vp8l.cuh); n_planes = 0: none
    Vp8lBatch vp8l;
    uint64_t vp8l_hand_bytes;  // size of vp8l.handoff, cleared before every pass
};
inline uint64_t alpha_handoff_size(uint64_t alpha_bytes) { return ((alpha_bytes >> 3) + 511) & ~(uint64_t)255; }

cudaError_t upload_tables_once();  // prediction tables -> __constant__ (per device, call with it current)
// resident blocks per SM of k_predict / k_loopfilter on the current device (an occupancy query; once per context)
void wavefront_blocks_per_sm(int* predict_blocks, int* predict_blocks_latency, int* filter_blocks);

// dequant + IWHT/IDCT + intra prediction, row-per-warp macroblock wavefront
cudaError_t launch_predict(const BatchDev& b, cudaStream_t s);
// simple / normal loop filter, same wavefront
cudaError_t launch_loopfilter(const BatchDev& b, cudaStream_t s);
// fancy / simple upsampling + YUV -> RGB(A)
cudaError_t launch_yuv2rgb(const BatchDev& b, cudaStream_t s);
// ALPH un-filter + merge into byte 3 of the RGBA output; adds its launch count to *launches
cudaError_t launch_alpha(const BatchDev& b, cudaStream_t s, int* launches);

// device entropy stage (entropy.cu): macroblock headers + coefficient tokens of frames
// [b.frame_first, +b.n_frames) from the staged bitstreams into d_mbs / d_levels (both cleared by
// the caller); d_status[frame] is set to WEBPB200_ERR_BITSTREAM for a frame that ran out of bits
void upload_entropy_tables_once();
void set_entropy_faults(int flags);  // test hook, see webpb200_debug_entropy_faults
// same_probs: every frame of the launch has the same token probabilities (one shared-memory table per block);
// rows_done (may be null): host-mapped array, [frame * 8 + partition] = number of macroblock rows that
// token partition has finished (plain store after a system-scope fence), so the host can start
// reconstructing the top of the frames while the bottom is still being decoded;
// d_sync (may be null): 2 * (frame_first + n_frames) zeroed words for the split layout (single-partition batches: mode warps
// and token warps in separate blocks, entropy.cu), without it every frame gets one block of 1 + partitions warps
cudaError_t launch_entropy(const BatchDev& b, const Vp8EntropyHdr* d_hdrs, const uint8_t* d_bits, webpb200_mb_desc* d_mbs,
                           int16_t* d_levels, int* d_status, int max_parts, bool same_probs, unsigned int* rows_done, unsigned int* d_sync,
                           cudaStream_t s);

struct CompositeArgs {  // one composite_frame call, reference extended.rs:30-129
    uint8_t* canvas;       // canvas_w * canvas_h * 4, device
    const uint8_t* frame;  // frame pixels, device (3 or 4 bytes per pixel)
    uint8_t* out;          // canvas copy handed to the caller after this frame, device
    uint32_t canvas_w, canvas_h;
    uint32_t fx, fy, fw, fh;
    uint32_t px, py, pw, ph;  // previous frame rectangle
    uint8_t clear[4];
    int has_clear, frame_has_alpha, use_blend, out_bpp;
};
cudaError_t launch_composite(const CompositeArgs& a, cudaStream_t s);

// CPU evaluation of the B_PRED lookup table (for the table's unit test; no GPU needed):
// edge = {P, A0..A7, L0..L3}, out = 16 predicted pixels, mode = IntraMode 1..9
void bpred_table_eval(int mode, const uint8_t edge[13], uint8_t out[16]);

}  // namespace webpb200
