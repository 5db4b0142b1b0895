// static_scratch.cu -- the prediction stage for callers that are not allowed to allocate.
//
// Who this is for: a service that hands each component a fixed slice of device memory at start-up
// and forbids cudaMalloc afterwards (an allocation synchronises the whole device, stalls every other
// stream of the process and can fail late, when memory is full or fragmented), or a caller running
// inside another runtime's memory pool where a raw allocation would bypass that pool's accounting.
// Such a caller owns its input and output arrays already; what it cannot provide on the spot is the
// small working memory the macroblock wavefront needs (ticket counter, the records through which a
// macroblock row hands its bottom pixel row to the row below, the row -> image map, one flag per
// macroblock).  Here that working memory is a
// fixed array inside the library image itself, so the call below touches no allocator at all.
//
// Consequences the caller must know (also stated at the declaration in include/webpb200.h):
//   * the array exists once per GPU that loads the library and stays until the process exits;
//   * it bounds the batch: at most WEBPB200_STATIC_MAX_ROWS macroblock rows,
//     WEBPB200_STATIC_MAX_MBS macroblocks and WEBPB200_STATIC_MAX_WIDTH_MBS macroblock columns
//     summed over the frames per call, larger batches are refused;
//   * all calls on a GPU share it, so two calls that overlap on that GPU (different streams,
//     different host threads) corrupt each other -- run them one after another;
//   * the first call on each GPU looks the array's device address up (a one-time set-up);
//     make that first call before you begin any stream capture.
#include <cuda_runtime.h>

#include <cstring>
#include <mutex>

#include "../../include/webpb200.h"
#include "kernels.cuh"

namespace {

using namespace webpb200;

constexpr uint64_t kMaxRows = WEBPB200_STATIC_MAX_ROWS;
constexpr uint64_t kMaxMbs = WEBPB200_STATIC_MAX_MBS;
constexpr uint64_t kMaxEdgeSlots = 2 * (uint64_t)WEBPB200_STATIC_MAX_WIDTH_MBS;
constexpr uint64_t up256(uint64_t v) { return (v + 255) & ~uint64_t(255); }
// same arithmetic as scratch_layout(rows, mbs, slots, with_filter = false) in kernels.cu, evaluated for
// the largest batch accepted: head + prediction edge records + three per-row tables + per-macroblock flags
constexpr uint64_t kScratchBytes = up256(256 + kEdgePredBytes * kMaxEdgeSlots) + 3 * up256(4 * kMaxRows) + up256(kMaxMbs);

// The working memory.  In host code this name is only a handle for cudaGetSymbolAddress; the
// address kernels must use is the one that call returns.
static __device__ __align__(256) unsigned char g_wavefront_scratch[kScratchBytes];

struct DeviceSlot {
    bool ready = false;
    unsigned char* scratch = nullptr;  // device address of g_wavefront_scratch on this GPU
    int sm_count = 0;
    int predict_blocks = 1, predict_blocks_latency = 1;
};
DeviceSlot g_slots[64];
std::mutex g_slots_mu;

// row -> image map, ticket order and edge-record slots, written straight into the static array
// (device code may name it)
__global__ void k_static_rowmap(const webpb200_frame_desc* frames, int n, unsigned long long rowmap_off, unsigned long long order_off,
                                unsigned long long recbase_off) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t* rowmap = reinterpret_cast<uint32_t*>(g_wavefront_scratch + rowmap_off);
    uint32_t* order = reinterpret_cast<uint32_t*>(g_wavefront_scratch + order_off);
    uint32_t* recbase = reinterpret_cast<uint32_t*>(g_wavefront_scratch + recbase_off);
    uint32_t slot = 0;  // image i owns 2 * mbw_i slots after those of the images before it
    for (int j = 0; j < i; j++) slot += 2u * frames[j].mbw;
    const uint32_t base = frames[i].row_base, mbw = frames[i].mbw;
    for (uint32_t r = 0; r < frames[i].mbh; r++) {
        rowmap[base + r] = (uint32_t)i;
        order[base + r] = base + r;  // rows handed out image by image: valid, and needs no sort
        recbase[base + r] = slot + (r & 1) * mbw;
    }
}

// One-time set-up for the current device: resolve the array's device address.
int slot_for_current_device(DeviceSlot** out) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return WEBPB200_ERR_NO_DEVICE;
    if (dev < 0 || dev >= 64) return WEBPB200_ERR_INVALID_PARAM;
    std::lock_guard<std::mutex> lock(g_slots_mu);
    DeviceSlot& s = g_slots[dev];
    if (!s.ready) {
        void* p = nullptr;
        if (cudaGetSymbolAddress(&p, g_wavefront_scratch) != cudaSuccess || p == nullptr) {
            cudaGetLastError();
            return WEBPB200_ERR_CUDA;
        }
        if (cudaDeviceGetAttribute(&s.sm_count, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return WEBPB200_ERR_CUDA;
        // prediction table -> __constant__ memory of this device (a copy, not an allocation)
        if (upload_tables_once() != cudaSuccess) {
            cudaGetLastError();
            return WEBPB200_ERR_CUDA;
        }
        int filter_blocks = 0;
        wavefront_blocks_per_sm(&s.predict_blocks, &s.predict_blocks_latency, &filter_blocks);
        s.scratch = static_cast<unsigned char*>(p);
        s.ready = true;
    }
    *out = &s;
    return 0;
}

}  // namespace

// Dequantisation + inverse transforms + intra prediction (the work of webpb200_reconstruct_device
// with stages = 1) without any allocation; see the notes at the top of this file and the
// declaration in include/webpb200.h.  Returns 0 or a negative WEBPB200_ERR_* like the other
// device-resident entry points.
extern "C" int webpb200_predict_device_static(const webpb200_frame_desc* h_frames, const webpb200_frame_desc* d_frames, int n,
                                              const webpb200_mb_desc* d_mbs, const int16_t* d_levels, uint8_t* d_planes,
                                              uint8_t* d_nonzero, void* stream) {
    if (n <= 0) return 0;
    if (!h_frames || !d_frames || !d_mbs || !d_levels || !d_planes) return -WEBPB200_ERR_INVALID_PARAM;
    uint64_t rows = 0, mbs = 0, rows_present = 0, slots = 0;
    BatchDev bd;
    std::memset(&bd, 0, sizeof(bd));
    for (int i = 0; i < n; i++) {
        // a frame without macroblocks has nothing to predict and would make the kernel prefetch past the arrays
        if (h_frames[i].mbw == 0 || h_frames[i].mbh == 0) return -WEBPB200_ERR_INVALID_PARAM;
        const uint64_t m = (uint64_t)h_frames[i].mbw * h_frames[i].mbh;
        rows_present += h_frames[i].mbh;
        slots += 2 * (uint64_t)h_frames[i].mbw;
        if ((uint64_t)h_frames[i].row_base + h_frames[i].mbh > rows) rows = (uint64_t)h_frames[i].row_base + h_frames[i].mbh;
        if ((uint64_t)h_frames[i].mb_base + m > mbs) mbs = (uint64_t)h_frames[i].mb_base + m;
        if (h_frames[i].width > bd.max_w) bd.max_w = h_frames[i].width;
        if (h_frames[i].height > bd.max_h) bd.max_h = h_frames[i].height;
    }
    // the fixed array cannot grow: refuse what does not fit instead of overrunning it
    if (rows > kMaxRows || mbs > kMaxMbs || slots > kMaxEdgeSlots) return -WEBPB200_ERR_STATIC_SCRATCH_LIMIT;
    // row ids must be dense (row_base of each frame = sum of the earlier frames' mbh): a hole
    // would be a ticket that points at nothing
    if (rows_present != rows) return -WEBPB200_ERR_INVALID_PARAM;

    DeviceSlot* slot = nullptr;
    int st = slot_for_current_device(&slot);
    if (st) return -st;
    bd.lay = scratch_layout(rows, mbs, slots, false);
    if (bd.lay.total > kScratchBytes) return -WEBPB200_ERR_STATIC_SCRATCH_LIMIT;

    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    // the array still holds the previous call's contents: the ticket counter, the watchdog word and
    // every edge record must start at zero (the tables and the flags are rewritten in full)
    if (cudaMemsetAsync(slot->scratch, 0, (size_t)bd.lay.zero_bytes, s) != cudaSuccess) return -WEBPB200_ERR_CUDA;
    k_static_rowmap<<<(n + 127) / 128, 128, 0, s>>>(d_frames, n, (unsigned long long)bd.lay.rowmap_off, (unsigned long long)bd.lay.order_off,
                                                    (unsigned long long)bd.lay.recbase_off);
    if (cudaGetLastError() != cudaSuccess) return -WEBPB200_ERR_CUDA;

    bd.frames = d_frames;
    bd.mbs = d_mbs;
    bd.levels = d_levels;
    bd.planes = d_planes;
    bd.scratch = slot->scratch;
    bd.n_frames = n;
    bd.total_rows = (uint32_t)rows;
    bd.sm_count = slot->sm_count;
    bd.predict_blocks = slot->predict_blocks;
    bd.predict_blocks_latency = slot->predict_blocks_latency;
    if (launch_predict(bd, s) != cudaSuccess) return -WEBPB200_ERR_CUDA;
    if (d_nonzero && cudaMemcpyAsync(d_nonzero, slot->scratch + bd.lay.mbinfo_off, (size_t)mbs, cudaMemcpyDeviceToDevice, s) != cudaSuccess)
        return -WEBPB200_ERR_CUDA;
    return 0;
}

// The calls above only queue work.  This waits for `stream` and reports whether the wavefront of the
// last webpb200_predict_device_static call on the current GPU ran to the end: 0, or
// -WEBPB200_ERR_CUDA if the stream failed or a macroblock row gave up waiting for the row above (which
// happens when two calls overlapped on the GPU and raced on the shared array); d_planes and d_nonzero
// are then incomplete.  Allocates nothing (one 4-byte copy from the array).
extern "C" int webpb200_predict_device_static_check(void* stream) {
    DeviceSlot* slot = nullptr;
    int st = slot_for_current_device(&slot);
    if (st) return -st;
    if (cudaStreamSynchronize(reinterpret_cast<cudaStream_t>(stream)) != cudaSuccess) return -WEBPB200_ERR_CUDA;
    unsigned int word = 0;
    if (cudaMemcpy(&word, slot->scratch + 8, 4, cudaMemcpyDeviceToHost) != cudaSuccess) return -WEBPB200_ERR_CUDA;
    return word ? -WEBPB200_ERR_CUDA : 0;
}
