// api.cu -- C ABI of libwebpb200.so (include/webpb200.h): contexts, the batched lossy path
// (host entropy decode -> packed arrays -> device kernels -> RGB(A)), the WebPDecoder-shaped
// single-image surface and animation compositing.  There is no CPU pixel path in this library:
// every decode call needs a CUDA device.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <deque>
#include <condition_variable>
#include <functional>
#include <memory>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/webpb200.h"
#include "host_container.h"
#include "host_vp8.h"
#include "host_vp8l.h"
#include "kernels.cuh"

using namespace webpb200;

namespace {

// A small pool of parked worker threads: a decode_batch call runs three or four short parallel
// loops (container walk, probe, header parse), and creating 15 threads for each costs more than
// the loops themselves.  One user at a time; a second concurrent caller falls back to plain threads.
class WorkerPool {
public:
    static WorkerPool& get() {
        static WorkerPool* p = new WorkerPool();  // leaked on purpose: workers may outlive static destruction
        return *p;
    }
    // runs job() on `extra` pool threads and on the caller; returns false if the pool is in use
    bool run(int extra, const std::function<void()>& job) {
        if (busy_.exchange(true)) return false;
        {
            std::unique_lock<std::mutex> lk(mu_);
            while ((int)threads_.size() < extra) threads_.emplace_back([this, id = (int)threads_.size()] { loop(id); });
            job_ = &job;
            want_ = extra;
            pending_ = extra;
            ++gen_;
        }
        cv_.notify_all();
        job();
        {
            std::unique_lock<std::mutex> lk(mu_);
            done_.wait(lk, [this] { return pending_ == 0; });
            job_ = nullptr;
        }
        busy_.store(false);
        return true;
    }

private:
    void loop(int id) {
        uint64_t seen = 0;
        for (;;) {
            const std::function<void()>* job = nullptr;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [&] { return gen_ != seen; });
                seen = gen_;
                if (id < want_) job = job_;
            }
            if (!job) continue;
            (*job)();
            std::unique_lock<std::mutex> lk(mu_);
            if (--pending_ == 0) done_.notify_one();
        }
    }
    std::mutex mu_;
    std::condition_variable cv_, done_;
    std::vector<std::thread> threads_;
    const std::function<void()>* job_ = nullptr;
    uint64_t gen_ = 0;
    int want_ = 0, pending_ = 0;
    std::atomic<bool> busy_{false};
};

template <class F>
void parallel_for(int n, int threads, F fn) {
    if (threads < 1) threads = 1;
    if (threads > n) threads = n;
    if (threads <= 1) {
        for (int i = 0; i < n; i++) fn(i);
        return;
    }
    std::atomic<int> next{0};
    auto work = [&]() {
        for (;;) {
            const int i = next.fetch_add(1);
            if (i >= n) return;
            fn(i);
        }
    };
    if (threads - 1 <= 63 && WorkerPool::get().run(threads - 1, work)) return;
    std::vector<std::thread> pool;
    for (int t = 1; t < threads; t++) pool.emplace_back(work);
    work();
    for (auto& t : pool) t.join();
}

struct DevBuf {
    uint8_t* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        n = (n + (size_t(1) << 20) - 1) & ~((size_t(1) << 20) - 1);
        cudaError_t e = cudaMalloc(&p, n);
        if (e == cudaSuccess) cap = n;
        return e;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};
struct PinBuf {
    uint8_t* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
        n = (n + (size_t(1) << 20) - 1) & ~((size_t(1) << 20) - 1);
        cudaError_t e = cudaMallocHost(&p, n);
        if (e == cudaSuccess) cap = n;
        return e;
    }
    void release() {
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
    }
};

// One unit of reconstruction work: a VP8 chunk, optionally with the ALPH chunk that goes with it.
struct WorkItem {
    const uint8_t* vp8 = nullptr;
    size_t vp8_len = 0;
    const uint8_t* alph = nullptr;
    size_t alph_len = 0;
    int check_w = 0, check_h = 0;  // container's size; 0 = do not compare (decoder.rs:814-857)
    int alph_w = 0, alph_h = 0;    // size the ALPH plane is coded for
    int bpp = 3;
    int status = 0;
    Vp8Dims dims{0, 0, 0, 0};
};

inline uint64_t align256(uint64_t v) { return (v + 255) & ~uint64_t(255); }

static_assert(sizeof(webpb200_frame_desc) == 136 && sizeof(webpb200_mb_desc) == 16 && sizeof(webpb200_segment) == 16,
              "boundary structs changed size: update include/webpb200.h and the bindings");

}  // namespace

struct webpb200_ctx {
    int device = 0;
    int sm_count = 0;
    int predict_blocks = 1, predict_blocks_latency = 1, filter_blocks = 1;  // resident blocks per SM of the wavefront kernels on this device
    cudaStream_t stream = nullptr;
    cudaStream_t stream2 = nullptr;      // reconstruction of finished bands while k_entropy still runs
    cudaStream_t stream3 = nullptr;      // the bands' downloads: a queued copy (other contexts share the copy engine) must not hold up the next band's kernels
    cudaEvent_t band_ev[16] = {};        // band k's pixels are converted
    unsigned int* h_rows = nullptr;      // mapped pinned: [frame * 8 + partition] rows finished (written by k_entropy)
    unsigned int* d_rows = nullptr;
    size_t rows_cap = 0;
    cudaEvent_t ev[15] = {};
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;  // the ALPH stage runs beside k_predict / k_loopfilter on stream2
    std::string err;
    uint64_t launches = 0;
    std::atomic<bool> in_flight{false};  // a webpb200_decode_batch_begin has not been ended yet
    // AUTO placement of the entropy stage: microseconds per macroblock of one device chain and of one host thread.
    // Start values measured on B200 with a 16-core host; every call that runs a stage refines its figure (moving average).
    double dev_us_per_mb = 33.0, host_us_per_mb = 3.0;
    double last_chain_mbs = 0;           // longest chain x oversubscription of the last device-entropy call

    PinBuf h_frames, h_mbs, h_levels, h_alpha, h_aux, h_bits, h_ehdr, h_status, h_vp8l_px, h_vp8l_planes, h_vp8l_aux;
    uint64_t last_h2d = 0, last_d2h = 0;  // bytes the last decode_batch copied each way
    int last_entropy = WEBPB200_ENTROPY_HOST;  // where its entropy stage ran
    float last_entropy_ms = 0;  // device time of k_entropy in the last decode_batch (0 if it ran on the host)
    int entropy_where = WEBPB200_ENTROPY_AUTO;  // WEBPB200_ENTROPY_*: where decode_batch runs the boolean decode
    DevBuf d_frames, d_mbs, d_levels, d_alpha, d_alpha_out, d_aux, d_planes, d_out, d_scratch, d_flush, d_canvas, d_frames_out, d_handoff, d_bits, d_ehdr, d_status, d_vp8l_px, d_vp8l_planes, d_vp8l_aux, d_vp8l_hand;
    // VP8L-compressed ALPH planes of the current batch (entropy-decoded on the host, transforms undone on the device)
    std::vector<int> vp8l_of_item;          // item -> plane index, -1 = none
    std::vector<uint64_t> vp8l_aux_off;     // plane -> word offset of its transform data in the aux arena
    uint64_t vp8l_px_words = 0, vp8l_aux_words = 0, vp8l_hand_units = 0;
    int vp8l_n = 0;

    // current batch
    int n = 0;
    std::vector<WorkItem> items;
    uint64_t total_mbs = 0, total_rows = 0, total_px = 0;
    uint64_t plane_bytes = 0, out_bytes = 0, alpha_bytes = 0;
    size_t aux_rowmap_off = 0, aux_order_off = 0, aux_recbase_off = 0, aux_lists_off = 0, aux_bytes = 0;
    BatchDev bd{};
    bool uploaded = false;

    bool fail(cudaError_t e, const char* what) {
        if (e == cudaSuccess) return false;
        err = std::string(what) + ": " + cudaGetErrorString(e);
        return true;
    }
};

namespace {

#define CU(c, call, what)                      \
    do {                                       \
        if ((c)->fail((call), what)) return -WEBPB200_ERR_CUDA; \
    } while (0)

// Ticket order of the wavefront kernels: macroblock row 0 of every image, then row 1 of every
// image, ...  A row only ever waits for the row above it in the same image, which in this order
// was handed out a whole batch of rows earlier and is (nearly always) far ahead or finished, so
// resident warps work instead of waiting.  (Image-major order keeps 2/3 of the warps spinning.)
// Frames [i0, i1) only: their rows are contiguous in row-id space, so the order of a chunk of the
// batch can be written in place at order[row_base(i0) ...].
void fill_row_order(const webpb200_frame_desc* frames, int i0, int i1, uint32_t* order) {
    uint32_t max_h = 0;
    for (int i = i0; i < i1; i++) max_h = std::max<uint32_t>(max_h, frames[i].mbh);
    size_t k = i0 < i1 ? frames[i0].row_base : 0;
    for (uint32_t r = 0; r < max_h; r++)
        for (int i = i0; i < i1; i++)
            if (frames[i].mbh > r) order[k++] = frames[i].row_base + r;
}

// row id -> image, and row id -> first edge-record slot of that row (kernels.cuh: image i owns 2 * mbw_i
// slots, its row y uses the mbw_i slots of parity y & 1)
void fill_row_tables(const webpb200_frame_desc* frames, int n, uint32_t* rowmap, uint32_t* recbase) {
    uint32_t slot = 0;
    for (int i = 0; i < n; i++) {
        for (uint32_t r = 0; r < frames[i].mbh; r++) {
            rowmap[frames[i].row_base + r] = (uint32_t)i;
            recbase[frames[i].row_base + r] = slot + (r & 1) * frames[i].mbw;
        }
        slot += 2u * frames[i].mbw;
    }
}

// Host side of a batch, part 1: probe sizes, lay the arenas out, size the pinned staging, build
// the row -> image map.  Nothing here depends on the entropy-coded data.
int layout_items(webpb200_ctx* c, std::vector<WorkItem>&& in_items, int upsampling, int threads, bool host_staging = true) {
    cudaSetDevice(c->device);
    c->items = std::move(in_items);
    c->n = (int)c->items.size();
    c->uploaded = false;
    const int n = c->n;
    auto& items = c->items;

    parallel_for(n, threads, [&](int i) {
        WorkItem& it = items[i];
        if (it.status) return;
        it.status = vp8_probe(it.vp8, it.vp8_len, &it.dims);
        if (it.status) return;
        if (it.dims.width == 0 || it.dims.height == 0) {
            it.status = WEBPB200_ERR_INCONSISTENT_SIZE;  // no pixels to reconstruct
            return;
        }
        if (it.check_w && (it.dims.width != it.check_w || it.dims.height != it.check_h)) it.status = WEBPB200_ERR_INCONSISTENT_SIZE;
    });

    // arenas: dead items keep zero-sized slots
    CU(c, c->h_frames.reserve(sizeof(webpb200_frame_desc) * (size_t)std::max(n, 1)), "pinned frames");
    webpb200_frame_desc* frames = reinterpret_cast<webpb200_frame_desc*>(c->h_frames.p);
    std::memset(frames, 0, sizeof(webpb200_frame_desc) * (size_t)n);
    uint64_t mbs = 0, rows = 0, px = 0, poff = 0, ooff = 0, aoff = 0;
    uint32_t max_w = 0, max_h = 0;
    for (int i = 0; i < n; i++) {
        WorkItem& it = items[i];
        webpb200_frame_desc& fd = frames[i];
        fd.mb_base = (uint32_t)mbs;
        fd.row_base = (uint32_t)rows;
        fd.y_off = poff;
        fd.u_off = poff;
        fd.v_off = poff;
        fd.out_off = ooff;
        fd.alpha_off = aoff;
        fd.out_bpp = (uint8_t)it.bpp;
        fd.upsampling = (uint8_t)upsampling;
        if (it.status) continue;  // width/height/mbw/mbh stay 0: kernels see an empty frame
        const uint64_t m = (uint64_t)it.dims.mbw * it.dims.mbh;
        if (mbs + m > 0xffffffffull) { it.status = WEBPB200_ERR_IMAGE_TOO_LARGE; continue; }
        fd.width = (uint16_t)it.dims.width; fd.height = (uint16_t)it.dims.height;
        fd.mbw = (uint16_t)it.dims.mbw; fd.mbh = (uint16_t)it.dims.mbh;
        fd.y_off = poff;
        fd.u_off = align256(fd.y_off + m * 256);
        fd.v_off = align256(fd.u_off + m * 64);
        poff = align256(fd.v_off + m * 64);
        ooff = align256(ooff + (uint64_t)it.dims.width * it.dims.height * it.bpp);
        if (it.alph && it.bpp == 4) {
            fd.alpha_mode = 1;
            aoff = align256(aoff + (uint64_t)it.dims.width * it.dims.height);
        }
        mbs += m;
        rows += it.dims.mbh;
        px += (uint64_t)it.dims.width * it.dims.height;
        max_w = std::max<uint32_t>(max_w, it.dims.width);
        max_h = std::max<uint32_t>(max_h, it.dims.height);
    }
    c->total_mbs = mbs; c->total_rows = rows; c->total_px = px;
    c->plane_bytes = poff; c->out_bytes = ooff; c->alpha_bytes = aoff;

    if (host_staging) {
        CU(c, c->h_mbs.reserve(sizeof(webpb200_mb_desc) * (size_t)std::max<uint64_t>(mbs, 1)), "pinned mbs");
        CU(c, c->h_levels.reserve(sizeof(int16_t) * WEBPB200_LEVELS_PER_MB * (size_t)std::max<uint64_t>(mbs, 1)), "pinned levels");
    }
    CU(c, c->h_alpha.reserve((size_t)std::max<uint64_t>(aoff, 1)), "pinned alpha");
    // VP8L-compressed ALPH planes (the compression method is the low two bits of the chunk's first byte,
    // extended.rs:259-266): a slot each in the pixel, aux and hand-off arenas, sized for the worst case
    c->vp8l_of_item.assign((size_t)n, -1);
    c->vp8l_aux_off.clear();
    c->vp8l_n = 0;
    c->vp8l_px_words = c->vp8l_aux_words = c->vp8l_hand_units = 0;
    {
        std::vector<uint64_t> px_off, hand_off;
        for (int i = 0; i < n; i++) {
            const WorkItem& it = items[i];
            if (it.status || !frames[i].alpha_mode || !it.alph || it.alph_len < 1 || (it.alph[0] & 3) != 1) continue;
            const uint64_t w = frames[i].width, h = frames[i].height;
            c->vp8l_of_item[i] = c->vp8l_n++;
            px_off.push_back(c->vp8l_px_words);
            c->vp8l_px_words += (w * h + 1 + 63) & ~uint64_t(63);
            c->vp8l_aux_off.push_back(c->vp8l_aux_words);
            c->vp8l_aux_words += (2 * ((w + 3) / 4) * ((h + 3) / 4) + 256 + 63) & ~uint64_t(63);
            hand_off.push_back(c->vp8l_hand_units);
            c->vp8l_hand_units += 2 * w;
        }
        if (c->vp8l_n) {
            CU(c, c->h_vp8l_px.reserve(4 * (size_t)c->vp8l_px_words), "pinned VP8L pixels");
            CU(c, c->h_vp8l_aux.reserve(4 * (size_t)c->vp8l_aux_words), "pinned VP8L transform data");
            CU(c, c->h_vp8l_planes.reserve(sizeof(Vp8lPlane) * (size_t)c->vp8l_n), "pinned VP8L planes");
            Vp8lPlane* planes = reinterpret_cast<Vp8lPlane*>(c->h_vp8l_planes.p);
            std::memset(planes, 0, sizeof(Vp8lPlane) * (size_t)c->vp8l_n);
            for (int i = 0; i < n; i++) {
                const int k = c->vp8l_of_item[i];
                if (k < 0) continue;
                planes[k].width = frames[i].width;
                planes[k].height = frames[i].height;
                for (int b = 0; b < 3; b++) planes[k].buf_off[b] = (uint64_t)b * c->vp8l_px_words + px_off[k];
                planes[k].hand_off = hand_off[k];
                planes[k].alpha_off = frames[i].alpha_off;
            }
        }
    }
    c->aux_rowmap_off = 0;
    c->aux_order_off = (size_t)align256(4 * rows);
    c->aux_recbase_off = c->aux_order_off + (size_t)align256(4 * rows);
    c->aux_lists_off = c->aux_recbase_off + (size_t)align256(4 * rows);
    c->aux_bytes = c->aux_lists_off + (size_t)align256(16 * (uint64_t)std::max(n, 1));
    CU(c, c->h_aux.reserve(c->aux_bytes), "pinned aux");
    fill_row_tables(frames, n, reinterpret_cast<uint32_t*>(c->h_aux.p + c->aux_rowmap_off), reinterpret_cast<uint32_t*>(c->h_aux.p + c->aux_recbase_off));
    BatchDev& bd = c->bd;
    std::memset(&bd, 0, sizeof(bd));
    bd.n_frames = n;
    bd.total_rows = (uint32_t)rows;
    bd.max_w = max_w;
    bd.max_h = max_h;
    bd.sm_count = c->sm_count;
    bd.predict_blocks = c->predict_blocks;
    bd.predict_blocks_latency = c->predict_blocks_latency;
    bd.filter_blocks = c->filter_blocks;
    bd.lay = scratch_layout(rows, mbs, edge_slots_of(frames, n));
    return 0;
}

// ALPH chunk of item i into the pinned alpha arena (extended.rs:253-308); host work in both entropy modes
int alpha_one(webpb200_ctx* c, int i, webpb200_frame_desc& fd) {
    WorkItem& it = c->items[i];
    int filter = 0;
    bool compressed = false;
    int st = parse_alph_header(it.alph, it.alph_len, &filter, &compressed);
    uint8_t* dst = c->h_alpha.p + fd.alpha_off;
    const size_t want = (size_t)it.alph_w * it.alph_h, have = (size_t)fd.width * fd.height;
    if (!st && want != have) st = WEBPB200_ERR_INCONSISTENT_SIZE;
    if (!st) {
        fd.alpha_filter = (uint8_t)filter;
        const int k = i < (int)c->vp8l_of_item.size() ? c->vp8l_of_item[i] : -1;
        if (compressed && k >= 0) {
            // prefix codes / LZ77 here, the inverse transforms on the device (vp8l.cu): the plane's slot of the pixel
            // arena gets the image in the transformed domain, its plane record the transforms to undo
            Vp8lPlane& P = reinterpret_cast<Vp8lPlane*>(c->h_vp8l_planes.p)[k];
            uint32_t* px = reinterpret_cast<uint32_t*>(c->h_vp8l_px.p) + P.buf_off[0];
            uint32_t* aux = reinterpret_cast<uint32_t*>(c->h_vp8l_aux.p);
            Vp8lPlan plan;
            st = vp8l_decode_entropy(it.alph + 1, it.alph_len - 1, it.alph_w, it.alph_h, px, &plan);
            P.n_stages = 0;
            if (!st) {
                uint64_t at = c->vp8l_aux_off[k];
                for (int q = 0; q < plan.n_stages; q++) {
                    Vp8lStage& S = P.st[q];
                    S.kind = (uint32_t)plan.st[q].kind;
                    S.bits = (uint32_t)plan.st[q].bits;
                    S.table_size = (uint32_t)plan.st[q].table_size;
                    S.w_in = (uint32_t)plan.st[q].w_in;
                    S.data_off = at;
                    if (!plan.st[q].data.empty()) std::memcpy(aux + at, plan.st[q].data.data(), 4 * plan.st[q].data.size());
                    at += plan.st[q].data.size();
                }
                P.n_stages = (uint32_t)plan.n_stages;
            }
        } else if (compressed) st = vp8l_decode_green(it.alph + 1, it.alph_len - 1, it.alph_w, it.alph_h, dst);
        else if (it.alph_len - 1 < want) st = WEBPB200_ERR_IO;
        else std::memcpy(dst, it.alph + 1, want);
    }
    return st;
}

// Part 2, per image (any host thread): the serial entropy decode into the pinned staging arrays.
void entropy_one(webpb200_ctx* c, int i) {
    webpb200_frame_desc* frames = reinterpret_cast<webpb200_frame_desc*>(c->h_frames.p);
    webpb200_mb_desc* hm = reinterpret_cast<webpb200_mb_desc*>(c->h_mbs.p);
    int16_t* hl = reinterpret_cast<int16_t*>(c->h_levels.p);
    {
        WorkItem& it = c->items[i];
        if (it.status) return;
        webpb200_frame_desc& fd = frames[i];
        const size_t m = (size_t)fd.mbw * fd.mbh;
        int st = vp8_entropy_decode(it.vp8, it.vp8_len, &fd, hm + fd.mb_base, hl + (size_t)fd.mb_base * WEBPB200_LEVELS_PER_MB);
        if (!st && fd.alpha_mode) st = alpha_one(c, i, fd);
        if (st) {  // keep the slot harmless for the kernels: all-DC, all-zero macroblocks
            it.status = st;
            std::memset(hm + fd.mb_base, 0, sizeof(webpb200_mb_desc) * m);
            std::memset(hl + (size_t)fd.mb_base * WEBPB200_LEVELS_PER_MB, 0, sizeof(int16_t) * WEBPB200_LEVELS_PER_MB * m);
            fd.filter_level = 0;
            fd.alpha_mode = 0;
        }
    }
}

// Part 3, after every image is decoded: the ALPH lists by filter method (the filter method is in
// the ALPH chunk header, so it is only known now).
void finish_items(webpb200_ctx* c) {
    const webpb200_frame_desc* frames = reinterpret_cast<const webpb200_frame_desc*>(c->h_frames.p);
    int* lists = reinterpret_cast<int*>(c->h_aux.p + c->aux_lists_off);
    BatchDev& bd = c->bd;
    const int n = c->n;
    for (int m = 0; m < 4; m++) bd.alpha_counts[m] = 0, bd.alpha_maxw[m] = 0, bd.alpha_maxh[m] = 0;
    for (int i = 0; i < n; i++) {
        const webpb200_frame_desc& fd = frames[i];
        if (fd.alpha_mode && fd.out_bpp == 4) {
            const int m = fd.alpha_filter & 3;
            lists[m * n + bd.alpha_counts[m]++] = i;
            bd.alpha_maxw[m] = std::max<uint32_t>(bd.alpha_maxw[m], fd.width);
            bd.alpha_maxh[m] = std::max<uint32_t>(bd.alpha_maxh[m], fd.height);
        }
    }
}

// The three parts in one go, tickets ordered over the whole batch.
int prepare_items(webpb200_ctx* c, std::vector<WorkItem>&& in_items, int upsampling, int threads) {
    int rc = layout_items(c, std::move(in_items), upsampling, threads);
    if (rc) return rc;
    parallel_for(c->n, threads, [&](int i) { entropy_one(c, i); });
    fill_row_order(reinterpret_cast<const webpb200_frame_desc*>(c->h_frames.p), 0, c->n,
                   reinterpret_cast<uint32_t*>(c->h_aux.p + c->aux_order_off));
    finish_items(c);
    return 0;
}

int reserve_device(webpb200_ctx* c) {
    cudaSetDevice(c->device);
    const int n = c->n;
    CU(c, c->d_frames.reserve(sizeof(webpb200_frame_desc) * (size_t)std::max(n, 1)), "device frames");
    CU(c, c->d_mbs.reserve(sizeof(webpb200_mb_desc) * (size_t)std::max<uint64_t>(c->total_mbs, 1)), "device mbs");
    CU(c, c->d_levels.reserve(sizeof(int16_t) * WEBPB200_LEVELS_PER_MB * (size_t)std::max<uint64_t>(c->total_mbs, 1)), "device levels");
    CU(c, c->d_alpha.reserve((size_t)std::max<uint64_t>(c->alpha_bytes, 1)), "device alpha");
    CU(c, c->d_aux.reserve(c->aux_bytes), "device aux");
    CU(c, c->d_planes.reserve((size_t)std::max<uint64_t>(c->plane_bytes, 1)), "device planes");
    CU(c, c->d_out.reserve((size_t)std::max<uint64_t>(c->out_bytes, 1)), "device out");
    CU(c, c->d_scratch.reserve((size_t)c->bd.lay.total), "device scratch");
    if (c->alpha_bytes) CU(c, c->d_handoff.reserve((size_t)alpha_handoff_size(c->alpha_bytes)), "device alpha handoff");
    if (c->alpha_bytes) CU(c, c->d_alpha_out.reserve((size_t)c->alpha_bytes), "device alpha planes");
    if (c->vp8l_n) {
        CU(c, c->d_vp8l_px.reserve(3 * 4 * (size_t)c->vp8l_px_words), "device VP8L pixels");
        CU(c, c->d_vp8l_aux.reserve(4 * (size_t)std::max<uint64_t>(c->vp8l_aux_words, 1)), "device VP8L transform data");
        CU(c, c->d_vp8l_planes.reserve(sizeof(Vp8lPlane) * (size_t)c->vp8l_n), "device VP8L planes");
        CU(c, c->d_vp8l_hand.reserve(8 * (size_t)std::max<uint64_t>(c->vp8l_hand_units, 1)), "device VP8L hand-off");
    }
    BatchDev& bd = c->bd;
    bd.frames = reinterpret_cast<const webpb200_frame_desc*>(c->d_frames.p);
    bd.mbs = reinterpret_cast<const webpb200_mb_desc*>(c->d_mbs.p);
    bd.levels = reinterpret_cast<const int16_t*>(c->d_levels.p);
    bd.alpha = c->d_alpha.p;
    bd.planes = c->d_planes.p;
    bd.out = c->d_out.p;
    bd.scratch = c->d_scratch.p;
    bd.alpha_lists = reinterpret_cast<const int*>(c->d_aux.p + c->aux_lists_off);
    bd.alpha_planar = c->alpha_bytes ? c->d_alpha_out.p : nullptr;
    bd.merge_alpha = 0;
    bd.alpha_handoff = c->alpha_bytes ? reinterpret_cast<uint16_t*>(c->d_handoff.p) : nullptr;
    bd.alpha_handoff_bytes = c->alpha_bytes ? alpha_handoff_size(c->alpha_bytes) : 0;
    std::memset(&bd.vp8l, 0, sizeof(bd.vp8l));
    bd.vp8l_hand_bytes = 0;
    if (c->vp8l_n) {
        bd.vp8l.planes = reinterpret_cast<const Vp8lPlane*>(c->d_vp8l_planes.p);
        bd.vp8l.n_planes = c->vp8l_n;
        bd.vp8l.max_stages = 4;  // a stage index beyond a plane's own count is skipped by the kernel
        bd.vp8l.max_height = bd.max_h;
        bd.vp8l.pixels = reinterpret_cast<uint32_t*>(c->d_vp8l_px.p);
        bd.vp8l.aux = reinterpret_cast<const uint32_t*>(c->d_vp8l_aux.p);
        bd.vp8l.handoff = reinterpret_cast<uint2*>(c->d_vp8l_hand.p);
        bd.vp8l.alpha = c->d_alpha.p;
        bd.vp8l.ticket = reinterpret_cast<unsigned int*>(c->d_scratch.p) + 4;
        bd.vp8l.abort_word = reinterpret_cast<unsigned int*>(c->d_scratch.p) + 2;
        bd.vp8l.sm_count = c->sm_count;
        bd.vp8l_hand_bytes = 8 * c->vp8l_hand_units;
    }
    return 0;
}

// the VP8L planes [p0, p1) of the batch: transformed-domain pixels, plane records and transform data -> device
int upload_vp8l(webpb200_ctx* c, int p0, int p1, cudaStream_t s) {
    if (p1 <= p0) return 0;
    const Vp8lPlane* planes = reinterpret_cast<const Vp8lPlane*>(c->h_vp8l_planes.p);
    const uint64_t w0 = planes[p0].buf_off[0], w1 = p1 < c->vp8l_n ? planes[p1].buf_off[0] : c->vp8l_px_words;
    CU(c, cudaMemcpyAsync(c->d_vp8l_px.p + 4 * w0, c->h_vp8l_px.p + 4 * w0, 4 * (size_t)(w1 - w0), cudaMemcpyHostToDevice, s), "H2D VP8L pixels");
    const uint64_t a0 = c->vp8l_aux_off[p0], a1 = p1 < c->vp8l_n ? c->vp8l_aux_off[p1] : c->vp8l_aux_words;
    CU(c, cudaMemcpyAsync(c->d_vp8l_aux.p + 4 * a0, c->h_vp8l_aux.p + 4 * a0, 4 * (size_t)(a1 - a0), cudaMemcpyHostToDevice, s), "H2D VP8L transform data");
    CU(c, cudaMemcpyAsync(c->d_vp8l_planes.p + sizeof(Vp8lPlane) * (size_t)p0, c->h_vp8l_planes.p + sizeof(Vp8lPlane) * (size_t)p0,
                          sizeof(Vp8lPlane) * (size_t)(p1 - p0), cudaMemcpyHostToDevice, s), "H2D VP8L planes");
    c->last_h2d += 4 * (w1 - w0) + 4 * (a1 - a0);
    return 0;
}

// undo the VP8L transforms of the planes [p0, p1) (alpha stage, before the un-filter kernels)
int run_vp8l(webpb200_ctx* c, const BatchDev& bd, int p0, int p1, cudaStream_t s) {
    if (p1 <= p0 || !bd.vp8l.n_planes) return 0;
    Vp8lBatch v = bd.vp8l;
    v.planes += p0;
    v.n_planes = p1 - p0;
    CU(c, cudaMemsetAsync(v.handoff, 0, (size_t)bd.vp8l_hand_bytes, s), "clear VP8L hand-off");
    int l = 0;
    CU(c, launch_vp8l_inverse(v, s, &l), "k_vp8l");
    c->launches += (uint64_t)l;
    return 0;
}

// rowmap, ticket order and edge-slot table -> scratch arena (they live next to the counters they are used with)
int upload_row_tables(webpb200_ctx* c, const uint8_t* aux, cudaMemcpyKind kind, cudaStream_t s) {
    CU(c, cudaMemsetAsync(c->d_scratch.p, 0, 256, s), "clear watchdog");
    if (!c->total_rows) return 0;
    const size_t bytes = 4 * (size_t)c->total_rows;
    CU(c, cudaMemcpyAsync(c->d_scratch.p + c->bd.lay.rowmap_off, aux + c->aux_rowmap_off, bytes, kind, s), "rowmap");
    CU(c, cudaMemcpyAsync(c->d_scratch.p + c->bd.lay.order_off, aux + c->aux_order_off, bytes, kind, s), "row order");
    CU(c, cudaMemcpyAsync(c->d_scratch.p + c->bd.lay.recbase_off, aux + c->aux_recbase_off, bytes, kind, s), "edge slots");
    return 0;
}

int upload(webpb200_ctx* c) {
    int rc = reserve_device(c);
    if (rc) return rc;
    const int n = c->n;
    cudaStream_t s = c->stream;
    CU(c, cudaMemcpyAsync(c->d_frames.p, c->h_frames.p, sizeof(webpb200_frame_desc) * (size_t)n, cudaMemcpyHostToDevice, s), "H2D frames");
    if (c->total_mbs) {
        CU(c, cudaMemcpyAsync(c->d_mbs.p, c->h_mbs.p, sizeof(webpb200_mb_desc) * (size_t)c->total_mbs, cudaMemcpyHostToDevice, s), "H2D mbs");
        CU(c, cudaMemcpyAsync(c->d_levels.p, c->h_levels.p, sizeof(int16_t) * WEBPB200_LEVELS_PER_MB * (size_t)c->total_mbs, cudaMemcpyHostToDevice, s), "H2D levels");
    }
    if (c->alpha_bytes) CU(c, cudaMemcpyAsync(c->d_alpha.p, c->h_alpha.p, (size_t)c->alpha_bytes, cudaMemcpyHostToDevice, s), "H2D alpha");
    if (int r = upload_vp8l(c, 0, c->vp8l_n, s)) return r;
    CU(c, cudaMemcpyAsync(c->d_aux.p, c->h_aux.p, c->aux_bytes, cudaMemcpyHostToDevice, s), "H2D aux");
    if (int r = upload_row_tables(c, c->d_aux.p, cudaMemcpyDeviceToDevice, s)) return r;
    c->uploaded = true;
    return 0;
}

// events: optional array of 5 events recorded before/after each stage (timing runs)
int reconstruct(webpb200_ctx* c, const BatchDev& bd, unsigned stages, cudaStream_t s, cudaEvent_t* events) {
    if (bd.total_rows == 0) return 0;
    // ticket counters and edge records start at zero; the watchdog word (offset 8) is cleared once
    // per batch, not per pass, so that a trip in any pass is still there when the host looks.  The
    // clears belong to the pass: a timing run records its first event before them.
    if (events) cudaEventRecord(events[0], s);
    // The ALPH stage (VP8L inverse transforms, un-filter into the planar alpha arena) touches nothing the VP8 kernels
    // touch, and its gradient filter and predictor transform are wavefronts of their own: it runs beside k_predict and
    // k_loopfilter on a second stream and joins before the colour conversion, which puts the planes into byte 3 of
    // the RGBA pixels it writes (stage 8 without stage 4 prepares the planes only).
    const bool alpha_stage = (stages & 8) && bd.alpha != nullptr &&
                             (bd.vp8l.n_planes > 0 || bd.alpha_counts[1] + bd.alpha_counts[2] + bd.alpha_counts[3] > 0);
    if (alpha_stage) {
        if (!c->stream2) CU(c, cudaStreamCreateWithFlags(&c->stream2, cudaStreamNonBlocking), "second stream");
        cudaStream_t as = c->stream2;
        CU(c, cudaEventRecord(c->ev_fork, s), "fork");
        CU(c, cudaStreamWaitEvent(as, c->ev_fork, 0), "fork");
        if (events) cudaEventRecord(events[12], as);
        if (int r = run_vp8l(c, bd, 0, bd.vp8l.n_planes, as)) return r;
        int l = 0;
        CU(c, launch_alpha(bd, as, &l), "k_alpha");
        c->launches += (uint64_t)l;
        if (events) cudaEventRecord(events[13], as);
        CU(c, cudaEventRecord(c->ev_join, as), "join");
    } else if (events) {
        cudaEventRecord(events[12], s);
        cudaEventRecord(events[13], s);
    }
    CU(c, cudaMemsetAsync(bd.scratch, 0, 8, s), "clear tickets");
    CU(c, cudaMemsetAsync(bd.scratch + bd.lay.edge_pred_off, 0, (size_t)(bd.lay.zero_bytes - bd.lay.edge_pred_off), s), "clear edge records");
    if (stages & 1) { CU(c, launch_predict(bd, s), "k_predict"); c->launches++; }
    if (events) cudaEventRecord(events[1], s);
    if (stages & 2) { CU(c, launch_loopfilter(bd, s), "k_loopfilter"); c->launches++; }
    if (events) cudaEventRecord(events[2], s);
    if (alpha_stage) CU(c, cudaStreamWaitEvent(s, c->ev_join, 0), "join");
    if (events) cudaEventRecord(events[14], s);
    if (stages & 4) {
        BatchDev rgb = bd;
        rgb.merge_alpha = (stages & (8 | 16)) && bd.alpha != nullptr;  // 16: the planes were un-filtered by the caller
        CU(c, launch_yuv2rgb(rgb, s), "k_yuv2rgb");
        c->launches++;
    }
    if (events) { cudaEventRecord(events[3], s); cudaEventRecord(events[4], s); }
    return 0;
}

// after a stream sync: did a wavefront kernel trip its watchdog?
int check_abort(webpb200_ctx* c) {
    if (!c->uploaded || c->bd.total_rows == 0) return 0;
    unsigned int word = 0;
    CU(c, cudaMemcpy(&word, c->bd.scratch + c->bd.lay.tickets_off + 8, 4, cudaMemcpyDeviceToHost), "read abort word");
    if (word) {
        c->err = "wavefront watchdog: a row waited too long for the row above";
        return -WEBPB200_ERR_CUDA;
    }
    return 0;
}

// D2H of the output arena into caller buffers; neighbouring images are sent as one copy when
// both sides are contiguous
int download_range(webpb200_ctx* c, uint8_t* const* outs, const size_t* out_lens, int i0, int i1) {
    const webpb200_frame_desc* frames = reinterpret_cast<const webpb200_frame_desc*>(c->h_frames.p);
    const int n = i1;
    int i = i0;
    while (i < n) {
        if (c->items[i].status) { i++; continue; }
        const size_t len = (size_t)frames[i].width * frames[i].height * frames[i].out_bpp;
        if (out_lens[i] != len) { c->items[i].status = WEBPB200_ERR_IMAGE_TOO_LARGE; i++; continue; }  // decoder.rs:670-672
        uint64_t src = frames[i].out_off;
        uint8_t* dst = outs[i];
        size_t run = len;
        int j = i + 1;
        while (j < n && !c->items[j].status && frames[j].out_off == src + run && outs[j] == dst + run &&
               out_lens[j] == (size_t)frames[j].width * frames[j].height * frames[j].out_bpp) {
            run += out_lens[j];
            j++;
        }
        CU(c, cudaMemcpyAsync(dst, c->d_out.p + src, run, cudaMemcpyDeviceToHost, c->stream), "D2H out");
        c->last_d2h += run;
        i = j;
    }
    return 0;
}

int download(webpb200_ctx* c, uint8_t* const* outs, const size_t* out_lens) {
    cudaSetDevice(c->device);
    int rc = download_range(c, outs, out_lens, 0, c->n);
    if (rc) return rc;
    CU(c, cudaStreamSynchronize(c->stream), "sync");
    return check_abort(c);
}

int items_from_files(const uint8_t* const* files, const size_t* lens, int n, int threads, std::vector<WorkItem>* out) {
    out->assign((size_t)n, WorkItem());
    parallel_for(n, threads, [&](int i) {
        WorkItem& it = (*out)[i];
        Container ct;
        it.status = parse_container(files[i], lens[i], &ct);
        if (it.status) return;
        if (ct.animated) { it.status = WEBPB200_ERR_INVALID_PARAM; return; }  // frames go through decode_animation / read_frame
        if (!ct.chunks[CK_VP8].present) { it.status = ct.chunks[CK_VP8L].present ? WEBPB200_ERR_NOT_LOSSY : WEBPB200_ERR_CHUNK_MISSING; return; }
        const Range r = ct.chunks[CK_VP8];
        if (r.end > lens[i]) { it.status = WEBPB200_ERR_IO; return; }
        it.vp8 = files[i] + r.start;
        it.vp8_len = (size_t)r.size();
        it.check_w = (int)ct.width;
        it.check_h = (int)ct.height;
        it.bpp = ct.has_alpha ? 4 : 3;
        if (ct.has_alpha) {
            const Range a = ct.chunks[CK_ALPH];
            if (!a.present) { it.status = WEBPB200_ERR_CHUNK_MISSING; return; }  // decoder.rs:707-711
            if (a.end > lens[i]) { it.status = WEBPB200_ERR_IO; return; }
            it.alph = files[i] + a.start;
            it.alph_len = (size_t)a.size();
            it.alph_w = (int)ct.width;
            it.alph_h = (int)ct.height;
        }
    });
    return 0;
}

std::mutex g_default_mu;
webpb200_ctx* g_default_ctx = nullptr;

// pinned allocations handed out by webpb200_host_alloc: input files that live in one of them can
// be sent to the device where they lie, without a staging copy
std::mutex g_pinned_mu;
std::map<uintptr_t, size_t> g_pinned;

// [lo, hi) inside one pinned allocation?
bool in_one_pinned_block(const uint8_t* lo, const uint8_t* hi) {
    std::lock_guard<std::mutex> lock(g_pinned_mu);
    auto it = g_pinned.upper_bound(reinterpret_cast<uintptr_t>(lo));
    if (it == g_pinned.begin()) return false;
    --it;
    return reinterpret_cast<uintptr_t>(hi) <= it->first + it->second;
}

webpb200_ctx* default_ctx(int* status) {
    std::lock_guard<std::mutex> lock(g_default_mu);
    if (!g_default_ctx) {
        *status = webpb200_ctx_create(0, &g_default_ctx);
        if (*status) g_default_ctx = nullptr;
    }
    return g_default_ctx;
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// contexts
// ------------------------------------------------------------------------------------------------
extern "C" int webpb200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

extern "C" int webpb200_device_pci_bus_id(int device, char* buf, int len) {
    if (!buf || len < 13) return -WEBPB200_ERR_INVALID_PARAM;
    if (cudaDeviceGetPCIBusId(buf, len, device) != cudaSuccess) {
        cudaGetLastError();
        return -WEBPB200_ERR_NO_DEVICE;
    }
    return 0;
}

extern "C" int webpb200_ctx_create(int device, webpb200_ctx** out) {
    *out = nullptr;
    if (device < 0 || device >= webpb200_device_count()) return WEBPB200_ERR_NO_DEVICE;
    if (cudaSetDevice(device) != cudaSuccess) return WEBPB200_ERR_NO_DEVICE;
    webpb200_ctx* c = new webpb200_ctx();
    c->device = device;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess || cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete c;
        return WEBPB200_ERR_CUDA;
    }
    c->sm_count = prop.multiProcessorCount;
    for (auto& e : c->ev) cudaEventCreate(&e);
    cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming);
    if (upload_tables_once() != cudaSuccess) {
        delete c;
        return WEBPB200_ERR_CUDA;
    }
    upload_entropy_tables_once();
    wavefront_blocks_per_sm(&c->predict_blocks, &c->predict_blocks_latency, &c->filter_blocks);
    if (const char* e = std::getenv("WEBPB200_ENTROPY")) c->entropy_where = std::strcmp(e, "device") == 0 ? WEBPB200_ENTROPY_DEVICE : std::strcmp(e, "host") == 0 ? WEBPB200_ENTROPY_HOST : WEBPB200_ENTROPY_AUTO;
    if (cudaDeviceSynchronize() != cudaSuccess) {
        delete c;
        return WEBPB200_ERR_CUDA;
    }
    *out = c;
    return 0;
}

extern "C" void webpb200_ctx_destroy(webpb200_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    if (c->stream2) cudaStreamSynchronize(c->stream2);
    if (c->stream3) cudaStreamSynchronize(c->stream3);
    for (PinBuf* b : {&c->h_frames, &c->h_mbs, &c->h_levels, &c->h_alpha, &c->h_aux, &c->h_bits, &c->h_ehdr, &c->h_status, &c->h_vp8l_px, &c->h_vp8l_planes, &c->h_vp8l_aux}) b->release();
    for (DevBuf* b : {&c->d_frames, &c->d_mbs, &c->d_levels, &c->d_alpha, &c->d_alpha_out, &c->d_aux, &c->d_planes, &c->d_out, &c->d_scratch, &c->d_flush, &c->d_canvas, &c->d_frames_out, &c->d_handoff, &c->d_bits, &c->d_ehdr, &c->d_status, &c->d_vp8l_px, &c->d_vp8l_planes, &c->d_vp8l_aux, &c->d_vp8l_hand})
        b->release();
    for (auto& e : c->ev) cudaEventDestroy(e);
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    if (c->ev_join) cudaEventDestroy(c->ev_join);
    cudaStreamDestroy(c->stream);
    if (c->stream2) cudaStreamDestroy(c->stream2);
    if (c->stream3) cudaStreamDestroy(c->stream3);
    for (auto& e : c->band_ev)
        if (e) cudaEventDestroy(e);
    if (c->h_rows) cudaFreeHost(c->h_rows);
    delete c;
}

extern "C" void* webpb200_host_alloc(size_t bytes) {
    void* p = nullptr;
    if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    std::lock_guard<std::mutex> lock(g_pinned_mu);
    g_pinned[reinterpret_cast<uintptr_t>(p)] = bytes ? bytes : 1;
    return p;
}
extern "C" void webpb200_host_free(void* p) {
    if (!p) return;
    {
        std::lock_guard<std::mutex> lock(g_pinned_mu);
        g_pinned.erase(reinterpret_cast<uintptr_t>(p));
    }
    cudaFreeHost(p);
}

extern "C" int webpb200_ctx_set_entropy(webpb200_ctx* c, int where) {
    if (!c || where < WEBPB200_ENTROPY_HOST || where > WEBPB200_ENTROPY_AUTO) return -WEBPB200_ERR_INVALID_PARAM;
    c->entropy_where = where;
    return 0;
}

extern "C" int webpb200_batch_output_device(webpb200_ctx* c, int image, void** d_ptr, size_t* len, uint32_t* width, uint32_t* height, int* bpp) {
    if (!c || !c->uploaded || image < 0 || image >= c->n) return -WEBPB200_ERR_INVALID_PARAM;
    if (c->items[image].status) return c->items[image].status;
    const webpb200_frame_desc& fd = reinterpret_cast<const webpb200_frame_desc*>(c->h_frames.p)[image];
    if (d_ptr) *d_ptr = c->d_out.p + fd.out_off;
    if (len) *len = (size_t)fd.width * fd.height * fd.out_bpp;
    if (width) *width = fd.width;
    if (height) *height = fd.height;
    if (bpp) *bpp = fd.out_bpp;
    return 0;
}

extern "C" int webpb200_last_transfer(const webpb200_ctx* c, uint64_t* h2d_bytes, uint64_t* d2h_bytes, int* entropy_where) {
    if (!c) return -WEBPB200_ERR_INVALID_PARAM;
    if (h2d_bytes) *h2d_bytes = c->last_h2d;
    if (d2h_bytes) *d2h_bytes = c->last_d2h;
    if (entropy_where) *entropy_where = c->last_entropy;
    return 0;
}

extern "C" float webpb200_last_entropy_ms(const webpb200_ctx* c) { return c ? c->last_entropy_ms : 0.f; }

extern "C" int webpb200_batch_download_packed(webpb200_ctx* c, int image, webpb200_mb_desc* mbs, int16_t* levels) {
    if (!c || !c->uploaded || image < 0 || image >= c->n) return -WEBPB200_ERR_INVALID_PARAM;
    cudaSetDevice(c->device);
    const webpb200_frame_desc& fd = reinterpret_cast<const webpb200_frame_desc*>(c->h_frames.p)[image];
    const size_t m = (size_t)fd.mbw * fd.mbh;
    CU(c, cudaStreamSynchronize(c->stream), "sync");
    if (m == 0) return 0;
    if (mbs) CU(c, cudaMemcpy(mbs, c->d_mbs.p + sizeof(webpb200_mb_desc) * (size_t)fd.mb_base, sizeof(webpb200_mb_desc) * m, cudaMemcpyDeviceToHost), "D2H mbs");
    if (levels) {
        CU(c, cudaMemcpy(levels, c->d_levels.p + sizeof(int16_t) * WEBPB200_LEVELS_PER_MB * (size_t)fd.mb_base,
                         sizeof(int16_t) * WEBPB200_LEVELS_PER_MB * m, cudaMemcpyDeviceToHost), "D2H levels");
        if (c->bd.levels_packed) {
            // k_entropy's packed records -> the dense boundary format the caller expects: slots in decode order
            // (Y2 first, then blocks 0..23) for the blocks whose nflags bit is set, everything else zero
            std::vector<webpb200_mb_desc> tmp;
            const webpb200_mb_desc* d = mbs;
            if (!d) {
                tmp.resize(m);
                CU(c, cudaMemcpy(tmp.data(), c->d_mbs.p + sizeof(webpb200_mb_desc) * (size_t)fd.mb_base, sizeof(webpb200_mb_desc) * m, cudaMemcpyDeviceToHost), "D2H mbs");
                d = tmp.data();
            }
            int16_t packed[WEBPB200_LEVELS_PER_MB];
            for (size_t i = 0; i < m; i++) {
                int16_t* rec = levels + i * WEBPB200_LEVELS_PER_MB;
                std::memcpy(packed, rec, sizeof(packed));
                std::memset(rec, 0, sizeof(packed));
                int slot = 0;
                for (int k = -1; k < 24; k++) {
                    const int blk = k < 0 ? 24 : k;
                    if ((d[i].nflags >> blk) & 1u) std::memcpy(rec + blk * 16, packed + 16 * slot++, 32);
                }
            }
        }
    }
    return 0;
}

extern "C" const char* webpb200_last_cuda_error(const webpb200_ctx* c) { return c ? c->err.c_str() : ""; }
extern "C" uint64_t webpb200_launch_count(const webpb200_ctx* c) { return c->launches; }

// ------------------------------------------------------------------------------------------------
// batched path
// ------------------------------------------------------------------------------------------------
extern "C" int webpb200_batch_prepare(webpb200_ctx* c, const uint8_t* const* files, const size_t* lens, int n, int* status,
                                      int upsampling, int host_threads) {
    if (!c || n < 0) return -WEBPB200_ERR_INVALID_PARAM;
    std::vector<WorkItem> items;
    items_from_files(files, lens, n, host_threads, &items);
    int rc = prepare_items(c, std::move(items), upsampling, host_threads);
    if (rc) return rc;
    int bad = 0;
    for (int i = 0; i < n; i++) {
        if (status) status[i] = c->items[i].status;
        bad += c->items[i].status != 0;
    }
    return bad;
}

extern "C" int webpb200_batch_upload(webpb200_ctx* c) { return upload(c); }

extern "C" int webpb200_batch_reconstruct(webpb200_ctx* c, unsigned stages) {
    if (!c->uploaded) return -WEBPB200_ERR_INVALID_PARAM;
    cudaSetDevice(c->device);
    return reconstruct(c, c->bd, stages, c->stream, nullptr);
}

extern "C" int webpb200_batch_sync(webpb200_ctx* c) {
    cudaSetDevice(c->device);
    CU(c, cudaStreamSynchronize(c->stream), "sync");
    return check_abort(c);
}

extern "C" int webpb200_batch_download(webpb200_ctx* c, uint8_t* const* outs, const size_t* out_lens) { return download(c, outs, out_lens); }

extern "C" int webpb200_batch_time(webpb200_ctx* c, unsigned stages, int iters, int flush_l2, float* ms_per_pass, float* kernel_ms) {
    if (!c->uploaded || iters < 1) return -WEBPB200_ERR_INVALID_PARAM;
    cudaSetDevice(c->device);
    const size_t flush_bytes = size_t(256) << 20;  // > 126 MB of L2
    if (flush_l2) CU(c, c->d_flush.reserve(flush_bytes), "flush buffer");
    double total = 0, per[4] = {0, 0, 0, 0};
    for (int it = 0; it < iters; it++) {
        if (flush_l2) CU(c, cudaMemsetAsync(c->d_flush.p, it & 0xff, flush_bytes, c->stream), "flush");
        int rc = reconstruct(c, c->bd, stages, c->stream, c->ev);
        if (rc) return rc;
        CU(c, cudaEventSynchronize(c->ev[4]), "event sync");
        float ms = 0;
        cudaEventElapsedTime(&ms, c->ev[0], c->ev[4]);
        total += ms;
        // events: 0 start, 1 after k_predict, 2 after k_loopfilter, 14 .. 3 = 4 k_yuv2rgb (after the ALPH stage has joined),
        // 12 .. 13 the ALPH stage on its own stream (beside the first two kernels: the four figures need not add up to the pass)
        static const int from[4] = {0, 1, 14, 12}, to[4] = {1, 2, 3, 13};
        for (int k = 0; k < 4; k++) {
            cudaEventElapsedTime(&ms, c->ev[from[k]], c->ev[to[k]]);
            per[k] += ms;
        }
    }
    CU(c, cudaStreamSynchronize(c->stream), "sync");
    if (int rc = check_abort(c)) return rc;
    if (ms_per_pass) *ms_per_pass = (float)(total / iters);
    if (kernel_ms)
        for (int k = 0; k < 4; k++) kernel_ms[k] = (float)(per[k] / iters);
    return 0;
}

extern "C" int webpb200_batch_size(const webpb200_ctx* c, int* n_images, uint64_t* n_macroblocks, uint64_t* n_pixels, uint64_t* in_bytes,
                                   uint64_t* out_bytes) {
    if (n_images) *n_images = c->n;
    if (n_macroblocks) *n_macroblocks = c->total_mbs;
    if (n_pixels) *n_pixels = c->total_px;
    if (in_bytes)
        *in_bytes = sizeof(webpb200_frame_desc) * (uint64_t)c->n + c->total_mbs * (sizeof(webpb200_mb_desc) + 2 * WEBPB200_LEVELS_PER_MB) +
                    c->alpha_bytes + c->aux_bytes;
    if (out_bytes) {
        uint64_t o = 0;
        const webpb200_frame_desc* f = reinterpret_cast<const webpb200_frame_desc*>(c->h_frames.p);
        for (int i = 0; i < c->n; i++) o += (uint64_t)f[i].width * f[i].height * f[i].out_bpp;
        *out_bytes = o;
    }
    return 0;
}

extern "C" int webpb200_batch_get_planes(webpb200_ctx* c, int image, uint8_t* y, uint8_t* u, uint8_t* v) {
    if (!c->uploaded || image < 0 || image >= c->n) return -WEBPB200_ERR_INVALID_PARAM;
    cudaSetDevice(c->device);
    const webpb200_frame_desc& fd = reinterpret_cast<const webpb200_frame_desc*>(c->h_frames.p)[image];
    const size_t m = (size_t)fd.mbw * fd.mbh;
    CU(c, cudaStreamSynchronize(c->stream), "sync");
    if (m == 0) return 0;
    CU(c, cudaMemcpy(y, c->d_planes.p + fd.y_off, m * 256, cudaMemcpyDeviceToHost), "D2H y");
    CU(c, cudaMemcpy(u, c->d_planes.p + fd.u_off, m * 64, cudaMemcpyDeviceToHost), "D2H u");
    CU(c, cudaMemcpy(v, c->d_planes.p + fd.v_off, m * 64, cudaMemcpyDeviceToHost), "D2H v");
    return 0;
}

extern "C" int webpb200_batch_get_packed(webpb200_ctx* c, int image, const webpb200_frame_desc** frame, const webpb200_mb_desc** mbs,
                                         const int16_t** levels) {
    if (image < 0 || image >= c->n) return -WEBPB200_ERR_INVALID_PARAM;
    const webpb200_frame_desc* f = reinterpret_cast<const webpb200_frame_desc*>(c->h_frames.p) + image;
    if (frame) *frame = f;
    if (mbs) *mbs = reinterpret_cast<const webpb200_mb_desc*>(c->h_mbs.p) + f->mb_base;
    if (levels) *levels = reinterpret_cast<const int16_t*>(c->h_levels.p) + (size_t)f->mb_base * WEBPB200_LEVELS_PER_MB;
    return 0;
}

namespace {
constexpr int kPreferHostEntropy = 1 << 30;  // internal: the cost model of the AUTO mode chose the host

// decode_batch with the entropy stage on the device: the host reads the frame headers and stages
// the compressed VP8 chunks; the boolean decode of macroblock headers and tokens runs in k_entropy,
// one warp per partition, all frames of the batch side by side; then the reconstruction kernels as
// before.  The 800 bytes per macroblock of levels are produced in HBM and never cross PCIe.
int decode_batch_device_entropy(webpb200_ctx* c, const std::vector<WorkItem>& items_in, uint8_t* const* outs, const size_t* out_lens,
                                int* status, int upsampling, int host_threads) {
    const bool trace = std::getenv("WEBPB200_TRACE") != nullptr;
    const auto t0 = std::chrono::steady_clock::now();
    auto ms_since = [&]() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); };
    std::vector<WorkItem> items = items_in;  // the caller keeps its list: the cost model may hand the batch to the host path
    const int n = (int)items.size();
    if (c->stream2) cudaStreamSynchronize(c->stream2);  // nothing of an earlier call that ended early may still be in flight
    if (c->stream3) cudaStreamSynchronize(c->stream3);
    const double t_items = ms_since();
    cudaStream_t s = c->stream;
    int rc = 0;

    // Bitstream arena.  If the caller's files already lie in one pinned block (webpb200_host_alloc)
    // and are reasonably dense there, that stretch of the block goes to the device as it is, and at
    // once: the upload then runs under the layout and header work below.  Otherwise the VP8 chunks
    // are staged back to back (16-byte aligned starts) through pinned memory further down.
    std::vector<uint64_t> bit_off((size_t)n + 1, 0);
    for (int i = 0; i < n; i++) bit_off[i + 1] = bit_off[i] + (items[i].status ? 0 : ((uint64_t)items[i].vp8_len + 15) & ~uint64_t(15));
    uint64_t bits_bytes = std::max<uint64_t>(bit_off[n], 16);
    const uint8_t* direct_base = nullptr;
    {
        const uint8_t *lo = nullptr, *hi = nullptr;
        for (int i = 0; i < n; i++) {
            const WorkItem& it = items[i];
            if (it.status) continue;
            if (!lo || it.vp8 < lo) lo = it.vp8;
            if (!hi || it.vp8 + it.vp8_len > hi) hi = it.vp8 + it.vp8_len;
        }
        if (lo && (uint64_t)(hi - lo) <= 2 * bit_off[n] && in_one_pinned_block(lo, hi)) {
            direct_base = lo;
            bits_bytes = (uint64_t)(hi - lo);
            for (int i = 0; i < n; i++) bit_off[i] = items[i].status ? 0 : (uint64_t)(items[i].vp8 - lo);
            cudaSetDevice(c->device);
            CU(c, c->d_bits.reserve((size_t)bits_bytes + 16), "device bitstreams");
            CU(c, cudaMemcpyAsync(c->d_bits.p, direct_base, (size_t)bits_bytes, cudaMemcpyHostToDevice, s), "H2D bitstreams");
        }
    }
    rc = layout_items(c, std::move(items), upsampling, host_threads, false);
    if (rc) return rc;
    const double t_layout = ms_since();
    webpb200_frame_desc* frames = reinterpret_cast<webpb200_frame_desc*>(c->h_frames.p);
    if (!direct_base) CU(c, c->h_bits.reserve((size_t)bits_bytes), "pinned bitstreams");
    CU(c, c->h_ehdr.reserve(sizeof(Vp8EntropyHdr) * (size_t)std::max(n, 1)), "pinned entropy headers");
    CU(c, c->h_status.reserve(sizeof(int) * (size_t)std::max(n, 1)), "pinned status");
    if ((rc = reserve_device(c)) != 0) return rc;
    if (!direct_base) CU(c, c->d_bits.reserve((size_t)bits_bytes + 16), "device bitstreams");
    CU(c, c->d_ehdr.reserve(sizeof(Vp8EntropyHdr) * (size_t)std::max(n, 1)), "device entropy headers");
    CU(c, c->d_status.reserve(sizeof(int) * 3 * (size_t)std::max(n, 1)), "device status");  // status[n] | k_entropy's sync words [2n]
    Vp8EntropyHdr* hdrs = reinterpret_cast<Vp8EntropyHdr*>(c->h_ehdr.p);

    // clear the outputs of the entropy stage while the host parses headers
    if (trace) cudaEventRecord(c->ev[8], s);
    if (c->total_mbs) {
        // descriptors only: the levels are written packed (a block that carries tokens clears its own slot,
        // the others are never touched), so the 800 bytes per macroblock need no clearing
        CU(c, cudaMemsetAsync(c->d_mbs.p, 0, sizeof(webpb200_mb_desc) * (size_t)c->total_mbs, s), "clear mbs");
    }
    CU(c, cudaMemsetAsync(c->d_status.p, 0, sizeof(int) * 3 * (size_t)std::max(n, 1), s), "clear status");
    if (trace) cudaEventRecord(c->ev[9], s);

    auto drop = [&](int i, int st) {  // the frame stays in the batch as all-DC, all-zero macroblocks nobody looks at
        c->items[i].status = st;
        hdrs[i].num_parts = 0;
        frames[i].filter_level = 0;
        frames[i].alpha_mode = 0;
    };
    // Host threads take the frames in order: header parse, then the compressed chunk into the pinned
    // arena.  This thread follows them and sends each finished slice of the arena on its way, so the
    // upload hides behind the staging.  (In AUTO mode the slices may turn out to be sent for nothing.)
    {
        const int nslices = std::max(1, std::min(16, n / 32));
        std::vector<int> slice_end((size_t)nslices);
        for (int k = 0; k < nslices; k++) slice_end[k] = (int)((int64_t)n * (k + 1) / nslices);
        std::unique_ptr<std::atomic<int>[]> left(new std::atomic<int>[(size_t)nslices]);
        for (int k = 0; k < nslices; k++) left[k].store(slice_end[k] - (k ? slice_end[k - 1] : 0));
        std::vector<int> slice_of((size_t)n);
        for (int k = 0, i = 0; k < nslices; k++)
            for (; i < slice_end[k]; i++) slice_of[i] = k;
        std::atomic<int> next{0};
        auto worker = [&]() {
            for (;;) {
                const int i = next.fetch_add(1);
                if (i >= n) return;
                WorkItem& it = c->items[i];
                Vp8EntropyHdr& eh = hdrs[i];
                std::memset(&eh, 0, sizeof(eh));
                if (!it.status) {
                    const int st = vp8_parse_header(it.vp8, it.vp8_len, &frames[i], &eh);
                    if (st) drop(i, st);
                    else if (!direct_base) std::memcpy(c->h_bits.p + bit_off[i], it.vp8, it.vp8_len);
                    eh.data_off = bit_off[i];
                }
                left[slice_of[i]].fetch_sub(1, std::memory_order_release);
            }
        };
        std::vector<std::thread> pool;
        if (direct_base) {  // nothing to upload behind the workers: this thread works too (parked pool threads, no thread creation)
            parallel_for(std::max(1, std::min(host_threads, n)), host_threads, [&](int) { worker(); });
        } else {
            for (int t = 0; t < std::max(1, std::min(host_threads, n)); t++) pool.emplace_back(worker);
        }
        cudaError_t e = cudaSuccess;
        for (int k = 0; k < nslices; k++) {
            while (left[k].load(std::memory_order_acquire) > 0) std::this_thread::yield();
            if (direct_base) continue;
            const uint64_t b0 = bit_off[k ? slice_end[k - 1] : 0], b1 = bit_off[slice_end[k]];
            if (b1 > b0 && e == cudaSuccess) e = cudaMemcpyAsync(c->d_bits.p + b0, c->h_bits.p + b0, (size_t)(b1 - b0), cudaMemcpyHostToDevice, s);
        }
        for (auto& t : pool) t.join();
        CU(c, e, "H2D bitstreams");
    }
    const double t_headers = ms_since();
    int max_parts = 1;
    for (int i = 0; i < n; i++) max_parts = std::max<int>(max_parts, hdrs[i].num_parts);
    bool same_probs = true;  // no frame updates its token probabilities differently from the others (usually: none does at all)
    {
        int first = -1;
        for (int i = 0; i < n && same_probs; i++) {
            if (!hdrs[i].num_parts) continue;
            if (first < 0) first = i;
            else same_probs = std::memcmp(hdrs[i].probs, hdrs[first].probs, sizeof(hdrs[i].probs)) == 0;
        }
    }
    if (c->entropy_where == WEBPB200_ENTROPY_AUTO) {
        // The device decodes every partition of every frame at once but each one slowly (one lane,
        // about 33 us per macroblock on B200); the host decodes fast (about 3 us per macroblock)
        // but only one frame per thread.  Many small frames: device.  A few large ones: host.
        // The two rates are per-context moving averages of what the last calls measured.
        double longest = 0, streams = 0;
        for (int i = 0; i < n; i++)
            if (hdrs[i].num_parts) {
                longest = std::max(longest, (double)frames[i].mbw * frames[i].mbh / hdrs[i].num_parts);
                streams += hdrs[i].num_parts;
            }
        // ALPH planes are decoded by the same host threads: beside k_entropy when the device decodes the VP8 tokens, on
        // top of the VP8 work when the host does (6 ns per pixel of a VP8L-compressed plane on one core of the boxes measured)
        const int workers = std::max(1, std::min(host_threads, n));
        double vp8l_px = 0;  // (raw ALPH planes cost the host a copy; the compression method is the chunk's first byte, extended.rs:266-283)
        for (int i = 0; i < n; i++)
            if (!c->items[i].status && c->items[i].alph && c->items[i].alph_len > 0 && (c->items[i].alph[0] & 3) == 1)
                vp8l_px += (double)frames[i].width * frames[i].height;
        const double est_alph_us = vp8l_px * 0.006 / workers;
        const double est_device_us = std::max(longest * c->dev_us_per_mb * std::max(1.0, streams / 1184.0), est_alph_us);
        const double est_host_us = (double)c->total_mbs * c->host_us_per_mb / workers + est_alph_us;
        if (est_device_us > est_host_us) return kPreferHostEntropy;
        c->last_chain_mbs = longest * std::max(1.0, streams / 1184.0);
    } else {
        double longest = 0, streams = 0;
        for (int i = 0; i < n; i++)
            if (hdrs[i].num_parts) {
                longest = std::max(longest, (double)frames[i].mbw * frames[i].mbh / hdrs[i].num_parts);
                streams += hdrs[i].num_parts;
            }
        c->last_chain_mbs = longest * std::max(1.0, streams / 1184.0);
    }
    fill_row_order(frames, 0, n, reinterpret_cast<uint32_t*>(c->h_aux.p + c->aux_order_off));
    const double t_host = ms_since();
    if (trace) std::fprintf(stderr, "  host: containers %.2f, layout %.2f, headers %.2f, staging %.2f ms\n", t_items, t_layout, t_headers, t_host);

    if (trace) cudaEventRecord(c->ev[10], s);
    CU(c, cudaMemcpyAsync(c->d_frames.p, frames, sizeof(webpb200_frame_desc) * (size_t)n, cudaMemcpyHostToDevice, s), "H2D frames");
    CU(c, cudaMemcpyAsync(c->d_ehdr.p, hdrs, sizeof(Vp8EntropyHdr) * (size_t)n, cudaMemcpyHostToDevice, s), "H2D entropy headers");
    if (int r2 = upload_row_tables(c, c->h_aux.p, cudaMemcpyHostToDevice, s)) return r2;
    c->uploaded = true;
    c->bd.levels_packed = 1;
    c->last_entropy = WEBPB200_ENTROPY_DEVICE;
    c->last_h2d = sizeof(webpb200_frame_desc) * (uint64_t)n + sizeof(Vp8EntropyHdr) * (uint64_t)n + bits_bytes + c->alpha_bytes + c->aux_bytes;
    // Band pipelining: when every frame has the same size and the caller's buffers are evenly spaced,
    // the frames are cut into horizontal bands of macroblock rows; k_entropy counts finished rows per
    // band in host-visible memory, and as soon as a band is complete in every frame this thread
    // launches its prediction, the loop filter of the band before (one band late: prediction reads
    // unfiltered neighbours, mod.rs:1289-1294), the colour conversion of the rows that are final
    // (the last rows above a band edge are still in the edge record of the loop filter) and
    // their copy to the host -- all on a second stream, under the still running k_entropy.
    constexpr int kMaxBands = 16;
    int nbands = 0;
    int band_start[kMaxBands + 1] = {0};
    size_t dpitch = 0, spitch = 0;
    {
        static const int want = [] { const char* e = std::getenv("WEBPB200_BANDS"); return e ? std::atoi(e) : 10; }();
        bool ok = want > 1 && n >= 2 && c->alpha_bytes == 0 && c->total_mbs >= 16384 && frames[0].mbh >= 8;
        for (int i = 0; ok && i < n; i++) {
            const webpb200_frame_desc& f = frames[i];
            ok = !c->items[i].status && f.width == frames[0].width && f.height == frames[0].height && f.out_bpp == 3 &&
                 (!outs || out_lens[i] == (size_t)f.width * f.height * 3);
        }
        if (ok && outs) {
            dpitch = (size_t)(outs[1] - outs[0]);
            spitch = (size_t)(frames[1].out_off - frames[0].out_off);
            ok = outs[1] > outs[0] && dpitch >= out_lens[0];
            for (int i = 0; ok && i < n; i++) ok = outs[i] == outs[0] + (size_t)i * dpitch && frames[i].out_off == frames[0].out_off + (uint64_t)i * spitch;
        }
        if (ok) {
            // bands shrink towards the bottom: what is left to do after k_entropy ends is the last band
            const int mbh = frames[0].mbh, target = std::min(want, kMaxBands);
            int row = 0;
            while (row < mbh) {
                const int left_bands = target - nbands;
                const int size = left_bands <= 1 ? mbh - row : std::max(1, (mbh - row + 3) / 4);
                band_start[nbands++] = row;
                row += std::min(size, mbh - row);
            }
            band_start[nbands] = mbh;
            if (!c->stream2) CU(c, cudaStreamCreateWithFlags(&c->stream2, cudaStreamNonBlocking), "second stream");
            if (!c->stream3) {
                CU(c, cudaStreamCreateWithFlags(&c->stream3, cudaStreamNonBlocking), "copy stream");
                for (auto& e : c->band_ev) CU(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming), "band events");
            }
            if (c->rows_cap < (size_t)n * 8) {
                if (c->h_rows) cudaFreeHost(c->h_rows);
                c->h_rows = nullptr;
                c->rows_cap = 0;
                CU(c, cudaHostAlloc(reinterpret_cast<void**>(&c->h_rows), sizeof(unsigned int) * 8 * (size_t)n, cudaHostAllocMapped), "row counters");
                CU(c, cudaHostGetDevicePointer(reinterpret_cast<void**>(&c->d_rows), c->h_rows, 0), "row counters");
                c->rows_cap = (size_t)n * 8;
            }
            std::memset(c->h_rows, 0, sizeof(unsigned int) * 8 * (size_t)n);
        }
    }
    if (!c->alpha_bytes) {  // the frame lists are complete without the ALPH headers: up they go before the long kernel
        finish_items(c);
        CU(c, cudaMemcpyAsync(c->d_aux.p + c->aux_lists_off, c->h_aux.p + c->aux_lists_off, c->aux_bytes - c->aux_lists_off, cudaMemcpyHostToDevice, s), "H2D lists");
    }
    // Let the clears and uploads drain before the long-lived entropy blocks are placed: launched
    // right behind them, the 1024 blocks usually ended up unevenly spread over the SMs and the kernel
    // ran 9.5 % longer (42.9 vs 39.2 ms, bimodal); the wait itself costs nothing that is not needed anyway.
    CU(c, cudaStreamSynchronize(s), "sync before k_entropy");
    CU(c, cudaEventRecord(c->ev[7], s), "event");  // every input of the reconstruction kernels is on the device
    cudaEventRecord(c->ev[5], s);
    // Which block layout k_entropy uses for single-partition batches.  A call made through webpb200_decode_batch_begin
    // expects other batches on the GPU: the split layout (mode warps in blocks of their own, half the registers per
    // chain) keeps the chains of three batches resident and is 10-20 % faster per step there, but a batch that has
    // the GPU to itself is 5 % slower in it (its token blocks spread a little less evenly): 36.6 against 35.0 ms.
    static const int split_env = [] { const char* e = std::getenv("WEBPB200_ENTROPY_SPLIT"); return e ? std::atoi(e) : -1; }();
    const bool split_layout = split_env >= 0 ? split_env != 0 : c->in_flight.load();
    if (c->total_mbs) {
        CU(c, launch_entropy(c->bd, reinterpret_cast<const Vp8EntropyHdr*>(c->d_ehdr.p), c->d_bits.p,
                             reinterpret_cast<webpb200_mb_desc*>(c->d_mbs.p), reinterpret_cast<int16_t*>(c->d_levels.p),
                             reinterpret_cast<int*>(c->d_status.p), max_parts, same_probs, nbands ? c->d_rows : nullptr,
                             split_layout ? reinterpret_cast<unsigned int*>(c->d_status.p) + std::max(n, 1) : nullptr, s), "k_entropy");
        c->launches++;
    }
    cudaEventRecord(c->ev[6], s);
    // Which frames ran out of bits is known when k_entropy has ended.  NOTHING that a copy engine executes may be
    // queued behind k_entropy: an engine takes the copies of all streams in the order they were enqueued, so a copy
    // that waits for this kernel would hold up the uploads and downloads of every other batch in flight for tens of
    // milliseconds (measured with three batches in flight: their calls fell into lockstep, one context's downloads
    // ran to the end before the next one's first band moved).  The status words are fetched after the kernel's
    // end has been seen (fetch_status), the frame lists went up before the launch, and what the ALPH planes need
    // goes up on a second stream.
    int* dev_status = reinterpret_cast<int*>(c->h_status.p);
    auto fetch_status = [&]() -> int {  // call with k_entropy finished
        CU(c, cudaMemcpyAsync(dev_status, c->d_status.p, sizeof(int) * (size_t)std::max(n, 1), cudaMemcpyDeviceToHost, s), "D2H status");
        CU(c, cudaStreamSynchronize(s), "sync");
        return 0;
    };
    // ALPH planes (extended.rs:253-308; VP8L or raw) are host work in both entropy modes: here they are decoded
    // on the host threads while k_entropy runs and uploaded beside it, together with the per-filter frame lists
    // (the filter method is in the ALPH chunk header); the descriptors they may have changed follow the kernel.
    if (c->alpha_bytes) {
        parallel_for(n, host_threads, [&](int i) {
            if (!c->items[i].status && frames[i].alpha_mode)
                if (const int st = alpha_one(c, i, frames[i])) drop(i, st);
        });
        finish_items(c);
        if (!c->stream2) CU(c, cudaStreamCreateWithFlags(&c->stream2, cudaStreamNonBlocking), "second stream");
        cudaStream_t up = c->stream2;
        CU(c, cudaMemcpyAsync(c->d_alpha.p, c->h_alpha.p, (size_t)c->alpha_bytes, cudaMemcpyHostToDevice, up), "H2D alpha");
        if (int r2 = upload_vp8l(c, 0, c->vp8l_n, up)) return r2;
        CU(c, cudaMemcpyAsync(c->d_aux.p + c->aux_lists_off, c->h_aux.p + c->aux_lists_off, c->aux_bytes - c->aux_lists_off, cudaMemcpyHostToDevice, up), "H2D lists");
        CU(c, cudaEventRecord(c->ev[11], up), "event");
        CU(c, cudaStreamSynchronize(s), "sync");  // k_entropy has ended (it reads the descriptors that are replaced now)
        CU(c, cudaMemcpyAsync(c->d_frames.p, frames, sizeof(webpb200_frame_desc) * (size_t)n, cudaMemcpyHostToDevice, s), "H2D frames");
        CU(c, cudaStreamWaitEvent(s, c->ev[11], 0), "wait uploads");
    }
    if (nbands) {
        cudaStream_t r = c->stream2;
        CU(c, cudaEventRecord(c->ev[4], s), "event");  // k_entropy finished
        CU(c, cudaStreamWaitEvent(r, c->ev[7], 0), "wait inputs");
        const BatchDev& B = c->bd;
        CU(c, cudaMemsetAsync(B.scratch + B.lay.edge_pred_off, 0, (size_t)(B.lay.zero_bytes - B.lay.edge_pred_off), r), "clear edge records");
        const int mbh = frames[0].mbh, w = frames[0].width, h = frames[0].height;
        const int q_all = h / 2 + 1;
        int q_done = 0, emitted = 0;
        cudaEvent_t bev[kMaxBands][4];
        int band_now = 0;
        double t_seen[kMaxBands];
        if (trace)
            for (int k = 0; k < nbands; k++)
                for (auto& e : bev[k]) cudaEventCreate(&e);
        struct BandCopy { size_t off, bytes; int ev; };
        std::deque<BandCopy> copies;  // converted, not yet handed to the copy engine
        auto pump_copies = [&](bool wait) -> int {
            while (!copies.empty()) {
                const BandCopy& bc = copies.front();
                if (wait) CU(c, cudaEventSynchronize(c->band_ev[bc.ev]), "band event");
                else if (cudaEventQuery(c->band_ev[bc.ev]) != cudaSuccess) break;
                CU(c, cudaMemcpy2DAsync(outs[0] + bc.off, dpitch, c->d_out.p + frames[0].out_off + bc.off, spitch, bc.bytes, (size_t)n, cudaMemcpyDeviceToHost, c->stream3), "D2H band");
                copies.pop_front();
            }
            return 0;
        };
        auto emit = [&](int q_end) -> int {  // convert row pairs [q_done, q_end) and send their pixel rows home
            if (q_end <= q_done) return 0;
            BatchDev bd = B;
            bd.q_first = q_done;
            bd.q_end = q_end;
            CU(c, launch_yuv2rgb(bd, r), "k_yuv2rgb");
            c->launches++;
            if (trace) cudaEventRecord(bev[band_now][3], r);
            const int y0 = q_done == 0 ? 0 : 2 * q_done - 1, y1 = std::min(h, 2 * q_end - 1);
            const size_t off = (size_t)y0 * w * 3, bytes = (size_t)(y1 - y0) * w * 3;
            if (!outs) {  // the images stay on the device (webpb200_batch_output_device)
                q_done = q_end;
                return 0;
            }
            // One 2-D copy for the band of every frame (a copy per frame is API-bound: 5 us x 1024 per band), on its own
            // stream, and handed to the copy engine only once the band's pixels exist (pump_copies): the engine takes
            // the copies of all streams in the order they were enqueued, so a copy that is still waiting for its event
            // holds up the finished bands of every other batch in flight (measured with three batches: the downloads of
            // one context ran to the end before the next context's first band moved).
            CU(c, cudaEventRecord(c->band_ev[emitted % 16], r), "band event");
            copies.push_back({off, bytes, emitted % 16});
            emitted++;
            c->last_d2h += bytes * (uint64_t)n;
            q_done = q_end;
            return 0;
        };
        auto rows_dev = [&](int r0, int r1) {  // macroblock rows [r0, r1) of every frame, in ticket order
            BatchDev bd = B;
            bd.order_first = (uint32_t)r0 * (uint32_t)n;
            bd.total_rows = (uint32_t)(r1 - r0) * (uint32_t)n;
            return bd;
        };
        const double t_launch = ms_since();
        cudaEvent_t probe[3];
        const bool copy_probe = trace && outs && std::getenv("WEBPB200_COPY_PROBE") != nullptr;
        if (copy_probe) {  // how fast is a D2H copy while k_entropy runs? (linear, then 2-D)
            for (auto& e : probe) cudaEventCreate(&e);
            cudaEventRecord(probe[0], r);
            cudaMemcpyAsync(outs[0], c->d_out.p + frames[0].out_off, (size_t)200 << 20, cudaMemcpyDeviceToHost, r);
            cudaEventRecord(probe[1], r);
            cudaMemcpy2DAsync(outs[0], dpitch, c->d_out.p + frames[0].out_off, spitch, 196608, (size_t)n, cudaMemcpyDeviceToHost, r);
            cudaEventRecord(probe[2], r);
        }
        int lf_done = 0;  // macroblock rows already loop-filtered
        int scan_from = 0;
        for (int k = 0; k < nbands; k++) {
            const int r0 = band_start[k], r1 = band_start[k + 1];
            // wait until every frame's decode frontier (the first row some partition has not finished) is past the band
            for (;;) {
                while (scan_from < n) {
                    const int P = hdrs[scan_from].num_parts;
                    const volatile unsigned int* slot = c->h_rows + (size_t)scan_from * 8;
                    int frontier = mbh;
                    for (int p = 0; p < P; p++) {
                        const int rows_p = p < mbh ? (mbh - p + P - 1) / P : 0, done = (int)slot[p];
                        if (done < rows_p) frontier = std::min(frontier, p + done * P);
                    }
                    if (frontier < r1) break;
                    scan_from++;
                }
                if (scan_from >= n || cudaEventQuery(c->ev[4]) != cudaErrorNotReady) break;
                if ((rc = pump_copies(false)) != 0) return rc;
                std::this_thread::sleep_for(std::chrono::microseconds(20));
            }
            scan_from = 0;
            if (trace) { t_seen[k] = ms_since(); band_now = k; cudaEventRecord(bev[k][0], r); }
            CU(c, cudaMemsetAsync(B.scratch, 0, 8, r), "clear tickets");
            CU(c, launch_predict(rows_dev(r0, r1), r), "k_predict");
            c->launches++;
            // the filter stays one macroblock row behind the prediction (which reads unfiltered pixels
            // of the row above), except at the very end
            const int lf_end = k + 1 < nbands ? r1 - 1 : r1;
            CU(c, launch_loopfilter(rows_dev(lf_done, lf_end), r), "k_loopfilter");
            c->launches++;
            lf_done = lf_end;
            // rows whose pixels are final and in the planes: the bottom 4 luma and chroma rows of the last filtered
            // macroblock row are still in its edge record (the row below filters and stores them), so the
            // last row pair converted is the one whose second chroma row is the row just above those
            if (trace) cudaEventRecord(bev[k][1], r);
            if ((rc = emit(lf_end < mbh ? 8 * lf_end - 4 : q_all)) != 0) return rc;
            if ((rc = pump_copies(false)) != 0) return rc;
            if (trace) cudaEventRecord(bev[k][2], outs ? c->stream3 : r);
        }
        const double t_lastband = ms_since();
        if ((rc = pump_copies(true)) != 0) return rc;
        CU(c, cudaStreamSynchronize(s), "sync");
        if ((rc = fetch_status()) != 0) return rc;
        CU(c, cudaStreamSynchronize(r), "sync");
        CU(c, cudaStreamSynchronize(c->stream3), "sync");
        cudaEventElapsedTime(&c->last_entropy_ms, c->ev[5], c->ev[6]);
        if (c->last_chain_mbs >= 256 && c->last_entropy_ms > 0) c->dev_us_per_mb = 0.5 * c->dev_us_per_mb + 0.5 * (1e3 * c->last_entropy_ms / c->last_chain_mbs);
        for (int i = 0; i < n; i++)
            if (dev_status[i]) c->items[i].status = dev_status[i];  // those pixels are not valid
        if ((rc = check_abort(c)) != 0) return rc;
        if (trace) {
            float ems = 0;
            cudaEventElapsedTime(&ems, c->ev[5], c->ev[6]);
            if (copy_probe) {
                float a = 0, b = 0;
                cudaEventElapsedTime(&a, probe[0], probe[1]);
                cudaEventElapsedTime(&b, probe[1], probe[2]);
                std::fprintf(stderr, "  copy probe under k_entropy: linear 210 MB %.2f ms (%.1f GB/s), 2-D %d x 196608 %.2f ms (%.1f GB/s)\n", a, 209.7 / a, n, b,
                             n * 196608e-6 / b);
                for (auto& e : probe) cudaEventDestroy(e);
            }
            for (int k = 0; k < nbands; k++) {
                float a = 0, b = 0, d = 0, y = 0;
                cudaEventElapsedTime(&y, c->ev[5], bev[k][3]);
                cudaEventElapsedTime(&a, c->ev[5], bev[k][0]);
                cudaEventElapsedTime(&b, c->ev[5], bev[k][1]);
                cudaEventElapsedTime(&d, c->ev[5], bev[k][2]);
                std::fprintf(stderr, "  band %d rows %d-%d: seen by host at %.1f ms; on the device (from k_entropy start) kernels %.1f-%.1f, yuv done %.1f, copy done %.1f\n", k,
                             band_start[k], band_start[k + 1], t_seen[k], a, b, y, d);
                for (auto& e : bev[k]) cudaEventDestroy(e);
            }
            {
                float a = 0, b = 0, d = 0;
                cudaEventElapsedTime(&a, c->ev[8], c->ev[9]);
                cudaEventElapsedTime(&b, c->ev[9], c->ev[10]);
                cudaEventElapsedTime(&d, c->ev[10], c->ev[5]);
                std::fprintf(stderr, "  device before k_entropy: clears %.2f ms, idle %.2f ms, uploads %.2f ms\n", a, b, d);
            }
            std::fprintf(stderr, "[webpb200] decode_batch(device entropy, %d bands) n=%d: host %.1f ms, launched at %.1f, k_entropy %.2f ms, last band seen at %.1f, total %.1f ms\n",
                         nbands, n, t_host, t_launch, ems, t_lastband, ms_since());
        }
        int bad = 0;
        for (int i = 0; i < n; i++) {
            if (status) status[i] = c->items[i].status;
            bad += c->items[i].status != 0;
        }
        return bad;
    }
    if ((rc = reconstruct(c, c->bd, c->alpha_bytes ? 0xF : 0x7, s, nullptr)) != 0) return rc;
    CU(c, cudaStreamSynchronize(s), "sync");
    if ((rc = fetch_status()) != 0) return rc;
    const double t_kernels = ms_since();
    for (int i = 0; i < n; i++)
        if (!c->items[i].status && dev_status[i]) c->items[i].status = dev_status[i];
    if (outs && (rc = download_range(c, outs, out_lens, 0, n)) != 0) return rc;
    CU(c, cudaStreamSynchronize(s), "sync");
    cudaEventElapsedTime(&c->last_entropy_ms, c->ev[5], c->ev[6]);
    if (c->last_chain_mbs >= 256 && c->last_entropy_ms > 0) c->dev_us_per_mb = 0.5 * c->dev_us_per_mb + 0.5 * (1e3 * c->last_entropy_ms / c->last_chain_mbs);
    if ((rc = check_abort(c)) != 0) return rc;
    if (trace) {
        float ems = 0;
        cudaEventElapsedTime(&ems, c->ev[5], c->ev[6]);
        std::fprintf(stderr, "[webpb200] decode_batch(device entropy) n=%d: host %.1f ms, +kernels %.1f ms (k_entropy %.2f ms), +download %.1f ms\n",
                     n, t_host, t_kernels, ems, ms_since());
    }
    int bad = 0;
    for (int i = 0; i < n; i++) {
        if (status) status[i] = c->items[i].status;
        bad += c->items[i].status != 0;
    }
    return bad;
}

}  // namespace

// Pipelined: the batch is cut into chunks of about kChunkMbs macroblocks; host threads entropy-
// decode images in order, and as soon as a chunk is complete its packed arrays go to the device,
// its kernels run and its pixels come back, all on the context's stream, while the host threads
// are already decoding the following chunks.  The copies and kernels hide behind the entropy decode.
namespace {
int decode_items_host_entropy(webpb200_ctx* c, std::vector<WorkItem>&& items, uint8_t* const* outs, const size_t* out_lens, int* status,
                              int upsampling, int host_threads);

// One batch of reconstruction work items (still images, or the frames of an animation) through
// whichever placement of the entropy stage the context asks for.  outs may be null (images stay on
// the device).  Returns the number of failed items, or a negative status.
int decode_items(webpb200_ctx* c, std::vector<WorkItem>&& items, uint8_t* const* outs, const size_t* out_lens, int* status, int upsampling,
                 int host_threads) {
    cudaSetDevice(c->device);
    c->last_d2h = 0;
    if (c->entropy_where != WEBPB200_ENTROPY_HOST) {
        const int r = decode_batch_device_entropy(c, items, outs, out_lens, status, upsampling, host_threads);
        if (r != kPreferHostEntropy) return r;
    }
    return decode_items_host_entropy(c, std::move(items), outs, out_lens, status, upsampling, host_threads);
}
}  // namespace

extern "C" int webpb200_decode_batch(webpb200_ctx* c, const uint8_t* const* files, const size_t* lens, int n, uint8_t* const* outs,
                                     const size_t* out_lens, int* status, int upsampling, int host_threads) {
    if (!c) return -WEBPB200_ERR_NO_DEVICE;
    if (n < 0) return -WEBPB200_ERR_INVALID_PARAM;
    std::vector<WorkItem> items;
    items_from_files(files, lens, n, host_threads, &items);
    return decode_items(c, std::move(items), outs, out_lens, status, upsampling, host_threads);
}

struct webpb200_pending {
    std::thread worker;
    int rc = 0;
    webpb200_ctx* ctx = nullptr;
};

extern "C" int webpb200_decode_batch_begin(webpb200_ctx* c, const uint8_t* const* files, const size_t* lens, int n, uint8_t* const* outs,
                                           const size_t* out_lens, int* status, int upsampling, int host_threads, webpb200_pending** pending) {
    if (!pending) return -WEBPB200_ERR_INVALID_PARAM;
    *pending = nullptr;
    if (!c) return -WEBPB200_ERR_NO_DEVICE;
    bool expected = false;
    if (!c->in_flight.compare_exchange_strong(expected, true)) return -WEBPB200_ERR_INVALID_PARAM;  // one call in flight per context
    webpb200_pending* p = new webpb200_pending();
    p->ctx = c;
    p->worker = std::thread([=]() { p->rc = webpb200_decode_batch(c, files, lens, n, outs, out_lens, status, upsampling, host_threads); });
    *pending = p;
    return 0;
}

extern "C" int webpb200_decode_batch_end(webpb200_pending* p) {
    if (!p) return -WEBPB200_ERR_INVALID_PARAM;
    p->worker.join();
    const int rc = p->rc;
    p->ctx->in_flight.store(false);
    delete p;
    return rc;
}

// Several contexts (GPUs), one host thread each, contiguous ranges balanced by macroblock count.
extern "C" int webpb200_decode_batch_multi(webpb200_ctx* const* ctxs, int n_ctx, const uint8_t* const* files, const size_t* lens, int n,
                                           uint8_t* const* outs, const size_t* out_lens, int* status, int upsampling, int host_threads,
                                           int* split) {
    if (!ctxs || n_ctx < 1 || n < 0) return -WEBPB200_ERR_INVALID_PARAM;
    for (int k = 0; k < n_ctx; k++)
        if (!ctxs[k]) return -WEBPB200_ERR_NO_DEVICE;
    // weights: macroblocks per file from the frame header (a bad file weighs nothing)
    std::vector<uint64_t> prefix((size_t)n + 1, 0);
    {
        std::vector<uint64_t> w((size_t)n, 0);
        parallel_for(n, host_threads, [&](int i) {
            Container ct;
            if (parse_container(files[i], lens[i], &ct) || ct.animated || !ct.chunks[CK_VP8].present) return;
            const Range r = ct.chunks[CK_VP8];
            Vp8Dims d{0, 0, 0, 0};
            if (r.end <= lens[i] && !vp8_probe(files[i] + r.start, (size_t)r.size(), &d)) w[i] = (uint64_t)d.mbw * d.mbh;
        });
        for (int i = 0; i < n; i++) prefix[i + 1] = prefix[i] + w[i] + 1;  // +1: files without macroblocks still spread out
    }
    std::vector<int> cut((size_t)n_ctx + 1, n);
    cut[0] = 0;
    for (int k = 1; k < n_ctx; k++) {
        const uint64_t target = prefix[n] * (uint64_t)k / (uint64_t)n_ctx;
        cut[k] = (int)(std::lower_bound(prefix.begin(), prefix.end(), target) - prefix.begin());
        cut[k] = std::max(cut[k], cut[k - 1]);
    }
    if (split)
        for (int k = 0; k <= n_ctx; k++) split[k] = cut[k];
    std::vector<int> rc((size_t)n_ctx, 0);
    const int per = std::max(1, host_threads / n_ctx);
    auto run = [&](int k) {
        const int i0 = cut[k], m = cut[k + 1] - cut[k];
        if (m <= 0) return;
        rc[k] = webpb200_decode_batch(ctxs[k], files + i0, lens + i0, m, outs ? outs + i0 : nullptr, out_lens ? out_lens + i0 : nullptr,
                                      status ? status + i0 : nullptr, upsampling, per);
    };
    std::vector<std::thread> th;
    for (int k = 1; k < n_ctx; k++) th.emplace_back(run, k);
    run(0);
    for (auto& t : th) t.join();
    int bad = 0;
    for (int k = 0; k < n_ctx; k++) {
        if (rc[k] < 0) return rc[k];
        bad += rc[k];
    }
    return bad;
}

namespace {
int decode_items_host_entropy(webpb200_ctx* c, std::vector<WorkItem>&& items, uint8_t* const* outs, const size_t* out_lens, int* status,
                              int upsampling, int host_threads) {
    const bool trace = std::getenv("WEBPB200_TRACE") != nullptr;
    const auto t0 = std::chrono::steady_clock::now();
    const int n = (int)items.size();
    int rc = layout_items(c, std::move(items), upsampling, host_threads);
    if (rc) return rc;
    if ((rc = reserve_device(c)) != 0) return rc;
    webpb200_frame_desc* frames = reinterpret_cast<webpb200_frame_desc*>(c->h_frames.p);
    cudaStream_t s = c->stream;
    c->last_entropy = WEBPB200_ENTROPY_HOST;
    c->last_entropy_ms = 0;
    c->last_d2h = 0;
    c->last_h2d = sizeof(webpb200_frame_desc) * (uint64_t)n + c->total_mbs * (sizeof(webpb200_mb_desc) + 2 * WEBPB200_LEVELS_PER_MB) + c->alpha_bytes + c->aux_bytes;

    // chunks by macroblock count; ticket order row-major inside each chunk
    const uint64_t kChunkMbs = 96 * 1024;
    std::vector<int> chunk_end;
    std::vector<int> chunk_of((size_t)n, 0);
    {
        uint64_t acc = 0;
        for (int i = 0; i < n; i++) {
            chunk_of[i] = (int)chunk_end.size();
            acc += (uint64_t)frames[i].mbw * frames[i].mbh;
            if (acc >= kChunkMbs || i == n - 1) {
                chunk_end.push_back(i + 1);
                acc = 0;
            }
        }
    }
    const int nchunks = (int)chunk_end.size();
    uint32_t* order = reinterpret_cast<uint32_t*>(c->h_aux.p + c->aux_order_off);
    for (int k = 0; k < nchunks; k++) fill_row_order(frames, k ? chunk_end[k - 1] : 0, chunk_end[k], order);
    if (int r2 = upload_row_tables(c, c->h_aux.p, cudaMemcpyHostToDevice, s)) return r2;
    c->uploaded = true;

    // host threads: images in order, one at a time each
    std::unique_ptr<std::atomic<int>[]> left(new std::atomic<int>[(size_t)std::max(nchunks, 1)]);
    for (int k = 0; k < nchunks; k++) left[k].store(chunk_end[k] - (k ? chunk_end[k - 1] : 0));
    std::atomic<int> next{0};
    auto worker = [&]() {
        for (;;) {
            const int i = next.fetch_add(1);
            if (i >= n) return;
            entropy_one(c, i);
            left[chunk_of[i]].fetch_sub(1, std::memory_order_release);
        }
    };
    int nthreads = std::max(1, std::min(host_threads, n));
    const auto t_workers = std::chrono::steady_clock::now();
    std::vector<std::thread> pool;
    for (int t = 0; t < nthreads; t++) pool.emplace_back(worker);

    const bool any_alpha = c->alpha_bytes != 0;
    int err = 0;
    for (int k = 0; k < nchunks && !err; k++) {
        while (left[k].load(std::memory_order_acquire) > 0) std::this_thread::sleep_for(std::chrono::microseconds(50));
        const int i0 = k ? chunk_end[k - 1] : 0, i1 = chunk_end[k];
        const uint64_t mb0 = frames[i0].mb_base;
        const uint64_t mb1 = i1 < n ? frames[i1].mb_base : c->total_mbs;
        const uint32_t row0 = frames[i0].row_base, row1 = i1 < n ? frames[i1].row_base : (uint32_t)c->total_rows;
        auto fail = [&](cudaError_t e, const char* what) { if (!err && c->fail(e, what)) err = -WEBPB200_ERR_CUDA; };
        fail(cudaMemcpyAsync(c->d_frames.p + sizeof(webpb200_frame_desc) * (size_t)i0, frames + i0, sizeof(webpb200_frame_desc) * (size_t)(i1 - i0), cudaMemcpyHostToDevice, s), "H2D frames");
        if (mb1 > mb0) {
            fail(cudaMemcpyAsync(c->d_mbs.p + sizeof(webpb200_mb_desc) * mb0, c->h_mbs.p + sizeof(webpb200_mb_desc) * mb0, sizeof(webpb200_mb_desc) * (mb1 - mb0), cudaMemcpyHostToDevice, s), "H2D mbs");
            const size_t lb = sizeof(int16_t) * WEBPB200_LEVELS_PER_MB;
            fail(cudaMemcpyAsync(c->d_levels.p + lb * mb0, c->h_levels.p + lb * mb0, lb * (mb1 - mb0), cudaMemcpyHostToDevice, s), "H2D levels");
        }
        if (any_alpha && !err) {
            // ALPH planes of this chunk: their filter methods are known now (entropy_one read the chunk
            // headers), so the un-filter runs per chunk too, ahead of the chunk's colour conversion (which merges the planes), and nothing is left for the end.
            // The chunk's index lists live at [4 * i0, 4 * i1) of the list area, method-major.
            const int nc = i1 - i0;
            int* lists = reinterpret_cast<int*>(c->h_aux.p + c->aux_lists_off) + (size_t)4 * i0;
            BatchDev bd = c->bd;
            bd.n_frames = nc;
            for (int m = 0; m < 4; m++) bd.alpha_counts[m] = 0, bd.alpha_maxw[m] = 0, bd.alpha_maxh[m] = 0;
            for (int i = i0; i < i1; i++) {
                const webpb200_frame_desc& fd = frames[i];
                if (!fd.alpha_mode || fd.out_bpp != 4) continue;
                const int m = fd.alpha_filter & 3;
                lists[m * nc + bd.alpha_counts[m]++] = i;
                bd.alpha_maxw[m] = std::max<uint32_t>(bd.alpha_maxw[m], fd.width);
                bd.alpha_maxh[m] = std::max<uint32_t>(bd.alpha_maxh[m], fd.height);
            }
            if (bd.alpha_counts[0] + bd.alpha_counts[1] + bd.alpha_counts[2] + bd.alpha_counts[3]) {
                const uint64_t a0 = frames[i0].alpha_off, a1 = i1 < n ? frames[i1].alpha_off : c->alpha_bytes;
                if (a1 > a0) fail(cudaMemcpyAsync(c->d_alpha.p + a0, c->h_alpha.p + a0, (size_t)(a1 - a0), cudaMemcpyHostToDevice, s), "H2D alpha");
                {   // this chunk's VP8L planes: upload what the host threads decoded, undo the transforms
                    int p0 = c->vp8l_n, p1 = 0;
                    for (int i = i0; i < i1; i++)
                        if (c->vp8l_of_item[i] >= 0) { p0 = std::min(p0, c->vp8l_of_item[i]); p1 = std::max(p1, c->vp8l_of_item[i] + 1); }
                    if (p1 > p0 && !err) {
                        if (int r = upload_vp8l(c, p0, p1, s)) err = r;
                        if (!err) if (int r = run_vp8l(c, c->bd, p0, p1, s)) err = r;
                    }
                }
                const size_t lo = c->aux_lists_off + sizeof(int) * 4 * (size_t)i0;
                fail(cudaMemcpyAsync(c->d_aux.p + lo, c->h_aux.p + lo, sizeof(int) * 4 * (size_t)nc, cudaMemcpyHostToDevice, s), "H2D lists");
                bd.alpha_lists = reinterpret_cast<const int*>(c->d_aux.p + lo);
                int l = 0;
                if (!err) fail(launch_alpha(bd, s, &l), "k_alpha");
                c->launches += (uint64_t)l;
            }
        }
        if (row1 > row0 && !err) {
            BatchDev bd = c->bd;
            bd.frame_first = i0;
            bd.n_frames = i1 - i0;
            bd.order_first = row0;
            bd.total_rows = row1 - row0;
            if (int r = reconstruct(c, bd, any_alpha ? 0x17 : 0x7, s, nullptr)) err = r;  // 0x10: merge the ALPH planes prepared above
        }
        if (!err && outs)
            if (int r = download_range(c, outs, out_lens, i0, i1)) err = r;
    }
    for (auto& t : pool) t.join();
    if (c->total_mbs >= 4096 && !c->alpha_bytes) {  // refine the host figure of the AUTO cost model (ALPH work would distort it)
        const double us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t_workers).count();
        c->host_us_per_mb = 0.5 * c->host_us_per_mb + 0.5 * (us * nthreads / (double)c->total_mbs);
    }
    if (err) {
        cudaStreamSynchronize(s);
        return err;
    }
    CU(c, cudaStreamSynchronize(s), "sync");
    if ((rc = check_abort(c)) != 0) return rc;
    if (any_alpha) {  // whole-batch lists, for the staged API that may follow (batch_reconstruct)
        finish_items(c);
        CU(c, cudaMemcpy(c->d_aux.p + c->aux_lists_off, c->h_aux.p + c->aux_lists_off, c->aux_bytes - c->aux_lists_off, cudaMemcpyHostToDevice), "H2D lists");
    }
    if (trace) std::fprintf(stderr, "[webpb200] decode_batch n=%d in %d chunks: %.1f ms\n", n, nchunks,
                            std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    int bad = 0;
    for (int i = 0; i < n; i++) {
        if (status) status[i] = c->items[i].status;
        bad += c->items[i].status != 0;
    }
    return bad;
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// host half alone (no CUDA calls): container walk + entropy decode into caller arrays
// ------------------------------------------------------------------------------------------------
extern "C" int webpb200_host_entropy_decode(const uint8_t* file, size_t len, webpb200_frame_desc* fd, webpb200_mb_desc* mbs, size_t mbs_cap,
                                            int16_t* levels, size_t levels_cap, uint8_t* alpha, size_t alpha_cap) {
    std::vector<WorkItem> items;
    const uint8_t* files[1] = {file};
    const size_t lens[1] = {len};
    items_from_files(files, lens, 1, 1, &items);
    WorkItem& it = items[0];
    if (it.status) return it.status;
    int st = vp8_probe(it.vp8, it.vp8_len, &it.dims);
    if (st) return st;
    if (it.dims.width != it.check_w || it.dims.height != it.check_h) return WEBPB200_ERR_INCONSISTENT_SIZE;
    std::memset(fd, 0, sizeof(*fd));
    fd->width = (uint16_t)it.dims.width; fd->height = (uint16_t)it.dims.height;
    fd->mbw = (uint16_t)it.dims.mbw; fd->mbh = (uint16_t)it.dims.mbh;
    fd->out_bpp = (uint8_t)it.bpp;
    fd->alpha_mode = it.alph != nullptr;
    if (!mbs || !levels) return 0;  // size query
    const size_t m = (size_t)it.dims.mbw * it.dims.mbh;
    if (mbs_cap < m || levels_cap < m * WEBPB200_LEVELS_PER_MB) return WEBPB200_ERR_INVALID_PARAM;
    st = vp8_entropy_decode(it.vp8, it.vp8_len, fd, mbs, levels);
    if (st) return st;
    if (it.alph && alpha) {
        int filter = 0;
        bool compressed = false;
        st = parse_alph_header(it.alph, it.alph_len, &filter, &compressed);
        if (st) return st;
        const size_t want = (size_t)it.alph_w * it.alph_h;
        if (alpha_cap < want) return WEBPB200_ERR_INVALID_PARAM;
        fd->alpha_filter = (uint8_t)filter;
        if (compressed) return vp8l_decode_green(it.alph + 1, it.alph_len - 1, it.alph_w, it.alph_h, alpha);
        if (it.alph_len - 1 < want) return WEBPB200_ERR_IO;
        std::memcpy(alpha, it.alph + 1, want);
    }
    return 0;
}

// The same through the hand-over structure of the device entropy stage (host_vp8.h), no ALPH part.
extern "C" void webpb200_debug_entropy_faults(int flags) { set_entropy_faults(flags); }

extern "C" int webpb200_debug_entropy_resumed(const uint8_t* file, size_t len, webpb200_frame_desc* fd, webpb200_mb_desc* mbs, size_t mbs_cap,
                                              int16_t* levels, size_t levels_cap) {
    std::vector<WorkItem> items;
    const uint8_t* files[1] = {file};
    const size_t lens[1] = {len};
    items_from_files(files, lens, 1, 1, &items);
    WorkItem& it = items[0];
    if (it.status) return it.status;
    int st = vp8_probe(it.vp8, it.vp8_len, &it.dims);
    if (st) return st;
    if (it.dims.width != it.check_w || it.dims.height != it.check_h) return WEBPB200_ERR_INCONSISTENT_SIZE;
    std::memset(fd, 0, sizeof(*fd));
    const size_t m = (size_t)it.dims.mbw * it.dims.mbh;
    if (mbs_cap < m || levels_cap < m * WEBPB200_LEVELS_PER_MB) return WEBPB200_ERR_INVALID_PARAM;
    return vp8_entropy_decode_resumed(it.vp8, it.vp8_len, fd, mbs, levels);
}

extern "C" void webpb200_debug_bpred_eval(int mode, const uint8_t* edge13, uint8_t* out16) { bpred_table_eval(mode, edge13, out16); }

// ------------------------------------------------------------------------------------------------
// device-resident seam
// ------------------------------------------------------------------------------------------------
extern "C" int webpb200_workspace_sizes(const webpb200_frame_desc* frames, int n, uint64_t* plane_bytes, uint64_t* out_bytes,
                                        uint64_t* scratch_bytes) {
    uint64_t p = 0, o = 0, rows = 0, mbs = 0, alpha_end = 0;
    for (int i = 0; i < n; i++) {
        const uint64_t m = (uint64_t)frames[i].mbw * frames[i].mbh;
        if (frames[i].alpha_mode) alpha_end = std::max(alpha_end, frames[i].alpha_off + (uint64_t)frames[i].width * frames[i].height);
        p = std::max(p, align256(frames[i].v_off + m * 64));
        o = std::max(o, align256(frames[i].out_off + (uint64_t)frames[i].width * frames[i].height * frames[i].out_bpp));
        rows = std::max<uint64_t>(rows, (uint64_t)frames[i].row_base + frames[i].mbh);
        mbs = std::max<uint64_t>(mbs, (uint64_t)frames[i].mb_base + m);
    }
    if (plane_bytes) *plane_bytes = p;
    if (out_bytes) *out_bytes = o;
    if (scratch_bytes)
        *scratch_bytes = scratch_layout(rows, mbs, edge_slots_of(frames, n)).total + align256(16 * (uint64_t)std::max(n, 1)) + (alpha_end ? alpha_handoff_size(alpha_end) + align256(alpha_end) : 0);
    return 0;
}

extern "C" int webpb200_reconstruct_device(const webpb200_frame_desc* hf, const webpb200_frame_desc* d_frames, int n,
                                           const webpb200_mb_desc* d_mbs, const int16_t* d_levels, const uint8_t* d_alpha,
                                           uint8_t* d_planes, uint8_t* d_out, void* d_scratch, unsigned stages, void* stream) {
    if (n <= 0) return 0;
    int st = 0;
    webpb200_ctx* c = default_ctx(&st);
    if (!c) return -st;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    uint64_t rows = 0, mbs = 0, alpha_end = 0;
    BatchDev bd;
    std::memset(&bd, 0, sizeof(bd));
    for (int i = 0; i < n; i++) {
        if (hf[i].alpha_mode) alpha_end = std::max(alpha_end, hf[i].alpha_off + (uint64_t)hf[i].width * hf[i].height);
        rows = std::max<uint64_t>(rows, (uint64_t)hf[i].row_base + hf[i].mbh);
        mbs = std::max<uint64_t>(mbs, (uint64_t)hf[i].mb_base + (uint64_t)hf[i].mbw * hf[i].mbh);
        bd.max_w = std::max<uint32_t>(bd.max_w, hf[i].width);
        bd.max_h = std::max<uint32_t>(bd.max_h, hf[i].height);
    }
    bd.lay = scratch_layout(rows, mbs, edge_slots_of(hf, n));
    // row tables and ALPH lists are built on the host and copied into the caller's scratch arena
    std::vector<uint32_t> rowmap((size_t)rows, 0), order((size_t)rows, 0), recbase((size_t)rows, 0);
    std::vector<int> lists((size_t)4 * n, 0);
    {
        size_t k = 0;  // rows in use may be fewer than `rows` if the caller left gaps; fill those with valid ids
        uint32_t max_h = 0;
        for (int i = 0; i < n; i++) max_h = std::max<uint32_t>(max_h, hf[i].mbh);
        for (uint32_t r = 0; r < max_h; r++)
            for (int i = 0; i < n; i++)
                if (hf[i].mbh > r) order[k++] = hf[i].row_base + r;
        rows = k;  // tickets handed out = rows that exist
    }
    fill_row_tables(hf, n, rowmap.data(), recbase.data());
    for (int i = 0; i < n; i++) {
        if (hf[i].alpha_mode && hf[i].out_bpp == 4) {
            const int m = hf[i].alpha_filter & 3;
            lists[(size_t)m * n + bd.alpha_counts[m]++] = i;
            bd.alpha_maxw[m] = std::max<uint32_t>(bd.alpha_maxw[m], hf[i].width);
            bd.alpha_maxh[m] = std::max<uint32_t>(bd.alpha_maxh[m], hf[i].height);
        }
    }
    uint8_t* scratch = static_cast<uint8_t*>(d_scratch);
    CU(c, cudaMemsetAsync(scratch, 0, 256, s), "clear watchdog");
    CU(c, cudaMemcpyAsync(scratch + bd.lay.rowmap_off, rowmap.data(), 4 * rowmap.size(), cudaMemcpyHostToDevice, s), "rowmap");
    CU(c, cudaMemcpyAsync(scratch + bd.lay.order_off, order.data(), 4 * order.size(), cudaMemcpyHostToDevice, s), "row order");
    CU(c, cudaMemcpyAsync(scratch + bd.lay.recbase_off, recbase.data(), 4 * recbase.size(), cudaMemcpyHostToDevice, s), "edge slots");
    CU(c, cudaMemcpyAsync(scratch + bd.lay.total, lists.data(), sizeof(int) * lists.size(), cudaMemcpyHostToDevice, s), "lists");
    CU(c, cudaStreamSynchronize(s), "sync");  // the host vectors go away
    bd.frames = d_frames; bd.mbs = d_mbs; bd.levels = d_levels; bd.alpha = d_alpha;
    bd.planes = d_planes; bd.out = d_out; bd.scratch = scratch;
    bd.n_frames = n; bd.total_rows = (uint32_t)rows; bd.sm_count = c->sm_count;
    bd.predict_blocks = c->predict_blocks; bd.predict_blocks_latency = c->predict_blocks_latency; bd.filter_blocks = c->filter_blocks;
    bd.alpha_lists = reinterpret_cast<const int*>(scratch + bd.lay.total);
    if (alpha_end) {
        bd.alpha_handoff = reinterpret_cast<uint16_t*>(scratch + bd.lay.total + align256(16 * (uint64_t)n));
        bd.alpha_handoff_bytes = alpha_handoff_size(alpha_end);
        bd.alpha_planar = reinterpret_cast<uint8_t*>(bd.alpha_handoff) + bd.alpha_handoff_bytes;
    }
    return reconstruct(c, bd, stages, s, nullptr);
}

// ------------------------------------------------------------------------------------------------
// WebPDecoder surface
// ------------------------------------------------------------------------------------------------
struct webpb200_decoder {
    Container ct;
    int upsampling = 0;
    bool has_bg = false;
    uint8_t bg[4] = {0, 0, 0, 0};
    size_t memory_limit = ~size_t(0);  // decoder.rs:314, 354
    // AnimationState, decoder.rs:240-263
    uint32_t next_frame = 0;
    uint64_t next_frame_start = 0;
    bool dispose_next = true;
    uint32_t pw = 0, ph = 0, px = 0, py = 0;
    bool canvas_live = false;
    DevBuf canvas, frame_out;
};

extern "C" int webpb200_open(const uint8_t* data, size_t len, int upsampling, webpb200_decoder** out) {
    *out = nullptr;
    webpb200_decoder* d = new webpb200_decoder();
    int st = parse_container(data, len, &d->ct);
    if (st) {
        delete d;
        return st;
    }
    d->upsampling = upsampling;
    d->next_frame_start = d->ct.first_anmf;
    *out = d;
    return 0;
}
extern "C" int webpb200_open_with_options(const uint8_t* data, size_t len, const webpb200_options* opt, webpb200_decoder** out) {
    int st = webpb200_open(data, len, opt ? opt->lossy_upsampling : 0, out);
    if (st || !opt) return st;
    if (opt->memory_limit) (*out)->memory_limit = (size_t)opt->memory_limit;
    if (opt->has_background_color) {
        st = webpb200_set_background_color(*out, opt->background_color);
        if (st) {
            webpb200_close(*out);
            *out = nullptr;
        }
    }
    return st;
}
extern "C" void webpb200_close(webpb200_decoder* d) {
    if (!d) return;
    d->canvas.release();
    d->frame_out.release();
    delete d;
}
extern "C" void webpb200_dimensions(const webpb200_decoder* d, uint32_t* w, uint32_t* h) { *w = d->ct.width; *h = d->ct.height; }
extern "C" int webpb200_has_alpha(const webpb200_decoder* d) { return d->ct.has_alpha; }
extern "C" int webpb200_is_animated(const webpb200_decoder* d) { return d->ct.animated; }
extern "C" int webpb200_is_lossy(const webpb200_decoder* d) { return d->ct.is_lossy; }
extern "C" uint32_t webpb200_num_frames(const webpb200_decoder* d) { return d->ct.num_frames; }
extern "C" uint32_t webpb200_loop_count(const webpb200_decoder* d) { return d->ct.loop_count; }
extern "C" uint64_t webpb200_loop_duration(const webpb200_decoder* d) { return d->ct.loop_duration; }
extern "C" size_t webpb200_output_buffer_size(const webpb200_decoder* d) {
    return (size_t)d->ct.width * d->ct.height * (d->ct.has_alpha ? 4 : 3);
}
extern "C" int webpb200_set_background_color(webpb200_decoder* d, const uint8_t rgba[4]) {
    if (d->ct.kind != 2) return WEBPB200_ERR_INVALID_PARAM;
    d->has_bg = true;
    std::memcpy(d->bg, rgba, 4);
    return 0;
}
extern "C" void webpb200_set_lossy_upsampling(webpb200_decoder* d, int upsampling) { d->upsampling = upsampling; }
extern "C" void webpb200_set_memory_limit(webpb200_decoder* d, size_t limit) { d->memory_limit = limit; }
extern "C" int webpb200_background_color_hint(const webpb200_decoder* d, uint8_t rgba[4]) {
    if (d->ct.kind != 2) return WEBPB200_ERR_INVALID_PARAM;
    std::memcpy(rgba, d->ct.background_hint, 4);
    return 0;
}
namespace {
// read_chunk, decoder.rs:620-640: the range is handed out where it lies in the caller's buffer
int metadata_chunk(webpb200_decoder* d, int kind, const uint8_t** data, size_t* len) {
    *data = nullptr;
    *len = 0;
    const Range r = d->ct.chunks[kind];
    if (!r.present) return 0;
    if (r.size() > (uint64_t)d->memory_limit) return WEBPB200_ERR_MEMORY_LIMIT;
    if (r.end > d->ct.len) return WEBPB200_ERR_IO;
    *data = d->ct.data + r.start;
    *len = (size_t)r.size();
    return 0;
}
}  // namespace
extern "C" int webpb200_icc_profile(webpb200_decoder* d, const uint8_t** data, size_t* len) { return metadata_chunk(d, CK_ICCP, data, len); }
extern "C" int webpb200_exif_metadata(webpb200_decoder* d, const uint8_t** data, size_t* len) { return metadata_chunk(d, CK_EXIF, data, len); }
extern "C" int webpb200_xmp_metadata(webpb200_decoder* d, const uint8_t** data, size_t* len) { return metadata_chunk(d, CK_XMP, data, len); }

namespace {

// One animation frame: reconstruct it as a batch of one and composite it on the device canvas.
// state_* mirror AnimationState; canvas lives in `canvas` (device).
int decode_and_composite(webpb200_ctx* c, const Container& ct, uint64_t pos, int upsampling, bool has_bg, const uint8_t bg[4],
                         bool dispose_next_in, uint32_t pw, uint32_t ph, uint32_t px, uint32_t py, DevBuf& canvas, bool& canvas_live,
                         DevBuf& frame_out, uint8_t* host_out, AnmfFrame* fr_out) {
    AnmfFrame fr;
    int st = parse_anmf(ct, pos, &fr);
    if (st) return st;
    if (fr.lossless) return WEBPB200_ERR_NOT_LOSSY;
    std::vector<WorkItem> items(1);
    WorkItem& it = items[0];
    it.vp8 = ct.data + fr.vp8.start;
    it.vp8_len = (size_t)fr.vp8.size();
    const bool frame_has_alpha = fr.alph.present;
    if (frame_has_alpha) {
        it.alph = ct.data + fr.alph.start;
        it.alph_len = (size_t)fr.alph.size();
        it.alph_w = (int)fr.w;
        it.alph_h = (int)fr.h;
        it.bpp = 4;
        // decoder.rs:814-857 does not compare the VP8 frame with the ANMF size; a mismatch would
        // index out of range there, so it is refused here
        it.check_w = (int)fr.w;
        it.check_h = (int)fr.h;
    } else {
        it.bpp = 3;
        it.check_w = (int)fr.w;
        it.check_h = (int)fr.h;  // decoder.rs:798-802
    }
    int rc = prepare_items(c, std::move(items), upsampling, 1);
    if (rc) return -rc;
    if (c->items[0].status) return c->items[0].status;
    if ((rc = upload(c)) != 0) return -rc;
    if ((rc = reconstruct(c, c->bd, 0xF, c->stream, nullptr)) != 0) return -rc;

    const size_t cbytes = (size_t)ct.width * ct.height * 4;
    const int out_bpp = ct.has_alpha ? 4 : 3;
    if (!canvas_live) {  // decoder.rs:872-882
        if (c->fail(canvas.reserve(cbytes), "canvas")) return WEBPB200_ERR_CUDA;
        if (has_bg) {
            std::vector<uint8_t> fill(cbytes);
            for (size_t i = 0; i < cbytes; i += 4) std::memcpy(&fill[i], bg, 4);
            if (c->fail(cudaMemcpy(canvas.p, fill.data(), cbytes, cudaMemcpyHostToDevice), "canvas fill")) return WEBPB200_ERR_CUDA;
        } else if (c->fail(cudaMemsetAsync(canvas.p, 0, cbytes, c->stream), "canvas clear")) {
            return WEBPB200_ERR_CUDA;
        }
        canvas_live = true;
    }
    if (c->fail(frame_out.reserve((size_t)ct.width * ct.height * out_bpp), "frame out")) return WEBPB200_ERR_CUDA;
    CompositeArgs a;
    std::memset(&a, 0, sizeof(a));
    a.canvas = canvas.p;
    a.frame = c->d_out.p + reinterpret_cast<const webpb200_frame_desc*>(c->h_frames.p)[0].out_off;
    a.out = frame_out.p;
    a.canvas_w = ct.width; a.canvas_h = ct.height;
    a.fx = fr.x; a.fy = fr.y; a.fw = fr.w; a.fh = fr.h;
    a.px = px; a.py = py; a.pw = pw; a.ph = ph;
    if (dispose_next_in) {  // decoder.rs:861-869
        if (has_bg) { a.has_clear = 1; std::memcpy(a.clear, bg, 4); }
        else if (frame_has_alpha) { a.has_clear = 1; }
    }
    a.frame_has_alpha = frame_has_alpha;
    a.use_blend = fr.blend;
    a.out_bpp = out_bpp;
    if (c->fail(launch_composite(a, c->stream), "k_composite")) return WEBPB200_ERR_CUDA;
    c->launches++;
    if (c->fail(cudaMemcpyAsync(host_out, frame_out.p, (size_t)ct.width * ct.height * out_bpp, cudaMemcpyDeviceToHost, c->stream), "D2H canvas"))
        return WEBPB200_ERR_CUDA;
    if (c->fail(cudaStreamSynchronize(c->stream), "sync")) return WEBPB200_ERR_CUDA;
    *fr_out = fr;
    return 0;
}

}  // namespace

extern "C" int webpb200_read_frame(webpb200_decoder* d, uint8_t* buf, size_t len, uint32_t* duration_ms) {
    if (!d->ct.animated) return WEBPB200_ERR_NOT_ANIMATED;
    if (len != webpb200_output_buffer_size(d)) return WEBPB200_ERR_IMAGE_TOO_LARGE;
    if (d->next_frame == d->ct.num_frames) return WEBPB200_ERR_NO_MORE_FRAMES;
    int st = 0;
    webpb200_ctx* c = default_ctx(&st);
    if (!c) return st;
    AnmfFrame fr;
    st = decode_and_composite(c, d->ct, d->next_frame_start, d->upsampling, d->has_bg, d->bg, d->dispose_next, d->pw, d->ph, d->px, d->py,
                              d->canvas, d->canvas_live, d->frame_out, buf, &fr);
    if (st) return st;
    d->pw = fr.w; d->ph = fr.h; d->px = fr.x; d->py = fr.y;
    d->dispose_next = fr.dispose;
    d->next_frame_start = fr.next;
    d->next_frame++;
    if (duration_ms) *duration_ms = fr.duration;
    return 0;
}

extern "C" int webpb200_reset_animation(webpb200_decoder* d) {
    if (!d->ct.animated) return WEBPB200_ERR_NOT_ANIMATED;
    d->next_frame = 0;
    d->next_frame_start = d->ct.first_anmf;
    d->dispose_next = true;
    return 0;
}

extern "C" int webpb200_read_image(webpb200_decoder* d, uint8_t* buf, size_t len) {
    if (len != webpb200_output_buffer_size(d)) return WEBPB200_ERR_IMAGE_TOO_LARGE;
    int st = 0;
    webpb200_ctx* c = default_ctx(&st);
    if (!c) return st;
    if (d->ct.animated) {  // decoder.rs:674-680: frame 1 on a fresh animation state
        DevBuf canvas, fout;
        bool live = false;
        AnmfFrame fr;
        st = decode_and_composite(c, d->ct, d->ct.first_anmf, d->upsampling, d->has_bg, d->bg, true, 0, 0, 0, 0, canvas, live, fout, buf, &fr);
        canvas.release();
        fout.release();
        return st;
    }
    const uint8_t* files[1] = {d->ct.data};
    const size_t lens[1] = {d->ct.len};
    uint8_t* outs[1] = {buf};
    const size_t out_lens[1] = {len};
    int status = 0;
    int rc = webpb200_decode_batch(c, files, lens, 1, outs, out_lens, &status, d->upsampling, 1);
    if (rc < 0) return -rc;
    return status;
}

extern "C" int webpb200_vp8_decode_frame(const uint8_t* vp8, size_t len, uint16_t* width, uint16_t* height, uint8_t* y, uint8_t* u, uint8_t* v) {
    Vp8Dims dims;
    int st = vp8_probe(vp8, len, &dims);
    if (st) return st;
    if (width) *width = (uint16_t)dims.width;
    if (height) *height = (uint16_t)dims.height;
    if (!y || !u || !v) return 0;
    webpb200_ctx* c = default_ctx(&st);
    if (!c) return st;
    std::vector<WorkItem> items(1);
    items[0].vp8 = vp8;
    items[0].vp8_len = len;
    items[0].bpp = 3;
    int rc = prepare_items(c, std::move(items), 0, 1);
    if (rc) return -rc;
    if (c->items[0].status) return c->items[0].status;
    if ((rc = upload(c)) != 0) return -rc;
    if ((rc = reconstruct(c, c->bd, 0x3, c->stream, nullptr)) != 0) return -rc;
    rc = webpb200_batch_get_planes(c, 0, y, u, v);
    return rc < 0 ? -rc : rc;
}

extern "C" int webpb200_decode_animation(webpb200_ctx* c, const uint8_t* file, size_t len, uint8_t* const* outs, size_t out_len,
                                         uint32_t* durations_ms, int max_frames, int upsampling, int host_threads) {
    webpb200_options opt;
    std::memset(&opt, 0, sizeof(opt));
    opt.lossy_upsampling = upsampling;
    return webpb200_decode_animation_opts(c, file, len, outs, out_len, durations_ms, max_frames, &opt, host_threads);
}

extern "C" int webpb200_decode_animation_opts(webpb200_ctx* c, const uint8_t* file, size_t len, uint8_t* const* outs, size_t out_len,
                                              uint32_t* durations_ms, int max_frames, const webpb200_options* opt, int host_threads) {
    if (!c) return -WEBPB200_ERR_NO_DEVICE;
    const int upsampling = opt ? opt->lossy_upsampling : 0;
    const bool has_bg = opt && opt->has_background_color;
    uint8_t bg[4] = {0, 0, 0, 0};
    if (has_bg) std::memcpy(bg, opt->background_color, 4);
    cudaSetDevice(c->device);
    Container ct;
    int st = parse_container(file, len, &ct);
    if (st) return -st;
    if (!ct.animated) return -WEBPB200_ERR_NOT_ANIMATED;
    const int out_bpp = ct.has_alpha ? 4 : 3;
    if (out_len != (size_t)ct.width * ct.height * out_bpp) return -WEBPB200_ERR_IMAGE_TOO_LARGE;
    const int nf = std::min<int>((int)ct.num_frames, max_frames);
    std::vector<AnmfFrame> frs((size_t)nf);
    std::vector<WorkItem> items((size_t)nf);
    uint64_t pos = ct.first_anmf;
    for (int f = 0; f < nf; f++) {
        st = parse_anmf(ct, pos, &frs[f]);
        if (st) return -st;
        if (frs[f].lossless) return -WEBPB200_ERR_NOT_LOSSY;
        WorkItem& it = items[f];
        it.vp8 = file + frs[f].vp8.start;
        it.vp8_len = (size_t)frs[f].vp8.size();
        it.check_w = (int)frs[f].w;
        it.check_h = (int)frs[f].h;
        it.bpp = 3;
        if (frs[f].alph.present) {
            it.alph = file + frs[f].alph.start;
            it.alph_len = (size_t)frs[f].alph.size();
            it.alph_w = (int)frs[f].w;
            it.alph_h = (int)frs[f].h;
            it.bpp = 4;
        }
        pos = frs[f].next;
    }
    // all frames reconstruct as one batch (every VP8 frame is a keyframe, lossy/mod.rs:513-518), through
    // the same pipelined path as still images with the pictures left on the device ...
    std::vector<int> fstatus((size_t)std::max(nf, 1), 0);
    int rc = decode_items(c, std::move(items), nullptr, nullptr, fstatus.data(), upsampling, host_threads);
    if (rc < 0) return rc;
    for (int f = 0; f < nf; f++)
        if (fstatus[f]) return -fstatus[f];
    // ... compositing is sequential over frames (decoder.rs:861-919)
    const size_t cbytes = (size_t)ct.width * ct.height * 4;
    CU(c, c->d_canvas.reserve(cbytes), "canvas");
    CU(c, c->d_frames_out.reserve(out_len * (size_t)nf), "canvas copies");
    if (has_bg) {  // decoder.rs:872-882: the canvas starts as the user's colour
        std::vector<uint8_t> fill(cbytes);
        for (size_t i = 0; i < cbytes; i += 4) std::memcpy(&fill[i], bg, 4);
        CU(c, cudaMemcpyAsync(c->d_canvas.p, fill.data(), cbytes, cudaMemcpyHostToDevice, c->stream), "canvas fill");
        CU(c, cudaStreamSynchronize(c->stream), "sync");  // `fill` goes away
    } else {
        CU(c, cudaMemsetAsync(c->d_canvas.p, 0, cbytes, c->stream), "canvas clear");
    }
    const webpb200_frame_desc* hf = reinterpret_cast<const webpb200_frame_desc*>(c->h_frames.p);
    bool dispose_next = true;
    uint32_t pw = 0, ph = 0, px = 0, py = 0;
    for (int f = 0; f < nf; f++) {
        CompositeArgs a;
        std::memset(&a, 0, sizeof(a));
        a.canvas = c->d_canvas.p;
        a.frame = c->d_out.p + hf[f].out_off;
        a.out = c->d_frames_out.p + out_len * (size_t)f;
        a.canvas_w = ct.width; a.canvas_h = ct.height;
        a.fx = frs[f].x; a.fy = frs[f].y; a.fw = frs[f].w; a.fh = frs[f].h;
        a.px = px; a.py = py; a.pw = pw; a.ph = ph;
        a.frame_has_alpha = frs[f].alph.present;
        if (dispose_next) {  // decoder.rs:861-869
            if (has_bg) { a.has_clear = 1; std::memcpy(a.clear, bg, 4); }
            else if (a.frame_has_alpha) a.has_clear = 1;
        }
        a.use_blend = frs[f].blend;
        a.out_bpp = out_bpp;
        CU(c, launch_composite(a, c->stream), "k_composite");
        c->launches++;
        pw = frs[f].w; ph = frs[f].h; px = frs[f].x; py = frs[f].y;
        dispose_next = frs[f].dispose;
        if (durations_ms) durations_ms[f] = frs[f].duration;
    }
    for (int f = 0; f < nf; f++)
        CU(c, cudaMemcpyAsync(outs[f], c->d_frames_out.p + out_len * (size_t)f, out_len, cudaMemcpyDeviceToHost, c->stream), "D2H canvases");
    CU(c, cudaStreamSynchronize(c->stream), "sync");
    return nf;
}
