/* include/webpb200.h -- C ABI of libwebpb200.so: B200-native lossy WebP (VP8) reconstruction.
 *
 * The reference (image-rs/image-webp v0.2.4) is a Rust crate with no FFI of its own; this
 * header is the C-ABI seam a Rust `extern "C"` block (or cgo / ctypes) binds to replace the
 * crate's lossy decode path.  Each entry point cites the reference item it stands in for.
 * Plain pointers and sizes only; no CUDA or torch types.  Return values, two conventions:
 *   - the WebPDecoder surface (webpb200_open ... webpb200_vp8_decode_frame) and the host-only helpers
 *     return 0 or a POSITIVE WEBPB200_ERR_* value mirroring DecodingError (decoder.rs:14-168);
 *   - the batched and device-resident entry points (webpb200_decode_batch, webpb200_batch_*,
 *     webpb200_decode_animation*, webpb200_reconstruct_device, webpb200_predict_device_static*) return
 *     0 / a non-negative count on success and a NEGATIVE -WEBPB200_ERR_* on a call-level failure.
 * Each declaration says which applies.  The library has NO CPU fallback: without a CUDA device every decode
 * call fails with WEBPB200_ERR_NO_DEVICE.
 */
#ifndef WEBPB200_H
#define WEBPB200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
    WEBPB200_OK = 0,
    WEBPB200_ERR_IO = 1,                 /* IoError / unexpected end of data      */
    WEBPB200_ERR_RIFF_SIGNATURE = 2,     /* ChunkHeaderInvalid("RIFF")            */
    WEBPB200_ERR_WEBP_SIGNATURE = 3,     /* WebpSignatureInvalid                  */
    WEBPB200_ERR_CHUNK_MISSING = 4,      /* ChunkMissing                          */
    WEBPB200_ERR_CHUNK_HEADER = 5,       /* ChunkHeaderInvalid(fourcc)            */
    WEBPB200_ERR_VP8_MAGIC = 6,          /* Vp8MagicInvalid                       */
    WEBPB200_ERR_UNSUPPORTED = 7,        /* UnsupportedFeature                    */
    WEBPB200_ERR_COLOR_SPACE = 8,        /* ColorSpaceInvalid                     */
    WEBPB200_ERR_BITSTREAM = 9,          /* BitStreamError                        */
    WEBPB200_ERR_INCONSISTENT_SIZE = 10, /* InconsistentImageSizes                */
    WEBPB200_ERR_IMAGE_TOO_LARGE = 11,   /* ImageTooLarge (also wrong buffer len) */
    WEBPB200_ERR_FRAME_OUTSIDE = 12,     /* FrameOutsideImage                     */
    WEBPB200_ERR_LOSSLESS_SIG = 13,      /* LosslessSignatureInvalid              */
    WEBPB200_ERR_VERSION = 14,           /* VersionNumberInvalid                  */
    WEBPB200_ERR_COLOR_CACHE = 15,       /* InvalidColorCacheBits                 */
    WEBPB200_ERR_HUFFMAN = 16,           /* HuffmanError                          */
    WEBPB200_ERR_TRANSFORM = 17,         /* TransformError                        */
    WEBPB200_ERR_ALPHA_PREPROC = 18,     /* InvalidAlphaPreprocessing             */
    WEBPB200_ERR_COMPRESSION = 19,       /* InvalidCompressionMethod              */
    WEBPB200_ERR_NO_MORE_FRAMES = 20,    /* NoMoreFrames                          */
    WEBPB200_ERR_CHUNK_SIZE = 21,        /* InvalidChunkSize                      */
    WEBPB200_ERR_NOT_ANIMATED = 22,      /* reference panics, decoder.rs:753      */
    WEBPB200_ERR_INVALID_PARAM = 23,     /* InvalidParameter                      */
    WEBPB200_ERR_NOMEM = 24,
    WEBPB200_ERR_NOT_LOSSY = 25,         /* VP8L still/animation frame: outside this library's path */
    WEBPB200_ERR_MEMORY_LIMIT = 26,      /* MemoryLimitExceeded (decoder.rs:154)  */
    WEBPB200_ERR_NO_DEVICE = 100,        /* no usable CUDA device                 */
    WEBPB200_ERR_CUDA = 101,             /* a CUDA call failed (see webpb200_last_cuda_error) */
    WEBPB200_ERR_STATIC_SCRATCH_LIMIT = 102 /* batch too large for webpb200_predict_device_static */
};

enum { WEBPB200_UPSAMPLING_BILINEAR = 0, WEBPB200_UPSAMPLING_SIMPLE = 1 }; /* decoder.rs:297-309 */

/* ---------------------------------------------------------------------------------------
 * Packed arrays at the host -> device boundary.  The host entropy decoder fills them; the
 * kernels read them.  All little-endian, naturally aligned.
 * ------------------------------------------------------------------------------------- */

/* One quantiser/loop-filter segment, common.rs:718-733 (quantisers already resolved per
 * lossy/mod.rs:397-451). */
typedef struct {
    int16_t ydc, yac, y2dc, y2ac, uvdc, uvac;
    int8_t lf_level;       /* Segment::loopfilter_level */
    uint8_t delta_values;  /* Segment::delta_values     */
    uint8_t pad[2];
} webpb200_segment;        /* 16 bytes */

/* One VP8 keyframe of the batch (Frame + the header fields reconstruction needs,
 * lossy/mod.rs:132-167, 262-268). */
typedef struct {
    uint16_t width, height, mbw, mbh;
    uint8_t filter_type;       /* 1 = simple, 0 = normal (mod.rs:164)  */
    uint8_t filter_level;      /* frame level 0..63                    */
    uint8_t sharpness;         /* 0..7                                 */
    uint8_t segments_enabled;
    uint8_t lf_adj_enabled;    /* loop_filter_adjustments_enabled      */
    int8_t ref_delta0;         /* ref_delta[0]  (mod.rs:1211)          */
    int8_t mode_delta0;        /* mode_delta[0] (mod.rs:1213)          */
    uint8_t out_bpp;           /* 3 = RGB, 4 = RGBA                    */
    uint8_t upsampling;        /* WEBPB200_UPSAMPLING_*                */
    uint8_t alpha_mode;        /* 0 none | 1 ALPH plane present (filter in alpha_filter) */
    uint8_t alpha_filter;      /* FilteringMethod 0..3 (extended.rs:245-251) */
    uint8_t pad0;
    webpb200_segment seg[4];
    uint32_t mb_base;          /* first macroblock in the MbDesc / level arrays      */
    uint32_t row_base;         /* first macroblock row in the row-progress arrays    */
    uint64_t y_off, u_off, v_off; /* byte offsets of the planes in the plane arena   */
    uint64_t out_off;          /* byte offset of the output image in the output arena */
    uint64_t alpha_off;        /* byte offset of the w*h alpha residual plane        */
} webpb200_frame_desc;         /* 136 bytes */

/* One macroblock as decoded by the host (MacroBlock, lossy/mod.rs:112-120). */
typedef struct {
    uint8_t ymode;      /* LumaMode: 0 DC 1 V 2 H 3 TM 4 B (common.rs:44-61)            */
    uint8_t uvmode;     /* ChromaMode: 0 DC 1 V 2 H 3 TM                                */
    uint8_t segment;    /* segmentid 0..3                                               */
    uint8_t flags;      /* bit0 = coeffs_skipped                                        */
    uint32_t nflags;    /* bit b: read_coefficients returned true for block b; 24 = Y2  */
    uint8_t bpred[8];   /* 16 IntraMode nibbles (common.rs:115-129): sub-block k in
                           bits 4*(k&1).. of byte k>>1                                 */
} webpb200_mb_desc;     /* 16 bytes */

/* Levels: int16 [macroblock][25][16], raw token values before dequantisation, de-zigzagged
 * (raster inside a 4x4 block); blocks 0-15 Y, 16-19 U, 20-23 V, 24 Y2. 800 bytes per MB. */
#define WEBPB200_LEVELS_PER_MB 400

/* ---------------------------------------------------------------------------------------
 * WebPDecoder surface (decoder.rs:312-940).  `data` is borrowed and must outlive the
 * decoder.  One decoder is single-threaded; separate decoders are independent.
 * ------------------------------------------------------------------------------------- */
typedef struct webpb200_decoder webpb200_decoder;

/* WebPDecodeOptions (decoder.rs:274-290) plus the two per-decoder settings the crate offers as setters, so
 * that one struct configures a decoder or a whole-animation call.  Zero-initialise for the defaults. */
typedef struct {
    int32_t lossy_upsampling;      /* WEBPB200_UPSAMPLING_*; default Bilinear                                  */
    uint8_t has_background_color;  /* 1: use background_color as set_background_color would (decoder.rs:568)   */
    uint8_t background_color[4];   /* RGBA                                                                      */
    uint8_t pad[3];
    uint64_t memory_limit;         /* set_memory_limit (decoder.rs:554); 0 = unlimited (the crate's usize::MAX) */
} webpb200_options;                /* 24 bytes */

/* WebPDecoder::new_with_options, decoder.rs:342-363 (container scan only, no pixel work).  Positive codes. */
int webpb200_open(const uint8_t* data, size_t len, int upsampling, webpb200_decoder** out);
/* the same with the full option set; a background colour on a file that is not an extended (VP8X) one is
 * WEBPB200_ERR_INVALID_PARAM, as set_background_color would answer */
int webpb200_open_with_options(const uint8_t* data, size_t len, const webpb200_options* opt, webpb200_decoder** out);
void webpb200_close(webpb200_decoder* d);
void webpb200_dimensions(const webpb200_decoder* d, uint32_t* w, uint32_t* h); /* decoder.rs:580 */
int webpb200_has_alpha(const webpb200_decoder* d);        /* decoder.rs:586 */
int webpb200_is_animated(const webpb200_decoder* d);      /* decoder.rs:591 */
int webpb200_is_lossy(const webpb200_decoder* d);         /* decoder.rs:599 */
uint32_t webpb200_num_frames(const webpb200_decoder* d);  /* decoder.rs:605 */
uint32_t webpb200_loop_count(const webpb200_decoder* d);  /* decoder.rs:610; 0 = forever */
uint64_t webpb200_loop_duration(const webpb200_decoder* d); /* decoder.rs:618 */
size_t webpb200_output_buffer_size(const webpb200_decoder* d); /* decoder.rs:659-664 */
int webpb200_set_background_color(webpb200_decoder* d, const uint8_t rgba[4]); /* decoder.rs:568 */
void webpb200_set_lossy_upsampling(webpb200_decoder* d, int upsampling);       /* decoder.rs:938 */
void webpb200_set_memory_limit(webpb200_decoder* d, size_t limit);             /* decoder.rs:554 */
/* background_color_hint, decoder.rs:559-565: 0 and the ANIM chunk's colour, or WEBPB200_ERR_INVALID_PARAM
 * where the crate returns None (not an extended file) */
int webpb200_background_color_hint(const webpb200_decoder* d, uint8_t rgba[4]);
/* icc_profile / exif_metadata / xmp_metadata, decoder.rs:643-655 (read_chunk 620-640): *data points into the
 * caller's file buffer (borrowed, not copied), NULL when the file has no such chunk; a chunk larger than the
 * memory limit is WEBPB200_ERR_MEMORY_LIMIT.  Positive codes. */
int webpb200_icc_profile(webpb200_decoder* d, const uint8_t** data, size_t* len);
int webpb200_exif_metadata(webpb200_decoder* d, const uint8_t** data, size_t* len);
int webpb200_xmp_metadata(webpb200_decoder* d, const uint8_t** data, size_t* len);
/* WebPDecoder::read_image, decoder.rs:669-741: tightly packed RGB8 / RGBA8 into buf;
 * len must equal output_buffer_size() else WEBPB200_ERR_IMAGE_TOO_LARGE. */
int webpb200_read_image(webpb200_decoder* d, uint8_t* buf, size_t len);
/* WebPDecoder::read_frame, decoder.rs:752-922: next composited canvas + its duration. */
int webpb200_read_frame(webpb200_decoder* d, uint8_t* buf, size_t len, uint32_t* duration_ms);
int webpb200_reset_animation(webpb200_decoder* d);        /* decoder.rs:929-935 */
/* vp8::Vp8Decoder::decode_frame, lossy/mod.rs:1247 (lib.rs:34-36): the loop-filtered,
 * macroblock-padded planes of one VP8 chunk payload.  y: mbw*16 x mbh*16, u/v: mbw*8 x mbh*8.
 * Pass NULL planes to query the sizes only. */
int webpb200_vp8_decode_frame(const uint8_t* vp8, size_t len, uint16_t* width, uint16_t* height,
                              uint8_t* y, uint8_t* u, uint8_t* v);

/* ---------------------------------------------------------------------------------------
 * Batched path (new; the reference decodes one image per call).  A context owns one CUDA
 * device's stream, pinned staging and device arenas; they grow on demand and are reused.
 * ------------------------------------------------------------------------------------- */
typedef struct webpb200_ctx webpb200_ctx;

int webpb200_device_count(void);                       /* 0 when no CUDA device is usable */
int webpb200_ctx_create(int device, webpb200_ctx** out);
/* PCI address of a device ("0000:1b:00.0"), so that a host process can place its threads and its pinned
 * buffers on the NUMA node the GPU hangs off (see sharding.bind_near_gpu). */
int webpb200_device_pci_bus_id(int device, char* buf, int len);
void webpb200_ctx_destroy(webpb200_ctx* c);
const char* webpb200_last_cuda_error(const webpb200_ctx* c);
/* Page-locked host memory for input/output buffers (copies to and from it run asynchronously
 * at full PCIe rate; pageable buffers work too, only slower).  NULL on failure. */
void* webpb200_host_alloc(size_t bytes);
void webpb200_host_free(void* p);

/* Where webpb200_decode_batch runs the boolean-entropy decode of macroblock headers and
 * coefficient tokens (lossy/mod.rs:615-671, 785-963; arithmetic_decoder.rs):
 *   HOST   one frame per host thread, packed levels copied to the device;
 *   DEVICE one warp per bitstream partition on the GPU, all frames of the batch side by side;
 *          only the compressed chunks cross PCIe.  Same results, bit for bit;
 *   AUTO   (the default) per batch, whichever a cost model expects to finish first: the device
 *          for many small frames or many partitions, the host for a few large frames.
 * The default can also be set with the environment variable WEBPB200_ENTROPY=host|device|auto. */
#define WEBPB200_ENTROPY_HOST 0
#define WEBPB200_ENTROPY_DEVICE 1
#define WEBPB200_ENTROPY_AUTO 2
int webpb200_ctx_set_entropy(webpb200_ctx* c, int where);
/* No counterpart in the reference (which fills host slices).  webpb200_decode_batch with outs = NULL
 * (out_lens is then ignored) leaves the decoded images in device memory; this gives the device
 * address, byte length (tightly packed RGB8 / RGBA8, as read_image would write it) and geometry of
 * image i.  The memory belongs to the context and is reused by its next batch call.  Returns the
 * image's status if it failed to decode. */
int webpb200_batch_output_device(webpb200_ctx* c, int image, void** d_ptr, size_t* len, uint32_t* width,
                                 uint32_t* height, int* bpp);
/* What the last webpb200_decode_batch on this context moved over PCIe (bytes each way) and
 * where its entropy stage ran (WEBPB200_ENTROPY_HOST or _DEVICE); for bench accounting. */
int webpb200_last_transfer(const webpb200_ctx* c, uint64_t* h2d_bytes, uint64_t* d2h_bytes, int* entropy_where);
/* Device time (CUDA events, ms) of the entropy kernel in the last webpb200_decode_batch; 0 if it ran on the host. */
float webpb200_last_entropy_ms(const webpb200_ctx* c);
/* After a decode_batch / batch_upload: copies one image's macroblock descriptors and levels as
 * the device holds them (mbw*mbh descriptors, mbw*mbh*400 levels) -- for parity tests of the
 * device entropy stage against webpb200_host_entropy_decode. */
int webpb200_batch_download_packed(webpb200_ctx* c, int image, webpb200_mb_desc* mbs, int16_t* levels);

/* N x (WebPDecoder::new + read_image) for still lossy files: container and frame-header parse on
 * `host_threads` threads, entropy decode on the device or on those threads (see
 * webpb200_ctx_set_entropy), reconstruction kernels, download into outs[i] (out_lens[i] must be
 * width*height*(3|4)); outs = NULL leaves the images on the device (webpb200_batch_output_device).
 * status[i] gets each file's own code; a bad file does not poison the batch, and outs[i] of a
 * failed file holds no valid image.  Returns the number of failed files, or a negative
 * WEBPB200_ERR_* for a batch-level failure (no device, CUDA error).  The stages overlap most when
 * the files lie in one block from webpb200_host_alloc and the outs of same-size images are evenly
 * spaced; neither is required. */
int webpb200_decode_batch(webpb200_ctx* c, const uint8_t* const* files, const size_t* lens, int n,
                          uint8_t* const* outs, const size_t* out_lens, int* status,
                          int upsampling, int host_threads);

/* webpb200_decode_batch without blocking the caller, so that two batches can be in flight on one GPU: the
 * boolean-entropy kernel of a batch is one long dependency chain per frame and leaves about half of the
 * GPU's issue slots idle, which the next batch's chains can use while this batch reconstructs and
 * downloads.  _begin starts the call on a library thread and returns at once; _end waits for it and returns
 * what webpb200_decode_batch would have (failed-file count, or a negative code).  Every argument array must
 * stay valid until _end.  One call in flight per context: use two contexts on the device (each owns its
 * streams and arenas) and alternate between them.  No counterpart in the reference (a WebPDecoder is
 * synchronous); independent decoders on separate threads are its equivalent (decoder.rs:312-331). */
typedef struct webpb200_pending webpb200_pending;
int webpb200_decode_batch_begin(webpb200_ctx* c, const uint8_t* const* files, const size_t* lens, int n,
                                uint8_t* const* outs, const size_t* out_lens, int* status,
                                int upsampling, int host_threads, webpb200_pending** pending);
int webpb200_decode_batch_end(webpb200_pending* pending);

/* The batch split over several contexts (normally one per GPU of the box; SURVEY 8(e): images are
 * independent, decoder.rs:312-331, so there is no exchange between the devices).  Files are dealt out as
 * contiguous index ranges with about equal macroblock counts (sizes come from a header probe), each
 * context runs webpb200_decode_batch on its range from its own host thread with host_threads / n_ctx
 * worker threads, and outs / status are filled exactly as one webpb200_decode_batch over everything would
 * fill them.  split (optional, n_ctx + 1 entries) receives the range boundaries.  Returns the number of
 * failed files, or a negative -WEBPB200_ERR_* if any context failed at call level. */
int webpb200_decode_batch_multi(webpb200_ctx* const* ctxs, int n_ctx, const uint8_t* const* files, const size_t* lens, int n,
                                uint8_t* const* outs, const size_t* out_lens, int* status, int upsampling,
                                int host_threads, int* split);

/* The same in stages, so the device-resident part can be timed alone:
 *   prepare  = host side of decode_frame (header + per-macroblock entropy decode,
 *              lossy/mod.rs:510-671, 785-963) into pinned packed arrays
 *   upload   = one async copy per array to the device
 *   reconstruct = everything after entropy decoding, on the device (see below)
 *   download = output arena back to outs[]                                             */
int webpb200_batch_prepare(webpb200_ctx* c, const uint8_t* const* files, const size_t* lens, int n,
                           int* status, int upsampling, int host_threads);
int webpb200_batch_upload(webpb200_ctx* c);
/* stages: bit0 dequant+IWHT/IDCT+intra prediction (mod.rs:673-783, 866, 898-952),
 * bit1 loop filter (mod.rs:966-1244), bit2 YUV->RGB(A) (mod.rs:184-241, yuv.rs),
 * bit3 alpha un-filter + merge (decoder.rs:718-734); 0xF = all.  Asynchronous.
 * The ALPH stage (bit3) un-filters into a planar alpha arena on a second stream beside bit0 / bit1, and
 * the colour conversion of the SAME call (bit2) puts the plane into byte 3 of the RGBA pixels it
 * writes: bit3 without bit2 prepares the planes only, bit2 without bit3 leaves byte 3 at 255. */
int webpb200_batch_reconstruct(webpb200_ctx* c, unsigned stages);
int webpb200_batch_sync(webpb200_ctx* c);
int webpb200_batch_download(webpb200_ctx* c, uint8_t* const* outs, const size_t* out_lens);
/* Runs reconstruct(stages) `iters` times back to back and returns the mean device time of
 * one pass in milliseconds, measured with CUDA events on the context's stream.  If
 * flush_l2 != 0 a buffer larger than L2 is overwritten between passes (outside the
 * timed spans).  kernel_ms (may be NULL) receives per-kernel mean times [4]: prediction, loop filter,
 * colour conversion, ALPH stage (which runs beside the first two: the four need not add up to the pass). */
int webpb200_batch_time(webpb200_ctx* c, unsigned stages, int iters, int flush_l2, float* ms_per_pass,
                        float* kernel_ms);
/* introspection for tests and bench */
int webpb200_batch_size(const webpb200_ctx* c, int* n_images, uint64_t* n_macroblocks,
                        uint64_t* n_pixels, uint64_t* in_bytes, uint64_t* out_bytes);
/* planes of image i as they are on the device now (after stage bit0: unfiltered, after
 * bit1: loop-filtered); macroblock-padded sizes as for webpb200_vp8_decode_frame. */
int webpb200_batch_get_planes(webpb200_ctx* c, int image, uint8_t* y, uint8_t* u, uint8_t* v);
/* host-side packed arrays of image i (pointers into pinned staging, valid until the next
 * prepare): lets tests diff the boundary arrays against the oracle. */
int webpb200_batch_get_packed(webpb200_ctx* c, int image, const webpb200_frame_desc** frame,
                              const webpb200_mb_desc** mbs, const int16_t** levels);
/* number of kernel launches issued by this context so far (bench.py's gpu_launches) */
uint64_t webpb200_launch_count(const webpb200_ctx* c);

/* The host half alone, without touching CUDA: container walk + everything Vp8Decoder reads
 * from the bitstream (lossy/mod.rs:510-671, 785-963) for one still lossy file, into caller
 * arrays; also the ALPH residual plane (extended.rs:253-308) when the file has one.  Call with
 * mbs = NULL to learn the sizes (fd->mbw * fd->mbh macroblocks, width*height alpha bytes). */
int webpb200_host_entropy_decode(const uint8_t* file, size_t len, webpb200_frame_desc* fd,
                                 webpb200_mb_desc* mbs, size_t mbs_cap, int16_t* levels, size_t levels_cap,
                                 uint8_t* alpha, size_t alpha_cap);
/* Evaluates the kernels' 4x4 sub-block prediction table on the CPU (table self-test):
 * edge13 = {P, A0..A7, L0..L3} (prediction.rs:290-337), mode = IntraMode 1..9. */
void webpb200_debug_bpred_eval(int mode, const uint8_t* edge13, uint8_t* out16);
/* webpb200_host_entropy_decode (VP8 part only) done the way the device entropy stage is fed: frame
 * header into the hand-over structure, then the macroblock loop from that structure alone with the
 * boolean decoder resumed from the exported state.  Same outputs and statuses: a CPU-side check of
 * the hand-over (csrc/entropy.h).  mbs_cap / levels_cap as in webpb200_host_entropy_decode. */
int webpb200_debug_entropy_resumed(const uint8_t* file, size_t len, webpb200_frame_desc* fd,
                                   webpb200_mb_desc* mbs, size_t mbs_cap, int16_t* levels, size_t levels_cap);
/* Test hook for the device entropy stage (process-wide, 0 = none).  Bit 0: in the layout that
 * webpb200_decode_batch_begin uses for single-partition batches (mode partitions and token partitions
 * in separate thread blocks), the mode blocks do nothing -- as if the hardware had never started them;
 * every token warp must then notice, claim its frame's mode partition and decode it itself.  The
 * results must not change. */
void webpb200_debug_entropy_faults(int flags);

/* Device-resident seam: reconstruct a batch whose packed arrays the caller already holds
 * in device memory (all pointers are device pointers; d_planes / d_out / d_scratch sized
 * from the frame descriptors by webpb200_workspace_sizes).  `stream` is a cudaStream_t
 * passed as void*.  Replaces, for n frames at once, the device half of
 * Vp8Decoder::decode_frame + Frame::fill_rgb/fill_rgba (+ the ALPH merge loop). */
int webpb200_workspace_sizes(const webpb200_frame_desc* frames, int n, uint64_t* plane_bytes,
                             uint64_t* out_bytes, uint64_t* scratch_bytes);
int webpb200_reconstruct_device(const webpb200_frame_desc* h_frames, const webpb200_frame_desc* d_frames,
                                int n, const webpb200_mb_desc* d_mbs, const int16_t* d_levels,
                                const uint8_t* d_alpha, uint8_t* d_planes, uint8_t* d_out,
                                void* d_scratch, unsigned stages, void* stream);

/* No counterpart in the reference (which allocates freely on the CPU heap).
 *
 * webpb200_reconstruct_device with stages = 1 (dequantisation, inverse transforms, intra
 * prediction into d_planes) for callers that must not allocate device memory: a process that
 * fixes its GPU memory budget at start-up, or one living inside another runtime's memory pool,
 * where a cudaMalloc would stall the device, dodge the pool's accounting or fail late.  The
 * wavefront's working memory (ticket counter, the records through which a macroblock row hands its
 * bottom edge to the row below, row map, per-macroblock flags) lives in a fixed array inside the
 * library instead of a caller-supplied scratch arena, and this call performs no allocation of any
 * kind.  The price:
 *   - every GPU that loads the library carries that array (about 1.4 MB) until the process exits;
 *   - a call may cover at most WEBPB200_STATIC_MAX_ROWS macroblock rows, WEBPB200_STATIC_MAX_MBS
 *     macroblocks and WEBPB200_STATIC_MAX_WIDTH_MBS macroblock columns summed over its frames
 *     (e.g. 256 images of 512x512); anything larger returns -WEBPB200_ERR_STATIC_SCRATCH_LIMIT and
 *     does nothing -- split the batch;
 *   - all calls on one GPU share the array and it keeps the last call's contents; the call clears
 *     what it needs cleared, but two calls that overlap in time on the same GPU (two streams, two
 *     threads) race on it and produce garbage.  Serialise them;
 *   - the first call on each GPU resolves the array's device address (one-time set-up, with
 *     that GPU current).  Make that first call before you start any stream capture.
 * h_frames / d_frames: the same descriptors in host and device memory (offsets as for
 * webpb200_reconstruct_device; every frame needs at least one macroblock); d_nonzero (optional, may
 * be NULL): receives one byte per macroblock, 1 where the loop filter must treat the macroblock as
 * having coefficients (MacroBlock::non_zero_dct).  Asynchronous on `stream` (a cudaStream_t passed
 * as void*).  Returns 0 or a negative WEBPB200_ERR_*, like webpb200_reconstruct_device.
 * webpb200_predict_device_static_check waits for `stream` and returns 0 if the last call on the
 * current GPU ran to completion, -WEBPB200_ERR_CUDA if its wavefront gave up (overlapping calls) or
 * the stream failed; it allocates nothing either. */
#define WEBPB200_STATIC_MAX_ROWS 8192
#define WEBPB200_STATIC_MAX_MBS 262144
#define WEBPB200_STATIC_MAX_WIDTH_MBS 8192
int webpb200_predict_device_static(const webpb200_frame_desc* h_frames, const webpb200_frame_desc* d_frames,
                                   int n, const webpb200_mb_desc* d_mbs, const int16_t* d_levels,
                                   uint8_t* d_planes, uint8_t* d_nonzero, void* stream);
int webpb200_predict_device_static_check(void* stream);

/* All frames of one animated lossy file, composited (extended.rs:30-129), as read_frame
 * would return them one after another: frames reconstruct as one batch, compositing runs
 * frame by frame on the device.  outs[f] = canvas after frame f. */
int webpb200_decode_animation(webpb200_ctx* c, const uint8_t* file, size_t len, uint8_t* const* outs,
                              size_t out_len, uint32_t* durations_ms, int max_frames, int upsampling,
                              int host_threads);
/* The same with webpb200_options: the background colour is honoured exactly as read_frame does after
 * set_background_color (initial canvas decoder.rs:872-882, clear colour 861-869).  Both return the number of
 * frames written (>= 0) or a negative -WEBPB200_ERR_*. */
int webpb200_decode_animation_opts(webpb200_ctx* c, const uint8_t* file, size_t len, uint8_t* const* outs,
                                   size_t out_len, uint32_t* durations_ms, int max_frames,
                                   const webpb200_options* opt, int host_threads);

#ifdef __cplusplus
}
#endif
#endif
