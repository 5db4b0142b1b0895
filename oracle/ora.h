/* oracle/ora.h -- TEST INFRASTRUCTURE, not product code.
 *
 * CPU oracle: a plain-C restatement of the image-webp (v0.2.4) decoder for the lossy
 * VP8 path, its container/ALPH/animation callers and the VP8L decoder that compressed
 * ALPH chunks need.  Each function cites the reference file:line it follows.
 *
 * Parity pin: checked against every lossy golden image of the reference's own tests
 * (tests/decode.rs:190-203 -> tests/golden/fixtures.npz) and its unit KATs; see
 * tests/test_oracle_golden.py.  The reference itself (Rust) cannot be built here.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library.  The product (image-webp_b200/) never does.
 */
#ifndef ORA_H
#define ORA_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* status codes: 0 ok; the rest mirror DecodingError variants (decoder.rs:14-168) */
enum {
    ORA_OK = 0,
    ORA_ERR_IO = 1,                 /* IoError (unexpected EOF)             */
    ORA_ERR_RIFF_SIGNATURE = 2,     /* ChunkHeaderInvalid("RIFF")           */
    ORA_ERR_WEBP_SIGNATURE = 3,     /* WebpSignatureInvalid                 */
    ORA_ERR_CHUNK_MISSING = 4,      /* ChunkMissing                         */
    ORA_ERR_CHUNK_HEADER = 5,       /* ChunkHeaderInvalid(other)            */
    ORA_ERR_VP8_MAGIC = 6,          /* Vp8MagicInvalid                      */
    ORA_ERR_UNSUPPORTED = 7,        /* UnsupportedFeature (non-keyframe)    */
    ORA_ERR_COLOR_SPACE = 8,        /* ColorSpaceInvalid                    */
    ORA_ERR_BITSTREAM = 9,          /* BitStreamError                       */
    ORA_ERR_INCONSISTENT_SIZE = 10, /* InconsistentImageSizes               */
    ORA_ERR_IMAGE_TOO_LARGE = 11,   /* ImageTooLarge (also wrong buf size)  */
    ORA_ERR_FRAME_OUTSIDE = 12,     /* FrameOutsideImage                    */
    ORA_ERR_LOSSLESS_SIG = 13,      /* LosslessSignatureInvalid             */
    ORA_ERR_VERSION = 14,           /* VersionNumberInvalid                 */
    ORA_ERR_COLOR_CACHE = 15,       /* InvalidColorCacheBits                */
    ORA_ERR_HUFFMAN = 16,           /* HuffmanError                         */
    ORA_ERR_TRANSFORM = 17,         /* TransformError                       */
    ORA_ERR_ALPHA_PREPROC = 18,     /* InvalidAlphaPreprocessing            */
    ORA_ERR_COMPRESSION = 19,       /* InvalidCompressionMethod             */
    ORA_ERR_NO_MORE_FRAMES = 20,    /* NoMoreFrames                         */
    ORA_ERR_CHUNK_SIZE = 21,        /* InvalidChunkSize                     */
    ORA_ERR_NOT_ANIMATED = 22,      /* reference panics (decoder.rs:753)    */
    ORA_ERR_INVALID_PARAM = 23,     /* InvalidParameter                     */
    ORA_ERR_NOMEM = 24,
    ORA_ERR_MEMORY_LIMIT = 26       /* MemoryLimitExceeded (decoder.rs:154)  */
};

/* ---- VP8 keyframe (lossy/mod.rs) ------------------------------------------------- */

typedef struct {
    int16_t ydc, yac, y2dc, y2ac, uvdc, uvac; /* common.rs:718-733 */
    int8_t quantizer_level, loopfilter_level;
    uint8_t delta_values;
} ora_segment;

/* One decoded macroblock, lossy/mod.rs:112-120, plus what the GPU boundary needs. */
typedef struct {
    uint8_t luma_mode;      /* 0 DC 1 V 2 H 3 TM 4 B */
    uint8_t chroma_mode;    /* 0 DC 1 V 2 H 3 TM     */
    uint8_t segmentid;
    uint8_t coeffs_skipped;
    uint8_t non_zero_dct;
    uint8_t bpred[16];      /* 0 DC 1 TM 2 VE 3 HE 4 LD 5 RD 6 VR 7 VL 8 HD 9 HU */
    uint32_t nflags;        /* bit b: read_coefficients returned true for block b (24 = Y2) */
} ora_mb;

typedef struct {
    int width, height, mbw, mbh;
    uint8_t *ybuf, *ubuf, *vbuf;       /* loop-filtered, MB-padded (mod.rs:546-549)   */
    uint8_t *ybuf0, *ubuf0, *vbuf0;    /* unfiltered copies (only if keep != 0)       */
    int version, for_display, pixel_type;
    int filter_type;                   /* 1 = simple (mod.rs:164)                     */
    int filter_level, sharpness_level;
    int segments_enabled, segments_update_map;
    int lf_adj_enabled;
    int ref_delta[4], mode_delta[4];
    int num_partitions;
    int prob_skip_false;               /* -1 = None */
    ora_segment segment[4];
    ora_mb* mbs;                       /* mbw*mbh, raster (only if keep != 0)         */
    int16_t* levels;                   /* [mb][25][16] raw token values, de-zigzagged;
                                          blocks 0-15 Y, 16-19 U, 20-23 V, 24 Y2       */
    int32_t* residue;                  /* [mb][24][16] post-IDCT residuals            */
} ora_frame;

/* Vp8Decoder::decode_frame, lossy/mod.rs:1247-1297.  data = VP8 chunk payload. */
int ora_vp8_decode(const uint8_t* data, size_t len, int keep, ora_frame* out);
void ora_frame_free(ora_frame* f);
/* Frame::fill_rgb / fill_rgba, lossy/mod.rs:184-241 (bpp 3|4; fancy 1 = Bilinear). */
void ora_frame_fill(const ora_frame* f, uint8_t* buf, int bpp, int fancy);

/* stage-level entry points for unit KATs and per-kernel diffs */
void ora_idct4x4(int32_t* block);            /* transform.rs:7-51   */
void ora_iwht4x4(int32_t* block);            /* transform.rs:54-86  */
void ora_fill_rgb_buffer_fancy(uint8_t* buf, const uint8_t* y, const uint8_t* u,
                               const uint8_t* v, int width, int height, int buffer_width,
                               int bpp);      /* yuv.rs:76-152 */
void ora_fill_rgb_buffer_simple(uint8_t* buf, size_t buf_len, const uint8_t* y,
                                const uint8_t* u, const uint8_t* v, int width,
                                int chroma_width, int buffer_width, int nrows_planes,
                                int bpp);     /* yuv.rs:293-382 */
void ora_yuv_to_rgb(uint8_t y, uint8_t u, uint8_t v, uint8_t* rgb); /* yuv.rs:59-71 */
/* predictors on a caller-made workspace (prediction.rs:159-499); mode = IntraMode */
void ora_predict_bmode(uint8_t* ws, int stride, int x0, int y0, int mode);
void ora_add_residue(uint8_t* ws, const int32_t* rb, int y0, int x0, int stride);
/* loop-filter primitives, loop_filter.rs:144-295; horizontal = contiguous 8 px */
void ora_lf_simple_h(int edge_limit, uint8_t* px);
void ora_lf_simple_v(int edge_limit, uint8_t* px, size_t point, size_t stride);
void ora_lf_subblock_h(int hev, int interior, int edge, uint8_t* px);
void ora_lf_subblock_v(int hev, int interior, int edge, uint8_t* px, size_t point, size_t stride);
void ora_lf_macroblock_h(int hev, int interior, int edge, uint8_t* px);
void ora_lf_macroblock_v(int hev, int interior, int edge, uint8_t* px, size_t point, size_t stride);

/* bool decoder KAT surface (arithmetic_decoder.rs) */
typedef struct ora_bool ora_bool;
ora_bool* ora_bool_new(const uint8_t* data, size_t len);
void ora_bool_free(ora_bool*);
int ora_bool_read_bool(ora_bool*, int prob);
int ora_bool_read_flag(ora_bool*);
int ora_bool_read_literal(ora_bool*, int n);
int ora_bool_past_eof(ora_bool*);

/* ---- VP8L (lossless/decoder/ *.rs) -- used for compressed ALPH and lossless files --- */
/* LosslessDecoder::decode_frame, lossless/decoder/mod.rs:67-152. buf = w*h*4 RGBA. */
int ora_vp8l_decode(const uint8_t* data, size_t len, uint32_t width, uint32_t height,
                    int implicit_dimensions, uint8_t* buf);

/* ---- extended.rs / alpha_blending.rs ---------------------------------------------- */
/* read_alpha_chunk, extended.rs:253-308: out = w*h alpha residuals, *filter = 0..3 */
int ora_read_alpha_chunk(const uint8_t* data, size_t len, int width, int height,
                         uint8_t* out, int* filter);
/* merge loop decoder.rs:718-734 with get_alpha_predictor extended.rs:131-191 */
void ora_alpha_merge(uint8_t* rgba, const uint8_t* alpha_data, int width, int height,
                     int filter);
/* composite_frame, extended.rs:30-129 */
void ora_composite_frame(uint8_t* canvas, uint32_t canvas_w, uint32_t canvas_h,
                         int has_clear, const uint8_t clear_color[4], const uint8_t* frame,
                         uint32_t fx, uint32_t fy, uint32_t fw, uint32_t fh,
                         int frame_has_alpha, int use_blend, uint32_t pw, uint32_t ph,
                         uint32_t px, uint32_t py);
/* do_alpha_blending, alpha_blending.rs:58-64 */
void ora_blend_pixel(const uint8_t src[4], const uint8_t dst[4], uint8_t out[4]);

/* ---- WebPDecoder (decoder.rs) ------------------------------------------------------- */
typedef struct ora_decoder ora_decoder;
/* new_with_options, decoder.rs:342-363; data is borrowed. upsampling: 0 Bilinear 1 Simple */
int ora_open(const uint8_t* data, size_t len, int upsampling, ora_decoder** out);
void ora_close(ora_decoder*);
void ora_dimensions(const ora_decoder*, uint32_t* w, uint32_t* h);
int ora_has_alpha(const ora_decoder*);
int ora_is_animated(const ora_decoder*);
int ora_is_lossy(const ora_decoder*);
uint32_t ora_num_frames(const ora_decoder*);
uint32_t ora_loop_count(const ora_decoder*);     /* 0 = forever */
uint64_t ora_loop_duration(const ora_decoder*);
size_t ora_output_buffer_size(const ora_decoder*);
int ora_set_background_color(ora_decoder*, const uint8_t color[4]);
/* set_memory_limit decoder.rs:554; background_color_hint 559-565 (returns 0 and fills rgba, or
 * ORA_ERR_INVALID_PARAM for "None"); icc_profile / exif_metadata / xmp_metadata 643-655 through read_chunk
 * 620-640: *data = NULL when the chunk is absent, ORA_ERR_MEMORY_LIMIT when it is larger than the limit */
void ora_set_memory_limit(ora_decoder*, size_t limit);
int ora_background_color_hint(const ora_decoder*, uint8_t rgba[4]);
int ora_metadata(ora_decoder*, int which /* 0 ICCP 1 EXIF 2 XMP */, const uint8_t** data, size_t* len);
void ora_set_lossy_upsampling(ora_decoder*, int upsampling);
int ora_read_image(ora_decoder*, uint8_t* buf, size_t len);
int ora_read_frame(ora_decoder*, uint8_t* buf, size_t len, uint32_t* duration_ms);
int ora_reset_animation(ora_decoder*);

/* CPU baseline helper: decode n files with `threads` worker threads, one image per
 * thread at a time (bench.py cpu_baseline / --impl reference).  outs[i] sized by caller. */
int ora_decode_batch(const uint8_t* const* files, const size_t* lens, int n,
                     uint8_t* const* outs, const size_t* out_lens, int* status, int threads);

#ifdef __cplusplus
}
#endif
#endif
