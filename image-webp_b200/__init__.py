"""image-webp_b200: B200-native lossy WebP (VP8) reconstruction behind the reference's decoder API.

Host-side mirror of image-rs/image-webp's decoder interface (src/decoder.rs) over the C ABI of
libwebpb200.so (include/webpb200.h).  Same names and argument meaning as the crate:

    WebPDecoder(data).dimensions() / has_alpha() / is_animated() / output_buffer_size()
    WebPDecoder.read_image(buf) / read_frame(buf) / reset_animation() / set_lossy_upsampling()
    DecodingError (status codes named after the crate's enum variants)

plus the batched entry point the reference does not have (BatchDecoder).  Pixels are produced
by the CUDA kernels only; if the shared library or a CUDA device is missing every decode call
raises -- there is no CPU fallback in this package.

The directory name carries a hyphen, so import it with
    importlib.import_module("image-webp_b200")
"""
import ctypes as C
import os
import weakref
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# WEBPB200_SO: developer override to A/B-test another build of the same library (tools/ab_build.sh)
_SO = os.environ.get("WEBPB200_SO") or os.path.join(_HERE, "libwebpb200.so")
_lib = None

BILINEAR, SIMPLE = 0, 1

ERROR_NAMES = {
    1: "IoError", 2: "ChunkHeaderInvalid(RIFF)", 3: "WebpSignatureInvalid", 4: "ChunkMissing",
    5: "ChunkHeaderInvalid", 6: "Vp8MagicInvalid", 7: "UnsupportedFeature", 8: "ColorSpaceInvalid",
    9: "BitStreamError", 10: "InconsistentImageSizes", 11: "ImageTooLarge", 12: "FrameOutsideImage",
    13: "LosslessSignatureInvalid", 14: "VersionNumberInvalid", 15: "InvalidColorCacheBits",
    16: "HuffmanError", 17: "TransformError", 18: "InvalidAlphaPreprocessing",
    19: "InvalidCompressionMethod", 20: "NoMoreFrames", 21: "InvalidChunkSize", 22: "NotAnimated",
    23: "InvalidParameter", 24: "OutOfMemory", 25: "NotLossy", 100: "NoCudaDevice", 101: "CudaError",
    102: "StaticScratchLimit",
}


class DecodingError(Exception):
    """Mirror of image_webp::DecodingError (decoder.rs:14-168); .code is the C-ABI status."""

    def __init__(self, code, detail=""):
        self.code = int(code)
        name = ERROR_NAMES.get(self.code, "status %d" % self.code)
        super().__init__(name + (": " + detail if detail else ""))


class Segment(C.Structure):
    _fields_ = [("ydc", C.c_int16), ("yac", C.c_int16), ("y2dc", C.c_int16), ("y2ac", C.c_int16),
                ("uvdc", C.c_int16), ("uvac", C.c_int16), ("lf_level", C.c_int8), ("delta_values", C.c_uint8),
                ("pad", C.c_uint8 * 2)]


class FrameDesc(C.Structure):
    _fields_ = [("width", C.c_uint16), ("height", C.c_uint16), ("mbw", C.c_uint16), ("mbh", C.c_uint16),
                ("filter_type", C.c_uint8), ("filter_level", C.c_uint8), ("sharpness", C.c_uint8),
                ("segments_enabled", C.c_uint8), ("lf_adj_enabled", C.c_uint8), ("ref_delta0", C.c_int8),
                ("mode_delta0", C.c_int8), ("out_bpp", C.c_uint8), ("upsampling", C.c_uint8),
                ("alpha_mode", C.c_uint8), ("alpha_filter", C.c_uint8), ("pad0", C.c_uint8),
                ("seg", Segment * 4), ("mb_base", C.c_uint32), ("row_base", C.c_uint32),
                ("y_off", C.c_uint64), ("u_off", C.c_uint64), ("v_off", C.c_uint64),
                ("out_off", C.c_uint64), ("alpha_off", C.c_uint64)]


class MbDesc(C.Structure):
    _fields_ = [("ymode", C.c_uint8), ("uvmode", C.c_uint8), ("segment", C.c_uint8), ("flags", C.c_uint8),
                ("nflags", C.c_uint32), ("bpred", C.c_uint8 * 8)]


assert C.sizeof(FrameDesc) == 136 and C.sizeof(MbDesc) == 16 and C.sizeof(Segment) == 16

MB_DTYPE = np.dtype([("ymode", "u1"), ("uvmode", "u1"), ("segment", "u1"), ("flags", "u1"),
                     ("nflags", "<u4"), ("bpred", "u1", (8,))])

# every symbol include/webpb200.h declares (tests check the library exports all of them)
SYMBOLS = [
    "webpb200_open", "webpb200_close", "webpb200_dimensions", "webpb200_has_alpha", "webpb200_is_animated",
    "webpb200_is_lossy", "webpb200_num_frames", "webpb200_loop_count", "webpb200_loop_duration",
    "webpb200_output_buffer_size", "webpb200_set_background_color", "webpb200_set_lossy_upsampling",
    "webpb200_read_image", "webpb200_read_frame", "webpb200_reset_animation", "webpb200_vp8_decode_frame",
    "webpb200_device_count", "webpb200_ctx_create", "webpb200_ctx_destroy", "webpb200_last_cuda_error",
    "webpb200_host_alloc", "webpb200_host_free", "webpb200_decode_batch", "webpb200_batch_prepare",
    "webpb200_batch_upload", "webpb200_batch_reconstruct", "webpb200_batch_sync", "webpb200_batch_download",
    "webpb200_batch_time", "webpb200_batch_size", "webpb200_batch_get_planes", "webpb200_batch_get_packed",
    "webpb200_launch_count", "webpb200_host_entropy_decode", "webpb200_debug_bpred_eval",
    "webpb200_workspace_sizes", "webpb200_reconstruct_device", "webpb200_decode_animation",
    "webpb200_predict_device_static", "webpb200_predict_device_static_check", "webpb200_open_with_options", "webpb200_set_memory_limit",
    "webpb200_background_color_hint", "webpb200_icc_profile", "webpb200_exif_metadata", "webpb200_xmp_metadata", "webpb200_decode_animation_opts", "webpb200_decode_batch_multi", "webpb200_decode_batch_begin", "webpb200_decode_batch_end", "webpb200_ctx_set_entropy", "webpb200_batch_download_packed",
    "webpb200_last_transfer", "webpb200_device_pci_bus_id", "webpb200_batch_output_device", "webpb200_last_entropy_ms", "webpb200_debug_entropy_resumed", "webpb200_debug_entropy_faults",
]


def build(force=False):
    """Compile the CUDA extension in-tree for sm_100a (csrc/Makefile); nvcc cross-compiles without a GPU."""
    src = os.path.join(_HERE, "csrc")
    deps = [os.path.join(src, f) for f in os.listdir(src) if f.endswith((".cu", ".cuh", ".cc", ".h"))]
    deps.append(os.path.join(_HERE, "..", "include", "webpb200.h"))
    if force or not os.path.exists(_SO) or any(os.path.getmtime(d) > os.path.getmtime(_SO) for d in deps):
        subprocess.check_call(["make", "-s", "-C", src])
    return _SO


def lib():
    """The loaded C-ABI library.  Raises if it is missing: nothing in this package decodes without it."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_SO):
        raise RuntimeError("libwebpb200.so is not built (run __graft_entry__.build()); there is no CPU fallback")
    L = C.CDLL(_SO)
    vp, sz, i32, u32 = C.c_void_p, C.c_size_t, C.c_int, C.c_uint32
    L.webpb200_open.argtypes = [vp, sz, i32, C.POINTER(vp)]
    L.webpb200_close.argtypes = [vp]
    L.webpb200_dimensions.argtypes = [vp, C.POINTER(u32), C.POINTER(u32)]
    for f in ("has_alpha", "is_animated", "is_lossy", "reset_animation"):
        getattr(L, "webpb200_" + f).argtypes = [vp]
    L.webpb200_num_frames.argtypes = [vp]; L.webpb200_num_frames.restype = u32
    L.webpb200_loop_count.argtypes = [vp]; L.webpb200_loop_count.restype = u32
    L.webpb200_loop_duration.argtypes = [vp]; L.webpb200_loop_duration.restype = C.c_uint64
    L.webpb200_output_buffer_size.argtypes = [vp]; L.webpb200_output_buffer_size.restype = sz
    L.webpb200_set_background_color.argtypes = [vp, vp]
    L.webpb200_set_lossy_upsampling.argtypes = [vp, i32]
    L.webpb200_open_with_options.argtypes = [vp, sz, vp, C.POINTER(vp)]
    L.webpb200_set_memory_limit.argtypes = [vp, sz]
    L.webpb200_background_color_hint.argtypes = [vp, vp]
    for f in ("icc_profile", "exif_metadata", "xmp_metadata"):
        getattr(L, "webpb200_" + f).argtypes = [vp, C.POINTER(vp), C.POINTER(sz)]
    L.webpb200_read_image.argtypes = [vp, vp, sz]
    L.webpb200_read_frame.argtypes = [vp, vp, sz, C.POINTER(u32)]
    L.webpb200_vp8_decode_frame.argtypes = [vp, sz, C.POINTER(C.c_uint16), C.POINTER(C.c_uint16), vp, vp, vp]
    L.webpb200_ctx_create.argtypes = [i32, C.POINTER(vp)]
    L.webpb200_ctx_destroy.argtypes = [vp]
    L.webpb200_last_cuda_error.argtypes = [vp]; L.webpb200_last_cuda_error.restype = C.c_char_p
    L.webpb200_host_alloc.argtypes = [sz]; L.webpb200_host_alloc.restype = vp
    L.webpb200_host_free.argtypes = [vp]
    pp, ps, pi = C.POINTER(vp), C.POINTER(sz), C.POINTER(i32)
    L.webpb200_decode_batch.argtypes = [vp, pp, ps, i32, pp, ps, pi, i32, i32]
    L.webpb200_decode_batch_begin.argtypes = [vp, pp, ps, i32, pp, ps, pi, i32, i32, C.POINTER(vp)]
    L.webpb200_decode_batch_end.argtypes = [vp]
    L.webpb200_decode_batch_multi.argtypes = [pp, i32, pp, ps, i32, pp, ps, pi, i32, i32, pi]
    L.webpb200_batch_prepare.argtypes = [vp, pp, ps, i32, pi, i32, i32]
    L.webpb200_batch_upload.argtypes = [vp]
    L.webpb200_batch_reconstruct.argtypes = [vp, C.c_uint]
    L.webpb200_batch_sync.argtypes = [vp]
    L.webpb200_batch_download.argtypes = [vp, pp, ps]
    L.webpb200_batch_time.argtypes = [vp, C.c_uint, i32, i32, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    L.webpb200_batch_size.argtypes = [vp, pi] + [C.POINTER(C.c_uint64)] * 4
    L.webpb200_batch_get_planes.argtypes = [vp, i32, vp, vp, vp]
    L.webpb200_batch_output_device.argtypes = [vp, i32, C.POINTER(vp), C.POINTER(sz), C.POINTER(u32), C.POINTER(u32), C.POINTER(i32)]
    L.webpb200_last_entropy_ms.argtypes = [vp]; L.webpb200_last_entropy_ms.restype = C.c_float
    L.webpb200_device_pci_bus_id.argtypes = [i32, C.c_char_p, i32]
    L.webpb200_last_transfer.argtypes = [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(i32)]
    L.webpb200_ctx_set_entropy.argtypes = [vp, i32]
    L.webpb200_batch_download_packed.argtypes = [vp, i32, vp, vp]
    L.webpb200_batch_get_packed.argtypes = [vp, i32, C.POINTER(C.POINTER(FrameDesc)), C.POINTER(C.POINTER(MbDesc)),
                                            C.POINTER(C.POINTER(C.c_int16))]
    L.webpb200_launch_count.argtypes = [vp]; L.webpb200_launch_count.restype = C.c_uint64
    L.webpb200_host_entropy_decode.argtypes = [vp, sz, C.POINTER(FrameDesc), vp, sz, vp, sz, vp, sz]
    L.webpb200_debug_bpred_eval.argtypes = [i32, vp, vp]
    L.webpb200_workspace_sizes.argtypes = [vp, i32] + [C.POINTER(C.c_uint64)] * 3
    L.webpb200_reconstruct_device.argtypes = [vp, vp, i32, vp, vp, vp, vp, vp, vp, C.c_uint, vp]
    L.webpb200_decode_animation.argtypes = [vp, vp, sz, pp, sz, C.POINTER(u32), i32, i32, i32]
    L.webpb200_decode_animation_opts.argtypes = [vp, vp, sz, pp, sz, C.POINTER(u32), i32, vp, i32]
    L.webpb200_predict_device_static.argtypes = [vp, vp, i32, vp, vp, vp, vp, vp]
    L.webpb200_predict_device_static_check.argtypes = [vp]
    _lib = L
    return L


def device_count():
    return lib().webpb200_device_count()


def _as_u8(data):
    if isinstance(data, np.ndarray):
        return np.ascontiguousarray(data, dtype=np.uint8).reshape(-1)
    return np.frombuffer(bytes(data), dtype=np.uint8)


def host_entropy_decode(data):
    """Host half only (no CUDA): returns (FrameDesc, mbs structured array, levels[mb,25,16], alpha or None)."""
    buf = _as_u8(data)
    fd = FrameDesc()
    st = lib().webpb200_host_entropy_decode(buf.ctypes.data, buf.size, C.byref(fd), None, 0, None, 0, None, 0)
    if st:
        raise DecodingError(st)
    m = fd.mbw * fd.mbh
    mbs = np.zeros(m, dtype=MB_DTYPE)
    levels = np.zeros((m, 25, 16), dtype=np.int16)
    alpha = np.zeros((fd.height, fd.width), dtype=np.uint8) if fd.alpha_mode else None
    st = lib().webpb200_host_entropy_decode(buf.ctypes.data, buf.size, C.byref(fd), mbs.ctypes.data, m,
                                            levels.ctypes.data, levels.size,
                                            alpha.ctypes.data if alpha is not None else None,
                                            alpha.size if alpha is not None else 0)
    if st:
        raise DecodingError(st)
    return fd, mbs, levels, alpha


def entropy_decode_resumed(data):
    """host_entropy_decode through the device stage's hand-over structure (host only): (FrameDesc, mbs, levels)."""
    buf = _as_u8(data)
    fd = FrameDesc()
    st = lib().webpb200_host_entropy_decode(buf.ctypes.data, buf.size, C.byref(fd), None, 0, None, 0, None, 0)
    if st:
        raise DecodingError(st)
    m = fd.mbw * fd.mbh
    mbs = np.zeros(m, dtype=MB_DTYPE)
    levels = np.zeros((m, 25, 16), dtype=np.int16)
    L = lib()
    L.webpb200_debug_entropy_resumed.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
    st = L.webpb200_debug_entropy_resumed(buf.ctypes.data, buf.size, C.byref(fd), mbs.ctypes.data, m, levels.ctypes.data, levels.size)
    if st:
        raise DecodingError(st)
    return fd, mbs, levels


class Options(C.Structure):
    """webpb200_options: WebPDecodeOptions (decoder.rs:274-290) + background colour + memory limit."""
    _fields_ = [("lossy_upsampling", C.c_int32), ("has_background_color", C.c_uint8), ("background_color", C.c_uint8 * 4),
                ("pad", C.c_uint8 * 3), ("memory_limit", C.c_uint64)]


def make_options(lossy_upsampling=BILINEAR, background_color=None, memory_limit=0):
    o = Options()
    o.lossy_upsampling = lossy_upsampling
    if background_color is not None:
        o.has_background_color = 1
        o.background_color[:] = background_color
    o.memory_limit = memory_limit
    return o


class WebPDecoder:
    """image_webp::WebPDecoder (decoder.rs:312-940) over the CUDA path."""

    def __init__(self, data, lossy_upsampling=BILINEAR, options=None):
        """WebPDecoder::new / new_with_options (decoder.rs:336-363): container scan only."""
        self._data = _as_u8(data)
        h = C.c_void_p()
        if options is not None:
            st = lib().webpb200_open_with_options(self._data.ctypes.data, self._data.size, C.byref(options), C.byref(h))
        else:
            st = lib().webpb200_open(self._data.ctypes.data, self._data.size, lossy_upsampling, C.byref(h))
        if st:
            raise DecodingError(st)
        self._h = h

    new = classmethod(lambda cls, data: cls(data))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().webpb200_close(self._h)
            self._h = None

    def dimensions(self):
        w, h = C.c_uint32(), C.c_uint32()
        lib().webpb200_dimensions(self._h, C.byref(w), C.byref(h))
        return w.value, h.value

    def has_alpha(self): return bool(lib().webpb200_has_alpha(self._h))
    def is_animated(self): return bool(lib().webpb200_is_animated(self._h))
    def is_lossy(self): return bool(lib().webpb200_is_lossy(self._h))
    def num_frames(self): return lib().webpb200_num_frames(self._h)
    def loop_count(self): return lib().webpb200_loop_count(self._h)
    def loop_duration(self): return lib().webpb200_loop_duration(self._h)
    def output_buffer_size(self): return lib().webpb200_output_buffer_size(self._h)

    def set_background_color(self, rgba):
        st = lib().webpb200_set_background_color(self._h, (C.c_uint8 * 4)(*rgba))
        if st:
            raise DecodingError(st)

    def set_lossy_upsampling(self, method):
        lib().webpb200_set_lossy_upsampling(self._h, method)

    def set_memory_limit(self, limit):
        """decoder.rs:554"""
        lib().webpb200_set_memory_limit(self._h, limit)

    def background_color_hint(self):
        """decoder.rs:559-565: the ANIM chunk's colour, None for a file that is not extended."""
        c = (C.c_uint8 * 4)()
        return None if lib().webpb200_background_color_hint(self._h, c) else tuple(c)

    def _metadata(self, fn):
        p, n = C.c_void_p(), C.c_size_t()
        st = fn(self._h, C.byref(p), C.byref(n))
        if st:
            raise DecodingError(st)
        return None if not p.value else C.string_at(p.value, n.value)

    def icc_profile(self): return self._metadata(lib().webpb200_icc_profile)      # decoder.rs:643
    def exif_metadata(self): return self._metadata(lib().webpb200_exif_metadata)  # decoder.rs:648
    def xmp_metadata(self): return self._metadata(lib().webpb200_xmp_metadata)    # decoder.rs:653

    def _new_buf(self):
        w, h = self.dimensions()
        return np.zeros((h, w, 4 if self.has_alpha() else 3), dtype=np.uint8)

    def read_image(self, buf=None):
        """decoder.rs:669-741.  buf: writable uint8 array of output_buffer_size() bytes (made if None)."""
        if buf is None:
            buf = self._new_buf()
        st = lib().webpb200_read_image(self._h, buf.ctypes.data, buf.size)
        if st:
            raise DecodingError(st)
        return buf

    def read_frame(self, buf=None):
        """decoder.rs:752-922.  Returns (canvas, duration_ms)."""
        if buf is None:
            buf = self._new_buf()
        dur = C.c_uint32()
        st = lib().webpb200_read_frame(self._h, buf.ctypes.data, buf.size, C.byref(dur))
        if st:
            raise DecodingError(st)
        return buf, dur.value

    def reset_animation(self):
        st = lib().webpb200_reset_animation(self._h)
        if st:
            raise DecodingError(st)


def vp8_decode_frame(vp8_payload):
    """vp8::Vp8Decoder::decode_frame (lossy/mod.rs:1247): (width, height, ybuf, ubuf, vbuf), planes MB-padded."""
    buf = _as_u8(vp8_payload)
    w, h = C.c_uint16(), C.c_uint16()
    st = lib().webpb200_vp8_decode_frame(buf.ctypes.data, buf.size, C.byref(w), C.byref(h), None, None, None)
    if st:
        raise DecodingError(st)
    mbw, mbh = (w.value + 15) // 16, (h.value + 15) // 16
    y = np.zeros((mbh * 16, mbw * 16), dtype=np.uint8)
    u = np.zeros((mbh * 8, mbw * 8), dtype=np.uint8)
    v = np.zeros((mbh * 8, mbw * 8), dtype=np.uint8)
    st = lib().webpb200_vp8_decode_frame(buf.ctypes.data, buf.size, C.byref(w), C.byref(h), y.ctypes.data, u.ctypes.data, v.ctypes.data)
    if st:
        raise DecodingError(st)
    return w.value, h.value, y, u, v


class PinnedBuffer:
    """Page-locked host memory exposed as a numpy uint8 array (webpb200_host_alloc).  The memory lives as long as
    any numpy view of it does: it is released when the last view and this object are gone, or by free()."""

    def __init__(self, nbytes):
        self.ptr = lib().webpb200_host_alloc(nbytes)
        if not self.ptr:
            raise DecodingError(101, "pinned allocation of %d bytes failed" % nbytes)
        self.nbytes = nbytes
        backing = (C.c_uint8 * nbytes).from_address(self.ptr)   # every view's base chain ends in this object
        self._fin = weakref.finalize(backing, _host_free, self.ptr)
        self.array = np.ctypeslib.as_array(backing)

    def free(self):
        """Release the memory now (the caller vouches that no view of it is used any more)."""
        if self.ptr:
            self.array = None
            self._fin()
            self.ptr = None


def _host_free(ptr):
    try:
        lib().webpb200_host_free(ptr)
    except Exception:  # interpreter shutdown
        pass


def device_pci_bus_id(device):
    buf = C.create_string_buffer(32)
    return buf.value.decode().lower() if lib().webpb200_device_pci_bus_id(device, buf, 32) == 0 else None


class DeviceImage:
    """A decoded image in device memory (borrowed from a BatchDecoder): uint8, (height, width, channels)."""

    def __init__(self, ptr, shape):
        self.ptr, self.shape = ptr, shape
        self.__cuda_array_interface__ = {"shape": shape, "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None}


class BufferList:
    """A list of byte buffers with its C pointer / length tables built once (what a `Vec<&[u8]>` is to
    the Rust API): pass it to BatchDecoder.decode instead of a plain list when the same buffers are
    used for many calls, so that the per-call Python work is O(1)."""

    def __init__(self, bufs):
        self.bufs = [_as_u8(b) for b in bufs]
        n = len(self.bufs)
        self.ptrs = (C.c_void_p * n)(*[b.ctypes.data for b in self.bufs])
        self.lens = (C.c_size_t * n)(*[b.size for b in self.bufs])

    def __len__(self):
        return len(self.bufs)


def decode_batch_multi(decoders, files, outs=None, lossy_upsampling=BILINEAR, host_threads=0):
    """One batch over several BatchDecoder contexts (one per GPU): webpb200_decode_batch_multi.
    Returns (outs, status list, split) where split[k]..split[k+1] is the index range context k decoded."""
    bufs = [_as_u8(f) for f in files]
    n = len(bufs)
    if outs is None:
        outs = []
        for b in bufs:
            try:
                outs.append(WebPDecoder(b)._new_buf())
            except DecodingError:
                outs.append(np.zeros(0, dtype=np.uint8))
    fp, fl = BatchDecoder._ptr_arrays(bufs)
    op, ol = BatchDecoder._ptr_arrays(outs)
    status = (C.c_int * max(n, 1))()
    split = (C.c_int * (len(decoders) + 1))()
    ctxs = (C.c_void_p * len(decoders))(*[d._h for d in decoders])
    rc = lib().webpb200_decode_batch_multi(ctxs, len(decoders), fp, fl, n, op, ol, status, lossy_upsampling,
                                           host_threads or (os.cpu_count() or 1), split)
    if rc < 0:
        raise DecodingError(-rc)
    return outs, list(status)[:n], list(split)


class BatchDecoder:
    """Batched still-image decode on one GPU (webpb200_ctx): N x (WebPDecoder::new + read_image)."""

    STAGE_PREDICT, STAGE_LOOPFILTER, STAGE_RGB, STAGE_ALPHA, STAGE_ALL = 1, 2, 4, 8, 15

    def __init__(self, device=0):
        h = C.c_void_p()
        st = lib().webpb200_ctx_create(device, C.byref(h))
        if st:
            raise DecodingError(st, "no usable CUDA device %d" % device)
        self._h = h
        self._keep = None

    def close(self):
        if getattr(self, "_h", None):
            try:
                lib().webpb200_ctx_destroy(self._h)
            except TypeError:  # interpreter shutdown: the module globals are already gone
                pass
            self._h = None

    __del__ = close

    def _check(self, rc):
        if rc < 0:
            raise DecodingError(-rc, (lib().webpb200_last_cuda_error(self._h) or b"").decode())
        return rc

    @staticmethod
    def _ptr_arrays(bufs):
        n = len(bufs)
        return (C.c_void_p * n)(*[b.ctypes.data for b in bufs]), (C.c_size_t * n)(*[b.size for b in bufs])

    def decode(self, files, outs=None, lossy_upsampling=BILINEAR, host_threads=0):
        """files: list of bytes/arrays.  Returns (outs, status list).  outs[i]: uint8 array (made if None)."""
        if isinstance(files, BufferList) and isinstance(outs, BufferList):
            n = len(files)
            status = (C.c_int * n)()
            self._check(lib().webpb200_decode_batch(self._h, files.ptrs, files.lens, n, outs.ptrs, outs.lens, status, lossy_upsampling,
                                                    host_threads or (os.cpu_count() or 1)))
            return outs.bufs, list(status)
        bufs = [_as_u8(f) for f in files]
        n = len(bufs)
        if outs is None:
            outs = []
            for b in bufs:
                try:
                    d = WebPDecoder(b)
                    outs.append(d._new_buf())
                except DecodingError:
                    outs.append(np.zeros(0, dtype=np.uint8))
        fp, fl = self._ptr_arrays(bufs)
        op, ol = self._ptr_arrays(outs)
        status = (C.c_int * n)()
        self._check(lib().webpb200_decode_batch(self._h, fp, fl, n, op, ol, status, lossy_upsampling, host_threads or (os.cpu_count() or 1)))
        return outs, list(status)

    def decode_begin(self, files, outs, lossy_upsampling=BILINEAR, host_threads=0):
        """webpb200_decode_batch_begin: files / outs are BufferList objects (they must outlive decode_end).
        Returns a handle for decode_end; one call in flight per BatchDecoder."""
        n = len(files)
        status = (C.c_int * max(n, 1))()
        h = C.c_void_p()
        self._check(lib().webpb200_decode_batch_begin(self._h, files.ptrs, files.lens, n, outs.ptrs, outs.lens, status, lossy_upsampling,
                                                      host_threads or (os.cpu_count() or 1), C.byref(h)))
        return (h, status, n, files, outs)

    def decode_end(self, handle):
        """webpb200_decode_batch_end: waits; returns the status list of the batch."""
        h, status, n, _, _ = handle
        self._check(lib().webpb200_decode_batch_end(h))
        return list(status)[:n]

    # staged API (bench + tests)
    def prepare(self, files, lossy_upsampling=BILINEAR, host_threads=0):
        bufs = [_as_u8(f) for f in files]
        self._keep = bufs  # the library borrows the file bytes until the next prepare
        n = len(bufs)
        fp, fl = self._ptr_arrays(bufs)
        status = (C.c_int * n)()
        self._check(lib().webpb200_batch_prepare(self._h, fp, fl, n, status, lossy_upsampling, host_threads or (os.cpu_count() or 1)))
        return list(status)

    def upload(self): self._check(lib().webpb200_batch_upload(self._h))
    def reconstruct(self, stages=15): self._check(lib().webpb200_batch_reconstruct(self._h, stages))
    def sync(self): self._check(lib().webpb200_batch_sync(self._h))

    def download(self, outs):
        op, ol = self._ptr_arrays(outs)
        self._check(lib().webpb200_batch_download(self._h, op, ol))

    def time(self, stages=15, iters=1, flush_l2=True):
        """Mean device milliseconds of one reconstruct pass (CUDA events) and per-kernel means."""
        ms = C.c_float()
        k = (C.c_float * 4)()
        self._check(lib().webpb200_batch_time(self._h, stages, iters, 1 if flush_l2 else 0, C.byref(ms), k))
        return ms.value, list(k)

    def size(self):
        n = C.c_int()
        v = [C.c_uint64() for _ in range(4)]
        lib().webpb200_batch_size(self._h, C.byref(n), *[C.byref(x) for x in v])
        return {"images": n.value, "macroblocks": v[0].value, "pixels": v[1].value, "in_bytes": v[2].value, "out_bytes": v[3].value}

    def planes(self, image):
        fd = self.packed(image)[0]
        y = np.zeros((fd.mbh * 16, fd.mbw * 16), dtype=np.uint8)
        u = np.zeros((fd.mbh * 8, fd.mbw * 8), dtype=np.uint8)
        v = np.zeros((fd.mbh * 8, fd.mbw * 8), dtype=np.uint8)
        self._check(lib().webpb200_batch_get_planes(self._h, image, y.ctypes.data, u.ctypes.data, v.ctypes.data))
        return y, u, v

    def packed(self, image):
        f, m, l = C.POINTER(FrameDesc)(), C.POINTER(MbDesc)(), C.POINTER(C.c_int16)()
        self._check(lib().webpb200_batch_get_packed(self._h, image, C.byref(f), C.byref(m), C.byref(l)))
        fd = f.contents
        n = fd.mbw * fd.mbh
        mbs = np.ctypeslib.as_array(C.cast(m, C.POINTER(C.c_uint8)), shape=(n * 16,)).view(MB_DTYPE).copy() if n else np.zeros(0, MB_DTYPE)
        levels = np.ctypeslib.as_array(l, shape=(n, 25, 16)).copy() if n else np.zeros((0, 25, 16), np.int16)
        return fd, mbs, levels

    ENTROPY_HOST, ENTROPY_DEVICE, ENTROPY_AUTO = 0, 1, 2

    def set_entropy(self, where):
        """Where decode() runs the boolean-entropy decode: ENTROPY_HOST (threads) or ENTROPY_DEVICE (k_entropy)."""
        self._check(lib().webpb200_ctx_set_entropy(self._h, where))

    def decode_to_device(self, files, lossy_upsampling=BILINEAR, host_threads=0):
        """Decode and leave the images in device memory: returns (list of objects exposing
        __cuda_array_interface__ with shape (h, w, channels), status list).  torch.as_tensor(x, device="cuda")
        or cupy.asarray(x) wraps one without a copy.  The memory is the context's: valid until its next call."""
        fl = files if isinstance(files, BufferList) else BufferList(files)
        n = len(fl)
        status = (C.c_int * n)()
        self._keep = fl
        self._check(lib().webpb200_decode_batch(self._h, fl.ptrs, fl.lens, n, None, None, status, lossy_upsampling,
                                                host_threads or (os.cpu_count() or 1)))
        out = []
        for i in range(n):
            p, ln, w, h, b = C.c_void_p(), C.c_size_t(), C.c_uint32(), C.c_uint32(), C.c_int()
            if status[i] or lib().webpb200_batch_output_device(self._h, i, C.byref(p), C.byref(ln), C.byref(w), C.byref(h), C.byref(b)):
                out.append(None)
                continue
            out.append(DeviceImage(p.value, (h.value, w.value, b.value)))
        return out, list(status)

    def last_transfer(self):
        """(h2d bytes, d2h bytes, 'host'|'device') of the last decode()."""
        a, b, w = C.c_uint64(), C.c_uint64(), C.c_int()
        self._check(lib().webpb200_last_transfer(self._h, C.byref(a), C.byref(b), C.byref(w)))
        return a.value, b.value, "device" if w.value == 1 else "host"

    def last_entropy_ms(self):
        """Device milliseconds of k_entropy in the last decode() (0.0 when the host did the entropy stage)."""
        return float(lib().webpb200_last_entropy_ms(self._h))

    def device_packed(self, image, frame_desc):
        """Macroblock descriptors and levels of one image as the device holds them (after decode/upload)."""
        n = frame_desc.mbw * frame_desc.mbh
        mbs = np.zeros(n, dtype=MB_DTYPE)
        levels = np.zeros((n, 25, 16), dtype=np.int16)
        self._check(lib().webpb200_batch_download_packed(self._h, image, mbs.ctypes.data, levels.ctypes.data))
        return mbs, levels

    def launch_count(self):
        return lib().webpb200_launch_count(self._h)

    def decode_animation(self, data, lossy_upsampling=BILINEAR, host_threads=0, outs=None, background_color=None):
        """All composited frames of an animated lossy file: (list of canvases, durations).
        outs: optional list of num_frames uint8 arrays of output_buffer_size() bytes (e.g. pinned).
        background_color: RGBA, as WebPDecoder.set_background_color before read_frame."""
        buf = _as_u8(data)
        d = WebPDecoder(buf)
        nf = d.num_frames()
        if outs is None:
            outs = [d._new_buf() for _ in range(nf)]
        op, _ = self._ptr_arrays(outs)
        dur = (C.c_uint32 * max(nf, 1))()
        opt = make_options(lossy_upsampling, background_color)
        rc = lib().webpb200_decode_animation_opts(self._h, buf.ctypes.data, buf.size, op, outs[0].size if outs else 0, dur, nf,
                                                  C.byref(opt), host_threads or (os.cpu_count() or 1))
        self._check(rc)
        return outs, list(dur)[:nf]
