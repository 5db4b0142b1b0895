"""ctypes binding of the CPU oracle (oracle/libwebp_oracle.so).  TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
import this module; the product package never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libwebp_oracle.so")


def build(force=False):
    """Compile the oracle with gcc (oracle/Makefile)."""
    srcs = [os.path.join(_HERE, f) for f in ("ora_vp8.c", "ora_vp8l.c", "ora_webp.c", "ora.h", "vp8_tables.h")]
    if force or not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs):
        subprocess.check_call(["make", "-s", "-C", _HERE, "CC=gcc"])
    return _SO


_lib = None


class Segment(C.Structure):
    _fields_ = [("ydc", C.c_int16), ("yac", C.c_int16), ("y2dc", C.c_int16), ("y2ac", C.c_int16),
                ("uvdc", C.c_int16), ("uvac", C.c_int16), ("quantizer_level", C.c_int8),
                ("loopfilter_level", C.c_int8), ("delta_values", C.c_uint8)]


class Mb(C.Structure):
    _fields_ = [("luma_mode", C.c_uint8), ("chroma_mode", C.c_uint8), ("segmentid", C.c_uint8),
                ("coeffs_skipped", C.c_uint8), ("non_zero_dct", C.c_uint8), ("bpred", C.c_uint8 * 16),
                ("nflags", C.c_uint32)]


class Frame(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("mbw", C.c_int), ("mbh", C.c_int),
                ("ybuf", C.POINTER(C.c_uint8)), ("ubuf", C.POINTER(C.c_uint8)), ("vbuf", C.POINTER(C.c_uint8)),
                ("ybuf0", C.POINTER(C.c_uint8)), ("ubuf0", C.POINTER(C.c_uint8)), ("vbuf0", C.POINTER(C.c_uint8)),
                ("version", C.c_int), ("for_display", C.c_int), ("pixel_type", C.c_int),
                ("filter_type", C.c_int), ("filter_level", C.c_int), ("sharpness_level", C.c_int),
                ("segments_enabled", C.c_int), ("segments_update_map", C.c_int), ("lf_adj_enabled", C.c_int),
                ("ref_delta", C.c_int * 4), ("mode_delta", C.c_int * 4), ("num_partitions", C.c_int),
                ("prob_skip_false", C.c_int), ("segment", Segment * 4),
                ("mbs", C.POINTER(Mb)), ("levels", C.POINTER(C.c_int16)), ("residue", C.POINTER(C.c_int32))]


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        u8p = C.POINTER(C.c_uint8)
        L.ora_open.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.POINTER(C.c_void_p)]
        L.ora_close.argtypes = [C.c_void_p]
        L.ora_dimensions.argtypes = [C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
        for f in ("ora_has_alpha", "ora_is_animated", "ora_is_lossy", "ora_reset_animation"):
            getattr(L, f).argtypes = [C.c_void_p]
        L.ora_num_frames.argtypes = [C.c_void_p]; L.ora_num_frames.restype = C.c_uint32
        L.ora_loop_count.argtypes = [C.c_void_p]; L.ora_loop_count.restype = C.c_uint32
        L.ora_loop_duration.argtypes = [C.c_void_p]; L.ora_loop_duration.restype = C.c_uint64
        L.ora_output_buffer_size.argtypes = [C.c_void_p]; L.ora_output_buffer_size.restype = C.c_size_t
        L.ora_set_background_color.argtypes = [C.c_void_p, C.c_void_p]
        L.ora_set_memory_limit.argtypes = [C.c_void_p, C.c_size_t]
        L.ora_background_color_hint.argtypes = [C.c_void_p, C.c_void_p]
        L.ora_metadata.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]
        L.ora_set_lossy_upsampling.argtypes = [C.c_void_p, C.c_int]
        L.ora_read_image.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.ora_read_frame.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_uint32)]
        L.ora_vp8_decode.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.POINTER(Frame)]
        L.ora_frame_free.argtypes = [C.POINTER(Frame)]
        L.ora_frame_fill.argtypes = [C.POINTER(Frame), C.c_void_p, C.c_int, C.c_int]
        L.ora_idct4x4.argtypes = [C.c_void_p]
        L.ora_iwht4x4.argtypes = [C.c_void_p]
        L.ora_fill_rgb_buffer_fancy.argtypes = [C.c_void_p] * 4 + [C.c_int] * 4
        L.ora_fill_rgb_buffer_simple.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_int] * 5
        L.ora_yuv_to_rgb.argtypes = [C.c_uint8, C.c_uint8, C.c_uint8, C.c_void_p]
        L.ora_predict_bmode.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
        L.ora_add_residue.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.ora_lf_simple_h.argtypes = [C.c_int, C.c_void_p]
        L.ora_lf_simple_v.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_size_t]
        L.ora_lf_subblock_h.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.ora_lf_subblock_v.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_size_t]
        L.ora_lf_macroblock_h.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.ora_lf_macroblock_v.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_size_t]
        L.ora_bool_new.argtypes = [C.c_void_p, C.c_size_t]; L.ora_bool_new.restype = C.c_void_p
        L.ora_bool_free.argtypes = [C.c_void_p]
        L.ora_bool_read_bool.argtypes = [C.c_void_p, C.c_int]
        L.ora_bool_read_flag.argtypes = [C.c_void_p]
        L.ora_bool_read_literal.argtypes = [C.c_void_p, C.c_int]
        L.ora_bool_past_eof.argtypes = [C.c_void_p]
        L.ora_vp8l_decode.argtypes = [C.c_void_p, C.c_size_t, C.c_uint32, C.c_uint32, C.c_int, C.c_void_p]
        L.ora_read_alpha_chunk.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_int)]
        L.ora_alpha_merge.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.ora_composite_frame.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_void_p, C.c_void_p] + \
            [C.c_uint32] * 4 + [C.c_int, C.c_int] + [C.c_uint32] * 4
        L.ora_blend_pixel.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.ora_decode_batch.argtypes = [C.POINTER(C.c_void_p), C.POINTER(C.c_size_t), C.c_int, C.POINTER(C.c_void_p),
                                       C.POINTER(C.c_size_t), C.POINTER(C.c_int), C.c_int]
        _lib = L
    return _lib


def stage_timing(on):
    """Switch the per-stage timers of the decoder on or off and clear them (bench.py's CPU baseline split)."""
    L = lib()
    C.c_int.in_dll(L, "ora_stage_timing").value = 1 if on else 0
    arr = (C.c_longlong * 4).in_dll(L, "ora_stage_ns")
    for i in range(4):
        arr[i] = 0


def stage_seconds():
    """Core-seconds spent per stage since stage_timing(True): entropy decode incl. dequantisation and inverse
    transforms, intra prediction, loop filter, YUV -> RGB(A) (summed over the worker threads)."""
    arr = (C.c_longlong * 4).in_dll(lib(), "ora_stage_ns")
    names = ["entropy+dequant+itx", "predict", "loopfilter", "rgb"]
    return {n: arr[i] / 1e9 for i, n in enumerate(names)}


class OracleError(Exception):
    def __init__(self, code):
        super().__init__("oracle status %d" % code)
        self.code = code


def _buf(data):
    a = np.frombuffer(bytes(data), dtype=np.uint8) if not isinstance(data, np.ndarray) else np.ascontiguousarray(data, dtype=np.uint8)
    return a


class WebPDecoder:
    """Same surface as the reference's WebPDecoder (decoder.rs:312-940)."""

    def __init__(self, data, upsampling="bilinear"):
        self._data = _buf(data)
        h = C.c_void_p()
        st = lib().ora_open(self._data.ctypes.data, self._data.size, 0 if upsampling == "bilinear" else 1, C.byref(h))
        if st:
            raise OracleError(st)
        self._h = h

    def __del__(self):
        if getattr(self, "_h", None):
            lib().ora_close(self._h)
            self._h = None

    def dimensions(self):
        w, h = C.c_uint32(), C.c_uint32()
        lib().ora_dimensions(self._h, C.byref(w), C.byref(h))
        return w.value, h.value

    def has_alpha(self): return bool(lib().ora_has_alpha(self._h))
    def is_animated(self): return bool(lib().ora_is_animated(self._h))
    def is_lossy(self): return bool(lib().ora_is_lossy(self._h))
    def num_frames(self): return lib().ora_num_frames(self._h)
    def loop_count(self): return lib().ora_loop_count(self._h)
    def loop_duration(self): return lib().ora_loop_duration(self._h)
    def output_buffer_size(self): return lib().ora_output_buffer_size(self._h)

    def set_background_color(self, rgba):
        c = (C.c_uint8 * 4)(*rgba)
        st = lib().ora_set_background_color(self._h, c)
        if st:
            raise OracleError(st)

    def set_lossy_upsampling(self, upsampling):
        lib().ora_set_lossy_upsampling(self._h, 0 if upsampling == "bilinear" else 1)

    def set_memory_limit(self, limit):
        lib().ora_set_memory_limit(self._h, limit)

    def background_color_hint(self):
        c = (C.c_uint8 * 4)()
        return None if lib().ora_background_color_hint(self._h, c) else tuple(c)

    def _metadata(self, which):
        p, n = C.c_void_p(), C.c_size_t()
        st = lib().ora_metadata(self._h, which, C.byref(p), C.byref(n))
        if st:
            raise OracleError(st)
        return None if not p.value else C.string_at(p.value, n.value)

    def icc_profile(self): return self._metadata(0)
    def exif_metadata(self): return self._metadata(1)
    def xmp_metadata(self): return self._metadata(2)

    def _shape(self):
        w, h = self.dimensions()
        return (h, w, 4 if self.has_alpha() else 3)

    def read_image(self, out=None):
        if out is None:
            out = np.zeros(self._shape(), dtype=np.uint8)
        st = lib().ora_read_image(self._h, out.ctypes.data, out.size)
        if st:
            raise OracleError(st)
        return out

    def read_frame(self, out=None):
        if out is None:
            out = np.zeros(self._shape(), dtype=np.uint8)
        dur = C.c_uint32()
        st = lib().ora_read_frame(self._h, out.ctypes.data, out.size, C.byref(dur))
        if st:
            raise OracleError(st)
        return out, dur.value

    def reset_animation(self):
        st = lib().ora_reset_animation(self._h)
        if st:
            raise OracleError(st)


class Vp8Frame:
    """Vp8Decoder::decode_frame with the intermediate arrays (keep=1)."""

    def __init__(self, vp8_payload, keep=True):
        self._data = _buf(vp8_payload)
        self.f = Frame()
        st = lib().ora_vp8_decode(self._data.ctypes.data, self._data.size, 1 if keep else 0, C.byref(self.f))
        if st:
            raise OracleError(st)
        f = self.f
        self.width, self.height, self.mbw, self.mbh = f.width, f.height, f.mbw, f.mbh
        ysz, csz = (f.mbh * 16, f.mbw * 16), (f.mbh * 8, f.mbw * 8)
        g = lambda p, shp: np.ctypeslib.as_array(p, shape=shp).copy()
        self.y, self.u, self.v = g(f.ybuf, ysz), g(f.ubuf, csz), g(f.vbuf, csz)
        if keep:
            self.y0, self.u0, self.v0 = g(f.ybuf0, ysz), g(f.ubuf0, csz), g(f.vbuf0, csz)
            n = f.mbw * f.mbh
            self.levels = g(f.levels, (n, 25, 16))
            self.residue = g(f.residue, (n, 24, 16))
            self.mbs = [f.mbs[i] for i in range(n)]

    def fill(self, bpp=3, fancy=True):
        out = np.zeros((self.height, self.width, bpp), dtype=np.uint8)
        lib().ora_frame_fill(C.byref(self.f), out.ctypes.data, bpp, 1 if fancy else 0)
        return out

    def __del__(self):
        if getattr(self, "f", None) is not None:
            lib().ora_frame_free(C.byref(self.f))
            self.f = None


def decode_batch(files, threads=1, has_alpha=False):
    """Full CPU decode of a list of byte strings, one image per worker thread at a time."""
    n = len(files)
    bufs = [_buf(f) for f in files]
    outs, out_lens = [], []
    for b in bufs:
        d = WebPDecoder(b)
        outs.append(np.zeros(d._shape(), dtype=np.uint8))
        out_lens.append(outs[-1].size)
        del d
    fp = (C.c_void_p * n)(*[b.ctypes.data for b in bufs])
    fl = (C.c_size_t * n)(*[b.size for b in bufs])
    op = (C.c_void_p * n)(*[o.ctypes.data for o in outs])
    ol = (C.c_size_t * n)(*out_lens)
    st = (C.c_int * n)()
    bad = lib().ora_decode_batch(fp, fl, n, op, ol, st, threads)
    return outs, list(st), bad
