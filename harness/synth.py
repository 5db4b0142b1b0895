"""Synthetic WebP inputs for tests and bench.py: ctypes binding of harness/vp8synth.c (the
bool-encoder VP8 keyframe writer) plus RIFF / VP8X / ALPH / ANIM container builders.

Container layouts follow what the reference parses: simple file decoder.rs:365-408, VP8X
header extended.rs:193-228, chunk walk decoder.rs:428-544, ALPH header extended.rs:253-308,
ANMF header decoder.rs:767-786.  Not part of the decode path.
"""
import ctypes as C
import os
import struct
import subprocess
from concurrent.futures import ThreadPoolExecutor

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libvp8synth.so")
_lib = None


class Params(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("seed", C.c_uint64),
                ("filter_type", C.c_int32), ("filter_level", C.c_int32), ("sharpness", C.c_int32),
                ("segments", C.c_int32), ("seg_quant", C.c_int32 * 4), ("seg_lf", C.c_int32 * 4),
                ("seg_probs", C.c_int32 * 3), ("lf_adj", C.c_int32), ("ref_delta", C.c_int32 * 4),
                ("mode_delta", C.c_int32 * 4), ("log2_parts", C.c_int32), ("yac_abs", C.c_int32),
                ("q_delta", C.c_int32 * 5), ("prob_skip", C.c_int32), ("p_skip_1024", C.c_int32),
                ("profile", C.c_int32), ("update_probs", C.c_int32), ("version", C.c_int32),
                ("density_1024", C.c_int32)]


PROFILES = {"realistic": 0, "uniform": 1, "extreme": 2, "all_tm": 3, "bpred_ld_vl": 4, "zero_tails": 5}


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in ("vp8synth.c", "vp8synth_tables.h")]
    if force or not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs):
        subprocess.check_call(["make", "-s", "-C", _HERE, "CC=gcc"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        L.vp8synth_bound.argtypes = [C.c_int, C.c_int]; L.vp8synth_bound.restype = C.c_size_t
        L.vp8synth_frame.argtypes = [C.POINTER(Params), C.c_void_p, C.c_size_t]; L.vp8synth_frame.restype = C.c_size_t
        L.vp8synth_kat_new.restype = C.c_void_p
        L.vp8synth_kat_bool.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.vp8synth_kat_literal.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.vp8synth_kat_opt_signed.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.vp8synth_kat_ymode.argtypes = [C.c_void_p, C.c_int]
        L.vp8synth_kat_finish.argtypes = [C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def vp8_frame(width, height, seed=0, filter_type=1, filter_level=20, sharpness=0, segments=0,
              seg_quant=(0, 0, 0, 0), seg_lf=(0, 0, 0, 0), seg_probs=(-1, -1, -1), lf_adj=0,
              ref_delta=(0, 0, 0, 0), mode_delta=(0, 0, 0, 0), log2_parts=0, yac_abs=30,
              q_delta=(0, 0, 0, 0, 0), prob_skip=-1, p_skip=0.0, profile="realistic",
              update_probs=0, version=0, density=1.0):
    """One VP8 keyframe payload (bytes) with random modes and coefficients."""
    p = Params()
    p.width, p.height, p.seed = width, height, seed
    p.filter_type, p.filter_level, p.sharpness = int(filter_type), filter_level, sharpness
    p.segments = segments
    p.seg_quant[:] = seg_quant
    p.seg_lf[:] = seg_lf
    p.seg_probs[:] = seg_probs
    p.lf_adj = lf_adj
    p.ref_delta[:] = ref_delta
    p.mode_delta[:] = mode_delta
    p.log2_parts, p.yac_abs = log2_parts, yac_abs
    p.q_delta[:] = q_delta
    p.prob_skip = prob_skip
    p.p_skip_1024 = int(round(p_skip * 1024))
    p.profile = PROFILES[profile] if isinstance(profile, str) else profile
    p.update_probs, p.version = update_probs, version
    p.density_1024 = int(round(density * 1024))
    cap = lib().vp8synth_bound(width, height)
    buf = np.empty(cap, dtype=np.uint8)
    n = lib().vp8synth_frame(C.byref(p), buf.ctypes.data, cap)
    if n == 0:
        raise ValueError("vp8synth_frame failed (bad parameters or bound too small)")
    return buf[:n].tobytes()


# ---- containers -------------------------------------------------------------------------------

def chunk(fourcc, payload):
    pad = b"\0" if len(payload) & 1 else b""
    return fourcc + struct.pack("<I", len(payload)) + payload + pad


def riff(chunks):
    body = b"WEBP" + b"".join(chunks)
    return b"RIFF" + struct.pack("<I", len(body)) + body


def webp_simple(vp8):
    """RIFF <size> WEBP 'VP8 ' <size> payload (decoder.rs:379-408)."""
    return riff([chunk(b"VP8 ", vp8)])


def _u24(v):
    return struct.pack("<I", v)[:3]


def vp8x_header(width, height, alpha=False, animation=False, icc=False, exif=False, xmp=False):
    flags = (0x10 if alpha else 0) | (0x02 if animation else 0) | (0x20 if icc else 0) | (0x08 if exif else 0) | (0x04 if xmp else 0)
    return chunk(b"VP8X", bytes([flags, 0, 0, 0]) + _u24(width - 1) + _u24(height - 1))


def alph_raw(alpha, filtering=0, preprocessing=0):
    """Uncompressed ALPH chunk from a (h, w) uint8 alpha plane: applies the forward of the
    un-filter in extended.rs:131-191 so that decoding gives `alpha` back."""
    a = np.asarray(alpha, dtype=np.uint8)
    h, w = a.shape
    ai = a.astype(np.int32)
    pred = np.zeros((h, w), dtype=np.int32)
    if filtering == 1:
        pred[:, 1:] = ai[:, :-1]
        pred[1:, 0] = ai[:-1, 0]
    elif filtering == 2:
        pred[1:, :] = ai[:-1, :]
        pred[0, 1:] = ai[0, :-1]
    elif filtering == 3:
        pred[1:, 1:] = np.clip(ai[1:, :-1] + ai[:-1, 1:] - ai[:-1, :-1], 0, 255)
        pred[0, 1:] = ai[0, :-1]
        pred[1:, 0] = ai[:-1, 0]
    res = ((ai - pred) & 0xff).astype(np.uint8)
    info = ((preprocessing & 3) << 4) | ((filtering & 3) << 2) | 0
    return chunk(b"ALPH", bytes([info]) + res.tobytes())


def webp_extended(vp8, width, height, alph_chunk=None, icc=None, exif=None, xmp=None):
    """VP8X + [ICCP] + [ALPH] + 'VP8 ' + [EXIF] + [XMP ] still image (metadata: raw payload bytes or None)."""
    parts = [vp8x_header(width, height, alpha=alph_chunk is not None, icc=icc is not None, exif=exif is not None, xmp=xmp is not None)]
    if icc is not None:
        parts.append(chunk(b"ICCP", icc))
    if alph_chunk is not None:
        parts.append(alph_chunk)
    parts.append(chunk(b"VP8 ", vp8))
    if exif is not None:
        parts.append(chunk(b"EXIF", exif))
    if xmp is not None:
        parts.append(chunk(b"XMP ", xmp))
    return riff(parts)


def anmf(x, y, w, h, duration, blend, dispose, subchunks):
    """ANMF chunk (decoder.rs:767-786): x/2, y/2, w-1, h-1, duration as LE24; flags bit1 = no blend, bit0 = dispose."""
    flags = (0 if blend else 2) | (1 if dispose else 0)
    hdr = _u24(x // 2) + _u24(y // 2) + _u24(w - 1) + _u24(h - 1) + _u24(duration) + bytes([flags])
    return chunk(b"ANMF", hdr + b"".join(subchunks))


def webp_animation(width, height, frames, alpha, loop_count=0, background=(0, 0, 0, 0)):
    """frames: list of dicts {x, y, w, h, duration, blend, dispose, vp8, alph (chunk bytes or None)}."""
    parts = [vp8x_header(width, height, alpha=alpha, animation=True),
             chunk(b"ANIM", bytes(background) + struct.pack("<H", loop_count))]
    for f in frames:
        sub = []
        if f.get("alph") is not None:
            sub.append(f["alph"])
        sub.append(chunk(b"VP8 ", f["vp8"]))
        parts.append(anmf(f["x"], f["y"], f["w"], f["h"], f["duration"], f["blend"], f["dispose"], sub))
    return riff(parts)


def smooth_alpha(width, height, seed):
    """smooth gradient + 5 % noise."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:height, 0:width]
    g = (xx * 255 // max(1, width - 1) + yy * 255 // max(1, height - 1)) // 2
    noise = rng.integers(0, 256, (height, width))
    mask = rng.random((height, width)) < 0.05
    return np.where(mask, noise, g).astype(np.uint8)


# ---- benchmark / test configurations -----------------------------------

def cfg2_image(i, width=512, height=512, profile="realistic"):
    """cfg2: simple filter, level 8..40, sharpness 0, no segments, 1 partition."""
    seed = 0xB200 + i
    r = np.random.default_rng(seed)
    return webp_simple(vp8_frame(width, height, seed=seed, filter_type=1, filter_level=int(r.integers(8, 41)),
                                 yac_abs=int(r.integers(10, 61)), prob_skip=200, p_skip=0.05, profile=profile))


def cfg3_image(i, width=3840, height=2160, profile="realistic"):
    """cfg3: normal filter, 4 segments (abs quant 20/35/50/70, LF 10/20/30/45), LF deltas, sharpness 3, 8 partitions."""
    seed = 0xB200 + i
    return webp_simple(vp8_frame(width, height, seed=seed, filter_type=0, filter_level=24, sharpness=3, segments=1,
                                 seg_quant=(20, 35, 50, 70), seg_lf=(10, 20, 30, 45), seg_probs=(128, 128, 128),
                                 lf_adj=2, ref_delta=(2, 0, 0, 0), mode_delta=(-3, 0, 0, 0), log2_parts=3,
                                 yac_abs=40, prob_skip=200, p_skip=0.05, profile=profile))


def cfg4_image(i, width=1920, height=1080, profile="realistic"):
    """cfg4: VP8X + ALPH (raw, filter method cycling 0..3) + VP8 normal filter; RGBA out."""
    seed = 0xB200 + i
    vp8 = vp8_frame(width, height, seed=seed, filter_type=0, filter_level=20, yac_abs=36, prob_skip=200,
                    p_skip=0.05, profile=profile)
    return webp_extended(vp8, width, height, alph_raw(smooth_alpha(width, height, seed), filtering=i % 4))


def make_batch(fn, n, threads=8, **kw):
    """Build n files with fn(i, **kw) on a thread pool (the C writer releases the GIL)."""
    with ThreadPoolExecutor(max_workers=threads) as ex:
        return list(ex.map(lambda i: fn(i, **kw), range(n)))
