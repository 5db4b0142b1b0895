"""The BASELINE.json configurations as input builders, shared by the parity tests at
BASELINE shape (tests/test_gpu_configs.py) and bench.py's `other_configs`.  Inputs only: nothing here decodes.

  cfg2   / cfg2u  512x512, simple filter, RGB ("realistic" / "uniform" mode profile)      -> synth.cfg2_image
  cfg3            3840x2160, normal filter, 4 segments, LF deltas, sharpness 3, 8 partitions -> synth.cfg3_image
  cfg4            1920x1080 VP8X + ALPH (VP8L-compressed, filter method cycling 0..3) + normal-filter VP8, RGBA
  cfg5            animation, 1920x1080 canvas, ALPH + VP8 per frame, blend on, dispose alternating;
                  the variant has every 4th frame as an offset sub-rectangle (even offsets)
"""
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np

from . import synth as S
from . import vp8l_writer as V

# how the VP8L stream of image i's ALPH chunk is shaped: every transform of reverse_transform.rs comes up
_ALPHA_TRANSFORMS = [
    lambda rng, w, h: (),
    lambda rng, w, h: (("predictor", 4, rng.integers(0, 14, (V.subsample(h, 4), V.subsample(w, 4)))),),
    lambda rng, w, h: (("subtract_green",), ("predictor", 5, 11)),
    lambda rng, w, h: (("color", 5, rng.integers(-128, 128, (V.subsample(h, 5), V.subsample(w, 5), 3))), ("predictor", 3, 12)),
]


def cfg4_alpha_chunk(i, width=1920, height=1080, variants=0):
    """ALPH chunk (VP8L-compressed) of cfg4 image i: smooth gradient + 5 % noise, filter method i % 4,
    transform mix (i // 4) % 4.  variants > 0 reuses the planes of image i % variants (bench: the VP8L
    writer is numpy code, a few seconds per 1080p plane)."""
    k = i % variants if variants else i
    rng = np.random.default_rng(0xA1FA + k)
    alpha = S.smooth_alpha(width, height, 0xB200 + k)
    return S.chunk(b"ALPH", V.alph_chunk_payload(alpha, filtering=k % 4, transforms=_ALPHA_TRANSFORMS[(k // 4) % 4](rng, width, height)))


def cfg4_image(i, width=1920, height=1080, alph=None, variants=0):
    seed = 0xB200 + i
    vp8 = S.vp8_frame(width, height, seed=seed, filter_type=0, filter_level=20, yac_abs=36, prob_skip=200, p_skip=0.05)
    return S.webp_extended(vp8, width, height, alph if alph is not None else cfg4_alpha_chunk(i, width, height, variants))


def cfg4_batch(lo, hi, threads=8, variants=16, width=1920, height=1080):
    """cfg4 images [lo, hi).  The ALPH chunks come in `variants` distinct planes (all four filter methods and
    all transform mixes for variants >= 16); the VP8 chunk is different for every image."""
    with ThreadPoolExecutor(max_workers=max(1, threads)) as ex:
        chunks = list(ex.map(lambda k: cfg4_alpha_chunk(k, width, height), range(min(variants, hi - lo) if variants else 0)))
        if variants:
            return list(ex.map(lambda i: cfg4_image(i, width, height, alph=chunks[i % len(chunks)]), range(lo, hi)))
        return list(ex.map(lambda i: cfg4_image(i, width, height), range(lo, hi)))


def libwebp_alpha_chunk(i, width=1920, height=1080):
    """An ALPH chunk made by libwebp's encoder (through PIL), for the cross-check that the writer above and
    libwebp's own streams decode alike.  None when PIL lacks WebP support."""
    try:
        import io
        from PIL import Image, features
        if not features.check("webp"):
            return None
    except Exception:
        return None
    rgba = np.zeros((height, width, 4), np.uint8)
    rgba[..., 3] = S.smooth_alpha(width, height, 0xB200 + i)
    buf = io.BytesIO()
    Image.fromarray(rgba, "RGBA").save(buf, "WEBP", quality=50, method=0)
    b = buf.getvalue()
    at = b.find(b"ALPH")
    n = int.from_bytes(b[at + 4:at + 8], "little")
    return S.chunk(b"ALPH", b[at + 8:at + 8 + n])


def cfg5_animation(n_frames=120, width=1920, height=1080, offsets=False, threads=8):
    """cfg5: n_frames ANMF frames, each ALPH (raw, filter k % 4) + VP8 normal filter, blend on, dispose
    alternating.  offsets=True makes every 4th frame a sub-rectangle at an even offset."""
    def frame(k):
        w, h, x, y = width, height, 0, 0
        if offsets and k % 4 == 3:
            w, h = (width // 2) & ~1, (height // 2) & ~1
            x, y = 2 * ((k * 37) % (width // 4)), 2 * ((k * 53) % (height // 4))
        vp8 = S.vp8_frame(w, h, seed=0xB200 + k, filter_type=0, filter_level=20, yac_abs=36, prob_skip=200, p_skip=0.05)
        return dict(x=x, y=y, w=w, h=h, duration=40, blend=True, dispose=bool(k & 1), vp8=vp8,
                    alph=S.alph_raw(S.smooth_alpha(w, h, k), filtering=k % 4))
    with ThreadPoolExecutor(max_workers=max(1, threads)) as ex:
        frames = list(ex.map(frame, range(n_frames)))
    return S.webp_animation(width, height, frames, alpha=True)


def still_batch(name, lo, hi, threads=None):
    """Files of configuration `name` for image indices [lo, hi)."""
    threads = threads or os.cpu_count() or 1
    if name == "cfg2":
        fn = S.cfg2_image
    elif name == "cfg2u":
        fn = lambda i: S.cfg2_image(i, profile="uniform")
    elif name == "cfg3":
        fn = S.cfg3_image
    elif name == "cfg4":
        return cfg4_batch(lo, hi, threads=threads)
    else:
        raise ValueError(name)
    with ThreadPoolExecutor(max_workers=threads) as ex:
        return list(ex.map(fn, range(lo, hi)))
