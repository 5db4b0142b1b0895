"""The synthetic bitstream writer: encoder known-answer bytes of the reference
(arithmetic_encoder.rs:177-240), and oracle vs libwebp (PIL) on streams whose features no
reference fixture covers (multi-partition, sharpness, LF deltas, delta segments ...)."""
import io

import numpy as np
import pytest

from util import SYNTH_CASES


@pytest.fixture(scope="module")
def synth():
    from harness import synth as S
    S.build()
    return S


def _kat(S, ops):
    L = S.lib()
    k = L.vp8synth_kat_new()
    for op in ops:
        getattr(L, "vp8synth_kat_" + op[0])(k, *op[1:])
    out = np.zeros(256, np.uint8)
    n = L.vp8synth_kat_finish(k, out.ctypes.data)
    return out[:n].tolist()


COMMON = [("bool", 0, 128), ("bool", 1, 10), ("bool", 0, 250), ("literal", 1, 1), ("literal", 3, 5),
          ("literal", 8, 64), ("literal", 8, 185)]


def test_encoder_short(synth):
    assert _kat(synth, COMMON) == [104, 101, 107, 128]


def test_encoder_hello(synth):
    ops = COMMON + [("literal", 8, 31), ("literal", 8, 134), ("opt_signed", 2, 0, 0)]
    # the reference writes Some(1) with its own (inverted) sign polarity: flag 1, magnitude 1, then a 1 bit
    ops += [("bool", 1, 128), ("literal", 2, 1), ("bool", 1, 128)]
    assert bytes(_kat(synth, ops)[:5]) == b"hello"


def test_encoder_tree(synth):
    assert _kat(synth, [("ymode", 3)]) == [233, 64, 0, 0]


def test_encoder_roundtrip_with_oracle_decoder(synth, oracle):
    """arithmetic_encoder.rs:205-231 with the oracle's decoder on the other end"""
    bits = [(1, 40), (1, 110), (0, 70), (0, 10), (1, 5)]
    data = np.array(_kat(synth, [("bool", b, p) for b, p in bits]), dtype=np.uint8)
    L = oracle.lib()
    d = L.ora_bool_new(data.ctypes.data, data.size)
    assert [L.ora_bool_read_bool(d, p) for _, p in bits] == [b for b, _ in bits]
    assert L.ora_bool_past_eof(d) == 0
    L.ora_bool_free(d)


@pytest.mark.parametrize("name", sorted(SYNTH_CASES))
def test_oracle_matches_libwebp(synth, oracle, name):
    """The reference's contract is dwebp-exactness (tests/CREDITS.md), so libwebp is a second
    opinion on features the fixtures never exercise.  Known exception: libwebp keeps
    de-quantised coefficients in 16 bits, the reference in 32/64 (transform.rs:8-11), so the
    'extreme' profile (|level| up to 2114 at q-index 127) may differ and is skipped."""
    from PIL import Image
    kw = SYNTH_CASES[name]
    f = synth.webp_simple(synth.vp8_frame(**kw))
    got = oracle.WebPDecoder(f).read_image()
    assert got.shape == (kw["height"], kw["width"], 3)
    if kw.get("profile") == "extreme":
        pytest.skip("outside libwebp's 16-bit coefficient range")
    im = Image.open(io.BytesIO(f))
    im.load()
    assert np.array_equal(got, np.asarray(im))


@pytest.mark.parametrize("filtering", [0, 1, 2, 3])
@pytest.mark.parametrize("size", [(37, 23), (1, 9), (9, 1), (64, 48)])
def test_raw_alpha_oracle_matches_libwebp(synth, oracle, filtering, size):
    """ALPH un-filter methods 0-3 on raw (uncompressed) alpha, extended.rs:131-191, 295-299"""
    from PIL import Image
    w, h = size
    alpha = synth.smooth_alpha(w, h, seed=filtering * 7 + w)
    vp8 = synth.vp8_frame(w, h, seed=w * h + filtering, filter_type=0, filter_level=10)
    f = synth.webp_extended(vp8, w, h, synth.alph_raw(alpha, filtering=filtering))
    got = oracle.WebPDecoder(f).read_image()
    assert got.shape == (h, w, 4)
    assert np.array_equal(got[:, :, 3], alpha)
    im = Image.open(io.BytesIO(f))
    im.load()
    assert np.array_equal(got, np.asarray(im.convert("RGBA")))
