"""Host half of the product (no GPU needed): the shared library loads and exports every symbol
include/webpb200.h declares, the entropy decoder's packed arrays equal the oracle's, truncated
streams get the same verdict, the kernels' B_PRED table equals the oracle's predictors."""
import ctypes as C
import re

import numpy as np
import pytest

from util import LOSSY_STILL, ROOT, SYNTH_CASES, chunk_payload, product, vp8_payload


@pytest.fixture(scope="module")
def W():
    w = product()
    w.build()
    return w


@pytest.fixture(scope="module")
def synth():
    from harness import synth as S
    S.build()
    return S


def test_library_exports_every_declared_symbol(W):
    header = open(ROOT + "/include/webpb200.h").read()
    declared = sorted(set(re.findall(r"\b(webpb200_[a-z0-9_]+)\s*\(", header)))
    assert declared == sorted(W.SYMBOLS)
    raw = C.CDLL(W._SO)
    for name in declared:
        assert hasattr(raw, name), name


def test_struct_sizes(W):
    assert C.sizeof(W.FrameDesc) == 136 and C.sizeof(W.MbDesc) == 16


def test_no_device_fails_loudly(W):
    """No CPU fallback: without a CUDA device decode calls return NoCudaDevice."""
    if W.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(W.DecodingError) as e:
        W.BatchDecoder(0)
    assert e.value.code == 100


def _compare_packed(W, oracle, data):
    fd, mbs, levels, alpha = W.host_entropy_decode(data)
    fr = oracle.Vp8Frame(vp8_payload(data), keep=True)
    f = fr.f
    assert (fd.width, fd.height, fd.mbw, fd.mbh) == (fr.width, fr.height, fr.mbw, fr.mbh)
    assert (fd.filter_type, fd.filter_level, fd.sharpness) == (f.filter_type, f.filter_level, f.sharpness_level)
    assert (fd.segments_enabled, fd.lf_adj_enabled) == (f.segments_enabled, f.lf_adj_enabled)
    assert (fd.ref_delta0, fd.mode_delta0) == (f.ref_delta[0], f.mode_delta[0])
    for i in range(4 if f.segments_enabled else 1):
        a, b = fd.seg[i], f.segment[i]
        assert (a.ydc, a.yac, a.y2dc, a.y2ac, a.uvdc, a.uvac) == (b.ydc, b.yac, b.y2dc, b.y2ac, b.uvdc, b.uvac)
    for i in range(4):
        assert fd.seg[i].lf_level == f.segment[i].loopfilter_level
        assert fd.seg[i].delta_values == f.segment[i].delta_values
    om = fr.mbs
    ym = np.array([m.luma_mode for m in om])
    assert (mbs["ymode"] == ym).all()
    assert (mbs["uvmode"] == np.array([m.chroma_mode for m in om])).all()
    assert (mbs["segment"] == np.array([m.segmentid for m in om])).all()
    assert ((mbs["flags"] & 1) == np.array([m.coeffs_skipped for m in om])).all()
    assert (mbs["nflags"] == np.array([m.nflags for m in om], dtype=np.uint32)).all()
    want = np.array([list(m.bpred) for m in om], dtype=np.uint8)
    mine = np.stack([(mbs["bpred"][:, k >> 1] >> (4 * (k & 1))) & 15 for k in range(16)], axis=1)
    assert (mine[ym == 4] == want[ym == 4]).all()
    assert (mine[ym != 4] == 0).all()
    assert np.array_equal(levels, fr.levels)
    return fd, alpha


@pytest.mark.parametrize("key", LOSSY_STILL)
def test_packed_arrays_on_fixtures(W, oracle, fixtures, key):
    data = fixtures["webp/" + key]
    fd, alpha = _compare_packed(W, oracle, data)
    if "_webp_a" in key:  # VP8L-compressed ALPH: the product's own VP8L decoder against the oracle's
        ch = np.frombuffer(chunk_payload(data, b"ALPH"), np.uint8)
        want = np.zeros((fd.height, fd.width), np.uint8)
        flt = C.c_int()
        assert oracle.lib().ora_read_alpha_chunk(ch.ctypes.data, ch.size, fd.width, fd.height, want.ctypes.data, C.byref(flt)) == 0
        assert fd.alpha_filter == flt.value
        assert np.array_equal(alpha, want)


@pytest.mark.parametrize("name", sorted(SYNTH_CASES))
def test_packed_arrays_on_synthetic(W, oracle, synth, name):
    _compare_packed(W, oracle, synth.webp_simple(synth.vp8_frame(**SYNTH_CASES[name])))


def test_truncated_streams_same_verdict(W, oracle, synth):
    """arithmetic_decoder.rs:249-276: one byte past the end of a partition is fine, more is BitStreamError"""
    vp8 = synth.vp8_frame(width=64, height=48, seed=77, filter_type=0, filter_level=10, log2_parts=1)
    for cut in list(range(len(vp8) - 1, 11, -5)) + [10, 9, 5, 3, 1]:
        try:
            W.host_entropy_decode(synth.webp_simple(vp8[:cut]))
            a = 0
        except W.DecodingError as e:
            a = e.code
        try:
            oracle.WebPDecoder(synth.webp_simple(vp8[:cut])).read_image()
            b = 0
        except oracle.OracleError as e:
            b = e.code
        assert a == b, cut


def test_container_errors(W, oracle, fixtures):
    good = bytes(fixtures["webp/gallery1/1"])
    cases = [good[:3], b"RIFX" + good[4:], good[:8] + b"WEBQ" + good[12:], good[:12] + b"ABCD" + good[16:],
             good[:23] + b"\x00" + good[24:],               # VP8 start code broken
             good[:20] + bytes([good[20] | 1]) + good[21:]]  # inter frame
    for data in cases:
        try:
            oracle.WebPDecoder(data)
            a = 0
        except oracle.OracleError as e:
            a = e.code
        h = C.c_void_p()
        buf = np.frombuffer(data, np.uint8)
        b = W.lib().webpb200_open(buf.ctypes.data, buf.size, 0, C.byref(h))
        if h:
            W.lib().webpb200_close(h)
        assert a == b and a != 0


def test_container_getters_match_oracle(W, oracle, fixtures):
    for key in LOSSY_STILL + ["animated/random_lossy", "animated/random_lossless", "gallery2/1_webp_ll"]:
        data = fixtures["webp/" + key]
        a, b = oracle.WebPDecoder(data), W.WebPDecoder(data)
        for name in ("dimensions", "has_alpha", "is_animated", "is_lossy", "num_frames", "loop_count", "loop_duration", "output_buffer_size"):
            assert getattr(a, name)() == getattr(b, name)(), (key, name)


def test_bpred_table_matches_oracle(W, oracle):
    """The kernels' sub-block prediction table, evaluated on the CPU, against the oracle's
    predictors (prediction.rs:238-499) on random edges."""
    rng = np.random.default_rng(1)
    for mode in range(1, 10):
        for _ in range(200):
            e = rng.integers(0, 256, 13, dtype=np.uint8)
            got = np.zeros(16, np.uint8)
            W.lib().webpb200_debug_bpred_eval(mode, e.ctypes.data, got.ctypes.data)
            ws = np.zeros(16 + 6 * 32, np.uint8)
            o = 16 + 32 + 4
            ws[o - 33] = e[0]
            ws[o - 32:o - 24] = e[1:9]
            for j in range(4):
                ws[o + 32 * j - 1] = e[9 + j]
            oracle.lib().ora_predict_bmode(ws.ctypes.data + 16, 32, 4, 1, mode)
            want = np.array([ws[o + 32 * y + x] for y in range(4) for x in range(4)])
            assert np.array_equal(got, want), mode


def _both_ways(W, data):
    """(status, mbs, levels) of the ordinary host entropy decode and of the one resumed from the hand-over
    structure the device entropy stage is fed with (csrc/entropy.h)."""
    out = []
    for fn in (lambda d: W.host_entropy_decode(d)[:3], W.entropy_decode_resumed):
        try:
            fd, mbs, levels = fn(data)
            out.append((0, mbs, levels))
        except W.DecodingError as e:
            out.append((e.code, None, None))
    return out


@pytest.mark.parametrize("key", LOSSY_STILL)
def test_header_handover_on_fixtures(W, fixtures, key):
    (sa, ma, la), (sb, mb, lb) = _both_ways(W, bytes(fixtures["webp/" + key]))
    assert sa == sb == 0 and np.array_equal(ma, mb) and np.array_equal(la, lb)


def test_header_handover_on_synthetic_and_truncated(W, synth):
    """Every synthetic feature case, and one multi-partition stream cut at many places: the decoder
    resumed from the exported header state gives the same arrays and the same verdicts."""
    for name, kw in SYNTH_CASES.items():
        (sa, ma, la), (sb, mb, lb) = _both_ways(W, synth.webp_simple(synth.vp8_frame(**kw)))
        assert sa == sb == 0 and np.array_equal(ma, mb) and np.array_equal(la, lb), name
    vp8 = synth.vp8_frame(width=96, height=64, seed=78, filter_type=0, filter_level=10, log2_parts=2, segments=1,
                          seg_quant=(20, 35, 50, 70), seg_lf=(10, 20, 30, 45), seg_probs=(128, 100, 200), prob_skip=150, p_skip=0.3)
    for cut in list(range(len(vp8), 10, -7)) + [10, 9, 4]:
        (sa, ma, la), (sb, mb, lb) = _both_ways(W, synth.webp_simple(vp8[:cut]))
        assert sa == sb, cut
        if sa == 0:
            assert np.array_equal(ma, mb) and np.array_equal(la, lb), cut


def test_metadata_memory_limit_and_options(W, oracle, synth, fixtures):
    """icc_profile / exif_metadata / xmp_metadata (decoder.rs:643-655), set_memory_limit (554, enforced in
    read_chunk 629), background_color_hint (559), new_with_options (342): product against the oracle."""
    vp8 = synth.vp8_frame(48, 32, seed=5)
    icc, exif, xmp = bytes(range(200)), b"Exif\0\0II*\0" + b"\x01" * 31, b"<x:xmpmeta/>"
    files = [synth.webp_extended(vp8, 48, 32, icc=icc, exif=exif, xmp=xmp), synth.webp_extended(vp8, 48, 32, exif=exif),
             synth.webp_simple(vp8), bytes(fixtures["webp/animated/random_lossy"]), bytes(fixtures["webp/gallery2/1_webp_a"])]
    for data in files:
        a, b = oracle.WebPDecoder(data), W.WebPDecoder(data)
        assert a.background_color_hint() == b.background_color_hint()
        for name in ("icc_profile", "exif_metadata", "xmp_metadata"):
            assert getattr(a, name)() == getattr(b, name)(), name
        for limit in (0, 30, 199, 200, 1 << 40):
            a.set_memory_limit(limit)
            b.set_memory_limit(limit)
            for name in ("icc_profile", "exif_metadata", "xmp_metadata"):
                try:
                    x = getattr(a, name)()
                except oracle.OracleError as e:
                    x = ("error", e.code)
                try:
                    y = getattr(b, name)()
                except W.DecodingError as e:
                    y = ("error", e.code)
                assert x == y, (name, limit)
    got = W.WebPDecoder(files[0])
    assert got.icc_profile() == icc and got.exif_metadata() == exif and got.xmp_metadata() == xmp
    got.set_memory_limit(100)
    with pytest.raises(W.DecodingError) as e:
        got.icc_profile()
    assert e.value.code == 26  # MemoryLimitExceeded
    # options: memory limit and background colour through new_with_options; the colour only on extended files
    d = W.WebPDecoder(files[0], options=W.make_options(memory_limit=100))
    with pytest.raises(W.DecodingError):
        d.icc_profile()
    assert d.xmp_metadata() == xmp
    with pytest.raises(W.DecodingError) as e:
        W.WebPDecoder(files[2], options=W.make_options(background_color=(1, 2, 3, 4)))
    assert e.value.code == 23  # InvalidParameter, as set_background_color answers (decoder.rs:568-577)
    W.WebPDecoder(files[3], options=W.make_options(background_color=(1, 2, 3, 4)))
    assert C.sizeof(W.Options) == 24


def test_return_code_signs(W):
    """The header's two conventions: decoder surface positive codes, batched / device-resident entry points negative."""
    h = C.c_void_p()
    junk = np.frombuffer(b"RIFX\0\0\0\0WEBP", np.uint8)
    assert W.lib().webpb200_open(junk.ctypes.data, junk.size, 0, C.byref(h)) == 2          # positive: RIFF signature
    assert W.lib().webpb200_decode_batch(None, None, None, 0, None, None, None, 0, 1) == -100  # negative: no context
    assert W.lib().webpb200_decode_animation(None, None, 0, None, 0, None, 0, 0, 1) == -100


def test_vp8l_longest_backward_reference(W, oracle, synth):
    """A backward reference whose length and distance symbols both have 15-bit codes and the most extra bits
    (10 and 18) needs 58 bits in one go.  The reference's bit reader counts 56..63 bits as available after a
    fill (`nbits |= 56`, decoder/mod.rs:718-738) and get_copy_distance (619-633) does not refill, so the
    reference decodes such a stream at most bit alignments and answers BitStreamError at the others.  Oracle and product follow it in both: same verdict at every alignment, same pixels where it decodes."""
    from harness import vp8l_writer as V
    w, h = 2048, 640
    n = w * h
    # explicit code lengths: a complete code 1, 2, ..., 14, 15, 15 over 16 symbols, the two symbols of interest at 15 bits
    green = [0] * 280
    for k, sym in enumerate([0, 7] + list(range(8, 21)) + [256 + 23]):
        green[sym] = min(k + 1, 15)
    distl = [0] * 40
    for k, sym in enumerate(list(range(0, 15)) + [39]):
        distl[sym] = min(k + 1, 15)
    length, dist = 4096, 1_048_456                   # length prefix 23 (10 extra bits), distance code prefix 39 (18 extra bits)
    vp8 = synth.vp8_frame(w, h, seed=3, filter_level=0)
    verdicts = []
    for shift in range(8):                            # the reference lands on every bit alignment of the reader's window
        rng = np.random.default_rng(5)
        g = rng.integers(0, 2, n).astype(np.uint32) * 7
        pos = 1_100_000 + shift
        g[pos - 8:pos] = 0                            # one-bit literals right before it: `shift` moves the bit position by one each
        g[pos:pos + length] = g[pos - dist:pos - dist + length]
        stream = V.encode((0xff000000 | (g << 8)).reshape(h, w), header=False, refs=[(pos, length, dist + 120)], force_lengths={0: green, 4: distl})
        payload = bytes([1]) + stream                 # ALPH header: compression 1, no filter
        f = synth.webp_extended(vp8, w, h, synth.chunk(b"ALPH", payload))
        want = np.zeros((h, w), np.uint8)
        flt = C.c_int()
        ch = np.frombuffer(payload, np.uint8)
        a = oracle.lib().ora_read_alpha_chunk(ch.ctypes.data, ch.size, w, h, want.ctypes.data, C.byref(flt))
        try:
            fd, mbs, levels, alpha = W.host_entropy_decode(f)
            b = 0
        except W.DecodingError as e:
            b = e.code
        assert a == b, shift
        verdicts.append(a)
        if a == 0:
            assert np.array_equal(want.reshape(-1), g.astype(np.uint8))
            assert np.array_equal(alpha.reshape(-1), want.reshape(-1)), shift
    assert sorted(set(verdicts)) == [0, 9], verdicts  # both happen; which alignments fail follows from the bits left over before the fill


def test_vp8l_host_decoder_all_transform_mixes(W, oracle, synth):
    """The product's host VP8L decoder (two-level prefix-code tables, literal-only fast path) against the oracle on
    noisy planes: every transform mix and ALPH filter of the cfg4 generator at a small size, with prefix codes longer
    than the root table (noise over a gradient gives 10-bit and longer codes) and with all three colour channels
    coded (subtract-green / colour transform mixes)."""
    from harness import configs
    w, h = 208, 120
    longest = 0
    for k in range(16):
        ch = configs.cfg4_alpha_chunk(k, w, h)
        payload = np.frombuffer(ch[8:], np.uint8)
        f = synth.webp_extended(synth.vp8_frame(w, h, seed=40 + k, filter_level=0), w, h, ch)
        want = np.zeros((h, w), np.uint8)
        flt = C.c_int()
        assert oracle.lib().ora_read_alpha_chunk(payload.ctypes.data, payload.size, w, h, want.ctypes.data, C.byref(flt)) == 0
        fd, mbs, levels, alpha = W.host_entropy_decode(f)
        assert fd.alpha_filter == flt.value
        assert np.array_equal(alpha.reshape(-1), want.reshape(-1)), k
        longest = max(longest, len(payload))
    assert longest > w * h // 2  # the planes are noisy enough not to collapse into a few short codes
