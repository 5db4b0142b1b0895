"""Parity of the CUDA path with the CPU oracle, through the C ABI (run on the B200 box).

Bit-exact everywhere: this is integer/byte work.  Three layers of evidence:
  * the reference's own golden images (tests/golden/fixtures.npz, from tests/reference/*.png),
  * the oracle on synthetic streams covering what no fixture does,
  * size-independent properties at batch scale (batch composition must not matter, repeated
    passes must agree, a checksum over the whole batch must equal the oracle's).
"""
import ctypes as C
import hashlib

import numpy as np
import pytest

from util import LOSSY_STILL, SYNTH_CASES, product, vp8_payload

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def W():
    w = product()
    w.lib()  # must already be built in-tree; raises otherwise (no fallback)
    assert w.device_count() > 0, "no CUDA device"
    return w


@pytest.fixture(scope="module")
def bd(W):
    b = W.BatchDecoder(0)
    yield b
    b.close()


@pytest.fixture(scope="module")
def synth():
    from harness import synth as S
    S.build()
    return S


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest()


# ---- the reference's golden images --------------------------------------------------------------

@pytest.mark.parametrize("key", LOSSY_STILL)
def test_golden_still(W, fixtures, key):
    d = W.WebPDecoder(fixtures["webp/" + key])
    w, h, c = (int(v) for v in fixtures["dims/" + key])
    assert d.dimensions() == (w, h) and d.has_alpha() == (c == 4) and d.is_lossy()
    img = d.read_image()
    assert sha(img) == fixtures["sha/" + key].tobytes()


@pytest.mark.parametrize("i", range(1, 6))
def test_golden_nofancy(W, fixtures, i):
    d = W.WebPDecoder(fixtures["webp/gallery1/%d" % i], lossy_upsampling=W.SIMPLE)
    assert sha(d.read_image()) == fixtures["sha/gallery1_nofancy/%d" % i].tobytes()
    d2 = W.WebPDecoder(fixtures["webp/gallery1/%d" % i])
    d2.set_lossy_upsampling(W.SIMPLE)  # decoder.rs:938
    assert sha(d2.read_image()) == fixtures["sha/gallery1_nofancy/%d" % i].tobytes()


def test_golden_animation(W, bd, fixtures):
    key = "animated/random_lossy"
    data = fixtures["webp/" + key]
    d = W.WebPDecoder(data)
    assert d.is_animated() and d.num_frames() == 4
    assert np.array_equal(d.read_image(), fixtures["full/%s-1" % key])  # frame 1, cursor untouched
    for i in range(1, 5):
        img, dur = d.read_frame()
        assert np.array_equal(img, fixtures["full/%s-%d" % (key, i)]), i
    with pytest.raises(W.DecodingError) as e:
        d.read_frame()
    assert e.value.code == 20  # NoMoreFrames
    d.reset_animation()
    assert np.array_equal(d.read_frame()[0], fixtures["full/%s-1" % key])
    frames, durs = bd.decode_animation(data)
    for i in range(4):
        assert np.array_equal(frames[i], fixtures["full/%s-%d" % (key, i + 1)]), i


def test_lossless_is_refused_not_faked(W, fixtures):
    """VP8L stills are outside this library's path: they must fail, not fall back to a CPU decode."""
    with pytest.raises(W.DecodingError) as e:
        W.WebPDecoder(fixtures["webp/gallery2/1_webp_ll"]).read_image()
    assert e.value.code == 25


def test_wrong_buffer_size(W, fixtures):
    d = W.WebPDecoder(fixtures["webp/regression/dark"])
    with pytest.raises(W.DecodingError) as e:
        d.read_image(np.zeros(7, np.uint8))
    assert e.value.code == 11  # decoder.rs:670-672


# ---- oracle on synthetic streams, stage by stage ------------------------------------------------

@pytest.mark.parametrize("name", sorted(SYNTH_CASES))
def test_stages_against_oracle(W, bd, oracle, synth, name):
    data = synth.webp_simple(synth.vp8_frame(**SYNTH_CASES[name]))
    fr = oracle.Vp8Frame(vp8_payload(data), keep=True)
    assert bd.prepare([data]) == [0]
    bd.upload()
    bd.reconstruct(bd.STAGE_PREDICT)
    bd.sync()
    y, u, v = bd.planes(0)
    assert np.array_equal(y, fr.y0) and np.array_equal(u, fr.u0) and np.array_equal(v, fr.v0), "prediction"
    bd.reconstruct(bd.STAGE_PREDICT | bd.STAGE_LOOPFILTER)
    bd.sync()
    y, u, v = bd.planes(0)
    assert np.array_equal(y, fr.y) and np.array_equal(u, fr.u) and np.array_equal(v, fr.v), "loop filter"
    for ups, fancy in ((W.BILINEAR, True), (W.SIMPLE, False)):
        assert bd.prepare([data], lossy_upsampling=ups) == [0]
        bd.upload()
        bd.reconstruct()
        out = np.zeros((fr.height, fr.width, 3), np.uint8)
        bd.download([out])
        assert np.array_equal(out, fr.fill(3, fancy)), "rgb fancy=%s" % fancy


def test_vp8_decode_frame_planes(W, oracle, synth):
    """vp8::Vp8Decoder::decode_frame (lib.rs:34-36): public MB-padded planes"""
    vp8 = synth.vp8_frame(**SYNTH_CASES["normal_sharp3_2parts_131x77"])
    fr = oracle.Vp8Frame(vp8, keep=False)
    w, h, y, u, v = W.vp8_decode_frame(vp8)
    assert (w, h) == (131, 77)
    assert np.array_equal(y, fr.y) and np.array_equal(u, fr.u) and np.array_equal(v, fr.v)


@pytest.mark.parametrize("filtering", [0, 1, 2, 3])
@pytest.mark.parametrize("size", [(37, 23), (1, 9), (9, 1), (64, 48), (131, 77), (24, 1100)])
def test_alpha_filters(W, bd, oracle, synth, filtering, size):
    """ALPH un-filter + merge, decoder.rs:718-734 / extended.rs:131-191, raw alpha, every method"""
    w, h = size
    alpha = np.random.default_rng(w * 31 + h + filtering).integers(0, 256, (h, w)).astype(np.uint8)
    vp8 = synth.vp8_frame(w, h, seed=w * h + filtering, filter_type=0, filter_level=10)
    f = synth.webp_extended(vp8, w, h, synth.alph_raw(alpha, filtering=filtering))
    want = oracle.WebPDecoder(f).read_image()
    assert np.array_equal(want[:, :, 3], alpha)
    outs, status = bd.decode([f])
    assert status == [0] and np.array_equal(outs[0], want)


def test_mixed_batch_with_bad_files(W, bd, oracle, synth, fixtures):
    """Independent images: a bad file gets its own status and does not poison the batch."""
    good = [synth.webp_simple(synth.vp8_frame(**SYNTH_CASES[k])) for k in sorted(SYNTH_CASES)]
    good += [bytes(fixtures["webp/" + k]) for k in ("gallery1/1", "gallery2/4_webp_a", "regression/dark")]
    w, h = 96, 80
    good.append(synth.webp_extended(synth.vp8_frame(w, h, seed=5, filter_type=0, filter_level=9), w, h,
                                    synth.alph_raw(synth.smooth_alpha(w, h, 3), filtering=3)))
    broken = good[3][:len(good[3]) // 2]                    # truncated mid-partition
    files = good[:5] + [broken, b"not a webp file at all"] + good[5:] + [bytes(fixtures["webp/gallery2/2_webp_ll"])]
    outs, status = bd.decode(files)
    for i, f in enumerate(files):
        try:
            want = oracle.WebPDecoder(f).read_image()
        except oracle.OracleError:
            want = None
        if i == len(files) - 1:
            assert status[i] == 25  # lossless still: refused
        elif want is None:
            assert status[i] != 0, i
        else:
            assert status[i] == 0 and np.array_equal(outs[i], want), i


def _anim_frames(synth, cw, ch, specs, alpha):
    frames = []
    for k, (x, y, w, h, blend, dispose) in enumerate(specs):
        vp8 = synth.vp8_frame(w, h, seed=100 + k, filter_type=k % 2, filter_level=5 + 7 * k)
        alph = None
        if alpha:
            a = np.random.default_rng(k).integers(0, 256, (h, w)).astype(np.uint8)
            a[: h // 3] = 0          # fully transparent band: blend keeps the canvas
            a[h // 3: h // 2] = 255  # opaque band
            alph = synth.alph_raw(a, filtering=k % 4)
        frames.append(dict(x=x, y=y, w=w, h=h, duration=40 + k, blend=blend, dispose=dispose, vp8=vp8, alph=alph))
    return synth.webp_animation(cw, ch, frames, alpha=alpha)


@pytest.mark.parametrize("alpha", [False, True])
def test_synthetic_animation(W, bd, oracle, synth, alpha):
    """composite_frame + do_alpha_blending (extended.rs:30-129, alpha_blending.rs): offsets,
    blend on/off, dispose on/off, lossy+ALPH frames -- none of which the fixtures exercise."""
    specs = [(0, 0, 64, 48, False, False), (10, 6, 30, 20, True, True), (0, 0, 64, 48, True, False),
             (32, 24, 32, 24, False, True), (2, 2, 17, 9, True, False), (0, 0, 64, 48, True, True)]
    data = _anim_frames(synth, 64, 48, specs, alpha)
    o = oracle.WebPDecoder(data)
    d = W.WebPDecoder(data)
    assert d.num_frames() == len(specs) and d.has_alpha() == alpha
    assert d.loop_duration() == o.loop_duration()
    want = []
    for _ in specs:
        a, da = o.read_frame()
        b, db = d.read_frame()
        assert da == db and np.array_equal(a, b)
        want.append(a)
    frames, durs = bd.decode_animation(data)
    for a, b in zip(want, frames):
        assert np.array_equal(a, b)
    # user background colour (decoder.rs:568, 861-882)
    o2, d2 = oracle.WebPDecoder(data), W.WebPDecoder(data)
    o2.set_background_color((10, 20, 30, 40))
    d2.set_background_color((10, 20, 30, 40))
    want_bg = []
    for _ in specs:
        want_bg.append(o2.read_frame()[0])
        assert np.array_equal(want_bg[-1], d2.read_frame()[0])
    # the same colour through webpb200_options on the whole-animation call
    frames_bg, _ = bd.decode_animation(data, background_color=(10, 20, 30, 40))
    for a, b in zip(want_bg, frames_bg):
        assert np.array_equal(a, b)


# ---- batch-scale properties ----------------------------------------------------------------------

def test_batch_properties(W, bd, oracle, synth):
    """256 cfg2-style images: whole-batch checksum equals the oracle's, a second pass gives the
    same bytes, and an image decoded alone equals the same image inside the batch."""
    n = 256
    files = synth.make_batch(synth.cfg2_image, n)
    outs, status = bd.decode(files)
    assert all(s == 0 for s in status)
    want, wst, bad = oracle.decode_batch(files, threads=8)
    assert bad == 0
    h1, h2 = hashlib.sha256(), hashlib.sha256()
    for a, b in zip(outs, want):
        h1.update(a.tobytes())
        h2.update(b.tobytes())
    assert h1.digest() == h2.digest()
    outs2, _ = bd.decode(files)
    assert all(np.array_equal(a, b) for a, b in zip(outs, outs2))
    for i in (0, 17, n - 1):
        alone, _ = bd.decode([files[i]])
        assert np.array_equal(alone[0], outs[i])
    assert bd.launch_count() > 0


def test_large_frame_normal_filter(W, bd, oracle, synth):
    """one cfg3-style frame (3840x2160, normal filter, 4 segments, 8 partitions, LF deltas)"""
    f = synth.cfg3_image(0)
    outs, status = bd.decode([f])
    assert status == [0]
    assert np.array_equal(outs[0], oracle.WebPDecoder(f).read_image())


def test_device_resident_seam(W, bd, oracle, synth):
    """webpb200_reconstruct_device on caller-owned device arrays (torch is only the allocator)."""
    torch = pytest.importorskip("torch")
    files = [synth.webp_simple(synth.vp8_frame(**SYNTH_CASES[k])) for k in ("simple_150x90", "normal_sharp3_2parts_131x77")]
    assert bd.prepare(files) == [0, 0]
    packed = [bd.packed(i) for i in range(2)]
    frames = (W.FrameDesc * 2)()
    for i, (fd, _, _) in enumerate(packed):
        C.memmove(C.byref(frames[i]), C.byref(fd), C.sizeof(W.FrameDesc))
    sizes = [C.c_uint64() for _ in range(3)]
    W.lib().webpb200_workspace_sizes(frames, 2, *[C.byref(s) for s in sizes])
    dev = torch.device("cuda:0")
    t_frames = torch.frombuffer(bytearray(bytes(frames)), dtype=torch.uint8).to(dev)
    t_mbs = torch.from_numpy(np.concatenate([p[1] for p in packed]).view(np.uint8)).to(dev)
    t_levels = torch.from_numpy(np.concatenate([p[2].reshape(-1) for p in packed])).to(dev)
    t_planes = torch.zeros(sizes[0].value, dtype=torch.uint8, device=dev)
    t_out = torch.zeros(sizes[1].value, dtype=torch.uint8, device=dev)
    t_scratch = torch.zeros(sizes[2].value, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    rc = W.lib().webpb200_reconstruct_device(frames, t_frames.data_ptr(), 2, t_mbs.data_ptr(), t_levels.data_ptr(), None,
                                             t_planes.data_ptr(), t_out.data_ptr(), t_scratch.data_ptr(), 15, None)
    assert rc == 0
    torch.cuda.synchronize()
    out = t_out.cpu().numpy()
    for i, f in enumerate(files):
        want = oracle.WebPDecoder(f).read_image()
        got = out[frames[i].out_off: frames[i].out_off + want.size].reshape(want.shape)
        assert np.array_equal(got, want)


def test_batch_split_over_contexts(W, oracle, synth, fixtures):
    """webpb200_decode_batch_multi: the batch dealt out over several contexts (one per GPU where the box has
    several, else two contexts on the one device), ranges balanced by macroblock count; every image equals
    the oracle's and a bad file keeps its own status."""
    files = [synth.webp_simple(synth.vp8_frame(64 + 16 * (i % 9), 48 + 16 * (i % 5), seed=500 + i, filter_type=i % 2, filter_level=10 + i % 40,
                                                 log2_parts=i % 3)) for i in range(40)]
    files.append(synth.cfg3_image(7, 1280, 720))      # one heavy file: the split must not be by count
    files.insert(5, bytes(fixtures["webp/gallery2/2_webp_a"]))
    files.insert(11, b"RIFF\x10\0\0\0WEBPjunkjunkjunk")
    ndev = max(1, min(W.device_count(), 4))
    ctxs = [W.BatchDecoder(k % ndev) for k in range(max(2, ndev))]
    outs, status, split = W.decode_batch_multi(ctxs, files, host_threads=8)
    assert split[0] == 0 and split[-1] == len(files) and all(a <= b for a, b in zip(split, split[1:]))
    assert split[1] > len(files) // 2   # the 720p frame at the end weighs as much as all the small ones
    for i, f in enumerate(files):
        if i == 11:
            assert status[i] != 0
            continue
        assert status[i] == 0, i
        assert np.array_equal(outs[i], oracle.WebPDecoder(f).read_image()), i
    for c in ctxs:
        c.close()


@pytest.mark.parametrize("size", [(37, 23), (1, 9), (9, 1), (200, 31), (64, 48), (131, 77), (33, 70)])
def test_vp8l_alpha_transforms_on_device(W, bd, oracle, synth, size):
    """VP8L-compressed ALPH planes: prefix codes on the host, inverse transforms on the device (vp8l.cu) --
    predictor (all fourteen modes, block sizes 2..5), colour, subtract-green, colour indexing with 1, 2 and 4
    bit packing and without, alone and chained, with every alpha filter; every pixel against the oracle
    (reverse_transform.rs:24-578, extended.rs:131-191, 284-294)."""
    from harness import vp8l_writer as V
    w, h = size
    rng = np.random.default_rng(w * 1000 + h)
    alpha = synth.smooth_alpha(w, h, 3)
    few = (rng.integers(0, 4, (h, w)) * 85).astype(np.uint8)

    def pal(a, flt):
        res = np.frombuffer(synth.alph_raw(a, filtering=flt)[9:9 + a.size], np.uint8)
        return 0xff000000 | (np.unique(res).astype(np.uint32) << 8)

    def blocks(bits, hi, lo=0, last=None):
        shape = (V.subsample(h, bits), V.subsample(w, bits)) + (() if last is None else (last,))
        return rng.integers(lo, hi, shape)

    cases = [(alpha, 0, ()), (alpha, 1, (("predictor", 2, blocks(2, 14)),)), (alpha, 2, (("predictor", 5, blocks(5, 14)),)),
             (alpha, 3, (("subtract_green",), ("predictor", 3, 11))), (alpha, 0, (("color", 2, blocks(2, 128, -128, 3)), ("predictor", 4, 12))),
             (alpha, 1, (("predictor", 3, 13), ("color", 3, (5, -7, 100)), ("subtract_green",))),
             (few, 0, (("palette", pal(few, 0)),)), (few, 2, (("palette", pal(few, 2)), ("predictor", 2, 11))),
             ((few > 100).astype(np.uint8) * 255, 0, (("palette", pal((few > 100).astype(np.uint8) * 255, 0)),)),
             (alpha, 3, (("palette", pal(alpha, 3)), ("predictor", 4, 5)))]
    files = []
    for k, (a, flt, tr) in enumerate(cases):
        vp8 = synth.vp8_frame(w, h, seed=900 + k, filter_type=k % 2, filter_level=15)
        files.append(synth.webp_extended(vp8, w, h, synth.chunk(b"ALPH", V.alph_chunk_payload(a, flt, tr))))
    for where in (bd.ENTROPY_HOST, bd.ENTROPY_DEVICE):
        bd.set_entropy(where)
        outs, status = bd.decode(files)
        assert status == [0] * len(files)
        for k, f in enumerate(files):
            want = oracle.WebPDecoder(f).read_image()
            assert np.array_equal(outs[k], want), (where, k)
            assert np.array_equal(W.WebPDecoder(f).read_image(), want), k
    bd.set_entropy(bd.ENTROPY_AUTO)
    # a repeated device-resident pass gives the same bytes (the uploaded pixels are never written by a kernel)
    bd.prepare(files)
    bd.upload()
    for _ in range(2):
        bd.reconstruct(15)
    bd.sync()
    again = [oracle.WebPDecoder(f).read_image() for f in files]
    outs2 = [np.zeros_like(a) for a in again]
    bd.download([o.reshape(-1) for o in outs2])
    for a, b in zip(again, outs2):
        assert np.array_equal(a, b)


def test_alpha_stage_bits_and_pinned_views(W, bd, oracle, synth):
    """Staged API: the ALPH merge happens in the colour conversion of the same reconstruct call (include/webpb200.h): without
    stage bit 3 byte 3 of every RGBA pixel stays 255, with it the plane is merged, whatever the filter method; bit 3 alone
    changes no pixel.  And a numpy view of a PinnedBuffer keeps the pinned memory alive after the object is gone."""
    w, h = 72, 40
    files, wants = [], []
    for filtering in range(4):
        alpha = np.random.default_rng(900 + filtering).integers(0, 256, (h, w)).astype(np.uint8)
        vp8 = synth.vp8_frame(w, h, seed=77 + filtering, filter_type=filtering & 1, filter_level=12)
        files.append(synth.webp_extended(vp8, w, h, synth.alph_raw(alpha, filtering=filtering)))
        wants.append(oracle.WebPDecoder(files[-1]).read_image())
    assert bd.prepare(files) == [0] * 4
    bd.upload()
    pin = W.PinnedBuffer(4 * w * h * 4)
    view = pin.array
    outs = [view[i * w * h * 4:(i + 1) * w * h * 4].reshape(h, w, 4) for i in range(4)]
    del pin                                            # the views own the memory now
    bd.reconstruct(bd.STAGE_PREDICT | bd.STAGE_LOOPFILTER | bd.STAGE_RGB)
    bd.download(outs)
    for o, want in zip(outs, wants):
        assert np.array_equal(o[:, :, :3], want[:, :, :3]) and (o[:, :, 3] == 255).all()
    bd.reconstruct(bd.STAGE_ALPHA)                     # prepares the planes only
    bd.download(outs)
    for o in outs:
        assert (o[:, :, 3] == 255).all()
    bd.reconstruct()
    bd.download(outs)
    for o, want in zip(outs, wants):
        assert np.array_equal(o, want)
