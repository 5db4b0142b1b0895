"""Device entropy stage (k_entropy) against the host entropy decoder and the CPU oracle.

The stage replaces, on the GPU, the boolean decode of macroblock headers and coefficient tokens
(reference lossy/mod.rs:615-671, 785-963, 1255-1281).  Its outputs are integers: every comparison
is bit-exact -- macroblock descriptors and levels against webpb200_host_entropy_decode, pixels
against the oracle, statuses of damaged streams against the host path.
"""
import numpy as np
import pytest

from util import LOSSY_STILL, SYNTH_CASES, product

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def W():
    w = product()
    w.lib()
    assert w.device_count() > 0, "no CUDA device"
    return w


@pytest.fixture(scope="module")
def synth():
    from harness import synth as S
    S.build()
    return S


@pytest.fixture()
def dev(W):
    b = W.BatchDecoder(0)
    b.set_entropy(b.ENTROPY_DEVICE)
    yield b
    b.close()


@pytest.fixture()
def host(W):
    b = W.BatchDecoder(0)
    b.set_entropy(b.ENTROPY_HOST)
    yield b
    b.close()


def _check_packed(W, dev, files, status):
    for i, f in enumerate(files):
        if status[i]:
            continue
        fd, mbs, levels, _ = W.host_entropy_decode(f)
        dm, dl = dev.device_packed(i, fd)
        assert np.array_equal(dm, mbs), "macroblock descriptors of image %d" % i
        assert np.array_equal(dl, levels), "levels of image %d" % i


def test_golden_files(W, dev, fixtures, oracle):
    files = [bytes(fixtures["webp/" + k]) for k in LOSSY_STILL]
    outs, status = dev.decode(files)
    assert status == [0] * len(files)
    _check_packed(W, dev, files, status)
    for f, o in zip(files, outs):
        assert np.array_equal(o, oracle.WebPDecoder(f).read_image())


def test_synthetic_features(W, dev, synth, oracle):
    """Every feature no fixture covers: partitions 1/2/4/8, segment maps, skip flags, probability
    updates, CAT6 extremes, all-skipped frames, 1x1, one-row and one-column frames."""
    files = [synth.webp_simple(synth.vp8_frame(**kw)) for kw in SYNTH_CASES.values()]
    outs, status = dev.decode(files)
    assert status == [0] * len(files)
    _check_packed(W, dev, files, status)
    for name, f, o in zip(SYNTH_CASES, files, outs):
        assert np.array_equal(o, oracle.WebPDecoder(f).read_image()), name


def test_alpha_and_mixed_partitions(W, dev, synth, oracle):
    w, h = 96, 80
    files = [synth.webp_extended(synth.vp8_frame(w, h, seed=20 + k, filter_type=0, filter_level=9, log2_parts=k % 4), w, h,
                                 synth.alph_raw(synth.smooth_alpha(w, h, k), filtering=k % 4)) for k in range(8)]
    files += [synth.webp_simple(synth.vp8_frame(200, 120, seed=40 + k, log2_parts=3 - k % 4)) for k in range(4)]
    outs, status = dev.decode(files)
    assert status == [0] * len(files)
    _check_packed(W, dev, files, status)
    for f, o in zip(files, outs):
        assert np.array_equal(o, oracle.WebPDecoder(f).read_image())


def test_damaged_streams_match_host(W, dev, host, synth):
    """Truncations at every kind of place (header, first partition, partition table, tokens): the
    per-image status must equal the host path's, and intact neighbours must be unaffected."""
    good = synth.webp_simple(synth.vp8_frame(160, 160, seed=4, filter_type=0, filter_level=12, log2_parts=3, prob_skip=180, p_skip=0.2))
    small = synth.webp_simple(synth.vp8_frame(64, 48, seed=9))
    cuts = [len(good) - 1, len(good) - 2, len(good) - 40, len(good) * 3 // 4, len(good) // 2, len(good) // 3, 400, 120, 60, 40, 31, 25]
    files = [good] + [good[:c] for c in cuts] + [small] + [small[:c] for c in (len(small) - 1, len(small) - 3, len(small) // 2, 50)]
    # RIFF sizes are left stale on purpose: the container walk decides what is readable, as in the reference
    o_dev, s_dev = dev.decode(files)
    o_host, s_host = host.decode(files)
    assert s_dev == s_host
    assert s_dev[0] == 0 and any(s_dev)
    for a, b, st in zip(o_dev, o_host, s_dev):
        if st == 0:
            assert np.array_equal(a, b)


def test_batch_scale_checksum(W, dev, host, synth):
    """256 cfg2-profile images: device-entropy pixels == host-entropy pixels, image by image."""
    files = synth.make_batch(synth.cfg2_image, 256)
    o_dev, s_dev = dev.decode(files)
    o_host, s_host = host.decode(files)
    assert s_dev == s_host == [0] * len(files)
    for i, (a, b) in enumerate(zip(o_dev, o_host)):
        assert np.array_equal(a, b), i
    _check_packed(W, dev, files[:4], s_dev[:4])


def _contiguous_outs(n, per):
    buf = np.zeros(n * per, dtype=np.uint8)
    return buf, [buf[i * per:(i + 1) * per] for i in range(n)]


@pytest.mark.parametrize("w,h,n,kw", [
    (512, 512, 128, dict(filter_type=1, filter_level=25, prob_skip=200, p_skip=0.05)),       # cfg2 shape, simple filter
    (200, 120, 320, dict(filter_type=0, filter_level=40, sharpness=2)),                      # ragged size, normal filter
    (333, 257, 96, dict(filter_type=0, filter_level=63, sharpness=7, log2_parts=2)),         # odd size, strongest filter, 4 partitions
    (256, 256, 96, dict(filter_type=0, filter_level=0)),                                     # no loop filter at all
])
def test_band_pipeline_matches_host_path(W, dev, host, synth, w, h, n, kw):
    """Same-size batches with evenly spaced output buffers take the band-pipelined route (bands of
    macroblock rows reconstructed and copied out while k_entropy still decodes the rows below).
    The band edges are where it could go wrong (loop filter one band late, colour conversion held
    back 3 rows): compare every pixel with the host-entropy path, which runs whole frames."""
    files = [synth.webp_simple(synth.vp8_frame(w, h, seed=1000 + i, **kw)) for i in range(n)]
    per = w * h * 3
    _, o_dev = _contiguous_outs(n, per)
    _, o_host = _contiguous_outs(n, per)
    _, s_dev = dev.decode(files, outs=o_dev)
    _, s_host = host.decode(files, outs=o_host)
    assert s_dev == s_host == [0] * n
    for i in range(n):
        assert np.array_equal(o_dev[i], o_host[i]), i
    # fancy and simple upsampling both
    _, s2 = dev.decode(files[:64] if n >= 128 else files, outs=o_dev[:64] if n >= 128 else o_dev, lossy_upsampling=W.SIMPLE)
    _, s3 = host.decode(files[:64] if n >= 128 else files, outs=o_host[:64] if n >= 128 else o_host, lossy_upsampling=W.SIMPLE)
    assert s2 == s3
    m = 64 if n >= 128 else n
    for i in range(m):
        assert np.array_equal(o_dev[i], o_host[i]), i


def test_band_pipeline_with_a_damaged_frame(W, dev, host, synth):
    """A frame that runs out of bits inside a token partition, in a batch that takes the banded route:
    its status must say so (as on the host path) and its neighbours must be intact."""
    n, w, h = 96, 256, 256
    files = [synth.webp_simple(synth.vp8_frame(w, h, seed=2000 + i, filter_type=0, filter_level=20)) for i in range(n)]
    files[17] = files[17][:len(files[17]) * 2 // 3]
    files[40] = files[40][:len(files[40]) - 5]
    per = w * h * 3
    _, o_dev = _contiguous_outs(n, per)
    _, o_host = _contiguous_outs(n, per)
    _, s_dev = dev.decode(files, outs=o_dev)
    _, s_host = host.decode(files, outs=o_host)
    assert s_dev == s_host and s_dev[17] != 0
    for i in range(n):
        if s_dev[i] == 0:
            assert np.array_equal(o_dev[i], o_host[i]), i


def test_inputs_in_pinned_memory_are_uploaded_in_place(W, dev, synth):
    """Files that lie in one block from webpb200_host_alloc skip the staging copy (the block goes to the
    device as it is, gaps included): same pixels as from ordinary memory."""
    files = [synth.webp_simple(synth.vp8_frame(96 + 16 * (i % 3), 64, seed=300 + i, log2_parts=i % 3)) for i in range(40)]
    want, st0 = dev.decode(files)
    pin = W.PinnedBuffer(sum(len(f) + 100 for f in files))
    views, off = [], 7  # odd start, ragged gaps
    for i, f in enumerate(files):
        pin.array[off:off + len(f)] = np.frombuffer(f, dtype=np.uint8)
        views.append(pin.array[off:off + len(f)])
        off += len(f) + (i * 13) % 90
    got, st1 = dev.decode(views)
    assert st0 == st1 == [0] * len(files)
    for a, b in zip(want, got):
        assert np.array_equal(a, b)
    pin.free()


def test_randomised_streams_device_vs_host(W, dev, host, synth):
    """120 frames with randomly drawn stream parameters (size incl. 1-pixel-wide and single-macroblock
    frames, 1/2/4/8 partitions, segments with and without a map, skip probabilities, probability
    updates, every coefficient profile, quantiser range), a third of them cut at a random place:
    descriptors, levels, statuses and pixels of the device entropy stage equal the host's."""
    rng = np.random.default_rng(20261107)
    profiles = ["realistic", "uniform", "extreme", "all_tm", "bpred_ld_vl", "zero_tails"]
    files = []
    for i in range(120):
        w = int(rng.choice([1, 15, 16, 17, 33, 64, 100, 160, 250]))
        h = int(rng.choice([1, 16, 31, 48, 65, 128, 200]))
        seg = int(rng.integers(0, 5))
        kw = dict(width=w, height=h, seed=5000 + i, filter_type=int(rng.integers(0, 2)), filter_level=int(rng.integers(0, 64)),
                  sharpness=int(rng.integers(0, 8)), log2_parts=int(rng.integers(0, 4)), yac_abs=int(rng.integers(0, 128)),
                  profile=str(rng.choice(profiles)), update_probs=int(rng.integers(0, 2)))
        if seg:
            kw.update(segments=seg, seg_quant=tuple(int(v) for v in rng.integers(0, 100, 4)), seg_lf=tuple(int(v) for v in rng.integers(0, 60, 4)),
                      seg_probs=tuple(int(v) for v in rng.integers(1, 255, 3)))
        if rng.integers(0, 2):
            kw.update(prob_skip=int(rng.integers(1, 255)), p_skip=float(rng.choice([0.0, 0.1, 0.5, 1.0])))
        try:
            f = synth.webp_simple(synth.vp8_frame(**kw))
        except ValueError:
            continue
        if i % 3 == 2:
            f = f[:int(rng.integers(21, len(f)))]
        files.append(f)
    assert len(files) > 100
    o_dev, s_dev = dev.decode(files)
    o_host, s_host = host.decode(files)
    assert s_dev == s_host
    assert sum(1 for s in s_dev if s == 0) > 60 and any(s_dev)
    for i, (a, b, st) in enumerate(zip(o_dev, o_host, s_dev)):
        if st == 0:
            assert np.array_equal(a, b), i
    _check_packed(W, dev, files, s_dev)


@pytest.mark.parametrize("where", ["device", "host"])
def test_images_left_in_device_memory(W, dev, host, synth, where):
    """outs = NULL: no download; webpb200_batch_output_device hands out the device copy of each image
    (wrapped here as torch tensors through __cuda_array_interface__, no copy)."""
    import torch
    bd = dev if where == "device" else host
    uniform = [synth.webp_simple(synth.vp8_frame(256, 256, seed=700 + i, filter_type=0, filter_level=15)) for i in range(80)]  # banded route
    w, h = 96, 80
    mixed = [synth.webp_extended(synth.vp8_frame(w, h, seed=20 + k, filter_type=0, filter_level=9), w, h,
                                 synth.alph_raw(synth.smooth_alpha(w, h, k), filtering=k % 4)) for k in range(4)]
    mixed += [synth.webp_simple(synth.vp8_frame(200, 120, seed=40)), b"junk"]
    for files in (uniform, mixed):
        want, st_want = host.decode(files)
        got, st = bd.decode_to_device(files)
        assert st == st_want
        assert bd.last_transfer()[1] == 0  # nothing was downloaded
        for g, wnt, s in zip(got, want, st):
            if s:
                assert g is None
                continue
            t = torch.as_tensor(g, device="cuda")
            assert tuple(t.shape) == g.shape and t.dtype == torch.uint8
            assert np.array_equal(t.cpu().numpy().reshape(-1), np.asarray(wnt).reshape(-1))


def test_uniform_size_random_parameters_and_ragged_count(W, dev, host, synth):
    """131 frames of one size (so the band pipeline and, for single-partition batches, the
    four-frames-per-block entropy layout are taken; 131 is not a multiple of 4) with per-frame random
    filter type / level / sharpness / segments / skip, two of them damaged and one not a WebP at all;
    then the same with random partition counts (one frame per block)."""
    rng = np.random.default_rng(77)
    w, h, n = 208, 176, 131
    for parts in (False, True):
        files = []
        for i in range(n):
            kw = dict(filter_type=int(rng.integers(0, 2)), filter_level=int(rng.integers(0, 64)), sharpness=int(rng.integers(0, 8)),
                      yac_abs=int(rng.integers(4, 100)), log2_parts=int(rng.integers(0, 4)) if parts else 0)
            if rng.integers(0, 2):
                kw.update(segments=1, seg_quant=tuple(int(v) for v in rng.integers(5, 90, 4)), seg_lf=tuple(int(v) for v in rng.integers(0, 60, 4)),
                          seg_probs=tuple(int(v) for v in rng.integers(1, 255, 3)))
            if rng.integers(0, 2):
                kw.update(prob_skip=int(rng.integers(1, 255)), p_skip=float(rng.choice([0.1, 0.6])))
            files.append(synth.webp_simple(synth.vp8_frame(w, h, seed=9000 + i, **kw)))
        files[5] = files[5][:len(files[5]) // 2]
        files[129] = files[129][:len(files[129]) - 3]
        files[64] = b"RIFF\\x10\\x00\\x00\\x00WEBPjunkjunkjunk"
        per = w * h * 3
        _, o_dev = _contiguous_outs(n, per)
        _, o_host = _contiguous_outs(n, per)
        _, s_dev = dev.decode(files, outs=o_dev)
        _, s_host = host.decode(files, outs=o_host)
        assert s_dev == s_host and s_dev[5] != 0 and s_dev[64] != 0
        for i in range(n):
            if s_dev[i] == 0:
                assert np.array_equal(o_dev[i], o_host[i]), (parts, i)
        good = [i for i in range(n) if s_dev[i] == 0]
        _check_packed(W, dev, [files[i] for i in good[:3]], [0, 0, 0]) if good[:3] == [0, 1, 2] else None


def test_two_batches_in_flight(W, synth, oracle):
    """webpb200_decode_batch_begin / _end: two contexts on one device, batches alternating between them so that
    two are always in flight; every image of every batch equals the oracle's, and a second begin on a busy
    context is refused."""
    n, w, h = 96, 256, 192
    batches = [[synth.webp_simple(synth.vp8_frame(w, h, seed=7000 + 100 * b + i, filter_type=(b + i) % 2, filter_level=10 + i % 30))
                for i in range(n)] for b in range(4)]
    ctxs = [W.BatchDecoder(0), W.BatchDecoder(0)]
    for c in ctxs:
        c.set_entropy(c.ENTROPY_DEVICE)
    per = w * h * 3
    bufs = [_contiguous_outs(n, per)[1] for _ in range(4)]
    lists = [(W.BufferList([np.frombuffer(f, np.uint8) for f in batches[b]]), W.BufferList(bufs[b])) for b in range(4)]
    pending = [None, None]
    for b in range(4):
        k = b % 2
        if pending[k] is not None:
            assert ctxs[k].decode_end(pending[k]) == [0] * n
        pending[k] = ctxs[k].decode_begin(lists[b][0], lists[b][1])
        if b == 1:  # both busy now: a third begin on a busy context must be refused, not queued
            with pytest.raises(W.DecodingError):
                ctxs[k].decode_begin(lists[b][0], lists[b][1])
    for k in range(2):
        assert ctxs[k].decode_end(pending[k]) == [0] * n
    for b in range(4):
        want, st, bad = oracle.decode_batch(batches[b], threads=8)
        assert bad == 0
        for i in range(n):
            assert np.array_equal(bufs[b][i], want[i].reshape(-1)), (b, i)
    for c in ctxs:
        c.close()


def test_split_layout_without_its_mode_blocks(W, synth, oracle):
    """The split layout of k_entropy (what webpb200_decode_batch_begin uses for single-partition batches: mode partitions
    and token partitions in separate thread blocks).  Nothing guarantees that a frame's mode block starts before its
    token block: with the test hook that makes every mode block do nothing, each token warp has to claim its frame's
    mode partition and decode it itself, and every image must still equal the oracle's."""
    n, w, h = 40, 192, 160
    files = [synth.webp_simple(synth.vp8_frame(w, h, seed=9100 + i, filter_type=i % 2, filter_level=5 + i)) for i in range(n)]
    want, st, bad = oracle.decode_batch(files, threads=8)
    assert bad == 0
    per = w * h * 3
    lib = W.lib()
    for faults in (0, 1):
        lib.webpb200_debug_entropy_faults(faults)
        try:
            bd = W.BatchDecoder(0)
            bd.set_entropy(bd.ENTROPY_DEVICE)
            outs = _contiguous_outs(n, per)[1]
            fl, ol = W.BufferList([np.frombuffer(f, np.uint8) for f in files]), W.BufferList(outs)
            assert bd.decode_end(bd.decode_begin(fl, ol)) == [0] * n
            for i in range(n):
                assert np.array_equal(outs[i], want[i].reshape(-1)), (faults, i)
            bd.close()
        finally:
            lib.webpb200_debug_entropy_faults(0)
