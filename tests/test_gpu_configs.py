"""Parity at the BASELINE.json shapes: every pixel of cfg3 / cfg4 / cfg5 inputs at their real
sizes against the CPU oracle, through the C ABI.  The small synthetic cases of test_gpu_parity.py cover the
features; these cover what only the real sizes reach -- the 8-partition device entropy stage at 240
macroblock columns, the band pipeline on 4K frames, VP8L-compressed ALPH planes at 1080p with every alpha
filter and every VP8L transform, and a 1080p animation with offset frames.

Reference paths: still decode + ALPH merge decoder.rs:669-741, animation decoder.rs:752-922, normal filter with
segments and deltas lossy/mod.rs:966-1244, ALPH chunk extended.rs:253-308 (+ lossless/decoder).
"""
import numpy as np
import pytest

from util import product

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def W():
    w = product()
    w.lib()
    assert w.device_count() > 0, "no CUDA device"
    return w


@pytest.fixture(scope="module")
def cfg():
    from harness import synth as S
    from harness import configs
    S.build()
    return configs


def _contiguous(sizes):
    buf = np.zeros(int(sum(sizes)), dtype=np.uint8)
    outs, at = [], 0
    for s in sizes:
        outs.append(buf[at:at + s])
        at += s
    return outs


@pytest.mark.parametrize("where", ["device", "host"])
def test_cfg3_4k_frames(W, oracle, cfg, where):
    """4 frames 3840x2160 (normal filter, 4 segments, LF deltas, sharpness 3, 8 DCT partitions), evenly spaced
    output buffers: with the entropy stage on the device this is the band-pipelined route with eight
    partition warps per frame; with it on the host, the chunked upload route."""
    files = cfg.still_batch("cfg3", 0, 4, threads=8)
    want, st, bad = oracle.decode_batch(files, threads=8)
    assert bad == 0
    bd = W.BatchDecoder(0)
    bd.set_entropy(bd.ENTROPY_DEVICE if where == "device" else bd.ENTROPY_HOST)
    outs = _contiguous([3840 * 2160 * 3] * 4)
    _, status = bd.decode(files, outs=outs)
    assert status == [0] * 4
    assert bd.last_transfer()[2] == where
    for i in range(4):
        assert np.array_equal(outs[i], want[i].reshape(-1)), i
    # and one frame alone (no band pipeline, AUTO placement) gives the same bytes
    bd.set_entropy(bd.ENTROPY_AUTO)
    alone, status = bd.decode([files[2]])
    assert status == [0] and np.array_equal(alone[0].reshape(-1), outs[2])
    bd.close()


def test_cfg4_1080p_vp8l_alpha(W, oracle, cfg):
    """8 images 1920x1080 VP8X + VP8L-compressed ALPH + normal-filter VP8, RGBA out: alpha filter methods 0..3
    twice over, the VP8L streams with no transform / predictor (14 modes) / subtract-green + predictor /
    colour + predictor; plus two whose ALPH chunk libwebp's own encoder made."""
    files = [cfg.cfg4_image(i) for i in range(8)]
    for i in range(2):
        alph = cfg.libwebp_alpha_chunk(i)
        if alph is not None:
            files.append(cfg.cfg4_image(100 + i, alph=alph))
    n = len(files)
    want, st, bad = oracle.decode_batch(files, threads=8)
    assert bad == 0
    bd = W.BatchDecoder(0)
    for where in (bd.ENTROPY_HOST, bd.ENTROPY_DEVICE):
        bd.set_entropy(where)
        outs, status = bd.decode(files)
        assert status == [0] * n
        for i in range(n):
            assert np.array_equal(outs[i].reshape(-1), want[i].reshape(-1)), (where, i)
    bd.close()


@pytest.mark.parametrize("offsets", [False, True])
def test_cfg5_animation_1080p(W, oracle, cfg, offsets):
    """12-frame 1080p animation (ALPH + VP8 per frame, blend on, dispose alternating; the variant has offset
    sub-rectangle frames): every composited canvas against the oracle's read_frame sequence."""
    data = cfg.cfg5_animation(12, offsets=offsets, threads=8)
    o = oracle.WebPDecoder(data)
    bd = W.BatchDecoder(0)
    frames, durs = bd.decode_animation(data)
    assert len(frames) == 12
    for k in range(12):
        want, dur = o.read_frame()
        assert dur == durs[k]
        assert np.array_equal(frames[k].reshape(-1), want.reshape(-1)), k
    bd.close()
    # frame by frame through the WebPDecoder surface: the first three
    d = W.WebPDecoder(data)
    o = oracle.WebPDecoder(data)
    for k in range(3):
        a, da = o.read_frame()
        b, db = d.read_frame()
        assert da == db and np.array_equal(a.reshape(-1), b.reshape(-1)), k
