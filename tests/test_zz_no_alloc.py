"""webpb200_predict_device_static: the prediction stage for callers that may not allocate.

Some users of this library run under a fixed device-memory budget (or inside another runtime's
memory pool) and must not call cudaMalloc once they are up; for them the wavefront's working
memory sits in a fixed array inside the library.  This test drives that entry point through the
C ABI and checks that the planes it writes are byte-identical to those of the regular entry point
(webpb200_reconstruct_device with a caller-supplied scratch arena), twice in a row with different
batches -- the second call finds the first call's leftovers in the shared array and must not be
disturbed by them.  It also checks that a batch beyond the stated limit is refused.

Named test_zz_* so that pytest collects it last; it passes on its own in a fresh process as well
as after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

from util import SYNTH_CASES, product, vp8_payload

pytestmark = pytest.mark.gpu


def _device_batch(W, bd, torch, files):
    """Host-prepare `files`, then put the packed arrays into caller-owned (torch) device memory."""
    n = len(files)
    assert bd.prepare(files) == [0] * n
    packed = [bd.packed(i) for i in range(n)]
    frames = (W.FrameDesc * n)()
    for i, (fd, _, _) in enumerate(packed):
        C.memmove(C.byref(frames[i]), C.byref(fd), C.sizeof(W.FrameDesc))
    sizes = [C.c_uint64() for _ in range(3)]
    W.lib().webpb200_workspace_sizes(frames, n, *[C.byref(s) for s in sizes])
    dev = torch.device("cuda:0")
    t = {
        "frames": torch.frombuffer(bytearray(bytes(frames)), dtype=torch.uint8).to(dev),
        "mbs": torch.from_numpy(np.concatenate([p[1] for p in packed]).view(np.uint8)).to(dev),
        "levels": torch.from_numpy(np.concatenate([p[2].reshape(-1) for p in packed])).to(dev),
    }
    n_mbs = sum(p[1].size for p in packed)
    return frames, t, [s.value for s in sizes], n_mbs


def _run_both(W, torch, frames, t, sizes, n_mbs):
    dev = torch.device("cuda:0")
    n = len(frames)
    # regular entry point: scratch arena supplied by the caller
    planes_a = torch.full((sizes[0],), 0xAA, dtype=torch.uint8, device=dev)
    out = torch.zeros(max(sizes[1], 1), dtype=torch.uint8, device=dev)
    scratch = torch.zeros(sizes[2], dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    rc = W.lib().webpb200_reconstruct_device(frames, t["frames"].data_ptr(), n, t["mbs"].data_ptr(), t["levels"].data_ptr(), None,
                                             planes_a.data_ptr(), out.data_ptr(), scratch.data_ptr(), 1, None)
    assert rc == 0
    torch.cuda.synchronize()
    # the entry point under test: every buffer below exists before the call
    planes_b = torch.full((sizes[0],), 0xAA, dtype=torch.uint8, device=dev)
    nonzero = torch.full((n_mbs,), 0x55, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    rc = W.lib().webpb200_predict_device_static(frames, t["frames"].data_ptr(), n, t["mbs"].data_ptr(), t["levels"].data_ptr(),
                                                planes_b.data_ptr(), nonzero.data_ptr(), None)
    assert rc == 0
    # the call only queues work: this waits for the stream and says whether the wavefront ran to its end
    assert W.lib().webpb200_predict_device_static_check(None) == 0
    torch.cuda.synchronize()
    return planes_a.cpu().numpy(), planes_b.cpu().numpy(), nonzero.cpu().numpy()


def test_static_scratch_matches_regular_entry_point_twice():
    torch = pytest.importorskip("torch")
    W = product()
    W.lib()
    assert W.device_count() > 0
    from harness import synth as S
    from oracle import oracle as O
    S.build()
    bd = W.BatchDecoder(0)
    batches = [
        [S.webp_simple(S.vp8_frame(**SYNTH_CASES[k])) for k in ("delta_segments_lfdeltas_8parts_skip_160x160", "uniform_17x33", "cfg2_like_512x512")],
        [S.webp_simple(S.vp8_frame(**SYNTH_CASES[k])) for k in ("normal_sharp3_2parts_131x77", "even_odd_254x255")],
    ]
    for files in batches:  # two calls in a row on the same shared array, different shapes
        frames, t, sizes, n_mbs = _device_batch(W, bd, torch, files)
        a, b, nonzero = _run_both(W, torch, frames, t, sizes, n_mbs)
        assert np.array_equal(a, b)  # integer work: exact, the tolerance the other GPU tests use
        # and both equal the oracle's unfiltered planes; the flags equal its non_zero_dct
        for i, f in enumerate(files):
            fr = O.Vp8Frame(vp8_payload(f), keep=True)
            fd = frames[i]
            y = b[fd.y_off: fd.y_off + fr.y0.size].reshape(fr.y0.shape)
            assert np.array_equal(y, fr.y0)
            want = np.array([m.non_zero_dct for m in fr.mbs], dtype=np.uint8)
            assert np.array_equal(nonzero[fd.mb_base: fd.mb_base + want.size], want)
    bd.close()


def test_static_scratch_rejects_oversized_batches():
    W = product()
    W.lib()
    assert W.device_count() > 0
    frames = (W.FrameDesc * 1)()
    frames[0].width, frames[0].height = 16383, 16383
    frames[0].mbw, frames[0].mbh = 1024, 1024  # 1 Mi macroblocks: beyond WEBPB200_STATIC_MAX_MBS
    dummy = C.c_void_p(256)                     # never dereferenced: the size check comes first
    rc = W.lib().webpb200_predict_device_static(frames, dummy, 1, dummy, dummy, dummy, None, None)
    assert rc == -102  # -WEBPB200_ERR_STATIC_SCRATCH_LIMIT (negative, like webpb200_reconstruct_device)
    frames[0].mbw, frames[0].mbh = 8, 8193      # too many macroblock rows
    rc = W.lib().webpb200_predict_device_static(frames, dummy, 1, dummy, dummy, dummy, None, None)
    assert rc == -102
    many = (W.FrameDesc * 9)()                  # 9 frames of 1024 macroblock columns: beyond WEBPB200_STATIC_MAX_WIDTH_MBS
    for i in range(9):
        many[i].width, many[i].height, many[i].mbw, many[i].mbh = 16383, 16, 1024, 1
        many[i].row_base, many[i].mb_base = i, 1024 * i
    rc = W.lib().webpb200_predict_device_static(many, dummy, 9, dummy, dummy, dummy, None, None)
    assert rc == -102
    frames[0].mbw, frames[0].mbh = 0, 4         # a frame without macroblocks is refused, not walked
    rc = W.lib().webpb200_predict_device_static(frames, dummy, 1, dummy, dummy, dummy, None, None)
    assert rc == -23  # -WEBPB200_ERR_INVALID_PARAM
