#!/usr/bin/env python3
"""Per-stage diff of the CUDA path against the CPU oracle on a handful of inputs (debug aid,
run on the GPU box: python tools/gpu_diag.py > gpurun_out/diag.log)."""
import importlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
W = importlib.import_module("image-webp_b200")
from harness import synth as S  # noqa: E402
from oracle import oracle as O  # noqa: E402


def vp8_payload(data):
    b = bytes(data)
    i = b.find(b"VP8 ")
    n = int.from_bytes(b[i + 4:i + 8], "little")
    return b[i + 8:i + 8 + n]


def first_bad(a, b, blk):
    d = np.argwhere(a != b)
    if len(d) == 0:
        return "ok"
    y, x = d[0][:2]
    return "%d bad, first at px(%d,%d) mb(%d,%d) got %s want %s" % (len(d), x, y, x // blk, y // blk, a[tuple(d[0])], b[tuple(d[0])])


def diag(bd, name, data, upsampling=0):
    fr = O.Vp8Frame(vp8_payload(data), keep=True)
    st = bd.prepare([data], lossy_upsampling=upsampling)
    if st[0]:
        print(name, "prepare status", st)
        return False
    bd.upload()
    bd.reconstruct(1); bd.sync()
    y, u, v = bd.planes(0)
    r1 = [first_bad(y, fr.y0, 16), first_bad(u, fr.u0, 8), first_bad(v, fr.v0, 8)]
    bd.reconstruct(3); bd.sync()
    y, u, v = bd.planes(0)
    r2 = [first_bad(y, fr.y, 16), first_bad(u, fr.u, 8), first_bad(v, fr.v, 8)]
    bd.reconstruct(15); bd.sync()
    d = O.WebPDecoder(data, upsampling="bilinear" if upsampling == 0 else "simple")
    want = d.read_image()
    out = np.zeros_like(want)
    bd.download([out])
    r3 = first_bad(out, want, 16)
    ok = all(r == "ok" for r in r1 + r2 + [r3])
    print("%-40s %s\n   predict Y:%s U:%s V:%s\n   filter  Y:%s U:%s V:%s\n   rgb     %s" % (name, "OK" if ok else "FAIL", *r1, *r2, r3))
    return ok


def main():
    print("devices", W.device_count())
    bd = W.BatchDecoder(0)
    fx = np.load(os.path.join(ROOT, "tests", "golden", "fixtures.npz"))
    ok = True
    ok &= diag(bd, "1x1", S.webp_simple(S.vp8_frame(1, 1, seed=10)))
    ok &= diag(bd, "16x16 one MB", S.webp_simple(S.vp8_frame(16, 16, seed=1, filter_level=0)))
    ok &= diag(bd, "48x16 row", S.webp_simple(S.vp8_frame(48, 16, seed=2, filter_level=0)))
    ok &= diag(bd, "48x48 nofilter", S.webp_simple(S.vp8_frame(48, 48, seed=3, filter_level=0)))
    ok &= diag(bd, "48x48 simple", S.webp_simple(S.vp8_frame(48, 48, seed=3, filter_type=1, filter_level=30)))
    ok &= diag(bd, "48x48 normal", S.webp_simple(S.vp8_frame(48, 48, seed=3, filter_type=0, filter_level=30)))
    ok &= diag(bd, "all_tm 64x48", S.webp_simple(S.vp8_frame(64, 48, seed=6, profile="all_tm", filter_level=0)))
    ok &= diag(bd, "bpred_ld_vl 80x80", S.webp_simple(S.vp8_frame(80, 80, seed=7, profile="bpred_ld_vl", filter_type=1, filter_level=50, sharpness=6)))
    ok &= diag(bd, "uniform 131x77", S.webp_simple(S.vp8_frame(131, 77, seed=2, profile="uniform", filter_type=0, filter_level=30, sharpness=3, log2_parts=1)))
    ok &= diag(bd, "extreme 64x64", S.webp_simple(S.vp8_frame(64, 64, seed=5, profile="extreme", yac_abs=127, filter_type=0, filter_level=40, sharpness=5)))
    ok &= diag(bd, "segments+deltas 160x160", S.webp_simple(S.vp8_frame(160, 160, seed=4, filter_type=0, filter_level=12, segments=2, seg_quant=(-10, 5, 20, -3), seg_lf=(3, -5, 10, 0), seg_probs=(90, 128, -1), lf_adj=2, ref_delta=(2, 0, 0, 0), mode_delta=(-3, 0, 0, 0), log2_parts=3, prob_skip=180, p_skip=0.2)))
    ok &= diag(bd, "cfg2 512x512", S.cfg2_image(0))
    ok &= diag(bd, "simple upsampling 150x90", S.webp_simple(S.vp8_frame(150, 90, seed=1, filter_type=1, filter_level=20)), upsampling=1)
    for k in ["gallery1/1", "gallery2/3_webp_a", "regression/dark"]:
        ok &= diag(bd, k, fx["webp/" + k])
    # batch
    t = time.time()
    files = S.make_batch(S.cfg2_image, 64)
    print("made 64 cfg2 files %.2fs" % (time.time() - t))
    t = time.time()
    outs, status = bd.decode(files)
    print("decode_batch 64: %.3fs status ok=%s" % (time.time() - t, all(s == 0 for s in status)))
    want, st2, bad = O.decode_batch(files, threads=8)
    nbad = sum(int((a != b).any()) for a, b in zip(outs, want))
    print("batch mismatching images:", nbad)
    ok &= nbad == 0
    ms, k = bd.time(15, iters=5)
    print("device ms/pass %.3f kernels %s -> %.1f MPix/s" % (ms, ["%.3f" % x for x in k], 64 * 512 * 512 / ms / 1e3))
    print("ALL OK" if ok else "SOME FAILED")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
