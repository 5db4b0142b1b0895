#!/usr/bin/env python3
"""Soak: many decode calls of different kinds on ONE context, in random order, each checked against
results computed once with the host-entropy path of a separate context.  Looks for state leaking from
one call into the next (band pipeline, second stream, worker pool, buffer reuse)."""
import importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
W = importlib.import_module("image-webp_b200")
from harness import synth as S

rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
seconds = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
ref = W.BatchDecoder(0); ref.set_entropy(ref.ENTROPY_HOST)
bd = W.BatchDecoder(0)

def uniform(n, w, h, **kw):
    return [S.webp_simple(S.vp8_frame(w, h, seed=int(rng.integers(1 << 30)), **kw)) for _ in range(n)]

cases = {}
cases["cfg2x192"] = uniform(192, 512, 512, filter_type=1, filter_level=20, prob_skip=200, p_skip=0.05)
cases["small_normal"] = uniform(300, 208, 176, filter_type=0, filter_level=33, sharpness=3)
cases["parts4"] = uniform(96, 333, 257, filter_type=0, filter_level=50, log2_parts=2)
cases["mixed"] = ([S.webp_simple(S.vp8_frame(int(rng.integers(1, 300)), int(rng.integers(1, 300)), seed=i, log2_parts=i % 4)) for i in range(40)]
                  + [S.webp_extended(S.vp8_frame(96, 80, seed=50 + k, filter_type=0, filter_level=9), 96, 80,
                                     S.alph_raw(S.smooth_alpha(96, 80, k), filtering=k % 4)) for k in range(8)] + [b"junk"])
cases["big"] = uniform(3, 1920, 1080, filter_type=0, filter_level=25)
want = {}
for k, files in cases.items():
    o, st = ref.decode(files)
    want[k] = ([np.array(x).reshape(-1).copy() for x in o], st)
print("cases ready", {k: len(v) for k, v in cases.items()}, flush=True)

t0 = time.time(); calls = 0
while time.time() - t0 < seconds:
    k = list(cases)[int(rng.integers(len(cases)))]
    files = cases[k]
    mode = int(rng.integers(0, 3))
    bd.set_entropy(mode)
    style = int(rng.integers(0, 3))
    exp, est = want[k]
    if style == 0:                      # separate output arrays
        o, st = bd.decode(files)
        got = [np.asarray(x).reshape(-1) for x in o]
    elif style == 1 and k != "mixed":   # one contiguous block (band pipeline when sizes agree)
        per = exp[0].size
        buf = np.zeros(per * len(files), np.uint8)
        outs = [buf[i * per:(i + 1) * per] for i in range(len(files))]
        _, st = bd.decode(files, outs=outs)
        got = outs
    else:                               # left on the device
        import torch
        o, st = bd.decode_to_device(files)
        got = [torch.as_tensor(x, device="cuda").cpu().numpy().reshape(-1) if x is not None else None for x in o]
    assert st == est, (k, mode, style, st, est)
    for i, (a, b) in enumerate(zip(got, exp)):
        if st[i] == 0:
            assert np.array_equal(a, b), (k, mode, style, i)
    calls += 1
print("soak ok: %d calls in %.0f s" % (calls, time.time() - t0))
