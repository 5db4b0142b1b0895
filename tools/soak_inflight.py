#!/usr/bin/env python3
"""Soak for calls in flight: K contexts on one GPU, webpb200_decode_batch_begin / _end going round them with batches
of different kinds (band pipeline or not, single- and multi-partition frames, ALPH planes, a bad file), every result
compared with what the host-entropy path of a separate context gave.  usage: tools/soak_inflight.py [seconds] [K] [seed]"""
import importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
W = importlib.import_module("image-webp_b200")
from harness import synth as S

seconds = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
K = int(sys.argv[2]) if len(sys.argv) > 2 else 3
rng = np.random.default_rng(int(sys.argv[3]) if len(sys.argv) > 3 else 1)
ref = W.BatchDecoder(0); ref.set_entropy(ref.ENTROPY_HOST)

def uniform(n, w, h, **kw):
    return [S.webp_simple(S.vp8_frame(w, h, seed=int(rng.integers(1 << 30)), **kw)) for _ in range(n)]

cases = {}
cases["cfg2x256"] = uniform(256, 512, 512, filter_type=1, filter_level=20, prob_skip=200, p_skip=0.05)
cases["small_normal"] = uniform(300, 208, 176, filter_type=0, filter_level=33, sharpness=3)
cases["parts4"] = uniform(96, 333, 257, filter_type=0, filter_level=50, log2_parts=2)
cases["alpha"] = [S.webp_extended(S.vp8_frame(160, 128, seed=50 + k, filter_type=0, filter_level=9), 160, 128,
                                  S.alph_raw(S.smooth_alpha(160, 128, k), filtering=k % 4)) for k in range(48)]
cases["with_junk"] = uniform(40, 256, 192, filter_type=1, filter_level=12) + [b"junk"] + uniform(8, 256, 192)
want, lists = {}, {}
for k, files in cases.items():
    o, st = ref.decode(files)
    want[k] = ([None if x is None else np.array(x).reshape(-1).copy() for x in o], st)
print("cases ready", {k: len(v) for k, v in cases.items()}, flush=True)

ctxs = [W.BatchDecoder(0) for _ in range(K)]
pending = [None] * K   # (handle, case, outs)
def finish(slot):
    h, k, outs = pending[slot]
    st = ctxs[slot].decode_end(h)
    exp, est = want[k]
    assert st == est, (k, st, est)
    for i, e in enumerate(exp):
        if est[i] == 0:
            assert np.array_equal(outs[i], e), (k, i)
    pending[slot] = None

t0 = time.time(); calls = 0
while time.time() - t0 < seconds:
    slot = calls % K
    if pending[slot] is not None: finish(slot)
    k = list(cases)[int(rng.integers(len(cases)))]
    files = cases[k]
    ctxs[slot].set_entropy(int(rng.integers(0, 3)))
    exp, est = want[k]
    sizes = [0 if e is None else e.size for e in exp]
    if int(rng.integers(0, 2)) and len(set(sizes)) == 1:   # one contiguous block at an even stride (band pipeline when it applies)
        pin = W.PinnedBuffer(sum(sizes))
        outs = [pin.array[i * sizes[0]:(i + 1) * sizes[0]] for i in range(len(files))]
    else:
        outs = [np.zeros(max(s, 1), np.uint8)[:s] for s in sizes]
    fl, ol = W.BufferList([np.frombuffer(f, np.uint8) for f in files]), W.BufferList(outs)
    pending[slot] = (ctxs[slot].decode_begin(fl, ol), k, outs)
    calls += 1
    if int(rng.integers(0, 4)) == 0: time.sleep(float(rng.random()) * 0.01)
for slot in range(K):
    if pending[slot] is not None: finish(slot)
print("in-flight soak ok: %d calls on %d contexts in %.0f s" % (calls, K, time.time() - t0))
