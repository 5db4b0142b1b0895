#!/usr/bin/env python3
"""End-to-end cfg2 steps with K batches in flight (webpb200_decode_batch_begin/_end on K contexts), with and without the
download: what bounds the pipelined step.  usage: tools/inflight.py [n_images] [steps]"""
import ctypes as C, importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
W = importlib.import_module("image-webp_b200")
from harness import synth as S
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
cores = os.cpu_count() or 1
CFG = os.environ.get("INFLIGHT_CFG", "cfg2")
files = S.make_batch(S.cfg3_image if CFG == "cfg3" else S.cfg2_image, n, threads=cores)
starts, total = [], 0
for f in files:
    starts.append(total); total += (len(f) + 63) & ~63
pin_in = W.PinnedBuffer(total)
pf = []
for f, o in zip(files, starts):
    pin_in.array[o:o + len(f)] = np.frombuffer(f, dtype=np.uint8); pf.append(pin_in.array[o:o + len(f)])
fl = W.BufferList(pf)
per = (3840 * 2160 if CFG == 'cfg3' else 512 * 512) * 3
KMAX = int(os.environ.get('INFLIGHT_KMAX', '4'))
ctxs = [W.BatchDecoder(0) for _ in range(KMAX)]
pins = [W.PinnedBuffer(n * per) for _ in range(KMAX)]
outs = [W.BufferList([p.array[i * per:(i + 1) * per] for i in range(n)]) for p in pins]
for c, o in zip(ctxs, outs):
    c.set_entropy(c.ENTROPY_DEVICE)
    c.decode(fl, outs=o, host_threads=cores)
lib = W.lib()
def begin(c, o):
    status = (C.c_int * n)(); h = C.c_void_p()
    rc = lib.webpb200_decode_batch_begin(c._h, fl.ptrs, fl.lens, n, o.ptrs if o else None, o.lens if o else None, status, W.BILINEAR, cores, C.byref(h))
    assert rc == 0, rc
    return h, status
def end(h):
    rc = lib.webpb200_decode_batch_end(h[0]); assert rc == 0, rc
    assert all(s == 0 for s in h[1])
STAGGER = float(os.environ.get('INFLIGHT_STAGGER_MS', '0'))
KS = [int(x) for x in os.environ.get('INFLIGHT_K', '1,2,3,4').split(',')]
DL = [bool(int(x)) for x in os.environ.get('INFLIGHT_DL', '1,0').split(',')]
for download in DL:
    for K in KS:
        pend = [None] * K
        total_steps = K * steps
        t0 = time.perf_counter()
        for k in range(total_steps):
            if pend[k % K] is not None: end(pend[k % K])
            pend[k % K] = begin(ctxs[k % K], outs[k % K] if download else None)
            if k < K - 1 and STAGGER: time.sleep(STAGGER / 1e3)
        for k in range(K): end(pend[k])
        ms = 1e3 * (time.perf_counter() - t0) / total_steps
        print("download %d  in flight %d: %.1f ms per step = %.2f GPix/s   k_entropy of the last calls: %s" %
              (download, K, ms, n * (per // 3) / ms / 1e6, " ".join("%.1f" % c.last_entropy_ms() for c in ctxs[:K])), flush=True)
