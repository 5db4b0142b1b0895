#!/usr/bin/env python3
"""k_entropy time on the cfg2 batch for the library named by WEBPB200_SO (A/B of build variants)."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["WEBPB200_TRACE"] = "1"
W = importlib.import_module("image-webp_b200")
from harness import synth as S
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
files = S.make_batch(S.cfg2_image, n, threads=os.cpu_count())
bd = W.BatchDecoder(0)
bd.set_entropy(bd.ENTROPY_DEVICE)
pin = W.PinnedBuffer(n * 512 * 512 * 3)
outs = [pin.array[i * 786432:(i + 1) * 786432] for i in range(n)]
import numpy as np
starts, total = [], 0
for f in files:
    starts.append(total + int(os.environ.get('SHIFT', '0'))); total += (len(f) + 63 + 64) & ~63
pin_in = W.PinnedBuffer(total)
pf = []
for f, o in zip(files, starts):
    pin_in.array[o:o + len(f)] = np.frombuffer(f, dtype=np.uint8); pf.append(pin_in.array[o:o + len(f)])
fl, ol = W.BufferList(pf if not os.environ.get("PAGEABLE") else files), W.BufferList(outs)
import time
for _ in range(4):
    t = time.perf_counter(); bd.decode(fl, outs=ol, host_threads=os.cpu_count()); dt = time.perf_counter() - t
print("python-side %.1f ms" % (dt * 1e3))
