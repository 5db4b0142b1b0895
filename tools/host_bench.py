#!/usr/bin/env python3
"""Single-thread speed of the host entropy stage (no GPU needed)."""
import ctypes as C, importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
W = importlib.import_module("image-webp_b200")
from harness import synth as S
n = 16
files = [np.frombuffer(f, np.uint8) for f in S.make_batch(S.cfg2_image, n)]
fd = W.FrameDesc(); m = 1024
mbs = np.zeros(m, W.MB_DTYPE); lv = np.zeros((m, 25, 16), np.int16)
L = W.lib()
best = 1e9
for rep in range(20):
    t = time.perf_counter()
    for f in files:
        st = L.webpb200_host_entropy_decode(f.ctypes.data, f.size, C.byref(fd), mbs.ctypes.data, m, lv.ctypes.data, lv.size, None, 0)
        assert st == 0
    best = min(best, time.perf_counter() - t)
print("host entropy decode: %.1f MPix/s per thread (%.2f us per macroblock)" % (n * 512 * 512 / best / 1e6, best / n / 1024 * 1e6))
