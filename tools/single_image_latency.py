#!/usr/bin/env python3
"""Latency of the WebPDecoder surface on ONE image (BASELINE configs[0]): product (GPU) and CPU oracle, per golden lossy file."""
import importlib, os, sys, time, glob
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
W = importlib.import_module("image-webp_b200")
from oracle import oracle as O
import numpy as np
fx = np.load(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "fixtures.npz"))
for key in sorted(k for k in fx.keys() if k.startswith("webp/")):
    path = key[5:]
    data = fx[key].tobytes()
    try:
        d = W.WebPDecoder(data)
        if not d.is_lossy() or d.is_animated(): continue
        w, h = d.dimensions()
        d.read_image()
    except Exception:
        continue
    best = 1e9
    for _ in range(20):
        t = time.perf_counter(); W.WebPDecoder(data).read_image(); best = min(best, time.perf_counter() - t)
    cpu = 1e9
    for _ in range(5):
        t = time.perf_counter(); O.WebPDecoder(data).read_image(); cpu = min(cpu, time.perf_counter() - t)
    print("%-40s %5dx%-5d  product %.2f ms   CPU oracle (1 core) %.2f ms" % (os.path.basename(path), w, h, best * 1e3, cpu * 1e3), flush=True)
