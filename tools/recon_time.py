#!/usr/bin/env python3
"""Device-resident reconstruct pass of the cfg2 batch: per-kernel CUDA-event times (for knob sweeps)."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
W = importlib.import_module("image-webp_b200")
from harness import synth as S
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
files = S.make_batch(S.cfg2_image, n, threads=os.cpu_count())
bd = W.BatchDecoder(0)
bd.prepare(files, host_threads=os.cpu_count()); bd.upload(); bd.sync()
bd.time(15, iters=3, flush_l2=False)
ms, k = bd.time(15, iters=10, flush_l2=False)
print("pass %.3f ms  predict %.3f  loopfilter %.3f  yuv2rgb %.3f" % (ms, k[0], k[1], k[2]))
