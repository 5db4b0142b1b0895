#!/usr/bin/env python3
"""One decode of 16 cfg4 images (4 per ALPH filter method), for an ncu launch list of the alpha kernels."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
W = importlib.import_module("image-webp_b200")
from harness import synth as S
files = S.make_batch(S.cfg4_image, 16, threads=os.cpu_count())
bd = W.BatchDecoder(0)
for _ in range(2):
    outs, st = bd.decode(files)
print("ok", st[:4])
