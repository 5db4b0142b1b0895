#!/usr/bin/env python3
"""Stage breakdown of webpb200_decode_batch on the cfg2 batch (WEBPB200_TRACE=1 prints it)."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["WEBPB200_TRACE"] = "1"
W = importlib.import_module("image-webp_b200")
from harness import synth as S
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
files = S.make_batch(S.cfg2_image, n, threads=os.cpu_count())
bd = W.BatchDecoder(0)
pin = W.PinnedBuffer(n * 512 * 512 * 3)
outs = [pin.array[i * 786432:(i + 1) * 786432] for i in range(n)]
th = os.cpu_count()
best = 1e9
for _ in range(4):
    t = time.perf_counter(); bd.prepare(files, host_threads=th); best = min(best, time.perf_counter() - t)
print("host stage alone (parse + entropy, %d threads): %.1f ms" % (th, best * 1e3), flush=True)
for th in (os.cpu_count(),):
    for _ in range(3):
        t = time.perf_counter(); bd.decode(files, outs=outs, host_threads=th); dt = time.perf_counter() - t
        print("threads %d: %.1f ms -> %.0f MPix/s" % (th, dt * 1e3, n * 512 * 512 / dt / 1e6), flush=True)
bd.set_entropy(bd.ENTROPY_DEVICE)
for _ in range(4):
    t = time.perf_counter(); bd.decode(files, outs=outs, host_threads=th); dt = time.perf_counter() - t
    print("device entropy: %.1f ms -> %.0f MPix/s" % (dt * 1e3, n * 512 * 512 / dt / 1e6), flush=True)
