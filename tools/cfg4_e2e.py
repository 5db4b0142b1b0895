import importlib, os, sys, time
sys.path.insert(0, '/root/repo')
os.environ["WEBPB200_TRACE"] = "1"
W = importlib.import_module("image-webp_b200")
from harness import configs
import numpy as np
n = int(os.environ.get("N", "256"))
files = configs.cfg4_batch(0, n, threads=os.cpu_count())
bd = W.BatchDecoder(0)
if os.environ.get("ENT"): bd.set_entropy(int(os.environ["ENT"]))
sizes = [1920 * 1080 * 4] * n
pin = W.PinnedBuffer(sum(sizes))
outs = [pin.array[i * sizes[0]:(i + 1) * sizes[0]] for i in range(n)]
for _ in range(3):
    t = time.perf_counter(); bd.decode(files, outs=outs, host_threads=os.cpu_count()); dt = time.perf_counter() - t
    print("cfg4 n=%d: %.1f ms" % (n, dt * 1e3), flush=True)
