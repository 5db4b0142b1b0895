import importlib, os, sys, time, threading
sys.path.insert(0, '/root/repo')
W = importlib.import_module("image-webp_b200")
from harness import synth as S
n = 1024
files = S.make_batch(S.cfg2_image, n, threads=os.cpu_count())
def mk(fs):
    bd = W.BatchDecoder(0); st = bd.prepare(fs, host_threads=os.cpu_count()); assert all(s == 0 for s in st); bd.upload(); bd.sync(); bd.time(15, iters=3, flush_l2=False); return bd
one = mk(files)
iters = 40
t = time.perf_counter(); ms, k = one.time(15, iters=iters, flush_l2=False); wall1 = (time.perf_counter() - t) / iters * 1e3
print("one context, 1024 images: %.3f ms per pass (events %.3f)" % (wall1, ms))
one.close()
for parts in (2, 3, 4):
    step = n // parts
    ctxs = [mk(files[i * step:(i + 1) * step]) for i in range(parts)]
    res = [None] * parts
    def run(i): res[i] = ctxs[i].time(15, iters=iters, flush_l2=False)
    th = [threading.Thread(target=run, args=(i,)) for i in range(parts)]
    t = time.perf_counter()
    for x in th: x.start()
    for x in th: x.join()
    wall = (time.perf_counter() - t) / iters * 1e3
    print("%d contexts x %d images concurrently: %.3f ms per pass of all (each context's own pass %.3f ms)" % (parts, step, wall, res[0][0]))
    for c in ctxs: c.close()
