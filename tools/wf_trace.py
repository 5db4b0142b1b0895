#!/usr/bin/env python3
"""Global-timer stamps of four consecutive macroblock rows of the prediction wavefront (debug build:
tools/ab_build.sh trace "-DWF_TRACE=60", run with WEBPB200_SO=ab/libwebpb200_trace.so).
usage: tools/wf_trace.py cfg3 [n]"""
import ctypes as C, importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
W = importlib.import_module("image-webp_b200")
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from recon_cfg import files_for
cfg = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1
files = files_for(cfg, n, os.cpu_count())
bd = W.BatchDecoder(0)
bd.prepare(files, host_threads=os.cpu_count()); bd.upload(); bd.sync()
bd.time(1, iters=2, flush_l2=False)
ms, k = bd.time(1, iters=1, flush_l2=False)
print("predict %.3f ms" % k[0])
tr = np.zeros((4, 4, 1024), dtype=np.int64)
assert W.lib().webpb200_debug_wf_trace(tr.ctypes.data_as(C.c_void_p)) == 0
mbw = bd.packed(0)[0].mbw
t0 = tr[0][0][0]
for r in range(4):
    t = tr[r][:, :mbw] - t0
    print("row +%d on SM %d: starts %d ns, ends %d ns; step (end to end) median %d ns; pre-part median %d, work-part median %d" %
          (r, tr[r][3][0], t[0][0], t[2][-1], np.median(np.diff(t[2])), np.median(t[1] - t[0]), np.median(t[2] - t[1])))
for r in range(1, 4):
    a, b = tr[r - 1][:, :mbw] - t0, tr[r][:, :mbw] - t0
    lag = b[2] - a[2]                       # end of step x in row r  -  end of step x in row r-1
    # record x+1 of row r-1 is written at a[2][x+1]; row r has it in hand (after its wait) at b[1][x]
    hand = b[1][:-1] - a[2][1:]
    print("row +%d vs +%d: lag median %d ns (first %s); hand-off (consumer has record x+1) - (producer wrote it): median %d, min %d, p90 %d, first 8 %s" %
          (r, r - 1, np.median(lag), lag[:4].tolist(), np.median(hand), hand.min(), np.percentile(hand, 90), hand[:8].tolist()))
bd.close()
print("raw timeline (ns from row+0 step 0 start): per row, steps 0..11: start / after-wait / end")
for r in range(4):
    t = tr[r][:, :mbw] - t0
    print(" row +%d:" % r, " ".join("%d/%d/%d" % (t[0][x], t[1][x], t[2][x]) for x in range(12)))
for r in range(4):
    t = tr[r][:, :mbw] - t0
    d = np.diff(t[2])
    big = np.flatnonzero(d > 4000)
    print(" row +%d: steps longer than 4 us:" % r, [(int(x + 1), int(d[x])) for x in big][:20])
allwork = np.concatenate([(tr[r][2, :mbw] - tr[r][1, :mbw]) for r in range(4)])
allpre = np.concatenate([(tr[r][1, :mbw] - tr[r][0, :mbw]) for r in range(4)])
print("work parts: median %d ns, >3us: %d of %d (sum of excess %d ns); pre parts >4us: %d (sum %d ns)" %
      (np.median(allwork), (allwork > 3000).sum(), allwork.size, (allwork[allwork > 3000] - np.median(allwork)).sum(), (allpre > 4000).sum(), allpre[allpre > 4000].sum()))
