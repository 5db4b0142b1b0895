#!/usr/bin/env python3
"""Which macroblocks of the traced rows are slow, pass after pass (debug build, see wf_trace.py)."""
import ctypes as C, importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
W = importlib.import_module("image-webp_b200")
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from recon_cfg import files_for
files = files_for("cfg3", 1, os.cpu_count())
bd = W.BatchDecoder(0)
bd.prepare(files, host_threads=os.cpu_count()); bd.upload(); bd.sync()
fd, mbs, lv = bd.packed(0)
mbw = fd.mbw
for p in range(4):
    bd.time(1, iters=1, flush_l2=False)
    tr = np.zeros((4, 4, 1024), dtype=np.int64)
    W.lib().webpb200_debug_wf_trace(tr.ctypes.data_as(C.c_void_p))
    for r in range(2):
        work = tr[r][2, :mbw] - tr[r][1, :mbw]
        slow = np.flatnonzero(work > 3000)
        print("pass %d row +%d slow work parts at mbx" % (p, r), slow.tolist(), "ns", work[slow].tolist())
for r in range(2):
    row = 60 + r
    m = mbs[row * mbw:(row + 1) * mbw]
    bp = np.stack([(m["bpred"][:, k >> 1] >> (4 * (k & 1))) & 15 for k in range(16)], axis=1)
    print("row", row, "ymode:", m["ymode"].tolist())
    np.save("gpurun_out/bpred_row%d.npy" % row, bp)
bd.close()
