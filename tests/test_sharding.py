"""The N>1 path on CPU: two gloo ranks shard a batch by image index, decode nothing on a GPU
(there is none here) but run the same barrier / aggregation code bench.py uses."""
import os
import socket
import subprocess
import sys

from util import ROOT, product

WORKER = r"""
import importlib, os, sys
sys.path.insert(0, %(root)r)
sh = importlib.import_module("image-webp_b200.sharding")
dist = sh.init_distributed("gloo")
rank, local, world = sh.env_rank_world()
lo, hi = sh.shard_range(11, rank, world)
sh.barrier(dist)
units, ms = sh.aggregate(dist, (hi - lo) * 100.0, 10.0 + rank)
wlo, whi = sh.weak_shard(4, rank)
print("RESULT", rank, lo, hi, units, ms, wlo, whi, flush=True)
"""


def test_shard_range_covers_everything():
    sh = __import__("importlib").import_module("image-webp_b200.sharding")
    for n in (0, 1, 7, 8, 1024, 1025):
        for world in (1, 2, 3, 8):
            spans = [sh.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def test_two_gloo_ranks():
    product()
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, "-c", WORKER % {"root": ROOT}], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = [p.communicate(timeout=120) for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    rows = sorted(line.split()[1:] for o, _ in outs for line in o.splitlines() if line.startswith("RESULT"))
    assert rows[0][:3] == ["0", "0", "6"] and rows[1][:3] == ["1", "6", "11"]   # 11 images: 6 + 5
    for r in rows:
        assert float(r[3]) == 1100.0 and float(r[4]) == 11.0                   # sum of units, max of times
    assert rows[0][5:] == ["0", "4"] and rows[1][5:] == ["4", "8"]
