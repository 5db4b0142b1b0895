#!/usr/bin/env python3
"""Device-resident reconstruct pass of one BASELINE configuration: per-kernel CUDA-event times.
usage: tools/recon_cfg.py cfg2|cfg2u|cfg3|cfg4|cfg4c [n_images] [iters]
Also the command that is run under ncu for the per-configuration captures in profiles/."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
W = importlib.import_module("image-webp_b200")
from harness import synth as S

DEFAULT_N = {"cfg2": 1024, "cfg2u": 1024, "cfg3": 32, "cfg4": 64, "cfg4c": 64}


def files_for(cfg, n, threads):
    if cfg == "cfg2":
        return S.make_batch(S.cfg2_image, n, threads=threads)
    if cfg == "cfg2u":
        return S.make_batch(lambda i: S.cfg2_image(i, profile="uniform"), n, threads=threads)
    if cfg == "cfg3":
        return S.make_batch(S.cfg3_image, n, threads=threads)
    if cfg == "cfg4":
        return S.make_batch(S.cfg4_image, n, threads=threads)
    if cfg == "cfg4c":  # VP8L-compressed ALPH (harness/configs.py), as bench.py's cfg4
        from harness import configs
        return configs.cfg4_batch(0, n, threads=threads)
    raise SystemExit("unknown configuration " + cfg)


def main():
    cfg = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
    n = int(sys.argv[2]) if len(sys.argv) > 2 else DEFAULT_N[cfg]
    iters = int(sys.argv[3]) if len(sys.argv) > 3 else 10
    cores = os.cpu_count() or 1
    files = files_for(cfg, n, cores)
    bd = W.BatchDecoder(0)
    st = bd.prepare(files, host_threads=cores)
    assert all(s == 0 for s in st)
    bd.upload(); bd.sync()
    size = bd.size()
    bd.time(15, iters=3, flush_l2=False)
    ms, k = bd.time(15, iters=iters, flush_l2=False)
    print("%s n=%d: pass %.3f ms = %.1f GPix/s  predict %.3f  loopfilter %.3f  yuv2rgb %.3f  alpha %.3f" %
          (cfg, n, ms, size["pixels"] / ms / 1e6, k[0], k[1], k[2], k[3]), flush=True)
    bd.close()


if __name__ == "__main__":
    main()
