#!/usr/bin/env python3
"""Throughput of the other BASELINE.json configurations (not bench.py lines; evidence for SURVEY 8d):
cfg2 with the 'uniform' mode profile, cfg3 (3840x2160, normal filter, 4 segments, 8 partitions),
cfg4 (1920x1080 VP8X + ALPH, RGBA) and cfg5 (animation, 1920x1080 frames with ALPH, blended).
Prints one JSON line per configuration: device-resident MPix/s (CUDA events), per-kernel ms,
end-to-end MPix/s through the C ABI, and the CPU oracle on all host cores for the same files."""
import importlib, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
W = importlib.import_module("image-webp_b200")
from harness import synth as S
from oracle import oracle as O

CORES = os.cpu_count() or 1


def run_still(name, files, bpp, check=2):
    bd = W.BatchDecoder(0)
    st = bd.prepare(files, host_threads=CORES)
    assert all(s == 0 for s in st)
    bd.upload(); bd.sync()
    size = bd.size()
    bd.time(15, iters=3, flush_l2=False)
    ms, k = bd.time(15, iters=10, flush_l2=False)
    pin = W.PinnedBuffer(size["out_bytes"])
    outs, off = [], 0
    for f in files:
        d = W.WebPDecoder(f); n = d.output_buffer_size()
        outs.append(pin.array[off:off + n]); off += n
    e2e_by_mode = {}
    for mode, label in ((bd.ENTROPY_HOST, "host"), (bd.ENTROPY_DEVICE, "device"), (bd.ENTROPY_AUTO, "auto")):
        bd.set_entropy(mode)
        bd.decode(files, outs=outs, host_threads=CORES)
        t = time.perf_counter()
        for _ in range(3):
            bd.decode(files, outs=outs, host_threads=CORES)
        e2e_by_mode[label] = (time.perf_counter() - t) / 3
    e2e = e2e_by_mode["auto"]
    t = time.perf_counter()
    want, wst, bad = O.decode_batch(files, threads=CORES)
    cpu = time.perf_counter() - t
    ok = all(np.array_equal(outs[i], want[i].reshape(-1)) for i in range(min(check, len(files))))
    px = size["pixels"]
    print(json.dumps({"config": name, "images": len(files), "MPix": px / 1e6, "device_MPix_s": px / ms / 1e3, "ms_per_pass": ms,
                      "kernel_ms": dict(zip(["predict", "loopfilter", "yuv2rgb", "alpha"], k)), "e2e_MPix_s": px / e2e / 1e6,
                      "e2e_ms_by_entropy_mode": {k: round(v * 1e3, 1) for k, v in e2e_by_mode.items()},
                      "cpu_oracle_MPix_s": px / cpu / 1e6, "cpu_cores": CORES, "bit_exact_vs_oracle": bool(ok)}), flush=True)
    pin.free(); bd.close()


def main():
    which = sys.argv[1:] or ["cfg2u", "cfg3", "cfg4", "cfg5"]
    if "cfg2u" in which:
        run_still("cfg2 uniform profile: 1024 x 512x512, simple filter, RGB",
                  S.make_batch(lambda i: S.cfg2_image(i, profile="uniform"), 1024, threads=CORES), 3)
    if "cfg3" in which:
        run_still("cfg3: 32 x 3840x2160 (one GPU's share of 256 on 8), normal filter, 4 segments, 8 partitions, RGB",
                  S.make_batch(S.cfg3_image, 32, threads=CORES), 3, check=1)
    if "cfg4" in which:
        run_still("cfg4: 64 x 1920x1080 VP8X + raw ALPH (filter 0..3 cycling), normal filter, RGBA",
                  S.make_batch(S.cfg4_image, 64, threads=CORES), 4)
    if "cfg4c" in which:
        # cfg4 as SURVEY 8(d) words it: the ALPH chunk VP8L-compressed by libwebp (PIL), spliced next to a synthetic VP8 chunk
        import io
        from PIL import Image
        sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
        from util import chunk_payload
        files = []
        for i in range(64):
            a = S.smooth_alpha(1920, 1080, i)
            rgba = np.zeros((1080, 1920, 4), np.uint8)
            rgba[..., 3] = a
            buf = io.BytesIO()
            Image.fromarray(rgba, "RGBA").save(buf, "WEBP", quality=50, method=0)
            alph = chunk_payload(buf.getvalue(), b"ALPH")
            vp8 = S.vp8_frame(1920, 1080, seed=0xB200 + i, filter_type=0, filter_level=20, yac_abs=36, prob_skip=200, p_skip=0.05)
            files.append(S.webp_extended(vp8, 1920, 1080, S.chunk(b"ALPH", alph)))
        run_still("cfg4c: 64 x 1920x1080 VP8X + VP8L-compressed ALPH (libwebp), normal filter, RGBA", files, 4)
    if "cfg5" in which:
        frames = []
        for k in range(120):
            vp8 = S.vp8_frame(1920, 1080, seed=0xB200 + k, filter_type=0, filter_level=20, yac_abs=36, prob_skip=200, p_skip=0.05)
            frames.append(dict(x=0, y=0, w=1920, h=1080, duration=40, blend=True, dispose=bool(k & 1), vp8=vp8,
                               alph=S.alph_raw(S.smooth_alpha(1920, 1080, k), filtering=k % 4)))
        data = S.webp_animation(1920, 1080, frames, alpha=True)
        bd = W.BatchDecoder(0)
        pin = W.PinnedBuffer(120 * 1920 * 1080 * 4)
        pouts = [pin.array[k * 1920 * 1080 * 4:(k + 1) * 1920 * 1080 * 4] for k in range(120)]
        bd.decode_animation(data, host_threads=CORES, outs=pouts)
        t = time.perf_counter()
        outs, durs = bd.decode_animation(data, host_threads=CORES, outs=pouts)
        e2e = time.perf_counter() - t
        outs = [o.reshape(1080, 1920, 4) for o in outs]
        t = time.perf_counter()
        o = O.WebPDecoder(data)
        ok = True
        for k in range(120):
            w, _ = o.read_frame()
            if k in (0, 1, 60, 119):
                ok &= bool(np.array_equal(w, outs[k]))
        cpu = time.perf_counter() - t
        px = 120 * 1920 * 1080
        print(json.dumps({"config": "cfg5: animation, 120 frames 1920x1080, ALPH + VP8 per frame, blend, dispose alternating",
                          "MPix": px / 1e6, "e2e_MPix_s": px / e2e / 1e6, "cpu_oracle_1core_MPix_s": px / cpu / 1e6,
                          "bit_exact_vs_oracle": ok}), flush=True)
        bd.close()


if __name__ == "__main__":
    main()
