#!/usr/bin/env python3
"""Build tests/golden/fixtures.npz from the reference's own test fixtures.

Run in the build container only (needs /root/reference and PIL for PNG reading):
    python tests/golden/make_golden.py

For every file the reference's integration tests decode (tests/decode.rs:190-203) this stores
  webp/<set>/<name>      the input file bytes
  dims/<set>/<name>      [width, height, channels] of the expected output
  sha/<set>/<name>[-i]   SHA-256 of the expected raw RGB/RGBA bytes (from tests/reference/*.png,
                         which are dwebp output, tests/CREDITS.md)
  rowcrc/<set>/<name>[-i] CRC-32 of each expected pixel row (to localise a mismatch)
  full/<set>/<name>[-i]  the expected bytes themselves for small images (<= 64 KiB)
The PNGs are not copied; the GPU box has no /root/reference, so tests read only this file.
"""
import hashlib, os, sys, zlib
import numpy as np
from PIL import Image

REF = "/root/reference/tests"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "fixtures.npz")

STILL = {
    "gallery1": ["1", "2", "3", "4", "5"],
    "gallery2": ["%d_webp_%s" % (i, k) for k in ("a", "ll") for i in range(1, 6)],
    "regression": ["color_index", "dark", "tiny", "lossless_indexed_1bit_palette",
                   "lossless_indexed_2bit_palette", "lossless_indexed_4bit_palette"],
}
ANIM = {"animated": [("random_lossless", 3), ("random_lossy", 4)]}
NOFANCY = {"gallery1": ["1", "2", "3", "4", "5"]}


def png_bytes(path):
    im = Image.open(path)
    assert im.mode in ("RGB", "RGBA"), (path, im.mode)
    a = np.asarray(im, dtype=np.uint8)
    return a


def add_expected(d, key, arr):
    raw = arr.tobytes()
    d["sha/" + key] = np.frombuffer(hashlib.sha256(raw).digest(), dtype=np.uint8)
    d["rowcrc/" + key] = np.array([zlib.crc32(arr[y].tobytes()) for y in range(arr.shape[0])], dtype=np.uint32)
    if len(raw) <= 65536:
        d["full/" + key] = arr.copy()


def main():
    d = {}
    for s, names in STILL.items():
        for n in names:
            key = "%s/%s" % (s, n)
            d["webp/" + key] = np.fromfile("%s/images/%s.webp" % (REF, key), dtype=np.uint8)
            a = png_bytes("%s/reference/%s.png" % (REF, key))
            d["dims/" + key] = np.array([a.shape[1], a.shape[0], a.shape[2]], dtype=np.int32)
            add_expected(d, key, a)
    for s, items in ANIM.items():
        for n, frames in items:
            key = "%s/%s" % (s, n)
            d["webp/" + key] = np.fromfile("%s/images/%s.webp" % (REF, key), dtype=np.uint8)
            for i in range(1, frames + 1):
                a = png_bytes("%s/reference/%s-%d.png" % (REF, key, i))
                d["dims/" + key] = np.array([a.shape[1], a.shape[0], a.shape[2]], dtype=np.int32)
                add_expected(d, "%s-%d" % (key, i), a)
            d["nframes/" + key] = np.array([frames], dtype=np.int32)
    for s, names in NOFANCY.items():
        for n in names:
            a = png_bytes("%s/reference/%s_nofancy/%s.png" % (REF, s, n))
            add_expected(d, "%s_nofancy/%s" % (s, n), a)
    np.savez_compressed(OUT, **d)
    print("wrote", OUT, os.path.getsize(OUT), "bytes,", len(d), "arrays")


if __name__ == "__main__":
    sys.exit(main())
