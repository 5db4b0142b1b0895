"""A small VP8L (WebP lossless) bitstream writer for tests and bench inputs: Huffman-coded literals,
optional backward references given explicitly, and any subset of the four transforms with the mode /
coefficient images chosen by the caller.  It exists so that every feature of the lossless ALPH path the
reference decodes (lossless/decoder/mod.rs, reverse_transform.rs, huffman.rs) can be forced, which
libwebp's encoder (PIL) does not allow.  Not part of the decode path.

Layout written (what LosslessDecoder::decode_frame reads, decoder/mod.rs:67-152):
  [signature 0x2f, width-1:14, height-1:14, alpha:1, version:3]   (omitted in an ALPH chunk: implicit size)
  transforms: {1, type:2, payload}* 0          read_transforms, decoder/mod.rs:177-261
  image stream: colour-cache flag (0), meta-code flag (0, ARGB role only), 5 prefix codes, symbols
"""
import heapq

import numpy as np

_CL_ORDER = [17, 18, 0, 1, 2, 3, 4, 5, 16, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15]


class Bits:
    """LSB-first bit sink (the order BitReader consumes, decoder/mod.rs:697-774)."""

    def __init__(self):
        self.chunks = []   # finished byte strings
        self.acc = 0
        self.n = 0

    def put(self, value, nbits):
        self.acc |= (int(value) & ((1 << nbits) - 1)) << self.n
        self.n += nbits
        if self.n >= 4096:
            k = self.n // 8
            self.chunks.append((self.acc & ((1 << (8 * k)) - 1)).to_bytes(k, "little"))
            self.acc >>= 8 * k
            self.n -= 8 * k

    def put_stream(self, codes, lengths):
        """Append many variable-length codes at once (numpy): codes[i] already bit-reversed for
        LSB-first emission, lengths[i] <= 32 bits each."""
        lengths = np.asarray(lengths, dtype=np.int64)
        codes = np.asarray(codes, dtype=np.uint64)
        keep = lengths > 0
        lengths, codes = lengths[keep], codes[keep]
        if lengths.size == 0:
            return
        # flush the accumulator to a whole number of bytes + a bit remainder that starts the stream
        k0 = self.n // 8
        if k0:
            self.chunks.append((self.acc & ((1 << (8 * k0)) - 1)).to_bytes(k0, "little"))
            self.acc >>= 8 * k0
            self.n -= 8 * k0
        head_bits = self.n
        head = self.acc
        self.acc, self.n = 0, 0
        off = np.cumsum(lengths) - lengths + head_bits
        total = int(off[-1] + lengths[-1])
        nwords = (total + 63) // 64 + 1
        words = np.zeros(nwords, dtype=np.uint64)
        widx = (off >> 6).astype(np.int64)
        sh = (off & 63).astype(np.uint64)
        lo = codes << sh                       # part that lands in word widx (overflow bits drop out)
        spill = (sh.astype(np.int64) + lengths) > 64
        hi = np.where(spill, codes >> (np.uint64(64) - np.where(spill, sh, np.uint64(1))), np.uint64(0))
        # offsets are sorted, so equal word indices form runs: OR of disjoint bits == sum
        starts = np.flatnonzero(np.diff(widx, prepend=-1))
        words[widx[starts]] += np.add.reduceat(lo, starts)
        if spill.any():
            si = np.flatnonzero(spill)
            w2 = widx[si] + 1
            st2 = np.flatnonzero(np.diff(w2, prepend=-1))
            words[w2[st2]] += np.add.reduceat(hi[si], st2)
        words[0] |= np.uint64(head)
        raw = words.tobytes()  # little-endian words -> LSB-first byte order
        k = total // 8
        self.chunks.append(raw[:k])
        rem = total - 8 * k
        if rem:
            self.acc = raw[k] & ((1 << rem) - 1)
            self.n = rem

    def bytes(self):
        out = b"".join(self.chunks)
        if self.n:
            k = (self.n + 7) // 8
            out += self.acc.to_bytes(k, "little")
        return out


def _reverse_bits(v, n):
    r = 0
    for _ in range(n):
        r = (r << 1) | (v & 1)
        v >>= 1
    return r


def canonical_codes(lengths):
    """Canonical prefix code (shorter first, then by symbol) as HuffmanTree::build_implicit assigns it
    (huffman.rs:46-177); returns the codes bit-reversed, ready for LSB-first emission."""
    lengths = list(lengths)
    maxl = max(lengths) if lengths else 0
    count = [0] * (maxl + 2)
    for l in lengths:
        count[l] += 1
    count[0] = 0
    nxt = [0] * (maxl + 2)
    code = 0
    for l in range(1, maxl + 1):
        code = (code + count[l - 1]) << 1
        nxt[l] = code
    out = [0] * len(lengths)
    for s, l in enumerate(lengths):
        if l:
            out[s] = _reverse_bits(nxt[l], l)
            nxt[l] += 1
    return out


def huffman_lengths(hist, max_len=15):
    """Code lengths for a histogram, limited to max_len (counts are flattened until the tree fits).
    At least two symbols get a length, so that the code is a complete binary tree."""
    hist = np.asarray(hist, dtype=np.int64).copy()
    used = np.flatnonzero(hist)
    n = len(hist)
    if len(used) == 0:
        used = np.array([0])
        hist[0] = 1
    if len(used) == 1:
        other = 1 if used[0] == 0 else 0
        hist[other] = 1
        used = np.flatnonzero(hist)
    while True:
        heap = [(int(hist[s]), int(s), None, None) for s in used]
        heapq.heapify(heap)
        uid = n
        while len(heap) > 1:
            a = heapq.heappop(heap)
            b = heapq.heappop(heap)
            heapq.heappush(heap, (a[0] + b[0], uid, a, b))
            uid += 1
        lengths = [0] * n
        stack = [(heap[0], 0)]
        while stack:
            node, d = stack.pop()
            if node[2] is None:
                lengths[node[1]] = max(d, 1)
            else:
                stack.append((node[2], d + 1))
                stack.append((node[3], d + 1))
        if max(lengths) <= max_len:
            return lengths
        hist[used] = (hist[used] + 1) // 2 + 1  # flatten and retry


def write_prefix_code(bw, lengths):
    """One prefix code as read_huffman_code reads it (decoder/mod.rs:345-382): the 'simple' form for one
    or two symbols below 256, else code lengths coded with a code-length code (385-452)."""
    used = [s for s, l in enumerate(lengths) if l]
    if len(used) <= 2 and all(s < 256 for s in used):
        if not used:
            used = [0]
        bw.put(1, 1)
        bw.put(len(used) - 1, 1)
        first8 = 1 if used[0] > 1 else 0
        bw.put(first8, 1)
        bw.put(used[0], 8 if first8 else 1)
        if len(used) == 2:
            bw.put(used[1], 8)
        return
    bw.put(0, 1)
    # code-length alphabet: literal lengths 0..15 only (no repeat codes), each with a 4-bit code where used
    # (a complete code over the lengths that occur; 16/17/18 unused)
    hist = [0] * 19
    for l in lengths:
        hist[l] += 1
    cl = huffman_lengths(hist, max_len=7)
    ncl = 19
    while ncl > 4 and cl[_CL_ORDER[ncl - 1]] == 0:
        ncl -= 1
    bw.put(ncl - 4, 4)
    for i in range(ncl):
        bw.put(cl[_CL_ORDER[i]], 3)
    bw.put(0, 1)  # max_symbol = alphabet size
    cc = canonical_codes(cl)
    lens = np.asarray(lengths, dtype=np.int64)
    bw.put_stream(np.asarray(cc, dtype=np.uint64)[lens], np.asarray(cl, dtype=np.int64)[lens])


def _prefix_of(value):
    """(prefix symbol, extra bits count, extra value) of a length / distance code (decoder get_copy_distance)."""
    v = value - 1
    if v < 4:
        return v, 0, 0
    nb = v.bit_length()
    second = (v >> (nb - 2)) & 1
    extra = nb - 2
    return 2 * (nb - 1) + second, extra, v & ((1 << extra) - 1)


def write_image_stream(bw, argb, argb_role, refs=None, force_lengths=None):
    """Entropy-coded image (decode_image_stream, decoder/mod.rs:159-174): no colour cache, one prefix-code
    group.  argb: 1-D uint32 array of 0xAARRGGBB pixels in scan order.  refs: optional list of
    (position, length, distance_code) backward references replacing the literals they cover.
    force_lengths: optional dict code index (0 green .. 4 distance) -> explicit code lengths (KATs)."""
    argb = np.asarray(argb, dtype=np.uint32).reshape(-1)
    bw.put(0, 1)                 # no colour cache
    if argb_role:
        bw.put(0, 1)             # no meta prefix image
    g = ((argb >> 8) & 0xff).astype(np.int64)
    r = ((argb >> 16) & 0xff).astype(np.int64)
    b = (argb & 0xff).astype(np.int64)
    a = ((argb >> 24) & 0xff).astype(np.int64)
    refs = sorted(refs or [])
    lit = np.ones(argb.size, dtype=bool)
    for pos, length, _ in refs:
        lit[pos:pos + length] = False
    hist = [np.bincount(g[lit], minlength=280), np.bincount(r[lit], minlength=256), np.bincount(b[lit], minlength=256),
            np.bincount(a[lit], minlength=256), np.zeros(40, dtype=np.int64)]
    for _, length, dcode in refs:
        hist[0][256 + _prefix_of(length)[0]] += 1
        hist[4][_prefix_of(dcode)[0]] += 1
    lengths = []
    for j in range(5):
        if force_lengths and j in force_lengths:
            lengths.append(list(force_lengths[j]))
        else:
            used = np.flatnonzero(hist[j])
            if len(used) <= 1 and (len(used) == 0 or used[0] < 256):  # one symbol: a single-node code, zero bits per use
                one = [0] * len(hist[j])
                one[int(used[0]) if len(used) else 0] = 1
                lengths.append(one)
            else:
                lengths.append(huffman_lengths(hist[j]))
        write_prefix_code(bw, lengths[j])
    codes = [np.asarray(canonical_codes(l), dtype=np.uint64) for l in lengths]
    lens = [np.asarray(l, dtype=np.int64) for l in lengths]
    single = [sum(1 for l in ls if l) <= 1 for ls in lengths]  # a one-symbol code takes no bits (huffman.rs single node)

    def emit_literals(lo, hi):
        if hi <= lo:
            return
        cs, ls = [], []
        for j, ch in ((0, g), (1, r), (2, b), (3, a)):
            v = ch[lo:hi]
            cs.append(codes[j][v])
            ls.append(np.zeros(hi - lo, dtype=np.int64) if single[j] else lens[j][v])
        bw.put_stream(np.stack(cs, axis=1).reshape(-1), np.stack(ls, axis=1).reshape(-1))

    at = 0
    for pos, length, dcode in refs:
        emit_literals(at, pos)
        p, eb, ev = _prefix_of(length)
        if not single[0]:
            bw.put(int(codes[0][256 + p]), int(lens[0][256 + p]))
        bw.put(ev, eb)
        p, eb, ev = _prefix_of(dcode)
        if not single[4]:
            bw.put(int(codes[4][p]), int(lens[4][p]))
        bw.put(ev, eb)
        at = pos + length
    emit_literals(at, argb.size)


# ---- forward transforms (the inverses are reverse_transform.rs:24-578) ------------------------------

def _ch(argb):
    """(h, w) uint32 ARGB -> int32 array (h, w, 4) in the reference's byte order R, G, B, A."""
    return np.stack([(argb >> 16) & 0xff, (argb >> 8) & 0xff, argb & 0xff, (argb >> 24) & 0xff], axis=-1).astype(np.int32)


def _argb(ch):
    ch = ch.astype(np.uint32) & 0xff
    return (ch[..., 3] << 24) | (ch[..., 0] << 16) | (ch[..., 1] << 8) | ch[..., 2]


def subsample(v, bits):
    return (v + (1 << bits) - 1) >> bits


def predictor_forward(argb, size_bits, modes):
    """Residuals of the predictor transform.  modes: (ceil(h/2^bits), ceil(w/2^bits)) array of 0..13.
    The inverse adds the prediction made from already-restored pixels, i.e. from the ORIGINAL image,
    so the forward direction is a plain vectorised subtraction (apply_predictor_transform, 24-78)."""
    h, w = argb.shape
    c = _ch(argb)
    flat = c.reshape(h * w, 4)
    idx = np.arange(h * w)
    L = flat[np.maximum(idx - 1, 0)]
    T = flat[np.maximum(idx - w, 0)]
    TL = flat[np.maximum(idx - w - 1, 0)]
    TR = flat[np.maximum(idx - w + 1, 0)]   # last column: the first pixel of the current row (memory order)
    a2 = lambda x, y: (x + y) >> 1
    black = np.zeros_like(flat)
    black[:, 3] = 255
    sel_l = np.abs(T - TL).sum(axis=1) < np.abs(L - TL).sum(axis=1)   # predictor 11: Manhattan distances to L + T - TL
    half = a2(L, T)
    d13 = half - TL
    preds = [black, L, T, TR, TL, a2(a2(L, TR), T), a2(L, TL), a2(L, T), a2(TL, T), a2(T, TR), a2(a2(L, TL), a2(T, TR)),
             np.where(sel_l[:, None], L, T), np.clip(L + T - TL, 0, 255),
             np.clip(half + np.where(d13 < 0, -((-d13) >> 1), d13 >> 1), 0, 255)]  # i16 division truncates towards zero
    ys, xs = np.divmod(idx, w)
    mode = np.asarray(modes)[ys >> size_bits, xs >> size_bits]
    pred = np.zeros_like(flat)
    for m in range(14):
        k = mode == m
        pred[k] = preds[m][k]
    pred[ys == 0] = L[ys == 0]          # top row: left
    pred[xs == 0] = T[xs == 0]          # left column: top
    pred[0] = black[0]                  # first pixel: 0xff000000
    return _argb((flat - pred).reshape(h, w, 4))


def color_forward(argb, size_bits, coeffs):
    """coeffs: (by, bx, 3) int8-valued array of (green_to_red, green_to_blue, red_to_blue) per block
    (apply_color_transform, reverse_transform.rs:353-391)."""
    h, w = argb.shape
    c = _ch(argb)
    ys, xs = np.mgrid[0:h, 0:w]
    k = np.asarray(coeffs, dtype=np.int32)[ys >> size_bits, xs >> size_bits]
    s8 = lambda v: ((v & 0xff) ^ 0x80) - 0x80
    delta = lambda t, v: (s8(t) * s8(v)) >> 5
    out = c.copy()
    out[..., 0] = c[..., 0] - delta(k[..., 0], c[..., 1])
    out[..., 2] = c[..., 2] - delta(k[..., 1], c[..., 1]) - delta(k[..., 2], c[..., 0])
    return _argb(out)


def subtract_green_forward(argb):
    c = _ch(argb)
    c[..., 0] -= c[..., 1]
    c[..., 2] -= c[..., 1]
    return _argb(c)


def encode(argb, transforms=(), header=True, refs=None, force_lengths=None):
    """VP8L bitstream of an (h, w) uint32 ARGB image.

    transforms: sequence applied by the ENCODER in this order (the decoder undoes them in reverse), each
      ("predictor", size_bits, modes)   modes: block array of 0..13, or an int for one mode everywhere
      ("color", size_bits, coeffs)      coeffs: block array (by, bx, 3) or a 3-tuple for every block
      ("subtract_green",)
      ("palette", table)                table: uint32 ARGB colours (<= 256) that cover the image
    header=False leaves the 5-byte VP8L header out (ALPH chunks carry implicit dimensions).
    refs / force_lengths: passed to the main image stream (see write_image_stream)."""
    argb = np.asarray(argb, dtype=np.uint32)
    h, w = argb.shape
    bw = Bits()
    if header:
        bw.put(0x2f, 8)
        bw.put(w - 1, 14)
        bw.put(h - 1, 14)
        bw.put(1, 1)
        bw.put(0, 3)
    cur = argb
    for t in transforms:
        kind = t[0]
        ch, cw = cur.shape
        bw.put(1, 1)
        if kind == "predictor":
            bits = int(t[1])
            by, bx = subsample(ch, bits), subsample(cw, bits)
            modes = np.broadcast_to(np.asarray(t[2], dtype=np.uint32), (by, bx)) if np.ndim(t[2]) == 0 else np.asarray(t[2], dtype=np.uint32)
            assert modes.shape == (by, bx)
            bw.put(0, 2)
            bw.put(bits - 2, 3)
            write_image_stream(bw, (0xff000000 | (modes << 8)).reshape(-1), False)
            cur = predictor_forward(cur, bits, modes)
        elif kind == "color":
            bits = int(t[1])
            by, bx = subsample(ch, bits), subsample(cw, bits)
            k = np.asarray(t[2], dtype=np.int64)
            k = np.broadcast_to(k, (by, bx, 3)) if k.ndim == 1 else k
            assert k.shape == (by, bx, 3)
            bw.put(1, 2)
            bw.put(bits - 2, 3)
            # transform pixel bytes R, G, B = red_to_blue, green_to_blue, green_to_red (reverse_transform.rs:375-377)
            img = 0xff000000 | ((k[..., 2] & 0xff) << 16) | ((k[..., 1] & 0xff) << 8) | (k[..., 0] & 0xff)
            write_image_stream(bw, img.astype(np.uint32).reshape(-1), False)
            cur = color_forward(cur, bits, k)
        elif kind == "subtract_green":
            bw.put(2, 2)
            cur = subtract_green_forward(cur)
        elif kind == "palette":
            table = np.asarray(t[1], dtype=np.uint32)
            n = len(table)
            assert 1 <= n <= 256
            bw.put(3, 2)
            bw.put(n - 1, 8)
            # the table is subtraction-coded per channel (adjust_color_map, decoder/mod.rs:263-268)
            tc = _ch(table.reshape(1, n))[0]
            dc = tc.copy()
            dc[1:] = tc[1:] - tc[:-1]
            write_image_stream(bw, _argb(dc), False)
            lut = {int(c): i for i, c in enumerate(table)}
            index = np.vectorize(lut.__getitem__, otypes=[np.uint32])(cur)
            pack = 3 if n <= 2 else 2 if n <= 4 else 1 if n <= 16 else 0
            if pack:
                per = 1 << pack
                bpp = 8 >> pack
                pw = subsample(cw, pack)
                padded = np.zeros((ch, pw * per), dtype=np.uint32)
                padded[:, :cw] = index
                g = np.zeros((ch, pw), dtype=np.uint32)
                for k2 in range(per):
                    g |= padded[:, k2::per] << (bpp * k2)
                index = g
            cur = 0xff000000 | (index << 8)
        else:
            raise ValueError(kind)
    bw.put(0, 1)
    write_image_stream(bw, cur.reshape(-1), True, refs=refs, force_lengths=force_lengths)
    return bw.bytes()


def alph_chunk_payload(alpha, filtering=0, transforms=(), preprocessing=0, **kw):
    """ALPH chunk payload with compression = 1: header byte (extended.rs:253-283) + a headerless VP8L
    stream whose GREEN channel carries the filtered alpha plane (284-294).  The forward alpha filter is
    the inverse of get_alpha_predictor (131-191)."""
    a = np.asarray(alpha, dtype=np.uint8)
    h, w = a.shape
    ai = a.astype(np.int32)
    pred = np.zeros((h, w), dtype=np.int32)
    if filtering == 1:
        pred[:, 1:] = ai[:, :-1]
        pred[1:, 0] = ai[:-1, 0]
    elif filtering == 2:
        pred[1:, :] = ai[:-1, :]
        pred[0, 1:] = ai[0, :-1]
    elif filtering == 3:
        pred[1:, 1:] = np.clip(ai[1:, :-1] + ai[:-1, 1:] - ai[:-1, :-1], 0, 255)
        pred[0, 1:] = ai[0, :-1]
        pred[1:, 0] = ai[:-1, 0]
    res = ((ai - pred) & 0xff).astype(np.uint32)
    info = ((preprocessing & 3) << 4) | ((filtering & 3) << 2) | 1
    return bytes([info]) + encode(0xff000000 | (res << 8), transforms=transforms, header=False, **kw)
