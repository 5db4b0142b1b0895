"""Known-answer tests taken from the reference's in-file unit tests, run on the oracle."""
import ctypes as C

import numpy as np
import pytest


def test_bool_decoder_hello(oracle):
    """arithmetic_decoder.rs:636-673"""
    L = oracle.lib()
    for data, tail in ((b"hel", []), (b"hello world", [31])):
        buf = np.frombuffer(data, dtype=np.uint8).copy()
        b = L.ora_bool_new(buf.ctypes.data, buf.size)
        assert L.ora_bool_read_flag(b) == 0
        assert L.ora_bool_read_bool(b, 10) == 1
        assert L.ora_bool_read_bool(b, 250) == 0
        assert L.ora_bool_read_literal(b, 1) == 1
        assert L.ora_bool_read_literal(b, 3) == 5
        assert L.ora_bool_read_literal(b, 8) == 64
        assert L.ora_bool_read_literal(b, 8) == 185
        for t in tail:
            assert L.ora_bool_read_literal(b, 8) == t
        assert L.ora_bool_past_eof(b) == 0
        L.ora_bool_free(b)


def test_bool_decoder_uninit(oracle):
    """arithmetic_decoder.rs:675-682"""
    L = oracle.lib()
    b = L.ora_bool_new(None, 0)
    L.ora_bool_read_flag(b)
    assert L.ora_bool_past_eof(b) == 1
    L.ora_bool_free(b)


def test_bool_decoder_one_byte_past_end(oracle):
    """arithmetic_decoder.rs:249-276: one zero byte past the end is tolerated, the next is EOF."""
    L = oracle.lib()
    buf = np.array([0x55], dtype=np.uint8)
    b = L.ora_bool_new(buf.ctypes.data, 1)
    for _ in range(4):   # at most 4 bits used: the window (8 bits) is still inside byte 0 + the zero byte
        L.ora_bool_read_flag(b)
    assert L.ora_bool_past_eof(b) == 0
    for _ in range(64):  # far beyond 2 bytes
        L.ora_bool_read_flag(b)
    assert L.ora_bool_past_eof(b) == 1
    L.ora_bool_free(b)


def test_yuv_conversions(oracle):
    """yuv.rs:574-581"""
    out = np.zeros(3, dtype=np.uint8)
    oracle.lib().ora_yuv_to_rgb(203, 40, 42, out.ctypes.data)
    assert out.tolist() == [80, 255, 40]


def test_fancy_grid(oracle):
    """yuv.rs:518-572"""
    L = oracle.lib()
    y = np.array([77, 162, 202, 185, 28, 13, 199, 182, 135, 147, 164, 135, 66, 27, 171, 130], dtype=np.uint8)
    u = np.array([34, 101, 123, 163], dtype=np.uint8)
    v = np.array([97, 167, 149, 23], dtype=np.uint8)
    uu = [34, 51, 84, 101, 56, 71, 101, 117, 101, 112, 136, 148, 123, 133, 153, 163]
    vv = [97, 115, 150, 167, 110, 115, 126, 131, 136, 117, 78, 59, 149, 118, 55, 23]
    got = np.zeros(48, dtype=np.uint8)
    L.ora_fill_rgb_buffer_fancy(got.ctypes.data, y.ctypes.data, u.ctypes.data, v.ctypes.data, 4, 4, 4, 3)
    want = np.zeros((16, 3), dtype=np.uint8)
    px = np.zeros(3, dtype=np.uint8)
    for i in range(16):
        L.ora_yuv_to_rgb(int(y[i]), uu[i], vv[i], px.ctypes.data)
        want[i] = px
    assert got.tolist() == want.reshape(-1).tolist()


def test_add_residue(oracle):
    """prediction.rs:706-719"""
    p = np.arange(1, 17, dtype=np.uint8)
    r = np.array([-1, -2, -3, -4, 250, 249, 248, 250, -10, -18, -192, -17, -3, 15, 18, 9], dtype=np.int32)
    oracle.lib().ora_add_residue(p.ctypes.data, r.ctypes.data, 0, 0, 4)
    assert p.tolist() == [0, 0, 0, 0, 255, 255, 255, 255, 0, 0, 0, 0, 10, 29, 33, 25]


def _predict(oracle, im, stride, x0, y0, mode, pad=16):
    buf = np.zeros(pad + len(im), dtype=np.uint8)
    buf[pad:] = im
    oracle.lib().ora_predict_bmode(buf.ctypes.data + pad, stride, x0, y0, mode)
    return buf[pad:].tolist()


def test_predict_bhepred(oracle):
    """prediction.rs:721-740"""
    im = [5, 0, 0, 0, 0, 4, 0, 0, 0, 0, 3, 0, 0, 0, 0, 2, 0, 0, 0, 0, 1, 0, 0, 0, 0]
    want = [5, 0, 0, 0, 0, 4, 4, 4, 4, 4, 3, 3, 3, 3, 3, 2, 2, 2, 2, 2, 1, 1, 1, 1, 1]
    assert _predict(oracle, im, 5, 1, 1, 3) == want


def test_predict_brdpred(oracle):
    """prediction.rs:742-761"""
    im = [5, 6, 7, 8, 9, 4, 0, 0, 0, 0, 3, 0, 0, 0, 0, 2, 0, 0, 0, 0, 1, 0, 0, 0, 0]
    want = [5, 6, 7, 8, 9, 4, 5, 6, 7, 8, 3, 4, 5, 6, 7, 2, 3, 4, 5, 6, 1, 2, 3, 4, 5]
    assert _predict(oracle, im, 5, 1, 1, 5) == want


def test_predict_bldpred(oracle):
    """prediction.rs:763-801 (x0 = 0: the predictor touches no left/corner pixel for LD results)"""
    im = [1, 2, 3, 4, 5, 6, 7, 8] + [0] * 64
    got = _predict(oracle, im, 8, 0, 1, 4)
    rows = [got[8 * (1 + r):8 * (1 + r) + 4] for r in range(4)]
    assert rows == [[2, 3, 4, 5], [3, 4, 5, 6], [4, 5, 6, 7], [5, 6, 7, 8]]


def test_predict_bvepred(oracle):
    """prediction.rs:803-838"""
    im = [1, 2, 3, 4, 5, 6, 7, 8, 9] + [0] * 72
    got = _predict(oracle, im, 9, 1, 1, 2)
    for r in range(4):
        assert got[9 * (1 + r) + 1:9 * (1 + r) + 5] == [2, 3, 4, 5]


def _dct4x4(block):
    """forward DCT of the reference's encoder (transform.rs:131-162), restated here only to
    drive the identity test transform.rs:168-183."""
    b = [int(x) for x in block]
    for i in range(4):
        a = (b[i * 4] + b[i * 4 + 3]) * 8
        bb = (b[i * 4 + 1] + b[i * 4 + 2]) * 8
        c = (b[i * 4 + 1] - b[i * 4 + 2]) * 8
        d = (b[i * 4] - b[i * 4 + 3]) * 8
        b[i * 4] = a + bb
        b[i * 4 + 2] = a - bb
        b[i * 4 + 1] = (c * 2217 + d * 5352 + 14500) >> 12
        b[i * 4 + 3] = (d * 2217 - c * 5352 + 7500) >> 12
    for i in range(4):
        a = b[i] + b[i + 12]
        bb = b[i + 4] + b[i + 8]
        c = b[i + 4] - b[i + 8]
        d = b[i] - b[i + 12]
        b[i] = (a + bb + 7) >> 4
        b[i + 8] = (a - bb + 7) >> 4
        b[i + 4] = ((c * 2217 + d * 5352 + 12000) >> 16) + (1 if d != 0 else 0)
        b[i + 12] = (d * 2217 - c * 5352 + 51000) >> 16
    return b


def test_dct_inverse(oracle):
    """transform.rs:168-183"""
    block = [38, 6, 210, 107, 42, 125, 185, 151, 241, 224, 125, 233, 227, 8, 57, 96]
    coef = np.array(_dct4x4(block), dtype=np.int32)
    oracle.lib().ora_idct4x4(coef.ctypes.data)
    assert coef.tolist() == block


RED = [0x52, 0x49, 0x46, 0x46, 0x3c, 0x00, 0x00, 0x00, 0x57, 0x45, 0x42, 0x50, 0x56, 0x50,
       0x38, 0x20, 0x30, 0x00, 0x00, 0x00, 0xd0, 0x01, 0x00, 0x9d, 0x01, 0x2a, 0x02, 0x00,
       0x02, 0x00, 0x02, 0x00, 0x34, 0x25, 0xa0, 0x02, 0x74, 0xba, 0x01, 0xf8, 0x00, 0x03,
       0xb0, 0x00, 0xfe, 0xf0, 0xc4, 0x0b, 0xff, 0x20, 0xb9, 0x61, 0x75, 0xc8, 0xd7, 0xff,
       0x20, 0x3f, 0xe4, 0x07, 0xfc, 0x80, 0xff, 0xf8, 0xf2, 0x00, 0x00, 0x00]


@pytest.mark.parametrize("n", [2, 3])
def test_single_color_image(oracle, n):
    """decoder.rs:987-1033: 2x2 and 3x3 red images given as inline bytes, every pixel equal"""
    b = list(RED)
    b[26] = n
    b[28] = n
    img = oracle.WebPDecoder(bytes(b)).read_image()
    assert img.shape == (n, n, 3)
    assert (img.reshape(-1, 3) == img.reshape(-1, 3)[0]).all()


def test_riff_overflow_does_not_crash(oracle):
    """decoder.rs:970-984 (add_with_overflow_size): must not crash; any status is fine"""
    b = bytes([0x52, 0x49, 0x46, 0x46, 0xaf, 0x37, 0x80, 0x47, 0x57, 0x45, 0x42, 0x50, 0x6c, 0x64,
               0x00, 0x00, 0xff, 0xff, 0xff, 0xff, 0xfb, 0x7e, 0x73, 0x00, 0x06, 0x00, 0x00, 0x00,
               0x00, 0x00, 0x00, 0x00, 0x05, 0x00, 0x00, 0x00, 0x65, 0x65, 0x65, 0x65, 0x65, 0x65,
               0x40, 0xfb, 0xff, 0xff, 0x65, 0x65, 0x65, 0x65, 0x65, 0x65, 0x65, 0x65, 0x65, 0x65,
               0x00, 0x00, 0x00, 0x00, 0x62, 0x00, 0x10, 0x00, 0x00, 0x00, 0x00, 0x00, 0x00, 0x49,
               0x49, 0x54, 0x55, 0x50, 0x4c, 0x54, 0x59, 0x50, 0x45, 0x33, 0x37, 0x44, 0x4d, 0x46])
    with pytest.raises(oracle.OracleError):
        oracle.WebPDecoder(b)


def _composite(oracle, canvas, cw, ch, clear, frame, fx, fy, fw, fh, alpha, blend, pw, ph, px, py):
    c = (C.c_uint8 * 4)(*(clear or [0, 0, 0, 0]))
    oracle.lib().ora_composite_frame(canvas.ctypes.data, cw, ch, 1 if clear is not None else 0, c, frame.ctypes.data,
                                     fx, fy, fw, fh, int(alpha), int(blend), pw, ph, px, py)


def test_dispose_clear_fullsize_rgb_frame(oracle):
    """extended.rs:316-347"""
    canvas = np.full(4 * 4 * 4, 0xAA, dtype=np.uint8)
    frame = np.full(4 * 4 * 3, 0xFF, dtype=np.uint8)
    _composite(oracle, canvas, 4, 4, [0, 0, 0, 0], frame, 0, 0, 4, 4, False, True, 4, 4, 0, 0)
    assert (canvas == 0xFF).all()


def test_dispose_clear_subframe_rgb_frame(oracle):
    """extended.rs:349-418"""
    canvas = np.full(8 * 8 * 4, 0xAA, dtype=np.uint8)
    frame = np.full(4 * 4 * 3, 0xFF, dtype=np.uint8)
    _composite(oracle, canvas, 8, 8, [0, 0, 0, 0], frame, 0, 0, 4, 4, False, True, 4, 4, 2, 2)
    c = canvas.reshape(8, 8, 4)
    for y in range(2, 6):
        for x in range(2, 6):
            if x >= 4 or y >= 4:
                assert c[y, x].tolist() == [0, 0, 0, 0]
    assert (c[:4, :4] == 0xFF).all()
    assert c[0, 7].tolist() == [0xAA] * 4


def test_alpha_blending_close_to_float(oracle):
    """alpha_blending.rs:109-129 (tolerance test): integer blend within 3 of the f64 formula on a sample"""
    L = oracle.lib()
    rng = np.random.default_rng(7)
    out = np.zeros(4, dtype=np.uint8)
    for _ in range(4000):
        s = rng.integers(0, 256, 4, dtype=np.uint8)
        d = rng.integers(0, 256, 4, dtype=np.uint8)
        L.ora_blend_pixel(s.ctypes.data, d.ctypes.data, out.ctypes.data)
        sa, da = s[3] / 255.0, d[3] / 255.0
        oa = sa + da * (1 - sa)
        if s[3] == 0:
            assert out.tolist() == d.tolist()
            continue
        for c in range(3):
            ref = (s[c] * sa + d[c] * da * (1 - sa)) / oa
            assert abs(float(out[c]) - ref) <= 3.0
        assert abs(float(out[3]) - oa * 255) <= 1.0
