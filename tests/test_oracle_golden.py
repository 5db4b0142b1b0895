"""The oracle against every golden image of the reference's integration tests
(reference tests/decode.rs:39-203).  Expected bytes come from tests/reference/*.png (dwebp
output) via tests/golden/fixtures.npz; comparison is exact, as decode.rs:85-115 does it."""
import hashlib
import zlib

import numpy as np
import pytest

STILL = (["gallery1/%d" % i for i in range(1, 6)]
         + ["gallery2/%d_webp_a" % i for i in range(1, 6)]
         + ["gallery2/%d_webp_ll" % i for i in range(1, 6)]
         + ["regression/" + n for n in ("color_index", "dark", "tiny", "lossless_indexed_1bit_palette",
                                        "lossless_indexed_2bit_palette", "lossless_indexed_4bit_palette")])


def check(fixtures, key, img):
    want = fixtures["sha/" + key].tobytes()
    got = hashlib.sha256(img.tobytes()).digest()
    if got != want:
        crc = fixtures["rowcrc/" + key]
        rows = [y for y in range(img.shape[0]) if zlib.crc32(img[y].tobytes()) != int(crc[y])]
        pytest.fail("%s: %d of %d rows differ (first %s)" % (key, len(rows), img.shape[0], rows[:5]))
    if "full/" + key in fixtures.files:
        assert np.array_equal(img, fixtures["full/" + key])


@pytest.mark.parametrize("key", STILL)
def test_still_images(oracle, fixtures, key):
    d = oracle.WebPDecoder(fixtures["webp/" + key])
    w, h, c = (int(v) for v in fixtures["dims/" + key])
    assert d.dimensions() == (w, h)
    assert d.has_alpha() == (c == 4)
    assert d.output_buffer_size() == w * h * c
    assert not d.is_animated()
    check(fixtures, key, d.read_image())


@pytest.mark.parametrize("i", range(1, 6))
def test_nofancy(oracle, fixtures, i):
    """reftest_nofancy!, decode.rs:169-191: UpsamplingMethod::Simple vs dwebp -nofancy."""
    d = oracle.WebPDecoder(fixtures["webp/gallery1/%d" % i], upsampling="simple")
    check(fixtures, "gallery1_nofancy/%d" % i, d.read_image())


@pytest.mark.parametrize("name", ["random_lossy", "random_lossless"])
def test_animated(oracle, fixtures, name):
    """decode.rs:118-151: read_image gives frame 1 without moving the cursor; read_frame walks all frames."""
    key = "animated/" + name
    d = oracle.WebPDecoder(fixtures["webp/" + key])
    n = int(fixtures["nframes/" + key][0])
    assert d.is_animated() and d.num_frames() == n
    assert d.is_lossy() == (name == "random_lossy")
    check(fixtures, key + "-1", d.read_image())
    for i in range(1, n + 1):
        img, dur = d.read_frame()
        check(fixtures, "%s-%d" % (key, i), img)
    with pytest.raises(oracle.OracleError) as e:
        d.read_frame()
    assert e.value.code == 20  # NoMoreFrames
    d.reset_animation()
    img, _ = d.read_frame()
    check(fixtures, key + "-1", img)


def test_wrong_buffer_size(oracle, fixtures):
    d = oracle.WebPDecoder(fixtures["webp/regression/dark"])
    with pytest.raises(oracle.OracleError) as e:
        d.read_image(np.zeros(5, dtype=np.uint8))
    assert e.value.code == 11  # ImageTooLarge, decoder.rs:670-672
