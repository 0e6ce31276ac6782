"""preprocess()'s zero-copy view of a PIL image's pixels (Pillow's Arrow export) against the copying
route (np.asarray): same pixels for every way an image can come into being, and no use of the
exporter on images that wrap foreign memory (Pillow 12.2 crashes on those). CPU only."""
import os

import numpy as np
import pytest
from PIL import Image

from triton_client_b200.image_client import _pil_pixels, _pil_to_array


def _same(img, c=3):
    x, y = _pil_pixels(img, c), _pil_to_array(img, c)
    k = y.shape[2]
    assert x.dtype == np.uint8 and x.shape[:2] == y.shape[:2] and x.shape[2] in (k, 4)
    assert np.array_equal(x[:, :, :k], y)
    return x


A = np.random.default_rng(0).integers(0, 256, (120, 160, 3), dtype=np.uint8)


def test_made_in_memory():
    base = Image.fromarray(A, "RGB")
    made = [base, base.convert("L"), base.convert("RGBA"), base.convert("P"), base.convert("CMYK"), base.convert("LA"),
            base.crop((3, 5, 100, 90)), base.resize((50, 40)), base.rotate(90), base.copy(), Image.new("RGB", (7, 3)),
            Image.frombytes("RGB", (160, 120), A.tobytes()), base.convert("RGBA").split()[0],
            # foreign memory: the exporter must not be used
            Image.frombuffer("RGB", (160, 120), A.tobytes(), "raw", "RGB", 0, 1),
            Image.frombuffer("L", (160, 120), A[:, :, 0].tobytes(), "raw", "L", 0, 1), Image.fromarray(A[:, :, 0], "L")]
    for img in made:
        _same(img)
        _same(img, 1)
    if hasattr(Image.Image, "__arrow_c_array__"):
        try:
            import pyarrow  # noqa: F401
        except ImportError:
            return
        assert _same(base).shape[2] == 4  # RGB arrives as Pillow stores it: RGBX, no copy


@pytest.mark.parametrize("fmt", ["PNG", "BMP", "JPEG", "TIFF", "PPM", "GIF", "WEBP"])
def test_opened_from_files(fmt, tmp_path):
    base = Image.fromarray(A, "RGB")
    for mode in ("RGB", "L", "RGBA"):
        fn = os.path.join(str(tmp_path), "x_%s.%s" % (mode, fmt.lower()))
        try:
            base.convert(mode).save(fn, fmt)
        except Exception:
            continue  # the format cannot hold this mode
        _same(Image.open(fn))  # lazily loaded, possibly memory-mapped


def test_too_large_for_one_block_takes_the_copying_route():
    big = Image.new("RGB", (5000, 4000))
    assert _same(big).shape == (4000, 5000, 3)
