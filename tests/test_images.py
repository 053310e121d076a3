"""CPU (-m "not gpu"): the dataset loader's image decoders (hn_decode_image_file: PNG reader in hn_dataset.cpp, JPEG reader in hn_jpeg.cpp)
against PIL (libpng / libjpeg-turbo) on images written here. PNG is lossless: exact. JPEG: the two decoders run different inverse DCTs
(float here, integer "islow" in libjpeg-turbo) and round the chroma interpolation differently: +-3 of 255 per channel, mean error < 0.6."""
import io
import os

import numpy as np
import pytest

PIL = pytest.importorskip("PIL.Image")


@pytest.fixture(scope="module")
def picture():
    rng = np.random.default_rng(4)
    h, w = 67, 93   # not a multiple of the 8 / 16 pixel MCU
    yy, xx = np.mgrid[0:h, 0:w]
    img = np.stack([127 + 120 * np.sin(xx / 9.0) * np.cos(yy / 13.0), 255.0 * xx / w, 255.0 * ((xx // 8 + yy // 8) % 2)], -1)
    img += rng.normal(0, 6, img.shape)
    return np.clip(img, 0, 255).astype(np.uint8)


def test_png_decoder_exact(hn, picture, tmp_path):
    rgba = np.concatenate([picture, (255 - picture[..., :1])], -1)
    for mode, arr in (("RGBA", rgba), ("RGB", picture), ("L", picture[..., 0]), ("LA", rgba[..., [0, 3]])):
        p = str(tmp_path / f"t_{mode}.png")
        PIL.fromarray(arr, mode).save(p)
        got = hn.decode_image_file(p)
        want = np.asarray(PIL.open(p).convert("RGBA"))
        assert np.array_equal(got, want), mode


@pytest.mark.parametrize("subsampling,quality,restart", [(0, 92, 0), (1, 90, 0), (2, 85, 0), (2, 60, 4), (0, 100, 0)])
def test_jpeg_decoder_vs_libjpeg(hn, picture, tmp_path, subsampling, quality, restart):
    p = str(tmp_path / "t.jpg")
    kw = dict(quality=quality, subsampling=subsampling)
    if restart:
        kw["restart_marker_blocks"] = restart
    PIL.fromarray(picture, "RGB").save(p, "JPEG", **kw)
    got = hn.decode_image_file(p)
    want = np.asarray(PIL.open(p).convert("RGB")).astype(int)
    d = np.abs(got[..., :3].astype(int) - want)
    assert got.shape == (picture.shape[0], picture.shape[1], 4) and np.all(got[..., 3] == 255)
    assert d.max() <= 3 and d.mean() < 0.6, (subsampling, quality, d.max(), d.mean())
    # and it is the picture that was saved (JPEG is lossy: loose bound)
    assert np.abs(got[..., :3].astype(int) - picture).mean() < 20   # noisy picture, subsampled chroma


def test_jpeg_grey_and_errors(hn, picture, tmp_path):
    p = str(tmp_path / "g.jpg")
    PIL.fromarray(picture[..., 0], "L").save(p, "JPEG", quality=90)
    got = hn.decode_image_file(p)
    want = np.asarray(PIL.open(p).convert("L")).astype(int)
    assert np.abs(got[..., 0].astype(int) - want).max() <= 2 and np.array_equal(got[..., 0], got[..., 1]) and np.array_equal(got[..., 0], got[..., 2])
    pp = str(tmp_path / "prog.jpg")
    PIL.fromarray(picture, "RGB").save(pp, "JPEG", quality=90, progressive=True)
    with pytest.raises(RuntimeError, match="progressive"):
        hn.decode_image_file(pp)
    with pytest.raises(RuntimeError):
        hn.decode_image_file(str(tmp_path / "missing.jpg"))
    (tmp_path / "junk.png").write_bytes(b"not an image at all")
    with pytest.raises(RuntimeError):
        hn.decode_image_file(str(tmp_path / "junk.png"))
    trunc = open(p, "rb").read()[:200]
    (tmp_path / "trunc.jpg").write_bytes(trunc)
    with pytest.raises(RuntimeError):
        hn.decode_image_file(str(tmp_path / "trunc.jpg"))
