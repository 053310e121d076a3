"""CPU (-m "not gpu"): the training error map in the oracle (oracle/orc_errmap.inc).
binary_search / sample_cdf_2d / image_idx: bit-exact against vectors from the reference's own host-callable functions
(tests/golden/ref_host_cdf.npz, made by tests/golden/make_golden.py cdf through oracle/ref_shim.cu).
construct_cdf_2d / construct_cdf_1d / the host image CDF / the bilinear deposit are __global__ or inline Testbed code in the
reference: checked against numpy closed forms."""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
f32 = np.float32


def bits_equal(a, b):
    return np.array_equal(np.ascontiguousarray(a, f32).view(np.uint32), np.ascontiguousarray(b, f32).view(np.uint32))


def test_cdf_sampling_pinned_on_reference_functions(orc):
    g = np.load(os.path.join(HERE, "golden", "ref_host_cdf.npz"))
    cx, cy, ci = g["cdf_x_cond_y"], g["cdf_y"], g["cdf_img"]
    # the committed CDFs are what the oracle builds from the committed error map
    ox, oy, oi, _ = orc.error_map_cdf(g["data"])
    assert bits_equal(ox, cx) and bits_equal(oy, cy) and bits_equal(oi, ci)
    for v, a, b in zip(g["bs_vals"], g["bs_img"], g["bs_row"]):
        assert orc.binary_search(float(v), ci) == a and orc.binary_search(float(v), cy[1]) == b
    n_img = cx.shape[0]
    for img in range(n_img):
        importance = 0
        for s, uv, pdf in zip(g["samples"], g[f"uv_{img}"], g[f"pdf_{img}"]):
            o_uv, o_pdf = orc.sample_cdf_2d(s, img, cx, cy)
            assert bits_equal(o_uv, uv) and bits_equal([o_pdf], [pdf])
            importance += s[0] >= 0.5
        assert importance > 100  # both halves of UNIFORM_SAMPLING_FRACTION are exercised
    for tag, cdf in (("cdf", ci), ("none", None)):
        for b, img, pdf in zip(g["ii_base"], g[f"ii_img_{tag}"], g[f"ii_pdf_{tag}"]):
            o_img, o_pdf = orc.image_idx(int(b), 4096, n_img, cdf)
            assert o_img == img and bits_equal([o_pdf], [pdf])
    assert len(set(g["ii_img_cdf"].tolist())) == n_img


def test_cdf_construction_closed_forms(orc):
    rng = np.random.default_rng(5)
    n, h, w = 4, 9, 13
    data = rng.random((n, h, w)).astype(f32)
    data[2] = 0.0   # an image without any error: everything falls back to the uniform floor
    cx, cy, ci, pi = orc.error_map_cdf(data)
    d = data.astype(np.float64) + 1e-10
    row = d.sum(2)
    want_x = 0.99 * np.cumsum(d, 2) / row[..., None] + 0.01 * (np.arange(w) + 1) / w
    want_y = 0.99 * np.cumsum(row, 1) / row.sum(1)[:, None] + 0.01 * (np.arange(h) + 1) / h
    mass = row.sum(1)
    want_i = 0.9 * np.cumsum(mass) / mass.sum() + 0.1 * (np.arange(n) + 1) / n
    want_p = 0.9 * mass / mass.sum() + 0.1 / n
    assert np.allclose(cx, want_x, atol=2e-6) and np.allclose(cy, want_y, atol=2e-6) and np.allclose(ci, want_i, atol=2e-6) and np.allclose(pi, want_p, atol=2e-6)
    assert np.allclose(cx[..., -1], 1.0, atol=1e-6) and np.allclose(cy[:, -1], 1.0, atol=1e-6) and abs(ci[-1] - 1.0) < 1e-6
    assert np.all(np.diff(cx, axis=2) > 0) and np.all(np.diff(cy, axis=1) > 0) and np.all(np.diff(ci) > 0)
    assert np.allclose(np.diff(cx[2], axis=1), 1.0 / w, atol=1e-6)  # empty image: uniform


def test_sampling_follows_the_error_and_pdf_integrates_to_one(orc):
    n_img, h, w = 2, 8, 8
    data = np.full((n_img, h, w), 1e-3, f32)
    data[0, 2, 5] = 10.0
    data[1, 6, 1] = 10.0; data[1] *= 3.0
    cx, cy, ci, pi = orc.error_map_cdf(data)
    n = 40000
    img, uv, pdf = orc.sample_training_pixels(n, n_img, seed=7, step=3, cdf_x_cond_y=cx, cdf_y=cy, cdf_img=ci)
    # image 1 carries 3x the error: pmf = 0.9 * 0.75 + 0.05
    assert abs((img == 1).mean() - pi[1]) < 0.02 and abs(pi[1] - 0.725) < 1e-3
    px = np.minimum((uv * [w, h]).astype(int), [w - 1, h - 1])
    for k, (hx, hy) in enumerate([(5, 2), (1, 6)]):
        m = img == k
        frac = ((px[m, 0] == hx) & (px[m, 1] == hy)).mean()
        assert 0.45 < frac < 0.55                       # half of the samples are uniform, the other half lands on the hot texel
    assert np.all((uv >= 0) & (uv <= 1))
    assert abs(np.mean(1.0 / pdf) - 1.0) < 0.05         # E_p[1/p] = 1: the pdf is the density the samples are drawn with... on the importance half
    # without CDFs: image by ray index, uv = the first two floats of the ray's pcg32 stream, pdf 1
    img0, uv0, pdf0 = orc.sample_training_pixels(1000, 5, seed=7, step=3)
    assert np.array_equal(img0, (np.arange(1000) * 5) // 1000) and np.all(pdf0 == 1.0)
    # a shard of the global ray stream (rank 1 of 2) is the second half of it
    img_a, uv_a, _ = orc.sample_training_pixels(500, 5, seed=7, step=3, ray_offset=500, n_rays_total=1000)
    assert np.array_equal(img_a, img0[500:]) and np.array_equal(uv_a, uv0[500:])


def test_deposit_is_bilinear_and_conserves_mass(orc):
    res, image_res = (6, 5), (64, 48)
    # a value at a texel centre goes to that texel alone; between four centres it is shared by the bilinear weights
    m = orc.error_map_deposit([0], [[(2 + 0.5) / 6, (3 + 0.5) / 5]], [2.0], 2, res, image_res)
    assert m[0, 3, 2] == 2.0 and m.sum() == 2.0
    m = orc.error_map_deposit([1], [[3.0 / 6, 2.0 / 5]], [4.0], 2, res, image_res)
    assert np.allclose(m[1, 1:3, 2:4], 1.0) and m[0].sum() == 0
    rng = np.random.default_rng(2)
    n = 5000
    img, uv, val = rng.integers(0, 2, n), rng.random((n, 2)).astype(f32), rng.random(n).astype(f32)
    m = orc.error_map_deposit(img, uv, val, 2, res, image_res)
    assert abs(m.sum() - val.sum()) < 1e-2 * val.sum() and np.all(m >= 0)
    assert abs(m[0].sum() - val[img == 0].sum()) < 1e-2 * val.sum()
