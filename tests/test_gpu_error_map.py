"""-m gpu: a-21 training error map through the C ABI (hn_op_error_map_cdf, hn_op_sample_training_pixels, hn_op_error_map_deposit,
hn_get_error_map, options sample_*_proportional_to_error) against the oracle (oracle/orc_errmap.inc, itself pinned on the reference's
host-callable sample_cdf_2d / image_idx / binary_search: tests/test_error_map.py)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
f32 = np.float32


def bits_equal(a, b):
    return np.array_equal(np.ascontiguousarray(a, f32).view(np.uint32), np.ascontiguousarray(b, f32).view(np.uint32))


@pytest.fixture(scope="module")
def tb(hn, root):
    t = hn.Testbed()
    t.reload_network_from_file(os.path.join(root, "configs", "nerf", "base.json"))
    return t


def test_cdf_kernels_match_oracle(tb, orc):
    rng = np.random.default_rng(21)
    for n, h, w in ((5, 23, 31), (1, 2, 2), (70, 40, 40)):
        data = (rng.random((n, h, w)) ** 3).astype(f32)
        data[0, 0] = 0.0
        if n > 2:
            data[2] = 0.0
        got, want = tb.op_error_map_cdf(data), orc.error_map_cdf(data)
        for name, a, b in zip(("cdf_x_cond_y", "cdf_y", "cdf_img", "pmf_img"), got, want):
            # sequential sums in the same order; the device code contracts a * b + c into FMAs (as the reference's nvcc build does), the oracle
            # does not: CDF values in [0, 1] agree to 2 ulp of 1.0
            assert np.abs(a.astype(np.float64) - b).max() <= 2.4e-7, (name, n, h, w)


def test_pixel_sampling_bit_exact(tb, orc):
    rng = np.random.default_rng(22)
    n_img, h, w = 6, 17, 19
    data = (rng.random((n_img, h, w)) ** 4).astype(f32)
    data[3, 4:6, 7:9] += 20.0
    cx, cy, ci, _ = orc.error_map_cdf(data)
    n = 20000
    for kw in (dict(), dict(cdf_x_cond_y=cx, cdf_y=cy), dict(cdf_img=ci), dict(cdf_x_cond_y=cx, cdf_y=cy, cdf_img=ci)):
        for off, total in ((0, n), (n, 4 * n)):
            img, uv, pdf = tb.op_sample_training_pixels(n, n_img, seed=99, step=12, ray_offset=off, n_rays_total=total, **kw)
            oimg, ouv, opdf = orc.sample_training_pixels(n, n_img, seed=99, step=12, ray_offset=off, n_rays_total=total, **kw)
            assert np.array_equal(img, oimg) and bits_equal(uv, ouv) and bits_equal(pdf, opdf), (sorted(kw), off)
    # with the CDFs the hot spot of image 3 draws about half of that image's rays
    m = img == 3
    px = (uv[m] * [w, h]).astype(int)
    frac = ((px[:, 0] >= 7) & (px[:, 0] < 9) & (px[:, 1] >= 4) & (px[:, 1] < 6)).mean()
    assert 0.22 < frac < 0.34   # importance half: 80 / (80 + 0.2 * 319) of the mass, uniform half: 4 / 323


def test_deposit_matches_oracle(tb, orc):
    rng = np.random.default_rng(23)
    n, n_img, res, image_res = 30000, 4, (29, 23), (640, 480)
    img, uv, val = rng.integers(0, n_img, n), rng.random((n, 2)).astype(f32), rng.random(n).astype(f32)
    uv[:4] = [[0, 0], [1, 1], [0, 1], [0.5, 0.5]]
    got, want = tb.op_error_map_deposit(img, uv, val, n_img, res, image_res), orc.error_map_deposit(img, uv, val, n_img, res, image_res)
    # float atomics in arbitrary order vs the oracle's ray order: ~30 values of O(1) per texel -> 1e-5 absolute
    assert np.abs(got - want).max() <= 1e-4 and abs(got.sum() - val.sum()) <= 1e-3 * val.sum()
    # the reference clamps the texel index against the IMAGE resolution - 2 (src/testbed_nerf.cu:1086): an image smaller than the map
    tiny = tb.op_error_map_deposit([0], [[0.9, 0.9]], [1.0], 1, (8, 8), (4, 4))
    assert np.allclose(tiny, orc.error_map_deposit([0], [[0.9, 0.9]], [1.0], 1, (8, 8), (4, 4)), atol=1e-7) and tiny[0, 2:4, 2:4].sum() == pytest.approx(1.0)


def test_training_fills_the_map_and_importance_sampling_trains(hn, root):
    imgs, cams, focal = hn.synthetic.sphere_dataset(8, 64)
    t = hn.Testbed()
    t.reload_network_from_file(os.path.join(root, "configs", "nerf", "base.json"))
    t.set_params(hn.synthetic.random_params(t.network_info(), 1337, 1e-4))
    t.set_density_grid(hn.synthetic.sphere_density_grid(0.35), 1)
    t.set_training_data(imgs, cams, focal)
    t.reset_trainer(3)
    first = t.train(1)
    em = t.error_map()
    st = t.train_stats()
    # resolution rule of train_nerf (:2987-2990): (n_steps_between * rays_per_batch / n_images)^(1/4) * 3.5, capped by the image
    assert em.shape[0] == 8 and em.shape[1] == em.shape[2] == min(int(np.sqrt(np.sqrt(128 * 4096 // 8)) * 3.5), 64)
    assert np.all(em >= 0) and em.sum() > 0
    # each ray deposits its mean loss once: the map's total is the step's summed loss (rays that found samples)
    assert em.sum() == pytest.approx(first * 4096, rel=2e-3)
    losses = [t.train(1) for _ in range(127)]           # step 128 closes the first period: CDFs are built, the next step starts a new map
    before = t.error_map().sum()
    assert before > em.sum()
    t.set_option("sample_focal_plane_proportional_to_error", 1.0)
    t.set_option("sample_image_proportional_to_error", 1.0)
    more = [t.train(1) for _ in range(64)]
    assert t.error_map().sum() < before                  # a new accumulation period began
    assert np.isfinite(more).all() and np.mean(more[-10:]) < 0.5 * first
    assert t.train_stats().step == 1 + 127 + 64
