"""CPU (-m "not gpu"): foveation, hidden-area mask, distortion grid, pos_to_uv and to_rgb of the oracle (oracle/orc_fov.inc, uv_to_ray in
orc_nerf.cpp) against vectors from the reference's own host-callable functions (tests/golden/ref_host_fov.npz, made by
tests/golden/make_golden.py fov through oracle/ref_shim.cu)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
f32 = np.float32


def bits_equal(a, b):
    return np.array_equal(np.ascontiguousarray(a, f32).view(np.uint32), np.ascontiguousarray(b, f32).view(np.uint32))


def load():
    sys.path.insert(0, os.path.join(HERE, "golden"))
    from make_golden import FOV_CASES
    return np.load(os.path.join(HERE, "golden", "ref_host_fov.npz")), FOV_CASES


def test_foveation_pinned(orc):
    g, cases = load()
    for tag, fov, mode, lp in cases:
        assert bits_equal(orc.foveation_coeffs(fov), g[f"{tag}_coeffs"]), tag
        w, un, de = orc.foveation_warp(fov, g["uv"])
        assert bits_equal(w, g[f"{tag}_warp"]) and bits_equal(de, g[f"{tag}_density"]), tag
        # unwarp takes a square root: libm's sqrtf and the reference's host sqrt are both correctly rounded
        assert bits_equal(un, g[f"{tag}_unwarp"]), tag
        if tag != "none":
            assert np.abs(w - g["uv"]).max() > 0.05, tag           # the case really warps
            back, _, _ = orc.foveation_warp(fov, un)               # warp(unwarp(y)) = y
            assert np.abs(back - g["uv"]).max() < 2e-5, tag
    assert bits_equal(orc.foveation_coeffs((0,) * 6), np.tile(np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 1, 0, 1], f32), (2, 1)))


def params_for(orc, g, fov, mode, lp, mask=None, dist=None):
    w, h = (int(v) for v in g["res"])
    p = orc.render_params(w, h, g["cam"], near=float(g["near"][0]), lens_mode=mode, lens_params=lp, aperture_size=0.0, focus_z=1.3, parallax_shift=g["parallax"],
                          foveation=fov, hidden_area_mask=mask, distortion=dist)
    p.focal[:] = g["focal"].tolist(); p.screen_center[:] = g["screen_center"].tolist()
    return p


def test_uv_to_ray_with_foveation_mask_and_distortion_pinned(orc):
    g, cases = load()
    for tag, fov, mode, lp in cases:
        for sub, use_mask, use_dist in (("plain", False, False), ("mask", True, False), ("dist", False, True), ("all", True, True)):
            p = params_for(orc, g, fov, mode, lp, g["mask"] if use_mask else None, g["distortion"] if use_dist else None)
            o, d, ok = orc.uv_to_ray(p, int(g["spp"][0]), g["uv"])
            assert np.array_equal(ok, g[f"{tag}_{sub}_valid"]), (tag, sub)
            v = ok.astype(bool)
            assert v.sum() > 100 and (not use_mask or (~v).sum() > 20), (tag, sub)
            if mode == 0:
                assert bits_equal(o[v], g[f"{tag}_{sub}_o"][v]) and bits_equal(d[v], g[f"{tag}_{sub}_d"][v]), (tag, sub)
            else:  # the iterative OpenCV undistortion: bit-exact as well (tests/test_oracle_pin.py pins it that way)
                assert bits_equal(o[v], g[f"{tag}_{sub}_o"][v]) and bits_equal(d[v], g[f"{tag}_{sub}_d"][v]), (tag, sub)
        # the grid and the foveation change the rays
        assert np.abs(g[f"{tag}_dist_d"] - g[f"{tag}_plain_d"]).max() > 1e-3
    assert np.abs(g["fov_mild_plain_d"] - g["none_plain_d"]).max() > 1e-2


def test_pos_to_uv_and_to_rgb_pinned(orc):
    g, cases = load()
    for tag, fov, mode, lp in cases:
        p = params_for(orc, g, fov, mode, lp)
        pts = g[f"{tag}_plain_o"] + g[f"{tag}_plain_d"]
        back = orc.pos_to_uv(p, pts)
        assert np.allclose(back, g[f"{tag}_pos_to_uv"], rtol=0, atol=2e-6), (tag, np.abs(back - g[f"{tag}_pos_to_uv"]).max())
        # pos_to_uv inverts uv_to_ray (perspective / OpenCV lens, no distortion grid): the round trip returns the pixel
        inside = (g["uv"] > 0.02).all(1) & (g["uv"] < 0.98).all(1)
        assert np.abs(back[inside] - g["uv"][inside]).max() < 5e-4, tag
    assert np.allclose(orc.to_rgb(g["rgb_dirs"]), g["rgb"], rtol=0, atol=1e-6)
    assert np.allclose(orc.to_rgb([[0, 0]]), 0)
