"""-m gpu: camera models of the ray init (uv_to_ray, common_device.cuh:416-483) through the C ABI against the oracle, which
is itself pinned on the reference's host-callable uv_to_ray (tests/test_oracle_pin.py::test_camera_models_pinned).
Perspective / OpenCV are plain float arithmetic -> bit-exact rays; the models that go through sin/cos/atan (f-theta,
lat-long, fisheye, equirectangular, and the disk sample of the depth of field) are compared to 2e-5, frames to the
north_star tolerance (max abs <= 1e-2 outside <= 0.2 % edge pixels, PSNR >= 50 dB)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CASES = [  # tag, lens_mode, lens_params, aperture, focus_z, parallax, exact
    ("parallax", 0, (0,) * 7, 0.0, 1.0, (0.03, -0.02, 0.6), True),
    ("dof", 0, (0,) * 7, 0.04, 1.6, (0, 0, 0), False),
    ("opencv", 1, (-0.12, 0.05, 0.002, -0.003, 0, 0, 0), 0.0, 1.0, (0, 0, 0), True),
    ("ftheta", 2, (0.0, 4.0e-3, 2.0e-8, -1.0e-11, 0.0, 800.0, 600.0), 0.0, 1.0, (0, 0, 0), False),
    ("latlong", 3, (0,) * 7, 0.0, 1.0, (0, 0, 0), False),
    ("fisheye", 4, (0.03, -0.01, 0.004, -0.001, 0, 0, 0), 0.0, 1.0, (0, 0, 0), False),
    ("equirect", 5, (0,) * 7, 0.0, 1.0, (0, 0, 0), False),
]


def psnr(a, b):
    mse = float(np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2))
    return 99.0 if mse == 0 else 10.0 * np.log10(1.0 / mse)


def apply_case(tb, case):
    _, mode, lp, ap, fz, par, _ = case
    tb.set_option("lens_mode", float(mode))
    tb.set_option("lens_params", list(lp))
    tb.set_option("aperture_size", ap)
    tb.set_option("focus_z", fz)
    tb.set_option("parallax_shift", list(par))


def case_kw(case):
    _, mode, lp, ap, fz, par, _ = case
    return dict(lens_mode=mode, lens_params=lp, aperture_size=ap, focus_z=fz, parallax_shift=par)


@pytest.fixture(scope="module")
def nerf(hn, root):
    tb = hn.Testbed()
    tb.reload_network_from_file(os.path.join(root, "configs", "nerf", "base.json"))
    info = tb.network_info()
    grid = hn.synthetic.sphere_density_grid(0.35)
    tb.set_density_grid(grid, 1)
    tb.snap_to_pixel_centers = True
    blob = hn.synthetic.random_params(info, 1337, 1.0)
    tb.set_params(blob)
    return tb, grid, blob


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_init_rays_camera_models(nerf, orc, hn, case):
    tb, grid, _ = nerf
    w, h, spp = 160, 120, 2
    cam = hn.synthetic.look_at_camera((1.2, 0.9, 1.4), (0.5, 0.5, 0.5))
    apply_case(tb, case)
    try:
        o, d, t, alive = tb.op_init_rays(cam, cam, w, h, spp)
    finally:
        apply_case(tb, ("", 0, (0,) * 7, 0.0, 1.0, (0, 0, 0), True))
    bits, _ = orc.bitfield_from_grid(grid, 1)
    p = orc.render_params(w, h, cam, spp=spp, snap=True, rel_focal=tb.get_option("relative_focal_length", 2), **case_kw(case))
    ro, rd, rt, ra = orc.init_rays(p, bits)
    assert ra.sum() > 200, case[0]
    if case[6]:
        assert np.array_equal(alive, ra)
        assert np.array_equal(o.view(np.uint32), ro.view(np.uint32)) and np.array_equal(d.view(np.uint32), rd.view(np.uint32))
        assert np.array_equal(t.view(np.uint32)[ra == 1], rt.view(np.uint32)[ra == 1])
    else:
        assert (alive != ra).mean() <= 2e-3, (case[0], (alive != ra).sum())
        assert np.allclose(o, ro, rtol=0, atol=2e-5) and np.allclose(d, rd, rtol=0, atol=2e-5), case[0]
        both = (alive == 1) & (ra == 1)
        assert np.mean(np.abs(t[both] - rt[both]) <= 1e-3) > 0.995, case[0]
    if case[0] == "ftheta":  # pixels outside the image circle are invalid rays: zero direction, never alive
        dead = np.all(rd == 0, axis=1)
        assert dead.sum() > 100 and not ra[dead].any() and np.array_equal(dead, np.all(d == 0, axis=1))


@pytest.mark.parametrize("case", [CASES[2], CASES[4], CASES[5]], ids=["opencv", "latlong", "fisheye"])
def test_nerf_frame_with_lens(nerf, orc, hn, case):
    tb, grid, blob = nerf
    w, h = 160, 120
    cam = hn.synthetic.look_at_camera((1.2, 0.9, 1.4), (0.5, 0.5, 0.5))
    tb.set_camera_cols(cam)
    plain = tb.render(w, h, 1, True)
    apply_case(tb, case)
    try:
        img = tb.render(w, h, 1, True)
        ref, _, _ = orc.render_reference_frame(tb, blob, grid, 1, [], cam, w, h, **case_kw(case))
    finally:
        apply_case(tb, ("", 0, (0,) * 7, 0.0, 1.0, (0, 0, 0), True))
    assert np.abs(img - plain).max() > 2e-3, "lens did not change the picture"
    bad = np.abs(img - ref).max(axis=2) > 1e-2
    assert bad.mean() <= 2e-3, (case[0], bad.sum())
    assert psnr(np.where(bad[..., None], ref, img), ref) >= 50.0


def test_hybrid_frame_depth_of_field_and_parallax(hn, orc, root):
    """Both ray inits (mesh :488-579 and NeRF :1428-1544) take the parallax shift and the aperture: the mesh and the NeRF
    must stay registered (same camera model on both passes) and agree with the oracle."""
    w, h = 192, 128
    mesh = hn.synthetic.torus_mesh(128, 64)
    aabb = [-1, -1, -1, 2, 2, 2]
    tb = hn.Testbed()
    tb.reload_network_from_file(os.path.join(root, "configs", "nerf", "base.json"))
    grid = hn.synthetic.sphere_density_grid(0.35)
    params = hn.synthetic.random_params(tb.network_info(), 1337, 1.0)
    tb.set_params(params)
    tb.set_density_grid(grid, 1)
    tb.add_mesh(mesh)
    tb.build_scene()
    tb.set_option("mesh_aabb", aabb)
    tb.snap_to_pixel_centers = True
    cam = hn.synthetic.look_at_camera((1.5, 1.3, 2.0), (0.5, 0.5, 0.5))
    tb.set_camera_cols(cam)
    plain = tb.render(w, h, 1, True)
    case = ("dof_parallax", 0, (0,) * 7, 0.03, 2.2, (0.02, 0.01, 0.5), False)
    apply_case(tb, case)
    img = tb.render(w, h, 1, True)
    ref, _, _ = orc.render_reference_frame(tb, params, grid, 1, [mesh], cam, w, h, mesh_aabb=aabb, **case_kw(case))
    assert np.abs(img - plain).max() > 0.05
    bad = np.abs(img - ref).max(axis=2) > 1e-2
    assert bad.mean() <= 3e-3, (bad.sum(), np.abs(img - ref).max())
    assert psnr(np.where(bad[..., None], ref, img), ref) >= 50.0


def rot_y_cols(deg):
    a = np.deg2rad(deg)
    c, s = np.float32(np.cos(a)), np.float32(np.sin(a))
    # column-major mat3: columns (c,0,-s), (0,1,0), (s,0,c)
    return np.array([c, 0, -s, 0, 1, 0, s, 0, c], np.float32)


def test_rotated_crop_box(nerf, orc, hn):
    """m_render_aabb_to_local (src/testbed_nerf.cu:1530-1532, nerf_device.cuh:461-495): the crop box test runs in the
    rotated frame. Pure float multiplies/adds in the same order on both sides -> ray set-up and samples bit-exact."""
    tb, grid, blob = nerf
    w, h = 128, 96
    cam = hn.synthetic.look_at_camera((1.2, 0.9, 1.4), (0.5, 0.5, 0.5))
    tb.set_camera_cols(cam)
    R = rot_y_cols(25.0)
    box = [0.05, 0.3, 0.35, 0.75, 0.7, 1.0]
    plain = tb.render(w, h, 1, True).copy()
    tb.set_option("render_aabb", box)
    tb.set_option("render_aabb_to_local", R.tolist())
    try:
        o, d, t, alive = tb.op_init_rays(cam, cam, w, h, 0)
        bits, _ = orc.bitfield_from_grid(grid, 1)
        p = orc.render_params(w, h, cam, rel_focal=tb.get_option("relative_focal_length", 2), render_aabb=box, render_aabb_to_local=R)
        ro, rd, rt, ra = orc.init_rays(p, bits)
        assert 300 < ra.sum() < w * h // 2
        assert np.array_equal(alive, ra) and np.array_equal(t.view(np.uint32)[ra == 1], rt.view(np.uint32)[ra == 1])
        counts, samples = tb.op_march_all(ro, rd, rt, ra, max_per_ray=48)
        rc, rs = orc.march_all(p, bits, ro, rd, rt, ra, max_per_ray=48)
        assert np.array_equal(counts, rc) and np.array_equal(samples.view(np.uint32), rs.view(np.uint32))
        img = tb.render(w, h, 1, True)
        ref, _, _ = orc.render_reference_frame(tb, blob, grid, 1, [], cam, w, h, render_aabb_to_local=R)
    finally:
        tb.set_option("render_aabb", [0, 0, 0, 1, 1, 1])
        tb.set_option("render_aabb_to_local", [1, 0, 0, 0, 1, 0, 0, 0, 1])
    assert np.abs(img - plain).max() > 0.05, "crop did not change the picture"
    assert np.abs(img - ref).max() <= 1e-2 and psnr(img, ref) >= 50.0
    # un-rotated crop of the same size gives a different picture (the rotation is really applied)
    tb.set_option("render_aabb", box)
    try:
        axis_aligned = tb.render(w, h, 1, True)
    finally:
        tb.set_option("render_aabb", [0, 0, 0, 1, 1, 1])
    assert np.abs(axis_aligned - img).max() > 0.05
