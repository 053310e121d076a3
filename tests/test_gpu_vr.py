"""-m gpu: the VR pieces of the camera model and the 2-D render modes through the C ABI - foveation (option "foveation"), hidden-area
mask, distortion grid, ERenderMode::Distortion / Slice, and the foveated blit (hn_unwarp_foveated) - against the oracle, whose
Foveation / uv_to_ray / pos_to_uv / to_rgb are pinned bit for bit on the reference's host-callable functions (tests/test_fov_oracle.py)."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
f32 = np.float32
HERE = os.path.dirname(os.path.abspath(__file__))

FOV = (0.5, 0.6, 0.45, 0.5, 0.2, 0.25)
OPENCV = (0.05, -0.02, 0.001, -0.002, 0, 0, 0)


def psnr(a, b):
    mse = float(np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2))
    return 99.0 if mse == 0 else 10.0 * np.log10(1.0 / mse)


def bits_equal(a, b):
    return np.array_equal(np.ascontiguousarray(a, f32).view(np.uint32), np.ascontiguousarray(b, f32).view(np.uint32))


@pytest.fixture()
def scene(hn, root):
    tb = hn.Testbed()
    tb.reload_network_from_file(os.path.join(root, "configs", "nerf", "base.json"))
    info = tb.network_info()
    grid = hn.synthetic.sphere_density_grid(0.35)
    tb.set_density_grid(grid, 1)
    tb.snap_to_pixel_centers = True
    blob = hn.synthetic.random_params(info, 1337, 1.0)
    tb.set_params(blob)
    rng = np.random.default_rng(77)
    mask = (rng.random((9, 12)) > 0.25).astype(np.uint8)
    dist = (rng.normal(size=(32, 32, 2)) * 0.02).astype(f32)
    return tb, grid, blob, mask, dist


def test_foveation_coefficients_equal_the_reference(scene, orc):
    tb = scene[0]
    sys.path.insert(0, os.path.join(HERE, "golden"))
    from make_golden import FOV_CASES
    g = np.load(os.path.join(HERE, "golden", "ref_host_fov.npz"))
    for tag, fov, _, _ in FOV_CASES:
        c = tb.op_foveation(fov)
        assert bits_equal(c, g[f"{tag}_coeffs"]) and bits_equal(c, orc.foveation_coeffs(fov)), tag
    assert bits_equal(tb.op_foveation((0,) * 6), orc.foveation_coeffs((0,) * 6))


@pytest.mark.parametrize("lens", [0, 1])
def test_init_rays_with_foveation_mask_and_distortion_bit_exact(scene, orc, hn, lens):
    tb, grid, _, mask, dist = scene
    w, h = 160, 120
    cam = hn.synthetic.look_at_camera((1.7, 1.2, 1.9), (0.5, 0.5, 0.5))
    tb.set_option("foveation", list(FOV))
    tb.set_hidden_area_mask(mask)
    tb.set_distortion_map(dist)
    tb.set_option("lens_mode", float(lens)); tb.set_option("lens_params", list(OPENCV))
    tb.snap_to_pixel_centers = False
    o, d, t, alive = tb.op_init_rays(cam, cam, w, h, 3)
    bits, _ = orc.bitfield_from_grid(grid, 1)
    p = orc.render_params(w, h, cam, spp=3, snap=False, rel_focal=tb.get_option("relative_focal_length", 2), foveation=FOV, hidden_area_mask=mask, distortion=dist,
                          lens_mode=lens, lens_params=OPENCV)
    ro, rd, rt, ra = orc.init_rays(p, bits)
    assert np.array_equal(alive, ra) and ra.sum() > 500
    assert bits_equal(o, ro) and bits_equal(d, rd) and bits_equal(t[ra == 1], rt[ra == 1])
    # the three pieces are really on: without them the rays differ and more of them live
    for name in ("foveation", "mask", "dist"):
        if name == "foveation": tb.set_option("foveation", [0] * 6)
        elif name == "mask": tb.set_hidden_area_mask(None)
        else: tb.set_distortion_map(None)
        o2, d2, _, a2 = tb.op_init_rays(cam, cam, w, h, 3)
        assert not np.array_equal(d2, d) or not np.array_equal(a2, alive), name
        tb.set_option("foveation", list(FOV)); tb.set_hidden_area_mask(mask); tb.set_distortion_map(dist)


def test_hybrid_frame_with_foveation_and_mask_vs_oracle(scene, orc, hn):
    tb, grid, blob, mask, dist = scene
    tris = hn.synthetic.torus_mesh(96, 48, R=0.45, r=0.12, center=(0.5, 0.2, 0.5), seed=3)
    tb.add_mesh(tris); tb.build_scene()
    cam = hn.synthetic.look_at_camera((1.5, 1.3, 1.7), (0.5, 0.4, 0.5))
    tb.set_camera_cols(cam)
    tb.set_option("foveation", list(FOV)); tb.set_hidden_area_mask(mask); tb.set_distortion_map(dist)
    w, h = 192, 128
    img, depth = tb.render(w, h, 1, True, return_depth=True)
    ref, rdepth, _ = orc.render_reference_frame(tb, blob, grid, 1, [tris], cam, w, h, foveation=FOV, hidden_area_mask=mask, distortion=dist)
    bad = np.abs(img - ref).max(axis=2) > 1e-2
    print(f"foveated hybrid frame: max {np.abs(img - ref).max():.3e}, PSNR {psnr(img, ref):.1f} dB, pixels > 1e-2: {int(bad.sum())}")
    assert bad.sum() <= 4 and psnr(img, ref) >= 50.0
    tb.set_hidden_area_mask(None)
    open_img = tb.render(w, h, 1, True)
    covered = open_img[..., :3].sum(-1) > 0
    blanked = covered & (img[..., :3].sum(-1) == 0)
    assert covered.mean() > 0.1 and 0.1 < blanked.sum() / covered.sum() < 0.45   # the mask (a quarter of its cells are 0) blanks its share in BOTH passes
    assert tb.frame_stats().n_bvh_stack_overflows == 0


@pytest.mark.parametrize("lens,use_dist", [(0, True), (1, False), (1, True)])
def test_distortion_mode_vs_oracle(scene, orc, hn, lens, use_dist):
    tb, grid, blob, mask, dist = scene
    cam = hn.synthetic.look_at_camera((1.4, 1.2, 1.6), (0.5, 0.5, 0.5))
    tb.set_camera_cols(cam)
    tb.set_option("lens_mode", float(lens)); tb.set_option("lens_params", list(OPENCV))
    tb.set_distortion_map(dist if use_dist else None)
    tb.set_option("foveation", list(FOV))
    tb.set_option("render_mode", 8.0)
    w, h = 160, 120
    img, depth = tb.render(w, h, 1, True, return_depth=True)
    ref, rdepth, _ = orc.render_reference_frame(tb, blob, grid, 1, [], cam, w, h, render_mode=8, lens_mode=lens, lens_params=OPENCV, foveation=FOV,
                                                distortion=dist if use_dist else None)
    # a hue: atan2f of the device vs libm moves it by ~1e-7, visible only where the hue crosses a sector edge at full brightness
    bad = np.abs(img - ref).max(axis=2) > 1e-3
    print(f"Distortion mode lens {lens} grid {use_dist}: max {np.abs(img - ref).max():.3e}, pixels > 1e-3: {int(bad.sum())}, mean brightness {img[..., :3].max(-1).mean():.3f}")
    assert bad.mean() <= 2e-3 and np.array_equal(depth, rdepth) and np.all(depth == 1.0)
    assert img[..., :3].max(-1).mean() > 0.02  # the picture is not black: lens / grid do move pixels


def test_slice_mode_vs_oracle(scene, orc, hn):
    tb, grid, blob, mask, dist = scene
    cam = hn.synthetic.look_at_camera((1.4, 1.2, 1.6), (0.5, 0.5, 0.5))
    tb.set_camera_cols(cam)
    tb.set_option("render_mode", 10.0)
    tb.set_option("aperture_size", 0.05)   # ignored in Slice mode (plane_z < 0 switches it off)
    w, h = 160, 120
    for plane in (1.7, 1.9):
        tb.set_option("focus_z", plane)
        img, depth = tb.render(w, h, 1, True, return_depth=True)
        ref, rdepth, _ = orc.render_reference_frame(tb, blob, grid, 1, [], cam, w, h, render_mode=10, focus_z=plane, aperture_size=0.05)
        print(f"Slice at z = {plane}: max {np.abs(img - ref).max():.3e}, PSNR {psnr(img, ref):.1f} dB, mean alpha {ref[..., 0].mean():.4f}")
        # alpha = 1 - exp(-0.01 sigma) and the sRGB -> linear step make a faint picture (~5e-4): the frame tolerance of 1e-2 would be
        # vacuous here, so the bound is 1 % of the picture's own maximum
        assert ref[..., :3].max() > 3e-4 and np.abs(img - ref).max() <= 1e-2 * ref[..., :3].max() and psnr(img, ref) >= 50.0
        assert np.array_equal(depth, rdepth) and np.all(depth == f32(plane))
    st = tb.frame_stats()
    assert st.n_samples == w * h


def test_foveated_blit(scene, orc, hn):
    tb, grid, blob, mask, dist = scene
    cam = hn.synthetic.look_at_camera((1.4, 1.2, 1.6), (0.5, 0.5, 0.5))
    tb.set_camera_cols(cam)
    w, h = 192, 144
    plain = tb.render(w, h, 1, True)
    tb.set_option("foveation", list(FOV))
    warped, wdepth = tb.render(w, h, 1, True, return_depth=True)
    out, odepth = tb.unwarp_foveated(warped, FOV, depth=wdepth)
    # the foveated frame is a different (warped) picture and the blit brings it back: compare the silhouettes (the random-weight NeRF is
    # noise-like inside, and the blit is a resampling, so pixel values are not expected to agree)
    sil = lambda im: im[..., :3].sum(-1) > 1e-3
    iou = lambda a, b: (a & b).sum() / max((a | b).sum(), 1)
    print(f"silhouette IoU with the plain render: warped {iou(sil(warped), sil(plain)):.3f}, unwarped {iou(sil(out), sil(plain)):.3f}")
    assert iou(sil(out), sil(plain)) > max(iou(sil(warped), sil(plain)), 0.85)
    # the kernel against a numpy restatement of the shader on the same warped frame
    ys, xs = np.meshgrid(np.arange(h), np.arange(w), indexing="ij")
    uv = np.stack([(xs.reshape(-1) + 0.5) / w, (ys.reshape(-1) + 0.5) / h], 1).astype(f32)
    _, un, _ = orc.foveation_warp(FOV, uv)
    sx, sy = un[:, 0].astype(np.float64) * w - 0.5, un[:, 1].astype(np.float64) * h - 0.5
    x0, y0 = np.floor(sx).astype(int), np.floor(sy).astype(int)
    wx, wy = (sx - x0)[:, None], (sy - y0)[:, None]
    cl = lambda v, n: np.clip(v, 0, n - 1)
    W = warped.astype(np.float64)
    want = ((1 - wx) * (1 - wy) * W[cl(y0, h), cl(x0, w)] + wx * (1 - wy) * W[cl(y0, h), cl(x0 + 1, w)]
            + (1 - wx) * wy * W[cl(y0 + 1, h), cl(x0, w)] + wx * wy * W[cl(y0 + 1, h), cl(x0 + 1, w)]).reshape(h, w, 4)
    assert np.abs(out - want).max() <= 2e-4   # float32 texel weights at |coordinate| ~ 200
    # another output size
    small = tb.unwarp_foveated(warped, FOV, out_res=(96, 72))
    assert small.shape == (72, 96, 4) and psnr(small, plain[::2, ::2]) > 15.0
