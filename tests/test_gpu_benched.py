"""-m gpu: parity ON THE BENCHED CONFIGURATION (BASELINE.json configs[1]/[2] at full size), not only on small stand-ins.

* the 1 048 576-triangle torus of bench.py: >= 100 k rays (camera rays of the bench view, sun shadow rays from their hit points,
  axis-aligned rays, rays starting inside the mesh box) through hn_op_trace against the oracle's BVH (itself checked against brute
  force on the small meshes of test_gpu_mesh.py): (mesh, ORIGINAL triangle index, t) bit-exact, and no traversal-stack overflow;
* a stratified subset of the real 1920x1080 hybrid frame (every 16th row and column) against the oracle rendering exactly those
  pixels of the full-size frame: max abs <= 1e-2, PSNR >= 50 dB, per-ray sample counts equal, total samples consistent;
* the same for configs/geometry/base.json and for the aabb_scale = 4 variant (3 cascades, cone angle 1/256: tolerance parity).
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
f32 = np.float32
W, H = 1920, 1080


def psnr(a, b):
    mse = float(np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2))
    return 99.0 if mse == 0 else 10.0 * np.log10(1.0 / mse)


@pytest.fixture(scope="module")
def bench_scene(hn, orc, root):
    tb = hn.Testbed()
    tb.reload_network_from_file(os.path.join(root, "configs", "nerf", "base.json"))
    info = tb.network_info()
    params = hn.synthetic.random_params(info, 1337, 1e-4)
    grid = hn.synthetic.sphere_density_grid(0.35)
    tb.set_params(params)
    tb.set_density_grid(grid, 1)
    mesh = hn.synthetic.torus_mesh(1024, 512)
    tb.add_mesh(mesh, normalize=True)
    tb.build_scene()
    tb.set_option("render_aabb", tb.get_option("mesh_aabb", 6))
    tb.set_option("compat_nearest_aabb_only", 1.0)
    tb.snap_to_pixel_centers = True
    cam = hn.synthetic.default_camera()
    tb.set_camera_cols(cam)
    nmesh = orc.normalize_mesh(mesh)
    scene = orc.Scene([nmesh])
    return tb, info, params, grid, cam, nmesh, scene


def test_trace_one_million_triangles_bit_exact(hn, orc, bench_scene):
    tb, info, params, grid, cam, nmesh, scene = bench_scene
    assert nmesh.shape[0] == 1048576
    assert np.array_equal(tb.scene_aabb(), scene.aabb())
    rng = np.random.default_rng(31)
    o_all, d_all, _, _ = tb.op_init_rays(cam, cam, W, H, 0)
    sel = np.arange(0, W * H, 17)                      # ~122 k camera rays of the bench view
    o_c, d_c = o_all[sel], d_all[sel]
    p = orc.render_params(8, 8, cam, compat=True, n_threads=os.cpu_count() or 8)
    m0, t0, tt0 = scene.trace(p, o_c, d_c, brute=False)
    hit = t0 >= 0
    # sun shadow rays from the hit points (prepare_shadow_rays: origin pushed off the surface)
    sun = f32(0.57735026919)
    pos = (o_c[hit] + tt0[hit, None] * d_c[hit]).astype(f32)
    o_s = (pos + f32(1e-3) * np.array([sun, sun, sun], f32)).astype(f32)[:30000]
    d_s = np.tile(np.array([[sun, sun, sun]], f32), (o_s.shape[0], 1))
    o_a = rng.uniform(0.1, 0.9, (4096, 3)).astype(f32); o_a[:, 2] = 2.5
    d_a = np.tile(np.array([[0, 0, -1]], f32), (4096, 1))
    o_i = rng.uniform(0.3, 0.7, (4096, 3)).astype(f32)   # start inside the mesh box
    d_i = rng.normal(size=(4096, 3)).astype(f32); d_i /= np.linalg.norm(d_i, axis=1, keepdims=True).astype(f32)
    o = np.concatenate([o_c, o_s, o_a, o_i], 0); d = np.concatenate([d_c, d_s, d_a, d_i], 0)
    assert o.shape[0] >= 100000
    mesh_i, tri_i, t = tb.op_trace(o, d)
    om, ot, ott = scene.trace(p, o, d, brute=False)
    assert np.array_equal(tri_i, ot)
    assert np.array_equal(mesh_i, om)
    assert np.array_equal(t.view(np.uint32), ott.view(np.uint32))
    assert (ot >= 0).sum() > 20000 and (ot < 0).sum() > 10000
    assert tb.frame_stats().n_bvh_stack_overflows == 0
    # a brute-force spot check on the big mesh itself (512 rays x 1 Mi triangles)
    k = rng.choice(o.shape[0], 512, replace=False)
    bm, bt, btt = scene.trace(p, o[k], d[k], brute=True)
    assert np.array_equal(bt, tri_i[k]) and np.array_equal(btt.view(np.uint32), t[k].view(np.uint32))


def _subset_check(hn, orc, tb, cfg, params, bits, scene, aabb, cam, step, tol=1e-2, counts_exact=True, **kw):
    import torch
    mask = np.zeros((H, W), np.uint8)
    mask[::step, ::step] = 1
    m = mask.astype(bool)
    p = orc.render_params(W, H, cam, mesh_aabb=aabb if aabb is not None else (0, 0, 0, 1, 1, 1), render_aabb=tb.get_option("render_aabb", 6),
                          compat=True, snap=True, rel_focal=tb.get_option("relative_focal_length", 2), pixel_mask=mask,
                          max_mip=int(tb.get_option("max_cascade")), cone=tb.get_option("cone_angle_constant"), train_aabb=tb.get_option("nerf_aabb", 6), **kw)
    o_frame, o_depth, o_steps, _ = orc.render_frame(p, cfg, params, bits, scene)
    frame = torch.zeros(H, W, 4, device="cuda"); depth = torch.zeros(H, W, device="cuda")
    tb.render_frame(cam, cam, W, H, frame.data_ptr(), depth.data_ptr(), 0, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    g = frame.cpu().numpy()
    err = float(np.abs(g[m] - o_frame[m]).max())
    ps = psnr(g[m], o_frame[m])
    st = tb.frame_stats()
    o_, d_, t_, alive_ = tb.op_init_rays(cam, cam, W, H, 0)
    counts, _ = tb.op_march_all(o_[m.ravel()], d_[m.ravel()], t_[m.ravel()], alive_[m.ravel()])
    print(f"{int(m.sum())} pixels of the {W}x{H} frame: max abs {err:.3e}, PSNR {ps:.1f} dB, samples {st.n_samples} vs oracle subset x{step * step} = {int(o_steps[m].sum()) * step * step}")
    assert err <= tol and ps >= 50.0
    if counts_exact:
        assert np.array_equal(counts, o_steps[m])   # no ray terminates early in these scenes, so marched == composited
    assert abs(int(o_steps[m].sum()) * step * step - int(st.n_samples)) <= 0.02 * st.n_samples
    assert st.n_bvh_stack_overflows == 0
    return st


def test_benched_hybrid_frame_subset_vs_oracle(hn, orc, bench_scene):
    tb, info, params, grid, cam, nmesh, scene = bench_scene
    bb = scene.aabb()
    aabb = np.concatenate([bb[:3] - f32(4.0), bb[3:] + f32(4.0)])
    bits, _ = orc.bitfield_from_grid(grid, 1)
    st = _subset_check(hn, orc, tb, orc.net_config_from_info(info), params, bits, scene, aabb, cam, 16)
    assert st.n_samples > 60_000_000


def test_geometry_config_1080p_subset_vs_oracle(hn, orc, root):
    tb = hn.Testbed()
    tb.reload_network_from_file(os.path.join(root, "configs", "geometry", "base.json"))
    info = tb.network_info()
    params = hn.synthetic.random_params(info, 1337, 1.0)
    grid = hn.synthetic.sphere_density_grid(0.35)
    tb.set_params(params); tb.set_density_grid(grid, 1); tb.snap_to_pixel_centers = True
    cam = hn.synthetic.default_camera()
    tb.set_camera_cols(cam)
    bits, _ = orc.bitfield_from_grid(grid, 1)
    _subset_check(hn, orc, tb, orc.net_config_from_info(info), params, bits, None, None, cam, 24, counts_exact=False)


def test_large_scene_aabb_scale_4_subset_vs_oracle(hn, orc, root):
    """SURVEY 8(d) config 2 'large' variant: aabb_scale 4 -> 3 cascades, cone angle 1/256 (logf/expf in the stepping: tolerance parity)."""
    tb = hn.Testbed()
    tb.set_aabb_scale(4)
    tb.reload_network_from_file(os.path.join(root, "configs", "nerf", "base.json"))
    info = tb.network_info()
    params = hn.synthetic.random_params(info, 1337, 1.0)
    grid = hn.synthetic.sphere_density_grid(0.6, 3)
    tb.set_params(params); tb.set_density_grid(grid, 3); tb.snap_to_pixel_centers = True
    cam = hn.synthetic.look_at_camera((0.5, 0.9, 3.4), (0.5, 0.5, 0.5))
    tb.set_camera_cols(cam)
    assert int(tb.get_option("max_cascade")) == 2 and abs(tb.get_option("cone_angle_constant") - 1 / 256) < 1e-9
    bits, _ = orc.bitfield_from_grid(grid, 3)
    _subset_check(hn, orc, tb, orc.net_config_from_info(info), params, bits, None, None, cam, 24, counts_exact=False)
