"""CPU (-m "not gpu"): pins the DEVICE-ONLY half of the oracle against the reference's own CUDA build.

tests/golden/ref_gpu_*.npz hold outputs of the unmodified reference (ngp::Testbed + tiny-cuda-nn compiled for sm_100a,
oracle/ref_gpu.mk) run on a B200 by tests/golden/make_ref_gpu_golden.py: hash-grid encoding (kernel_grid), SH (kernel_sh),
NerfNetwork::inference_mixed_precision, occupancy bits (grid_to_bitfield / bitfield_max_pool), mesh_raytrace_kernel and
whole frames through Testbed::render_frame (NeRF-only and hybrid). The oracle must reproduce them:

  bit-exact   hash-grid encoding, occupancy bitfield
  <= 1 ulp    SH coefficients (fp16; the reference contracts FMAs under --use_fast_math)
  tolerance   raw network output |oracle - ref| <= 1e-3 + 1e-2 |ref| (the reference accumulates in fp16 inside wmma,
              fully_fused_mlp.cu:68,333,436; the oracle and the product in fp32) - both differences are printed
  frames      per-pixel max abs <= 1e-2 and PSNR >= 50 dB (north_star); hybrid frames: the same away from mesh silhouettes
              and shadow edges (the reference traces with fast-math rays, hit/miss flips there are its own rounding)
"""
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")
f32 = np.float32


def ulp16(a, b):
    def key(x):
        u = np.ascontiguousarray(x, np.float16).view(np.uint16).astype(np.int32)
        return np.where(u & 0x8000, -(u & 0x7FFF), u)
    return np.abs(key(a) - key(b))


def psnr(a, b):
    mse = float(np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2))
    return 99.0 if mse == 0 else 10.0 * np.log10(1.0 / mse)


@pytest.fixture(scope="module")
def synthetic():
    import importlib.util
    root = os.path.dirname(HERE)
    spec = importlib.util.spec_from_file_location("hn_synthetic", os.path.join(root, "surface-irradiance-estimation-from-neural-radiance-fields_b200", "synthetic.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


CONFIGS = {
    # tag: (L, F, log2_T, base_res, density_hidden, rgb_hidden, blob feature range)
    "nerf_init": (8, 4, 19, 16, 1, 2, 1e-4),
    "nerf_contrast": (8, 4, 19, 16, 1, 2, 1.0),
    "geom_contrast": (16, 2, 19, 16, 2, 2, 1.0),
}


def oracle_net(orc, synthetic, tag, gold):
    L, F, T, base, dh, rh, rng_range = CONFIGS[tag]
    # per_level_scale as Testbed::reset_network derives it (float exp/log, src/testbed.cu:3958-3966)
    pls = float(np.exp(np.log(f32(2048.0) / f32(base)) / f32(L - 1)))
    cfg = orc.net_config(L, F, T, base, pls, dh, rh)
    if "level_scales" in gold.files:  # the scales kernel_grid computed on the device (MUFU.EX2 + FFMA), stored with the vectors
        cfg = orc.with_device_scales(cfg, gold["level_scales"])

    class Info:
        n_levels, n_features_per_level, density_hidden_layers, rgb_hidden_layers = L, F, dh, rh
        n_params = orc.lib.orc_n_params(orc.C.byref(cfg))
        grid_param_offset = orc.lib.orc_grid_param_offset(orc.C.byref(cfg))

    return cfg, synthetic.random_params(Info, 1337, rng_range)


@pytest.mark.parametrize("tag", list(CONFIGS))
def test_hash_grid_bit_exact_vs_reference_kernel_grid(orc, synthetic, tag):
    g = np.load(os.path.join(GOLD, f"ref_gpu_{tag}.npz"))
    if tag == "geom_contrast" and "level_scales" not in g.files:
        pytest.skip("golden predates the device scale table")
    cfg, params = oracle_net(orc, synthetic, tag, g)
    enc = orc.encode(cfg, params, g["pos"])
    assert np.array_equal(enc.view(np.uint16), g["ref_encode"].view(np.uint16)), f"max ulp {int(ulp16(enc, g['ref_encode']).max())}"


@pytest.mark.parametrize("tag", ["nerf_init"])
def test_sh_within_one_ulp_of_reference_kernel_sh(orc, tag):
    g = np.load(os.path.join(GOLD, f"ref_gpu_{tag}.npz"))
    sh = orc.sh4(g["dir01"])
    d = ulp16(sh, g["ref_sh"])
    assert d.max() <= 1 and (d > 0).mean() < 1e-3


@pytest.mark.parametrize("tag", list(CONFIGS))
def test_network_vs_reference_inference(orc, synthetic, tag):
    g = np.load(os.path.join(GOLD, f"ref_gpu_{tag}.npz"))
    if tag == "geom_contrast" and "level_scales" not in g.files:
        pytest.skip("golden predates the device scale table")
    cfg, params = oracle_net(orc, synthetic, tag, g)
    out = orc.network(cfg, params, g["coords"]).astype(f32)
    ref = g["ref_network"].astype(f32)
    d = np.abs(out - ref)
    print(f"{tag}: |oracle - ref_gpu| max {d.max():.3e} (rgb {d[:, :3].max():.3e}, sigma {d[:, 3].max():.3e})")
    assert np.all(d <= 1e-3 + 1e-2 * np.abs(ref))


@pytest.mark.parametrize("tag", list(CONFIGS))
def test_bitfield_bit_exact_vs_reference(orc, synthetic, tag):
    g = np.load(os.path.join(GOLD, f"ref_gpu_{tag}.npz"))
    bits, _ = orc.bitfield_from_grid(synthetic.sphere_density_grid(0.35), 1)
    assert np.array_equal(bits, g["bitfield_packed"])


def display(frame):
    """product / reference epilogue for spp = 1, linear output, background (0,0,0,1): alpha -> 1, clamp"""
    out = frame.copy()
    out[..., 3] = out[..., 3] + (1.0 - out[..., 3]) * 1.0
    return np.clip(out, 0.0, 1.0)


@pytest.mark.parametrize("tag", ["nerf_init", "nerf_contrast"])
def test_nerf_frame_vs_reference_render_frame(orc, synthetic, tag):
    """configs[0]: 256x256 view of nerf/base.json; a stratified quarter of the pixels (every 2nd row and column)."""
    g = np.load(os.path.join(GOLD, f"ref_gpu_{tag}.npz"))
    cfg, params = oracle_net(orc, synthetic, tag, g)
    w, h = (int(v) for v in g["frame_res"])
    mask = np.zeros((h, w), np.uint8)
    mask[::2, ::2] = 1
    bits, _ = orc.bitfield_from_grid(synthetic.sphere_density_grid(0.35), 1)
    p = orc.render_params(w, h, synthetic.default_camera(), pixel_mask=mask)
    frame, depth, steps, _ = orc.render_frame(p, cfg, params, bits, None)
    m = mask.astype(bool)
    ours, ref = display(frame)[m], g["frame_display"].astype(f32)[m]
    err = np.abs(ours - ref).max()
    print(f"{tag}: oracle vs reference frame on {int(m.sum())} pixels: max abs {err:.3e}, PSNR {psnr(ours, ref):.1f} dB")
    assert err <= 1e-2 and psnr(ours, ref) >= 50.0
    # same pixels covered
    hit_ref = g["frame_depth"][m] < 1000
    assert (hit_ref != (depth[m] < 1000)).sum() <= 2


def edge_mask(depth, radius=1):
    """pixels within `radius` of a depth discontinuity (silhouettes) - where a last-bit difference in the ray flips hit/miss"""
    d = np.minimum(depth, 10.0)
    e = np.zeros(d.shape, bool)
    gy = np.abs(np.diff(d, axis=0)) > 0.05
    gx = np.abs(np.diff(d, axis=1)) > 0.05
    e[:-1] |= gy; e[1:] |= gy; e[:, :-1] |= gx; e[:, 1:] |= gx
    for _ in range(radius - 1):
        e2 = e.copy()
        e2[:-1] |= e[1:]; e2[1:] |= e[:-1]; e2[:, :-1] |= e[:, 1:]; e2[:, 1:] |= e[:, :-1]
        e = e2
    return e


@pytest.mark.parametrize("tag", ["hybrid_torus", "hybrid_three"])
def test_hybrid_frame_vs_reference_render_frame(orc, synthetic, tag):
    """Geometry mode: meshes (load_mesh normalisation, GeometryBvh nearest-box rule) + sun shadow rays + NeRF with the per-pixel
    depth test, against Testbed::render_frame of the reference on the same scene files."""
    g = np.load(os.path.join(GOLD, f"ref_gpu_{tag}.npz"))
    cfg, params = oracle_net(orc, synthetic, "nerf_contrast", g)
    w, h = (int(v) for v in g["res"])
    meshes = [orc.normalize_mesh(g[f"mesh_in_{i}"], g["centers"][i]) for i in range(len(g["centers"]))]
    scene = orc.Scene(meshes)
    bb = scene.aabb()
    aabb = np.concatenate([bb[:3] - f32(4.0), bb[3:] + f32(4.0)])
    bits, _ = orc.bitfield_from_grid(synthetic.sphere_density_grid(0.35), 1)
    # the Geometry-mode render box is the mesh box inflated by 4 (load_scene :3180-3189), the NeRF stays in the unit cube
    p = orc.render_params(w, h, g["cam"], mesh_aabb=aabb, render_aabb=aabb, compat=True)
    frame, depth, steps, _ = orc.render_frame(p, cfg, params, bits, scene)
    ours, ref = display(frame), g["frame_display"].astype(f32)
    d = np.abs(ours - ref).max(-1)
    edges = edge_mask(g["frame_depth"]) | edge_mask(depth)
    bad = d > 1e-2
    print(f"{tag}: oracle vs reference hybrid frame: PSNR {psnr(ours, ref):.1f} dB, pixels > 1e-2: {int(bad.sum())} of {w * h} "
          f"({int((bad & ~edges).sum())} away from depth edges), max {d.max():.3e}")
    assert psnr(ours, ref) >= 50.0
    assert (bad & ~edges).sum() <= 0.002 * w * h  # shadow-edge flips; everything else within 1e-2
    assert bad.sum() <= 0.01 * w * h


@pytest.mark.parametrize("tag", ["hybrid_torus", "hybrid_three"])
def test_mesh_trace_vs_reference_mesh_raytrace_kernel(orc, synthetic, tag):
    g = np.load(os.path.join(GOLD, f"ref_gpu_{tag}.npz"))
    meshes = [orc.normalize_mesh(g[f"mesh_in_{i}"], g["centers"][i]) for i in range(len(g["centers"]))]
    scene = orc.Scene(meshes)
    o, d = g["trace_o"], g["trace_d"]
    p = orc.render_params(8, 8, synthetic.default_camera(), compat=True, n_threads=8)
    mesh, tri, t = scene.trace(p, o, d, brute=False)
    pos_ref, dir_ref = g["ref_trace_pos"], g["ref_trace_dir"]
    hit_ref = np.any(dir_ref != d, axis=1)  # the kernel replaces the direction by the face normal on a hit
    hit = tri >= 0
    flips = int((hit != hit_ref).sum())
    both = hit & hit_ref
    t_ref = np.einsum("ij,ij->i", pos_ref - o, d)
    err = np.abs(t_ref - t)[both]
    print(f"{tag}: {int(hit_ref.sum())} reference hits, {flips} hit/miss flips, max |t - t_ref| {err.max():.3e}")
    assert flips <= 0.005 * o.shape[0]
    assert np.all(err <= 1e-5 * np.maximum(1.0, t[both]) + 2e-6)
    # normals of common hits agree
    n = np.zeros_like(d)
    for i in np.nonzero(both)[0][:512]:
        tr = meshes[mesh[i]][tri[i]].reshape(3, 3)
        nn = np.cross(tr[1] - tr[0], tr[2] - tr[0])
        n[i] = nn / np.linalg.norm(nn)
        assert np.abs(n[i] - dir_ref[i]).max() < 1e-4


def test_load_mesh_normalisation_bit_exact_vs_reference(orc):
    """Testbed::load_mesh's normalisation + centre offset: the triangles the reference holds after loading the .obj files are,
    as a set, bit-identical to the oracle's normalisation (GeometryBvh::build_mesh permutes the meshes themselves)."""
    g = np.load(os.path.join(GOLD, "ref_gpu_hybrid_three.npz"))
    key = lambda a: a[np.lexsort(a.T[::-1])]  # noqa: E731
    mine = [key(orc.normalize_mesh(g[f"mesh_in_{i}"], g["centers"][i])) for i in range(3)]
    matched = set()
    for i in range(3):
        t = key(g[f"ref_tris_{i}"])
        hits = [j for j, m in enumerate(mine) if np.array_equal(t.view(np.uint32), m.view(np.uint32))]
        assert len(hits) == 1
        matched.add(hits[0])
    assert matched == {0, 1, 2}
