"""CPU (-m "not gpu"): a-16 irradiance probes in the oracle (oracle/orc_probe.inc).

The pieces the reference ships (halton<base>, cylindrical_to_dir_nerf, the three probe ray set-ups) are checked against known
values here and against the reference's own CUDA kernels on the GPU box (tests/test_gpu_probe.py::test_live_*). The probe texture and
the mesh shading from it are this project's definition (the reference only declares them): checked against closed forms."""
import os

import numpy as np
import pytest

f32 = np.float32


@pytest.fixture(scope="module")
def synthetic():
    import importlib.util
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("hn_synthetic", os.path.join(root, "surface-irradiance-estimation-from-neural-radiance-fields_b200", "synthetic.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_halton_and_cylindrical_known_values(orc):
    assert [orc.halton(2, i) for i in range(1, 8)] == [0.5, 0.25, 0.75, 0.125, 0.625, 0.375, 0.875]
    assert np.allclose([orc.halton(3, i) for i in range(1, 5)], [1 / 3, 2 / 3, 1 / 9, 4 / 9], atol=1e-7)
    assert np.allclose([orc.halton(5, i) for i in (1, 5, 6)], [0.2, 0.04, 0.24], atol=1e-7)
    assert orc.halton(2, 0) == 0.0
    # p = (0, .) is the +z pole, p = (1, .) the -z pole, p = (0.5, 0.5) the +x axis (phi = 0)
    assert np.allclose(orc.cylindrical_to_dir([0.0, 0.3]), [0, 0, 1], atol=1e-6)
    assert np.allclose(orc.cylindrical_to_dir([1.0, 0.3]), [0, 0, -1], atol=1e-6)
    assert np.allclose(orc.cylindrical_to_dir([0.5, 0.5]), [1, 0, 0], atol=1e-6)
    assert np.allclose(orc.cylindrical_to_dir([0.5, 0.75]), [0, 1, 0], atol=1e-6)


def test_probe_ray_setups(orc):
    nt, nph = 12, 20
    o, d, texel = orc.probe_rays(0, nt, nph)
    assert np.array_equal(texel, np.arange(nt * nph)) and np.allclose(o, 0.5)
    assert np.allclose(np.linalg.norm(d, axis=1), 1.0, atol=1e-6)
    # ray idx = theta + n_theta * phi: cos(theta) = 1 - 2 theta / n_theta
    assert np.allclose(d[:, 2].reshape(nph, nt)[3], 1.0 - 2.0 * np.arange(nt) / nt, atol=1e-6)
    # read_probe is the inverse of the lattice (nudged to the texel centre)
    tex = np.zeros((nph, nt, 4), f32)
    tex[..., 0] = np.arange(nt * nph, dtype=f32).reshape(nph, nt)
    oc, dc, _ = orc.probe_rays(0, nt * 2, nph * 2)                    # the finer lattice's odd nodes are the coarse texel centres
    centres = dc.reshape(nph * 2, nt * 2, 3)[1::2, 1::2].reshape(-1, 3)
    assert np.array_equal(orc.read_probe(tex, centres)[:, 0], np.arange(nt * nph, dtype=f32))
    # mode 1: from a point outside looking in - the pole of the lattice points back at the origin's direction
    org = np.array([2.0, 0.5, 0.5], f32)
    o1, d1, _ = orc.probe_rays(1, nt, nph, origin=org)
    assert np.allclose(o1, org) and np.allclose(d1[0], -org / np.linalg.norm(org), atol=1e-6)
    # mode 2: n_origin^2 rays per texel, origins = centre + halton(2,3,5) - 0.5, same direction as the texel's mode-0 ray
    no = 3
    o2, d2, t2 = orc.probe_rays(2, nt, nph, no)
    assert o2.shape[0] == nt * nph * no * no and np.array_equal(np.bincount(t2), np.full(nt * nph, no * no))
    assert np.allclose(d2, d[t2], atol=1e-6)
    offs = np.unique(np.round(o2 - 0.5, 6), axis=0)
    want = np.array([[orc.halton(2, i) - 0.5, orc.halton(3, i) - 0.5, orc.halton(5, i) - 0.5] for i in range(no * no)])
    assert offs.shape[0] == no * no and np.allclose(np.sort(offs, 0), np.sort(np.round(want, 6), 0), atol=1e-6)


def test_probe_texture_of_a_uniform_medium(orc, synthetic):
    """hash features ~ 1e-4 -> sigma = exp(~0) ~ 1, colour = logistic(~0) ~ 0.5 everywhere inside the r = 0.35 occupancy sphere: every
    ray from the centre sees alpha = 1 - exp(-0.35) and the premultiplied linear colour srgb_to_linear(0.5 alpha)"""
    cfg = orc.net_config(8, 4, 19, 16, 2.0, 1, 2)

    class Info:
        n_levels, n_features_per_level, density_hidden_layers, rgb_hidden_layers = 8, 4, 1, 2
        n_params = orc.lib.orc_n_params(orc.C.byref(cfg))
        grid_param_offset = orc.lib.orc_grid_param_offset(orc.C.byref(cfg))

    params = synthetic.random_params(Info, 1337, 1e-4)
    bits, _ = orc.bitfield_from_grid(synthetic.sphere_density_grid(0.35), 1)
    p = orc.render_params(8, 8, synthetic.default_camera())
    tex, n = orc.render_probe(p, cfg, params, bits, 0, 6, 8)
    a = 1.0 - np.exp(-0.35)
    assert np.allclose(tex[..., 3], a, atol=2e-2) and n > 6 * 8 * 150
    lin = ((0.5 * a + 0.055) / 1.055) ** 2.4
    assert np.allclose(tex[..., :3], lin, atol=1e-2)
    # Halton-jittered centres (offsets up to +-0.5: some origins lie outside the occupied sphere) see other chords, averaged per texel
    tex2, n2 = orc.render_probe(p, cfg, params, bits, 2, 4, 4, 2)
    assert tex2.shape == (4, 4, 4) and np.all(tex2[..., 3] > 0.005) and np.all(tex2[..., 3] < 0.6) and n2 > n // 4


def test_mesh_lit_by_a_probe_closed_forms(orc, synthetic):
    """a floor quad under a constant probe (L = 1 from every direction, nothing in the way): cosine sampling returns exactly albedo,
    and restricting the light to the cap the NeRF subtends gives the same number from both samplers (K large)."""
    quad = np.array([[-3, 0.0, -3, 3, 0.0, -3, 3, 0.0, 3], [-3, 0.0, -3, 3, 0.0, 3, -3, 0.0, 3]], f32)
    scene = orc.Scene([quad])
    cam = synthetic.look_at_camera((0.5, 2.0, 0.5001), (0.5, 0.0, 0.5))
    base = dict(mesh_aabb=(-8, -8, -8, 8, 8, 8), n_threads=4, sun_dir=(0, -1, 0))   # sun from below: the direct term is its ambient only
    ones = np.ones((16, 8, 4), f32)
    p0 = orc.render_params(8, 8, cam, **base)
    dark, _, _, _ = orc.render_frame(p0, None, None, None, scene)
    p1 = orc.render_params(8, 8, cam, probe=ones, env_samples=64, env_sampling=0, **base)
    lit, _, _, _ = orc.render_frame(p1, None, None, None, scene)
    add = lit[4, 4, :3] - dark[4, 4, :3]
    brdf_base = np.array(orc.DEFAULT_BRDF[7:10], f32)
    assert np.allclose(add, brdf_base * brdf_base, atol=1e-5)        # albedo = base colour^2, estimator weight 1 per sample
    # probe that is bright only towards the NeRF's bounding sphere as seen from the floor point under it
    nt, nph = 64, 128
    o, d, _ = orc.probe_rays(0, nt, nph)
    hitp = np.array([0.5, 0.0, 0.5], f32)
    to_c = np.array([0.5, 0.5, 0.5], f32) - hitp
    dist, radius = np.linalg.norm(to_c), 0.5 * np.sqrt(3.0)
    assert dist < radius  # this point is inside the bounding sphere: cap sampling must fall back to the cosine sampler
    p2 = orc.render_params(8, 8, cam, probe=ones, env_samples=64, env_sampling=1, **base)
    lit2, _, _, _ = orc.render_frame(p2, None, None, None, scene)
    assert np.allclose(lit2[4, 4, :3], lit[4, 4, :3], atol=1e-6)
    # far floor: the cap sampler applies; a probe = indicator of that cap gives the same irradiance from both samplers
    quad_far = quad.copy(); quad_far[:, 1::3] = -3.0
    scene_far = orc.Scene([quad_far])
    cam_far = synthetic.look_at_camera((0.5, -1.0, 0.5001), (0.5, -3.0, 0.5))
    hit_far = np.array([0.5, -3.0, 0.5], f32)
    axis = (np.array([0.5, 0.5, 0.5], f32) - hit_far); dist = np.linalg.norm(axis); axis /= dist
    cos_max = np.sqrt(1.0 - (radius / dist) ** 2)
    ind = (d @ axis >= cos_max).astype(f32).reshape(nph, nt)
    tex = np.repeat(ind[..., None], 4, axis=2)
    res = []
    for sampling in (0, 1):
        pp = orc.render_params(16, 16, cam_far, probe=tex, env_samples=1024, env_sampling=sampling, **base)
        f, _, _, _ = orc.render_frame(pp, None, None, None, scene_far)
        pd = orc.render_params(16, 16, cam_far, **base)
        f0, _, _, _ = orc.render_frame(pd, None, None, None, scene_far)
        res.append((f[8, 8, :3] - f0[8, 8, :3]) / (brdf_base * brdf_base))
    # closed form for a cap centred on the normal: (1/pi) int cos = sin^2(theta_max)
    want = 1.0 - cos_max ** 2
    assert abs(res[1][0] - want) < 0.03 * want + 0.01 and abs(res[0][0] - want) < 0.15 * want + 0.02
