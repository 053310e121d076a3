"""-m gpu: BVH traversal, mesh pass and the hybrid frame against the oracle (brute-force triangle search = ground truth)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def psnr(a, b):
    mse = float(np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2))
    return 99.0 if mse == 0 else 10.0 * np.log10(1.0 / mse)


def rays_towards(rng, n, centre=(0.5, 0.5, 0.5), spread=0.5, dist=2.5):
    o = rng.normal(size=(n, 3)).astype(np.float32)
    o = o / np.linalg.norm(o, axis=1, keepdims=True) * np.float32(dist) + np.asarray(centre, np.float32)
    tgt = np.asarray(centre, np.float32) + (rng.random((n, 3), dtype=np.float32) - 0.5) * np.float32(2 * spread)
    d = tgt - o
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    return o.astype(np.float32), d


@pytest.mark.parametrize("compat", [False, True])
def test_trace_matches_brute_force(hn, orc, compat):
    rng = np.random.default_rng(21)
    m0 = hn.synthetic.torus_mesh(96, 48)                                    # 9216 triangles
    m1 = hn.synthetic.torus_mesh(40, 20, R=0.2, r=0.08, center=(0.9, 0.6, 0.4), seed=9)
    m2 = np.array([[0, 0, 0.2, 1, 0, 0.2, 0, 1, 0.2], [1, 0, 0.2, 1, 1, 0.2, 0, 1, 0.2],   # two coplanar quads -> exact t ties
                   [0, 0, 0.2, 1, 0, 0.2, 0, 1, 0.2], [1, 0, 0.2, 1, 1, 0.2, 0, 1, 0.2]], np.float32)
    meshes = [m0, m1, m2]
    tb = hn.Testbed()
    for m in meshes:
        tb.add_mesh(m)
    tb.build_scene()
    tb.set_option("compat_nearest_aabb_only", float(compat))
    scene = orc.Scene(meshes)
    assert np.array_equal(tb.scene_aabb(), scene.aabb())
    o, d = rays_towards(rng, 20000)
    # axis-aligned rays (zero direction components) and rays starting inside a mesh box
    o[:64] = [0.3, 0.4, 3.0]; d[:64] = [0, 0, -1]
    o[64:128] = rng.random((64, 3), dtype=np.float32)
    p = orc.render_params(8, 8, hn.synthetic.default_camera(), compat=compat, n_threads=8)
    mesh, tri, t = tb.op_trace(o, d)
    bm, bt, btt = scene.trace(p, o, d, brute=True)
    assert np.array_equal(tri, bt)
    assert np.array_equal(mesh, bm)
    assert np.array_equal(t.view(np.uint32), btt.view(np.uint32))
    assert (bt >= 0).sum() > 5000 and (bt < 0).sum() > 100
    # the oracle's own BVH agrees with its brute force (it is what the 1M-triangle CPU baseline runs)
    om, ot, ott = scene.trace(p, o, d, brute=False)
    assert np.array_equal(ot, bt) and np.array_equal(ott.view(np.uint32), btt.view(np.uint32))


def _parse_obj_triangles(path):
    """a plain OBJ reader for the test (v / f with optional vt / vn indices, polygons fanned): [n][9] float32"""
    verts, tris = [], []
    for line in open(path):
        t = line.split()
        if not t:
            continue
        if t[0] == "v":
            verts.append([float(x) for x in t[1:4]])
        elif t[0] == "f":
            idx = [int(x.split("/")[0]) for x in t[1:]]
            idx = [i - 1 if i > 0 else len(verts) + i for i in idx]
            for k in range(1, len(idx) - 1):
                tris.append(verts[idx[0]] + verts[idx[k]] + verts[idx[k + 1]])
    return np.asarray(tris, np.float32)


@pytest.mark.parametrize("compat", [False, True])
def test_reference_fixture_meshes(hn, orc, root, compat):
    """bunny.obj x3 with the centres of the reference's data/geometry/geometrypaths.json: the product's OBJ loader + load_mesh normalisation +
    BVH against the oracle's brute-force search on triangles parsed independently here. The fixture is read from the reference tree, or from
    oracle/_ref/fixtures/ (`make -C oracle ref` puts it there; git-ignored, it travels to the GPU box with the built checker)."""
    cands = ["/root/reference/data/geometry/objs/bunny.obj", os.path.join(root, "oracle", "_ref", "fixtures", "bunny.obj")]
    path = next((p for p in cands if os.path.exists(p)), None)
    if path is None:
        pytest.skip("bunny.obj neither under /root/reference nor under oracle/_ref/fixtures (make -C oracle ref)")
    centers = [(0.0, 0.0, 0.0), (1.0, 1.0, 1.0), (-1.0, 1.0, 1.0)]
    tb = hn.Testbed()
    for c in centers:
        tb._check(hn.testbed._lib.hn_load_mesh_file(tb._h, path.encode(), hn.testbed._ptr(np.asarray(c, np.float32))))
    tb.build_scene()
    tb.set_option("compat_nearest_aabb_only", float(compat))
    raw = _parse_obj_triangles(path)
    assert raw.shape[0] == 4968
    meshes = [orc.normalize_mesh(raw, c) for c in centers]
    scene = orc.Scene(meshes)
    assert np.array_equal(tb.scene_aabb(), scene.aabb())          # load_mesh normalisation bit-identical
    rng = np.random.default_rng(31)
    o, d = rays_towards(rng, 30000, centre=(0.5, 1.0, 1.0), spread=1.6, dist=4.0)
    o[:64] = [0.5, 0.5, 4.0]; d[:64] = [0, 0, -1]
    p = orc.render_params(8, 8, hn.synthetic.default_camera(), compat=compat, n_threads=8)
    mesh, tri, t = tb.op_trace(o, d)
    bm, bt, btt = scene.trace(p, o, d, brute=True)
    assert np.array_equal(tri, bt) and np.array_equal(mesh, bm) and np.array_equal(t.view(np.uint32), btt.view(np.uint32))
    assert (bt >= 0).sum() > 3000 and len(set(bm[bt >= 0].tolist())) == 3     # all three bunnies are hit
    assert tb.frame_stats().n_bvh_stack_overflows == 0


def hybrid_setup(hn, root, mesh, mesh_aabb, blob_scale=1.0):
    tb = hn.Testbed()
    tb.reload_network_from_file(os.path.join(root, "configs", "nerf", "base.json"))
    info = tb.network_info()
    grid = hn.synthetic.sphere_density_grid(0.35)
    params = hn.synthetic.random_params(info, 1337, blob_scale)
    tb.set_params(params)
    tb.set_density_grid(grid, 1)
    tb.add_mesh(mesh)
    tb.build_scene()
    tb.set_option("mesh_aabb", mesh_aabb)
    tb.snap_to_pixel_centers = True
    return tb, params, grid


@pytest.mark.parametrize("compat,clip", [(False, False), (True, False), (False, True)])
def test_hybrid_frame(hn, orc, root, compat, clip):
    """configs[2] at parity size: NeRF sphere interpenetrating a torus mesh, sun shadow rays, depth-merged blend."""
    w, h = 192, 128
    mesh = hn.synthetic.torus_mesh(128, 64)
    aabb = [-1, -1, -1, 2, 2, 2]
    tb, params, grid = hybrid_setup(hn, root, mesh, aabb)
    tb.set_option("compat_nearest_aabb_only", float(compat))
    tb.set_option("depth_merge_clip", float(clip))
    cam = hn.synthetic.look_at_camera((1.5, 1.3, 2.0), (0.5, 0.5, 0.5))
    tb.set_camera_cols(cam)
    img, depth = tb.render(w, h, 1, True, return_depth=True)
    ref, rdepth, _ = orc.render_reference_frame(tb, params, grid, 1, [mesh], cam, w, h, mesh_aabb=aabb, compat=compat, clip=clip)
    diff = np.abs(img - ref).max(axis=2)
    # a pixel whose NeRF depth is within rounding of the mesh depth can flip the per-pixel depth test: allow <= 0.1 % of them
    bad = diff > 1e-2
    assert bad.mean() <= 1e-3, (bad.sum(), diff.max())
    assert psnr(np.where(bad[..., None], ref, img), ref) >= 50.0
    both = (depth < 16000) & (rdepth < 16000) & ~bad
    assert np.mean(np.abs(depth[both] - rdepth[both]) <= 1e-3 * np.abs(rdepth[both]) + 1e-3) > 0.99
    assert (rdepth < 16000).mean() > 0.05


def test_mesh_only_frame(hn, orc, root):
    w, h = 160, 120
    mesh = hn.synthetic.torus_mesh(64, 32)
    tb = hn.Testbed()
    tb.add_mesh(mesh)
    tb.add_mesh(hn.synthetic.torus_mesh(32, 16, R=0.15, r=0.05, center=(0.2, 0.9, 0.3), seed=2))
    tb.build_scene()
    tb.snap_to_pixel_centers = True
    tb.set_option("floor_enable", 1.0)
    cam = hn.synthetic.look_at_camera((1.9, 1.6, 2.4), (0.5, 0.5, 0.5))
    tb.set_camera_cols(cam)
    img, depth = tb.render(w, h, 1, True, return_depth=True)
    bb = tb.scene_aabb()
    aabb = np.concatenate([bb[:3] - np.float32(4), bb[3:] + np.float32(4)])
    meshes = [mesh, hn.synthetic.torus_mesh(32, 16, R=0.15, r=0.05, center=(0.2, 0.9, 0.3), seed=2)]
    ref, rdepth, _ = orc.render_reference_frame(tb, None, None, 1, meshes, cam, w, h, mesh_aabb=aabb, floor_y=float(aabb[1] + np.float32(0.001)))
    assert np.abs(img - ref).max() <= 1e-3  # same IEEE operation order on both sides; only logf/sqrtf ulp differences remain
    assert np.array_equal(depth < 16000, rdepth < 16000)
    assert np.abs(depth - rdepth)[rdepth < 16000].max() <= 1e-5


def test_mesh_distance_queries(hn, orc):
    """TriangleBvh::unsigned_distance / signed_distance_watertight (src/triangle_bvh.cu:195-297) through hn_op_mesh_distance:
    the BVH-pruned walk against the oracle's brute force over every triangle (per-triangle functions pinned bit-exact on the
    reference's). Distances are the minimum of identical per-triangle values -> bit-exact; the sign of the watertight rule
    must agree everywhere except within 1e-5 of the surface."""
    mesh = hn.synthetic.torus_mesh(96, 48)
    small = hn.synthetic.torus_mesh(32, 16, R=0.15, r=0.05, center=(0.2, 0.9, 0.3), seed=2)
    tb = hn.Testbed()
    tb.add_mesh(mesh)
    tb.add_mesh(small)
    tb.build_scene()
    rng = np.random.default_rng(9)
    pos = rng.uniform(-0.2, 1.2, (4096, 3)).astype(np.float32)
    pos[:256] = mesh.reshape(-1, 3, 3)[rng.integers(0, mesh.reshape(-1, 9).shape[0], 256)].mean(axis=1) + rng.normal(0, 0.01, (256, 3)).astype(np.float32)
    scene = orc.Scene([mesh, small])
    want, _ = scene.distance(pos, 0)
    got = tb.op_mesh_distance(pos, 0)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert 0.0 < want.min() < 0.01 and want.max() > 0.3
    swant, _ = scene.distance(pos, 1)
    sgot = tb.op_mesh_distance(pos, 1)
    assert np.array_equal(np.abs(sgot).view(np.uint32), want.view(np.uint32))
    far = want > 1e-5
    assert np.array_equal(np.sign(sgot[far]), np.sign(swant[far]))
    assert (swant < 0).sum() > 50 and (swant > 0).sum() > 50  # both sides of the surface occur (the sign convention follows the mesh winding)
    # bounded search: nothing within reach -> max_distance
    out = tb.op_mesh_distance(np.array([[30.0, 30.0, 30.0]], np.float32), 0, max_distance=5.0)
    assert out[0] == np.float32(5.0)
    # raystab sign (32 Fibonacci stab rays, pcg32 offset per element): same rays on both sides; a ray grazing an edge can
    # flip one hit/miss through sin/cos rounding -> allow 0.2 % of the points
    rwant, _ = scene.distance(pos, 2)
    rgot = tb.op_mesh_distance(pos, 2)
    assert np.array_equal(np.abs(rgot).view(np.uint32), want.view(np.uint32))
    assert (np.sign(rgot) != np.sign(rwant)).mean() <= 2e-3
    assert (rwant < 0).sum() > 50 and (rwant > 0).sum() > 1000  # closed tori: inside is negative whatever the winding
    with pytest.raises(RuntimeError):
        tb.op_mesh_distance(pos[:4], 3)
