"""-m gpu: a-21 training step — loss and parameter gradients against the fp32 oracle (loss within 1e-2
relative, gradient cosine >= 0.999), the data-parallel split of a batch, and a short optimisation run on a teacher scene."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
f32 = np.float32


def cosine(a, b):
    a = a.astype(np.float64); b = b.astype(np.float64)
    return float(a @ b / (np.linalg.norm(a) * np.linalg.norm(b) + 1e-300))


def make_rays(rng, n):
    o = rng.normal(size=(n, 3)).astype(f32)
    o = o / np.linalg.norm(o, axis=1, keepdims=True) * f32(1.2) + f32(0.5)
    tgt = f32(0.5) + (rng.random((n, 3), dtype=f32) - f32(0.5)) * f32(0.6)
    d = tgt - o
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(f32)
    t0 = np.zeros(n, f32)
    for i in range(n):
        with np.errstate(divide="ignore"):
            a = (0.0 - o[i]) / d[i]; b = (1.0 - o[i]) / d[i]
        t0[i] = max(np.minimum(a, b).max(), 0.0) + 1e-4
    return o, d, t0


def nerf_setup(hn, root, blob_scale, config="nerf/base.json", grid_value=1.0):
    tb = hn.Testbed()
    tb.reload_network_from_file(os.path.join(root, "configs", *config.split("/")))
    info = tb.network_info()
    grid = hn.synthetic.sphere_density_grid(0.35, value=grid_value)
    params = hn.synthetic.random_params(info, 1337, blob_scale)
    tb.set_params(params)
    tb.set_density_grid(grid, 1)
    tb.snap_to_pixel_centers = True
    return tb, params, grid


@pytest.mark.parametrize("loss_type,config,grid_value", [(1, "nerf/base.json", 1.0), (0, "nerf/base.json", 1.0), (4, "geometry/base.json", 1.0),
                                                         (5, "nerf/base.json", 1.0), (1, "nerf/base.json", 0.02)])
@pytest.mark.parametrize("mlp_impl", [1, 0])
def test_gradients_match_oracle(hn, orc, root, loss_type, config, grid_value, mlp_impl):
    # grid_value 0.02: mean density 0.004 < NERF_MIN_OPTICAL_THICKNESS -> the L1 density regulariser of the reference is active
    # mlp_impl 1 = the tcgen05 training kernels (forward with stored activations, backward dX, split-K dW), 0 = cuBLAS GEMMs
    tb, params, grid = nerf_setup(hn, root, 0.5, config, grid_value)
    tb.set_option("train_mlp_impl", float(mlp_impl))
    rng = np.random.default_rng(5)
    n = 96
    o, d, t0 = make_rays(rng, n)
    targets = rng.random((n, 3), dtype=f32)
    bgs = rng.random((n, 3), dtype=f32)
    loss, grads, steps = tb.op_train_gradients(o, d, t0, targets, bgs, loss_type=loss_type, loss_scale=128.0)
    info = tb.network_info()
    cfg = orc.net_config_from_info(info)
    bits = orc.bitfield_from_grid(grid, 1)[0]
    p = orc.render_params(8, 8, hn.synthetic.default_camera(), density_grid_mean=float(np.maximum(grid, 0).mean()))
    wl, wg, ws = orc.train_gradients(p, cfg, params, bits, o, d, t0, targets, bgs, loss_type=loss_type, loss_scale=128.0)
    assert ws.sum() > 5000
    assert np.mean(steps == ws) >= 0.97            # the T < 1e-4 early stop can flip by a sample
    assert abs(loss - wl) <= 1e-2 * abs(wl)
    n_mlp = int(info.grid_param_offset)
    n_density = n_mlp - (64 * 32 + (info.rgb_hidden_layers - 1) * 64 * 64 + 16 * 64)
    assert cosine(grads[:n_density], wg[:n_density]) >= 0.999
    assert cosine(grads[n_density:n_mlp], wg[n_density:n_mlp]) >= 0.999
    assert cosine(grads[n_mlp:], wg[n_mlp:]) >= 0.999
    # magnitudes agree too, and untouched table entries stay exactly zero on both sides
    assert abs(np.linalg.norm(grads) / np.linalg.norm(wg) - 1.0) <= 2e-2
    assert np.all(grads[n_mlp:][wg[n_mlp:] == 0] == 0)   # (fp16 gradient underflow may zero a few more on the GPU side)
    assert np.mean((grads[n_mlp:] == 0) == (wg[n_mlp:] == 0)) > 0.999


def test_own_mlp_kernels_agree_with_cublas_path(hn, root):
    """same rays, same weights: the tcgen05 forward / backward / dW kernels against the cuBLAS GEMM path (both fp16 in, fp32 accumulate)"""
    rng = np.random.default_rng(11)
    n = 200                                       # ~55 k samples (below the entry point's cap of 65 536): ~430 tiles and a ragged last one
    o, d, t0 = make_rays(rng, n)
    targets = rng.random((n, 3), dtype=f32); bgs = rng.random((n, 3), dtype=f32)
    res = {}
    for config in ("nerf/base.json", "geometry/base.json"):
        for impl in (0, 1):
            tb, params, grid = nerf_setup(hn, root, 0.5, config)
            tb.set_option("train_mlp_impl", float(impl))
            res[impl] = tb.op_train_gradients(o, d, t0, targets, bgs, loss_type=1, loss_scale=128.0)
        (l0, g0, s0), (l1, g1, s1) = res[0], res[1]
        assert np.array_equal(s0, s1) and s0.sum() > 30000
        assert abs(l0 - l1) <= 1e-4 * abs(l0)
        n_mlp = int(tb.network_info().grid_param_offset)
        print(f"{config}: |g_tc - g_cublas| / |g_cublas|: MLP {np.linalg.norm(g1[:n_mlp] - g0[:n_mlp]) / np.linalg.norm(g0[:n_mlp]):.2e}, "
              f"grid {np.linalg.norm(g1[n_mlp:] - g0[n_mlp:]) / np.linalg.norm(g0[n_mlp:]):.2e}")
        assert cosine(g1[:n_mlp], g0[:n_mlp]) >= 0.99999 and cosine(g1[n_mlp:], g0[n_mlp:]) >= 0.9999
        assert abs(np.linalg.norm(g1) / np.linalg.norm(g0) - 1.0) <= 1e-3


def teacher_dataset(hn, root, n_views=12, res=64):
    teacher, params, grid = nerf_setup(hn, root, 1.0)
    rng = np.random.default_rng(9)
    cams, imgs = [], []
    for i in range(n_views):
        eye = rng.normal(size=3)
        eye = eye / np.linalg.norm(eye) * 1.6 + 0.5
        cam = hn.synthetic.look_at_camera(tuple(eye), (0.5, 0.5, 0.5))
        teacher.set_camera_cols(cam)
        imgs.append(teacher.render(res, res, 1, True).copy())
        cams.append(np.asarray(cam, f32).reshape(12))
    focal = float(teacher.get_option("relative_focal_length", 2)[1]) * res
    return np.stack(imgs), np.stack(cams), focal, grid


def test_training_reduces_loss_and_improves_psnr(hn, root):
    imgs, cams, focal = hn.synthetic.sphere_dataset(16, 64)
    tb, _, _ = nerf_setup(hn, root, 1e-4)
    tb.set_training_data(imgs, cams, focal)
    tb.reset_trainer(7)
    target0 = imgs[0][..., :3] * imgs[0][..., 3:4]    # over the render's black background

    def psnr_view0():
        tb.set_camera_cols(cams[0])
        img = tb.render(64, 64, 1, True)
        mse = float(np.mean((img[..., :3] - target0) ** 2))
        return 10 * np.log10(1.0 / max(mse, 1e-12))

    before = psnr_view0()
    first = tb.train(1)
    losses = [tb.train(1) for _ in range(120)]
    after = psnr_view0()
    st = tb.train_stats()
    assert st.step == 121 and st.n_samples_last > 50000 and st.rays_per_batch % 128 == 0
    assert np.isfinite(losses).all()
    assert np.mean(losses[-10:]) < 0.5 * first, (first, losses[-10:])
    assert after > before + 3.0, (before, after)
    # the snapshot writer sees the trained parameters
    path = "/tmp/hn_trained.ingp"
    tb.save_snapshot(path)
    tb2 = hn.Testbed()
    tb2.load_snapshot(path)
    tb2.snap_to_pixel_centers = True
    tb2.set_camera_cols(cams[0]); tb.set_camera_cols(cams[0])
    # the snapshot stores the density grid as fp16 (the reference's format): cells at the occupancy threshold may flip on reload
    diff = np.abs(tb2.render(64, 64, 1, True) - tb.render(64, 64, 1, True))
    assert diff.max() <= 1e-2 and np.mean(diff > 1e-6) < 0.05


def test_data_parallel_split_sums_to_the_full_batch(hn, root):
    """Two ranks draw disjoint halves of the same ray stream; their gradient buffers sum to the one-rank gradient of the
    whole batch (what the NCCL all-reduce in bench.py --workload config5 relies on)."""
    import ctypes as C
    imgs, cams, focal, grid = teacher_dataset(hn, root, n_views=6, res=48)

    def grads_of(rank, world, n_rays):
        tb, _, _ = nerf_setup(hn, root, 0.3)
        tb.set_training_data(imgs, cams, focal)
        tb.set_tile_shard(rank, world)
        tb.reset_trainer(11)
        loss = tb.train_accumulate(n_rays)
        ptr, n = tb.train_gradient_buffer()
        out = np.empty(n, f32)
        rc = hn.testbed.cudart_memcpy_d2h(out, ptr)
        assert rc == 0
        return loss, out

    l0, g0 = grads_of(0, 2, 1024)
    l1, g1 = grads_of(1, 2, 1024)
    lf, gf = grads_of(0, 1, 2048)
    assert abs((l0 + l1) - lf) <= 1e-4 * abs(lf) + 1e-7
    assert cosine(g0 + g1, gf) >= 0.99999
    assert np.abs((g0 + g1) - gf).max() <= 1e-3 * np.abs(gf).max()
    assert cosine(g0, g1) < 0.999   # the halves really are different rays


def test_density_grid_update_matches_oracle(hn, orc, root):
    """update_density_grid_nerf — same cells, same jitter (pcg32), density network, max-splat, decayed maximum."""
    tb, params, grid = nerf_setup(hn, root, 0.5)
    n = 128 ** 3
    cfg = orc.net_config_from_info(tb.network_info())
    p = orc.render_params(8, 8, hn.synthetic.default_camera())
    start = (grid * np.float32(0.02)).astype(f32)          # occupied sphere (0.02 > threshold) so the non-uniform half finds cells
    tb.set_density_grid(start, 1)
    tb.update_density_grid(0.95, n // 4, n // 4)
    got = tb.get_density_grid(1)
    want = orc.update_density_grid(p, cfg, params, start, 1, 0, n // 4, n // 4, 0.95, 1337, 0)
    touched = want != start * np.float32(0.95)
    assert touched.mean() > 0.2
    assert np.array_equal(got == start * np.float32(0.95), ~touched) or np.mean((got != start * np.float32(0.95)) == touched) > 0.9999
    assert np.allclose(got, want, rtol=2e-2, atol=1e-6), np.abs(got - want).max()
    # second round continues the rng stream and the ema step
    tb.update_density_grid(0.95, n // 4, 0)
    got2 = tb.get_density_grid(1)
    want2 = orc.update_density_grid(p, cfg, params, want, 1, 1, n // 4, 0, 0.95, 1337, 2)
    assert np.allclose(got2, want2, rtol=2e-2, atol=1e-6)
    # the bitfield follows the grid (a-20 path, bit-exact given the same grid): compare through the oracle on the GPU's own grid
    bits = tb.get_occupancy_bitfield()
    wbits, _ = orc.bitfield_from_grid(got2, 1)
    assert np.array_equal(bits, wbits)


def test_mark_untrained_matches_oracle(hn, orc, root):
    imgs, _, focal = hn.synthetic.sphere_dataset(2, 32)
    # two cameras INSIDE the unit cube: cells behind them and outside their frusta are never seen
    cams = np.stack([np.asarray(hn.synthetic.look_at_camera((0.5, 0.5, 0.95), (0.5, 0.5, 0.5)), f32).reshape(12),
                     np.asarray(hn.synthetic.look_at_camera((0.1, 0.6, 0.4), (0.6, 0.5, 0.5)), f32).reshape(12)])
    tb, _, _ = nerf_setup(hn, root, 0.5)
    tb.set_training_data(imgs, cams, focal)
    tb.mark_untrained_density_grid()
    got = tb.get_density_grid(1)
    want = orc.mark_untrained_grid(1, cams, 32, 32, focal)
    assert set(np.unique(got)) <= {-1.0, 0.0}
    assert np.mean(got == want) > 0.9995
    assert 0.01 < np.mean(want < 0) < 0.99


def test_snapshot_with_optimizer_state_round_trip(hn, root):
    """save_snapshot(path, include_optimizer_state=True): Ema > ExponentialDecay > Adam nesting of tcnn's serialisers; a fresh
    Testbed resumes from it (step counter, Adam moments) and keeps converging."""
    import msgpack
    import zlib
    imgs, cams, focal = hn.synthetic.sphere_dataset(8, 48)
    tb, _, _ = nerf_setup(hn, root, 1e-4)
    tb.set_training_data(imgs, cams, focal)
    tb.reset_trainer(3)
    for _ in range(40):
        last = tb.train(1)
    path = "/tmp/hn_opt_state.ingp"
    tb.save_snapshot(path, include_optimizer_state=True)   # the pyngp signature (python_api.cu:437)
    import gzip
    blob = msgpack.unpackb(gzip.decompress(open(path, "rb").read()), raw=False, strict_map_key=False)
    snap = blob["snapshot"]
    n = int(tb.network_info().n_params)
    assert snap["training_step"] == 40 and abs(snap["loss"] - last) < 1e-6
    opt = snap["optimizer"]
    assert len(opt["weights_ema_binary"]) == 2 * n                          # Ema
    assert "learning_rate" in opt["nested"]                                 # ExponentialDecay
    adam = opt["nested"]["nested"]
    assert adam["current_step"] == 40 and len(adam["first_moments_binary"]) == 4 * n and len(adam["param_steps_binary"]) == 4 * n
    m2 = np.frombuffer(adam["second_moments_binary"], np.float32)
    assert np.isfinite(m2).all() and (m2 > 0).sum() > 10000
    tb2 = hn.Testbed()
    tb2.load_snapshot(path)
    tb2.set_training_data(imgs, cams, focal)
    assert tb2.training_step == 40
    resumed = [tb2.train(1) for _ in range(10)]
    assert tb2.training_step == 50 and np.isfinite(resumed).all()
    assert np.mean(resumed) < 2.0 * last + 1e-3


def _write_png(path, rgba8, filter_mode):
    """Minimal PNG writer (RGBA8, non-interlaced); filter_mode None = per-row cycling through the five filter types."""
    import struct
    import zlib
    h, w, _ = rgba8.shape
    rows = rgba8.astype(np.int32).reshape(h, w * 4)
    raw = bytearray()
    prev = np.zeros(w * 4, np.int32)
    for y in range(h):
        cur = rows[y]
        ft = (y % 5) if filter_mode is None else filter_mode
        left = np.concatenate([np.zeros(4, np.int32), cur[:-4]])
        upleft = np.concatenate([np.zeros(4, np.int32), prev[:-4]])
        if ft == 0: pred = np.zeros_like(cur)
        elif ft == 1: pred = left
        elif ft == 2: pred = prev
        elif ft == 3: pred = (left + prev) // 2
        else:
            p = left + prev - upleft
            pa, pb, pc = np.abs(p - left), np.abs(p - prev), np.abs(p - upleft)
            pred = np.where((pa <= pb) & (pa <= pc), left, np.where(pb <= pc, prev, upleft))
        raw.append(ft)
        raw += ((cur - pred) % 256).astype(np.uint8).tobytes()
        prev = cur

    def chunk(tag, data):
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xffffffff)
    comp = zlib.compress(bytes(raw), 6)
    half = len(comp) // 2   # two IDAT chunks: the reader has to concatenate them
    open(path, "wb").write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 6, 0, 0, 0)) + chunk(b"IDAT", comp[:half]) +
                           chunk(b"IDAT", comp[half:]) + chunk(b"IEND", b""))


def test_load_training_data_transforms_json_png(hn, root, tmp_path):
    """transforms.json + PNG frames -> the same training set as the in-memory path (cameras through
    nerf_matrix_to_ngp, focal from camera_angle_x, sRGB bytes with straight alpha), checked through the first step's loss."""
    import json
    n, res = 6, 40
    imgs, cams, focal = hn.synthetic.sphere_dataset(n, res)
    srgb = np.where(imgs[..., :3] < 0.0031308, 12.92 * imgs[..., :3], 1.055 * np.power(np.maximum(imgs[..., :3], 1e-12), 1 / 2.4) - 0.055)
    rgba8 = np.concatenate([np.round(np.clip(srgb, 0, 1) * 255), np.round(imgs[..., 3:4] * 255)], -1).astype(np.uint8)
    frames = []
    for i in range(n):
        _write_png(str(tmp_path / f"r_{i}.png"), rgba8[i], None if i % 2 == 0 else i % 5)
        c = cams[i].reshape(4, 3).astype(np.float64)                   # ngp columns {right, down, forward, pos}
        col = np.stack([c[:, 2], c[:, 0], c[:, 1]], 1)                  # undo the axis cycle: old[c][0] = new[c][2], old[c][1] = new[c][0], old[c][2] = new[c][1]
        col[1] *= -1; col[2] *= -1
        col[3] = (col[3] - 0.5) / 0.33
        m = np.eye(4); m[:3, :4] = col.T
        frames.append({"file_path": f"r_{i}", "transform_matrix": m.tolist()})
    meta = {"camera_angle_x": float(2 * np.arctan(0.5 * res / focal)), "aabb_scale": 1, "frames": frames}
    (tmp_path / "transforms.json").write_text(json.dumps(meta))

    def first_loss(loader):
        tb, _, _ = nerf_setup(hn, root, 0.3)
        tb.set_option("train_auto_grid", 0.0)
        loader(tb)
        tb.reset_trainer(5)
        return tb.train_accumulate(1024), tb

    # the byte path, and the float path fed with the very same quantised pixels (decoded like read_rgba does)
    l_png, tb_png = first_loss(lambda tb: tb.load_training_data(str(tmp_path)))
    lin = np.where(rgba8[..., :3] / 255.0 <= 0.04045, rgba8[..., :3] / 255.0 / 12.92, np.power((rgba8[..., :3] / 255.0 + 0.055) / 1.055, 2.4))
    q = np.concatenate([lin, rgba8[..., 3:4] / 255.0], -1).astype(f32)

    def float_loader(tb):
        tb.set_training_data(q, cams, focal)
    l_f32, tb_f = first_loss(float_loader)
    st = tb_png.train_stats()
    assert st.n_samples_last == tb_f.train_stats().n_samples_last and st.n_samples_last > 10000   # same cameras -> same rays
    # alpha is 0 or 1 in this set, so blending in sRGB (byte path) or in linear space (float path) gives the same targets
    assert abs(l_png - l_f32) <= 2e-5 * max(abs(l_f32), 1e-6) + 1e-7, (l_png, l_f32)
    with pytest.raises(RuntimeError):
        tb_png.load_training_data(str(tmp_path / "missing"))


def test_per_image_cameras_mark_untrained_and_rays(hn, orc, root):
    """TrainingImageMetadata per image (hn_train_set_image_metadata): mark_untrained_density_grid projects every cell corner with the image's
    own focal length, principal point and lens (src/testbed_nerf.cu:111-139), the training rays go through it (:806-808)."""
    imgs, _, focal = hn.synthetic.sphere_dataset(3, 32)
    cams = np.stack([np.asarray(hn.synthetic.look_at_camera((0.5, 0.5, 0.95), (0.5, 0.5, 0.5)), f32).reshape(12),
                     np.asarray(hn.synthetic.look_at_camera((0.1, 0.6, 0.4), (0.6, 0.5, 0.5)), f32).reshape(12),
                     np.asarray(hn.synthetic.look_at_camera((0.9, 0.2, 0.3), (0.5, 0.5, 0.6)), f32).reshape(12)])
    tb, _, _ = nerf_setup(hn, root, 0.5)
    tb.set_training_data(imgs, cams, focal)
    f = np.array([[focal, focal], [2.5 * focal, 2.0 * focal], [1.5 * focal, 1.5 * focal]], f32)      # a tele view among them
    pp = np.array([[0.5, 0.5], [0.4, 0.55], [0.5, 0.5]], f32)
    lm = np.array([0, 0, 1], np.int32)                                                                 # the third image through an OpenCV lens
    lp = np.zeros((3, 7), f32); lp[2, :4] = [0.08, -0.03, 0.002, -0.001]
    tb.set_image_metadata(f, pp, lm, lp)
    tb.mark_untrained_density_grid()
    got = tb.get_density_grid(1)
    want = orc.mark_untrained_grid_per_image(1, cams, 32, 32, f, pp, lm, lp)
    shared = orc.mark_untrained_grid(1, cams, 32, 32, focal)
    assert np.mean(got == want) > 0.9995 and np.mean(want != shared) > 0.01   # the narrower frusta see fewer cells
    # a lat-long image sees everything
    tb.set_image_metadata(f, pp, np.array([0, 3, 0], np.int32), lp)
    tb.mark_untrained_density_grid()
    assert np.all(tb.get_density_grid(1) == 0.0)
    with pytest.raises(RuntimeError):
        tb.set_image_metadata(f[:2])
    # rays: with all images given the SAME intrinsics explicitly a step equals the shared-pinhole step (same rays, same samples); other cameras change it
    def first_loss(meta):
        t, _, _ = nerf_setup(hn, root, 0.3)
        t.set_option("train_auto_grid", 0.0)
        t.set_training_data(imgs, cams, focal)
        if meta is not None:
            t.set_image_metadata(*meta)
        t.reset_trainer(5)
        return t.train_accumulate(1024), t.train_stats().n_samples_last
    base = first_loss(None)
    same = first_loss((np.full((3, 2), focal, f32), np.full((3, 2), 0.5, f32), np.zeros(3, np.int32), np.zeros((3, 7), f32)))
    tele = first_loss((f, pp, lm, lp))
    # (the loss is a sum of float atomics: equal up to the order of the additions)
    assert same[1] == base[1] and abs(same[0] - base[0]) <= 1e-5 * base[0]
    assert tele[1] != base[1] and abs(tele[0] - base[0]) > 1e-4 * base[0] and np.isfinite(tele[0])


def test_loader_per_frame_intrinsics_and_lens(hn, root, tmp_path):
    """read_focal_length / read_lens per frame over the file-level values (src/nerf_loader.cu:175-271,678-696): the loaded set trains
    exactly like the same images with hn_train_set_image_metadata called by hand."""
    import json
    n, res = 4, 32
    imgs, cams, focal = hn.synthetic.sphere_dataset(n, res)
    rgba8 = np.concatenate([np.round(np.clip(imgs[..., :3], 0, 1) * 255), np.round(imgs[..., 3:4] * 255)], -1).astype(np.uint8)
    frames = []
    for i in range(n):
        _write_png(str(tmp_path / f"r_{i}.png"), rgba8[i], None)
        c = cams[i].reshape(4, 3).astype(np.float64)
        col = np.stack([c[:, 2], c[:, 0], c[:, 1]], 1)
        col[1] *= -1; col[2] *= -1
        col[3] = (col[3] - 0.5) / 0.33
        m = np.eye(4); m[:3, :4] = col.T
        frames.append({"file_path": f"r_{i}.png", "transform_matrix": m.tolist()})
    frames[1].update({"fl_x": 60.0, "fl_y": 55.0, "cx": 14.0, "cy": 18.0, "w": res, "h": res})
    frames[2].update({"k1": 0.05, "k2": -0.01, "p1": 0.001, "p2": 0.002})
    frames[3].update({"is_fisheye": True, "k1": 0.02, "k2": 0.01, "k3": 0.0, "k4": 0.001, "camera_angle_x": 1.1})
    meta = {"fl_x": float(focal), "aabb_scale": 1, "k1": 0.01, "frames": frames}   # a file-level OpenCV lens that frames inherit
    (tmp_path / "transforms.json").write_text(json.dumps(meta))
    lin = np.where(rgba8[..., :3] / 255.0 <= 0.04045, rgba8[..., :3] / 255.0 / 12.92, np.power((rgba8[..., :3] / 255.0 + 0.055) / 1.055, 2.4))
    q = np.concatenate([lin, rgba8[..., 3:4] / 255.0], -1).astype(f32)
    fl3 = 0.5 * res / np.tan(0.5 * 1.1)
    f = np.array([[focal, focal], [60.0, 55.0], [focal, focal], [fl3, fl3]], f32)
    pp = np.array([[0.5, 0.5], [14.0 / res, 18.0 / res], [0.5, 0.5], [0.5, 0.5]], f32)
    lm = np.array([1, 1, 1, 4], np.int32)
    lp = np.zeros((n, 7), f32)
    lp[0, 0] = 0.01; lp[1, 0] = 0.01; lp[2, :4] = [0.05, -0.01, 0.001, 0.002]; lp[3, :4] = [0.02, 0.01, 0.0, 0.001]

    def first_loss(loader):
        tb, _, _ = nerf_setup(hn, root, 0.3)
        tb.set_option("train_auto_grid", 0.0)
        loader(tb)
        tb.reset_trainer(5)
        return tb.train_accumulate(1024), tb.train_stats().n_samples_last

    def by_hand(tb):
        tb.set_training_data(q, cams, focal)
        tb.set_image_metadata(f, pp, lm, lp)
    l_file, n_file = first_loss(lambda tb: tb.load_training_data(str(tmp_path)))
    l_hand, n_hand = first_loss(by_hand)
    # the loader derives frame 3's focal length from camera_angle_x in float32 (through degrees, like the reference), this test in
    # float64: the two sets may differ by a sample on a cell boundary
    assert abs(n_file - n_hand) <= 4 and n_file > 5000
    assert abs(l_file - l_hand) <= 1e-3 * max(abs(l_hand), 1e-6), (l_file, l_hand)
    l_plain, n_plain = first_loss(lambda tb: tb.set_training_data(q, cams, focal))
    assert abs(n_plain - n_file) > 100 and abs(l_plain - l_file) > 10 * abs(l_file - l_hand) + 1e-6      # the per-frame cameras are really in use


def test_load_training_data_jpeg_frames(hn, root, tmp_path):
    """the same set as JPEG files (what real transforms.json sets ship): decoded by hn_jpeg.cpp, trained like the float path fed with
    libjpeg-turbo's decode of the same files (the two decoders agree to +-3 / 255: tests/test_images.py)"""
    import json
    PIL = pytest.importorskip("PIL.Image")
    n, res = 6, 40
    imgs, cams, focal = hn.synthetic.sphere_dataset(n, res)
    srgb = np.where(imgs[..., :3] < 0.0031308, 12.92 * imgs[..., :3], 1.055 * np.power(np.maximum(imgs[..., :3], 1e-12), 1 / 2.4) - 0.055)
    rgb8 = np.round(np.clip(srgb * imgs[..., 3:4], 0, 1) * 255).astype(np.uint8)   # over black: JPEG has no alpha
    frames, decoded = [], []
    for i in range(n):
        p = tmp_path / f"r_{i}.jpg"
        PIL.fromarray(rgb8[i], "RGB").save(str(p), "JPEG", quality=95, subsampling=(0 if i % 2 else 2))
        decoded.append(np.asarray(PIL.open(str(p)).convert("RGB")))
        c = cams[i].reshape(4, 3).astype(np.float64)
        col = np.stack([c[:, 2], c[:, 0], c[:, 1]], 1)
        col[1] *= -1; col[2] *= -1
        col[3] = (col[3] - 0.5) / 0.33
        m = np.eye(4); m[:3, :4] = col.T
        frames.append({"file_path": f"r_{i}.jpg", "transform_matrix": m.tolist()})
    (tmp_path / "transforms.json").write_text(json.dumps({"camera_angle_x": float(2 * np.arctan(0.5 * res / focal)), "aabb_scale": 1, "frames": frames}))

    def first_loss(loader):
        tb, _, _ = nerf_setup(hn, root, 0.3)
        tb.set_option("train_auto_grid", 0.0)
        tb.set_option("train_random_bg", 0.0)
        loader(tb)
        tb.reset_trainer(5)
        return tb.train_accumulate(1024), tb.train_stats().n_samples_last

    d8 = np.stack(decoded).astype(np.float64) / 255.0
    lin = np.where(d8 <= 0.04045, d8 / 12.92, np.power((d8 + 0.055) / 1.055, 2.4))
    q = np.concatenate([lin, np.ones((n, res, res, 1))], -1).astype(f32)
    l_jpg, n_jpg = first_loss(lambda tb: tb.load_training_data(str(tmp_path)))
    l_f32, n_f32 = first_loss(lambda tb: tb.set_training_data(q, cams, focal))
    assert n_jpg == n_f32 and n_jpg > 10000
    assert abs(l_jpg - l_f32) <= 2e-2 * max(abs(l_f32), 1e-6), (l_jpg, l_f32)


@pytest.mark.parametrize("depth_loss_type", [3, 0])
def test_depth_supervision_gradients_match_oracle(hn, orc, root, depth_loss_type):
    """depth supervision of compute_loss_kernel_train_nerf (src/testbed_nerf.cu:1063-1067,1149-1154): the composited ray depth against a per-ray
    target distance, lambda x its gradient added to dL/dsigma. Same bars as the colour gradients (loss 1e-2, cosine >= 0.999)."""
    tb, params, grid = nerf_setup(hn, root, 0.5)
    rng = np.random.default_rng(6)
    n = 96
    o, d, t0 = make_rays(rng, n)
    targets = rng.random((n, 3), dtype=f32)
    bgs = rng.random((n, 3), dtype=f32)
    target_depth = (1.0 + rng.random(n, dtype=f32)).astype(f32)
    target_depth[::7] = -1.0                       # rays without a depth measurement
    lam = 0.5
    loss, grads, steps = tb.op_train_gradients_depth(o, d, t0, targets, bgs, target_depth, lam, depth_loss_type)
    info = tb.network_info()
    cfg = orc.net_config_from_info(info)
    bits = orc.bitfield_from_grid(grid, 1)[0]
    p = orc.render_params(8, 8, hn.synthetic.default_camera(), density_grid_mean=float(np.maximum(grid, 0).mean()))
    wl, wg, ws = orc.train_gradients_depth(p, cfg, params, bits, o, d, t0, targets, bgs, target_depth, lam, depth_loss_type)
    assert np.mean(steps == ws) >= 0.97 and abs(loss - wl) <= 1e-2 * abs(wl)
    n_mlp = int(info.grid_param_offset)
    assert cosine(grads[:n_mlp], wg[:n_mlp]) >= 0.999 and cosine(grads[n_mlp:], wg[n_mlp:]) >= 0.999
    assert abs(np.linalg.norm(grads) / np.linalg.norm(wg) - 1.0) <= 2e-2
    # the depth term really contributes: without it the gradients are others (the colour loss is unchanged)
    l0, g0, _ = tb.op_train_gradients(o, d, t0, targets, bgs)
    assert abs(l0 - loss) <= 1e-5 * abs(l0) and cosine(g0, grads) < 0.995
    # lambda 0 or no targets = the plain step
    l1, g1, _ = tb.op_train_gradients_depth(o, d, t0, targets, bgs, np.full(n, -1.0, f32), lam, depth_loss_type)
    assert cosine(g1, g0) > 0.99999


def test_depth_supervised_training_pulls_the_surface(hn, root):
    """the dataset path: depth images + option depth_supervision_lambda; a few steps with a strong depth term move the rendered depth towards the targets"""
    imgs, cams, focal = hn.synthetic.sphere_dataset(8, 48)
    tb, _, _ = nerf_setup(hn, root, 1e-4)
    tb.set_training_data(imgs, cams, focal)
    inside = imgs[..., 3] > 0.5
    depth = np.where(inside, 1.35, 0.0).astype(f32)      # cameras sit ~1.6 from the centre of an r = 0.25 sphere: its front is ~1.35 away
    with pytest.raises(RuntimeError):
        tb.set_training_depth(depth[:3])
    tb.set_training_depth(depth)
    tb.reset_trainer(4)
    tb.set_option("depth_supervision_lambda", 1.0)
    plain = hn.Testbed()
    plain.reload_network_from_file(os.path.join(root, "configs", "nerf", "base.json"))
    plain.set_params(hn.synthetic.random_params(plain.network_info(), 1337, 1e-4)); plain.set_density_grid(hn.synthetic.sphere_density_grid(0.35), 1)
    plain.set_training_data(imgs, cams, focal); plain.reset_trainer(4)
    la = [tb.train(1) for _ in range(40)]
    lb = [plain.train(1) for _ in range(40)]
    assert np.isfinite(la).all() and la[-1] < la[0]
    # the two runs see the same rays; the depth term changes the parameters
    tb.snap_to_pixel_centers = plain.snap_to_pixel_centers = True
    tb.set_camera_cols(cams[0]); plain.set_camera_cols(cams[0])
    assert np.abs(tb.render(48, 48, 1, True) - plain.render(48, 48, 1, True)).max() > 1e-3
