"""CPU (-m "not gpu"): the a-21 training oracle (oracle/orc_train.inc) checks itself — its analytic gradients against central
finite differences of its own loss, and the loss functions against the reference's host-callable ones when oracle/_ref is built."""
import ctypes as C
import importlib.util
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
f32 = np.float32


@pytest.fixture(scope="module")
def synthetic():
    root = os.path.dirname(HERE)
    spec = importlib.util.spec_from_file_location("hn_synthetic", os.path.join(root, "surface-irradiance-estimation-from-neural-radiance-fields_b200", "synthetic.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def make_rays(rng, n):
    o = rng.normal(size=(n, 3)).astype(f32)
    o = o / np.linalg.norm(o, axis=1, keepdims=True) * f32(1.2) + f32(0.5)
    tgt = f32(0.5) + (rng.random((n, 3), dtype=f32) - f32(0.5)) * f32(0.5)
    d = tgt - o
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(f32)
    # start t = entry into the unit cube
    t0 = np.zeros(n, f32)
    for i in range(n):
        with np.errstate(divide="ignore"):
            a = (0.0 - o[i]) / d[i]; b = (1.0 - o[i]) / d[i]
        t0[i] = max(np.minimum(a, b).max(), 0.0) + 1e-4
    return o, d, t0


def setup(orc, synthetic, scale=0.5):
    cfg = orc.net_config(8, 4, 19, 16, 2.0, 1, 2)

    class Info:
        n_levels, n_features_per_level, density_hidden_layers, rgb_hidden_layers = 8, 4, 1, 2
        n_params = orc.lib.orc_n_params(C.byref(cfg))
        grid_param_offset = orc.lib.orc_grid_param_offset(C.byref(cfg))

    params = synthetic.random_params(Info, 1337, scale)
    bits, _ = orc.bitfield_from_grid(synthetic.sphere_density_grid(0.35), 1)
    p = orc.render_params(8, 8, synthetic.default_camera())
    return cfg, Info, params, bits, p


@pytest.mark.parametrize("loss_type", [0, 1])
def test_oracle_gradients_match_finite_differences(orc, synthetic, loss_type):
    cfg, info, params, bits, p = setup(orc, synthetic)
    rng = np.random.default_rng(3)
    o, d, t0 = make_rays(rng, 6)
    targets = rng.random((6, 3), dtype=f32)
    bgs = rng.random((6, 3), dtype=f32)
    loss, grads, steps = orc.train_gradients(p, cfg, params, bits, o, d, t0, targets, bgs, loss_type=loss_type, loss_scale=1.0)
    assert steps.sum() > 300 and np.isfinite(loss) and loss > 0
    n_mlp = int(info.grid_param_offset)
    # parameters with the largest gradients in each region: density MLP, rgb MLP, coarse and fine grid levels
    picks = []
    picks += list(np.argsort(-np.abs(grads[:3072]))[:3])
    picks += list(3072 + np.argsort(-np.abs(grads[3072:n_mlp]))[:3])
    g = grads[n_mlp:]
    picks += list(n_mlp + np.argsort(-np.abs(g))[:4])
    fine = n_mlp + 4 * (4096 + 32768 + 262144 + 524288 * 3)          # inside level 6
    picks += list(fine + np.argsort(-np.abs(grads[fine:fine + 4 * 524288]))[:2])
    for idx in picks:
        w = params.copy()
        base = np.float32(w[idx])
        delta = max(abs(float(base)) * 2 ** -5, 2 ** -9)
        hi, lo = np.float16(base + delta), np.float16(base - delta)
        w[idx] = hi
        lp = orc.train_gradients(p, cfg, w, bits, o, d, t0, targets, bgs, loss_type=loss_type, loss_scale=1.0)[0]
        w[idx] = lo
        lm = orc.train_gradients(p, cfg, w, bits, o, d, t0, targets, bgs, loss_type=loss_type, loss_scale=1.0)[0]
        fd = (lp - lm) / (float(hi) - float(lo))
        # the oracle's "loss" is mean over channels (/3), its gradients follow the reference's convention (sum over channels of dL/drgb): factor 3
        an = float(grads[idx]) / 3.0
        assert abs(fd - an) <= 0.05 * max(abs(an), abs(fd)) + 2e-6, (idx, fd, an)


def test_loss_functions_pinned_against_reference_host(orc):
    ref = orc.ref
    if ref is None or not hasattr(ref, "ref_loss_and_gradient"):
        pytest.skip("oracle/_ref not built here (authoring container only); vectors in tests/golden/ref_host_env.npz cover the CPU suite")
    rng = np.random.default_rng(1)
    for _ in range(200):
        t, pr = rng.random(3, dtype=f32), (rng.random(3, dtype=f32) * 1.4 - 0.2).astype(f32)
        for lt in range(7):
            out = np.zeros(6, f32)
            ref.ref_loss_and_gradient(orc.P(t), orc.P(pr), lt, orc.P(out))
            got = np.zeros(6, f32)
            orc.lib.orc_loss_and_gradient(orc.P(t), orc.P(pr), lt, orc.P(got))
            assert np.array_equal(out.view(np.uint32), got.view(np.uint32))


def test_loss_functions_pinned_against_golden(orc):
    g = np.load(os.path.join(HERE, "golden", "ref_host_env.npz"))
    for lt, tag in ((0, "l2"), (1, "huber"), (2, "relativel2"), (3, "l1"), (4, "mape"), (5, "smape"), (6, "logl1")):
        for i in range(g["loss_target"].shape[0]):
            got = np.zeros(6, f32)
            orc.lib.orc_loss_and_gradient(orc.P(g["loss_target"][i].copy()), orc.P(g["loss_pred"][i].copy()), lt, orc.P(got))
            assert np.array_equal(got.view(np.uint32), g[f"loss_{tag}"][i].view(np.uint32))


def test_pcg32_pinned_against_golden(orc):
    g = np.load(os.path.join(HERE, "golden", "ref_host_env.npz"))
    orc.lib.orc_pcg32.argtypes = [C.c_uint64, C.c_uint64, C.c_uint32, C.c_void_p]
    for a, sd in enumerate(g["pcg_seeds"]):
        for b, ad in enumerate(g["pcg_advances"]):
            out = np.zeros(8, np.uint32)
            orc.lib.orc_pcg32(int(sd), int(ad), 8, orc.P(out))
            assert np.array_equal(out, g["pcg_out"][a, b])


def test_density_input_gradient_finite_differences(orc):
    """Normals mode (NerfNetwork::input_gradient, src/testbed_nerf.cu:2106-2107): the oracle's analytic gradient of the raw
    density w.r.t. the warped position against central differences of an all-fp32 forward pass of the same weights. The
    step stays inside one cell of the finest level (1/2048), away from the cell faces where the interpolant has kinks."""
    import ctypes as C
    from importlib import import_module
    import os, sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    synthetic = import_module("surface-irradiance-estimation-from-neural-radiance-fields_b200.synthetic") if False else None
    import importlib.util
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("hn_synthetic", os.path.join(root, "surface-irradiance-estimation-from-neural-radiance-fields_b200", "synthetic.py"))
    synthetic = importlib.util.module_from_spec(spec); spec.loader.exec_module(synthetic)
    L = orc.lib
    L.orc_density_input_gradient.restype = C.c_float
    L.orc_density_raw_f32.restype = C.c_float
    for layers, nl, nf in ((1, 8, 4), (2, 16, 2)):
        scale = float(np.exp(np.float32(np.log(np.float32(128.0 if nl == 8 else 2048.0 / 16)) / np.float32(nl - 1))))
        cfg = orc.net_config(nl, nf, 19, 16, scale, layers, 2)

        class Info:
            n_levels, n_features_per_level, density_hidden_layers, rgb_hidden_layers = nl, nf, layers, 2
            n_params = L.orc_n_params(C.byref(cfg))
            grid_param_offset = L.orc_grid_param_offset(C.byref(cfg))
        params = synthetic.random_params(Info, 11, 1.0)
        rng = np.random.default_rng(5)
        ok = 0
        for _ in range(40):
            # centre of a finest-level cell: fract(scale_max * p + 0.5) = 0.5 -> all coarser levels are away from a face by >= 2^-12
            cell = rng.integers(100, 1900, 3)
            wp = ((cell.astype(np.float64) + 0.5 - 0.5) / 2047.0 + 0.13 / 2047.0).astype(np.float32)
            g = np.zeros(3, np.float32)
            L.orc_density_input_gradient(C.byref(cfg), orc.P(params), orc.P(wp), orc.P(g))
            h = np.float32(2e-5)
            fd = np.zeros(3, np.float64)
            for d in range(3):
                a = wp.copy(); b = wp.copy(); a[d] += h; b[d] -= h
                fd[d] = (float(L.orc_density_raw_f32(C.byref(cfg), orc.P(params), orc.P(a))) - float(L.orc_density_raw_f32(C.byref(cfg), orc.P(params), orc.P(b)))) / float(a[d] - b[d])
            nrm = np.linalg.norm(fd)
            if nrm < 1e-3:
                continue
            cos = float(np.dot(fd, g) / (nrm * np.linalg.norm(g)))
            if cos > 0.99 and abs(np.linalg.norm(g) / nrm - 1) < 0.1:
                ok += 1
        assert ok >= 24, (layers, ok)  # of 40: central differences of a kinked interpolant, some points straddle a cell face of some level
