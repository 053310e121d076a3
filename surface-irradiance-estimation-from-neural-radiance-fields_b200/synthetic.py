"""Synthetic inputs of the named shapes: seeded parameter blobs, a 128^3 sphere occupancy grid in Morton
order, the reference's default camera, and a procedural 1 Mi-triangle mesh. numpy only; used by tests/ and bench.py."""
from __future__ import annotations

import math

import numpy as np


def morton_invert(x: np.ndarray) -> np.ndarray:
    x = x & 0x49249249
    x = (x | (x >> 2)) & 0xC30C30C3
    x = (x | (x >> 4)) & 0x0F00F00F
    x = (x | (x >> 8)) & 0xFF0000FF
    x = (x | (x >> 16)) & 0x0000FFFF
    return x


def sphere_density_grid(radius: float = 0.35, n_cascades: int = 1, value: float = 1.0) -> np.ndarray:
    """float32[128^3 * n_cascades], Morton order inside each cascade; cell = value if its centre (in that cascade's box
    [0.5 - 2^(c-1), 0.5 + 2^(c-1)]) lies within `radius` of (0.5, 0.5, 0.5)."""
    n = 128
    idx = np.arange(n ** 3, dtype=np.uint32)
    x, y, z = morton_invert(idx), morton_invert(idx >> 1), morton_invert(idx >> 2)
    c = (np.stack([x, y, z], 1).astype(np.float32) + 0.5) / n
    out = []
    for casc in range(n_cascades):
        scale = float(2 ** casc)
        p = (c - 0.5) * scale + 0.5
        out.append((np.linalg.norm(p - 0.5, axis=1) < radius).astype(np.float32) * value)
    return np.concatenate(out)


def random_params(info, seed: int = 1337, feature_range: float = 1e-4) -> np.ndarray:
    """fp16 blob [density MLP | rgb MLP | hash grid]: MLP weights Xavier-uniform (tcnn gpu_matrix.h:284-296), hash features
    U(-feature_range, feature_range) (tcnn grid.h:1059-1062 uses 1e-4; 1.0 gives the 'contrast' blob)."""
    rng = np.random.default_rng(seed)
    L, F = info.n_levels, info.n_features_per_level

    def xavier(rows, cols):
        s = math.sqrt(6.0 / (rows + cols))
        return rng.uniform(-s, s, size=(rows, cols)).astype(np.float32)

    parts = []
    for in_w, hidden in ((L * F, info.density_hidden_layers), (32, info.rgb_hidden_layers)):
        parts.append(xavier(64, in_w).ravel())
        for _ in range(hidden - 1):
            parts.append(xavier(64, 64).ravel())
        parts.append(xavier(16, 64).ravel())
    n_grid = int(info.n_params - info.grid_param_offset)
    parts.append(rng.uniform(-feature_range, feature_range, size=n_grid).astype(np.float32))
    blob = np.concatenate(parts).astype(np.float16)
    assert blob.size == info.n_params
    return blob


def default_camera() -> np.ndarray:
    """reset_camera default (src/testbed.cu:541-568) as 4 columns x vec3: right, down, forward, position."""
    return np.array([1, 0, 0, 0, -1, 0, 0, 0, -1, 0.5, 0.5, 2.0], dtype=np.float32)


def look_at_camera(eye, target, up=(0.0, 1.0, 0.0)) -> np.ndarray:
    eye, target, up = (np.asarray(v, dtype=np.float64) for v in (eye, target, up))
    fwd = target - eye
    fwd /= np.linalg.norm(fwd)
    right = np.cross(fwd, up)
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    return np.concatenate([right, down, fwd, eye]).astype(np.float32)


DEFAULT_REL_FOCAL = 0.5 / math.tan(0.5 * 50.625 * math.pi / 180.0)  # fov 50.625 deg on axis 1


def torus_mesh(nu: int = 1024, nv: int = 512, R: float = 0.33, r: float = 0.12, center=(0.5, 0.5, 0.5), noise: float = 0.01,
               seed: int = 7, tilt: float = 0.6) -> np.ndarray:
    """Tessellated torus with seeded radial vertex noise: nu*nv*2 triangles (1024 x 512 x 2 = 1 048 576), float32 [n, 9]."""
    rng = np.random.default_rng(seed)
    u = np.linspace(0.0, 2 * np.pi, nu, endpoint=False)
    v = np.linspace(0.0, 2 * np.pi, nv, endpoint=False)
    uu, vv = np.meshgrid(u, v, indexing="ij")
    rr = r * (1.0 + noise * rng.standard_normal(uu.shape))
    x = (R + rr * np.cos(vv)) * np.cos(uu)
    y = rr * np.sin(vv)
    z = (R + rr * np.cos(vv)) * np.sin(uu)
    ct, st = math.cos(tilt), math.sin(tilt)
    y, z = ct * y - st * z, st * y + ct * z
    p = np.stack([x, y, z], -1) + np.asarray(center)
    p00 = p
    p10 = np.roll(p, -1, axis=0)
    p01 = np.roll(p, -1, axis=1)
    p11 = np.roll(p10, -1, axis=1)
    t1 = np.stack([p00, p10, p11], 2)
    t2 = np.stack([p00, p11, p01], 2)
    tris = np.concatenate([t1.reshape(-1, 9), t2.reshape(-1, 9)], 0)
    return np.ascontiguousarray(tris, dtype=np.float32)


def load_obj(path: str) -> np.ndarray:
    """Tiny .obj reader for the reference's fixtures (v / f records) -> float32 [n, 9]."""
    verts, tris = [], []
    with open(path) as f:
        for line in f:
            if line.startswith("v "):
                verts.append([float(t) for t in line.split()[1:4]])
            elif line.startswith("f "):
                idx = [int(t.split("/")[0]) for t in line.split()[1:]]
                idx = [i - 1 if i > 0 else len(verts) + i for i in idx]
                for k in range(2, len(idx)):
                    tris.append(verts[idx[0]] + verts[idx[k - 1]] + verts[idx[k]])
    return np.asarray(tris, dtype=np.float32)


def envmap(width: int = 1024, height: int = 512, seed: int = 3, lobe=(0.3, 0.7), lobe_strength: float = 6.0) -> np.ndarray:
    """Lat-long RGBA32F environment map [height][width][4] (SURVEY config 4): seeded smooth noise (periodic in x) plus one
    bright lobe centred at (u, v) = `lobe`. Alpha = 1."""
    rng = np.random.default_rng(seed)
    u = (np.arange(width, dtype=np.float32) + 0.5) / width
    v = (np.arange(height, dtype=np.float32) + 0.5) / height
    U, V = np.meshgrid(u, v)
    img = np.zeros((height, width, 4), np.float32)
    for c in range(3):
        acc = np.full((height, width), 0.35, np.float32)
        for f in (1, 2, 3, 5):
            ph = rng.uniform(0, 2 * np.pi, 2)
            amp = np.float32(0.12 / f)
            acc += amp * np.sin(2 * np.pi * f * U + ph[0]).astype(np.float32) * np.cos(np.pi * f * V + ph[1]).astype(np.float32)
        img[..., c] = acc
    du = np.minimum(np.abs(U - lobe[0]), 1.0 - np.abs(U - lobe[0]))
    g = np.exp(-((du / 0.05) ** 2 + ((V - lobe[1]) / 0.08) ** 2)).astype(np.float32)
    img[..., :3] += np.float32(lobe_strength) * g[..., None] * np.array([1.0, 0.9, 0.75], np.float32)
    img[..., :3] = np.maximum(img[..., :3], 0.0)
    img[..., 3] = 1.0
    return img


def sphere_dataset(n_views: int = 16, res: int = 64, radius: float = 0.3, seed: int = 9, fov_deg: float = 50.625):
    """Analytic training set (SURVEY config 5 stand-in for a rendered one): views of a patterned sphere centred in the unit
    cube. Returns (images [n][res][res][4] linear RGB + straight alpha, cameras [n][12], focal length in pixels)."""
    rng = np.random.default_rng(seed)
    focal = 0.5 * res / np.tan(0.5 * np.deg2rad(fov_deg))
    imgs = np.zeros((n_views, res, res, 4), np.float32)
    cams = np.zeros((n_views, 12), np.float32)
    c = np.array([0.5, 0.5, 0.5], np.float64)
    ys, xs = np.meshgrid(np.arange(res) + 0.5, np.arange(res) + 0.5, indexing="ij")
    for i in range(n_views):
        eye = rng.normal(size=3)
        eye = eye / np.linalg.norm(eye) * 1.5 + c
        cam = np.asarray(look_at_camera(tuple(eye), tuple(c)), np.float64).reshape(4, 3)
        cams[i] = cam.reshape(12).astype(np.float32)
        d = ((xs / res - 0.5) * res / focal)[..., None] * cam[0] + ((ys / res - 0.5) * res / focal)[..., None] * cam[1] + cam[2]
        d /= np.linalg.norm(d, axis=-1, keepdims=True)
        oc = cam[3] - c
        b = (d * oc).sum(-1)
        disc = b * b - ((oc * oc).sum() - radius * radius)
        hit = disc > 0
        t = -b - np.sqrt(np.maximum(disc, 0))
        p = cam[3] + t[..., None] * d
        n = (p - c) / radius
        shade = 0.35 + 0.65 * np.clip(n @ np.array([0.5, 0.7, 0.5]), 0, 1)
        col = 0.5 + 0.5 * np.sin(9.0 * p + np.array([0.0, 2.0, 4.0]))
        imgs[i, ..., :3] = np.where(hit[..., None], col * shade[..., None], 0.0)
        imgs[i, ..., 3] = hit
    return imgs, cams, float(focal)
