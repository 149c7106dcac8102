"""Deterministic synthetic scenes / cameras / ray batches of the shapes BASELINE.json names.

numpy only (shared by bench.py, the tests and the oracle); nothing here touches the GPU.
Follows SURVEY.md section 8(d): analytic Gaussian-blob density written into the cascaded
128^3 occupancy grids in morton order (reference layout: networks.py:263-268,
raymarching.cu:222-223), hemisphere cameras, and *unnormalised* ray directions exactly as
the reference's datasets/ray_utils.py:34-35 (get_ray_directions) and :62-68 (get_rays).
"""
from __future__ import annotations

import dataclasses
import math

import numpy as np

GRID_SIZE = 128
MAX_SAMPLES = 1024          # models/rendering.py:8
NEAR_DISTANCE = 0.01        # models/rendering.py:9
DENSITY_THRESHOLD = 0.01 * MAX_SAMPLES / 3 ** 0.5   # train_ml.py:175


def cascades_for_scale(scale: float) -> int:
    """networks.py:258: max(1 + ceil(log2(2*scale)), 1)."""
    return max(1 + int(np.ceil(np.log2(2 * scale))), 1)


def exp_step_factor_for_scale(scale: float) -> float:
    """train_ml.py:101-102: 1/256 for scale > 0.5 else 0."""
    return 1.0 / 256 if scale > 0.5 else 0.0


def expand_bits(v: np.ndarray) -> np.ndarray:
    v = v.astype(np.uint32)
    v = (v * np.uint32(0x00010001)) & np.uint32(0xFF0000FF)
    v = (v * np.uint32(0x00000101)) & np.uint32(0x0F00F00F)
    v = (v * np.uint32(0x00000011)) & np.uint32(0xC30C30C3)
    v = (v * np.uint32(0x00000005)) & np.uint32(0x49249249)
    return v


def morton3d(x, y, z) -> np.ndarray:
    """30-bit morton code (raymarching.cu:44-50)."""
    return expand_bits(x) | (expand_bits(y) << np.uint32(1)) | (expand_bits(z) << np.uint32(2))


@dataclasses.dataclass
class Scene:
    scale: float
    cascades: int
    exp_step_factor: float
    mus: np.ndarray          # (J,3)
    sig: np.ndarray          # (J,)
    amp: np.ndarray          # (J,)
    slab: bool               # ground slab |z+0.3|<0.02 out to radius scale/2
    density_grid: np.ndarray  # (C, G^3) f32, morton order
    threshold: float          # min(mean(grid>0), DENSITY_THRESHOLD) (networks.py:405-406)

    def density(self, xyz: np.ndarray) -> np.ndarray:
        """Analytic sigma(x) at arbitrary points (N,3) -> (N,) f32."""
        xyz = np.asarray(xyz, np.float32)
        out = np.zeros(len(xyz), np.float32)
        for mu, s, a in zip(self.mus, self.sig, self.amp):
            d2 = ((xyz - mu) ** 2).sum(1)
            out += a * np.exp(-d2 / (2 * s * s)) * (d2 < (4 * s) ** 2)
        if self.slab:
            r = np.sqrt(xyz[:, 0] ** 2 + xyz[:, 1] ** 2)
            out += 50.0 * ((np.abs(xyz[:, 2] + 0.3) < 0.02) & (r < self.scale / 2))
        return out

    @staticmethod
    def color(xyz: np.ndarray) -> np.ndarray:
        """Analytic colour 0.5+0.5*sin(8x)."""
        return (0.5 + 0.5 * np.sin(8.0 * np.asarray(xyz, np.float32))).astype(np.float32)


def make_scene(scale: float = 0.5, seed: int = 1234, jitter_seed: int | None = None) -> Scene:
    """Bounded (scale<=0.5) or unbounded analytic scene.  `jitter_seed` perturbs the blob set
    (sub-NeRF i of a Rad-NeRF zoo sees the same scene with ~90% overlap)."""
    rng = np.random.RandomState(seed)
    mus = rng.uniform(-0.3, 0.3, (24, 3)).astype(np.float32)
    sig = rng.uniform(0.02, 0.045, 24).astype(np.float32)
    amp = np.full(24, 200.0, np.float32)
    unbounded = scale > 0.5
    if unbounded:
        dirs = rng.normal(size=(64, 3)); dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
        rad = rng.uniform(1.0, scale, 64)
        far_mu = (dirs * rad[:, None]).astype(np.float32)
        far_sig = (rng.uniform(0.015, 0.05, 64) * np.maximum(rad, 1.0)).astype(np.float32)
        mus = np.concatenate([mus, far_mu]); sig = np.concatenate([sig, far_sig])
        amp = np.concatenate([amp, np.full(64, 200.0, np.float32)])
    if jitter_seed is not None:
        jr = np.random.RandomState(10_000 + jitter_seed)
        mus = mus + (jr.normal(size=mus.shape) * 0.1 * sig[:, None]).astype(np.float32)
        keep = jr.uniform(size=len(mus)) > 0.05
        keep[0] = True
        mus, sig, amp = mus[keep], sig[keep], amp[keep]
    C = cascades_for_scale(scale)
    G = GRID_SIZE
    grid = np.zeros((C, G, G, G), np.float32)   # [c, x, y, z]
    for c in range(C):
        s = min(2.0 ** (c - 1), scale)           # networks.py:387
        # cell centre of integer coordinate k: ((k+0.5)/G*2-1)*s
        ctr = ((np.arange(G, dtype=np.float32) + 0.5) / G * 2 - 1) * s
        cell = 2 * s / G
        for mu, sg, a in zip(mus, sig, amp):
            r = 4 * sg
            lo = np.floor((mu - r + s) / cell).astype(int); hi = np.ceil((mu + r + s) / cell).astype(int)
            lo = np.clip(lo, 0, G); hi = np.clip(hi, 0, G)
            if np.any(hi <= lo):
                continue
            X = ctr[lo[0]:hi[0], None, None] - mu[0]
            Y = ctr[None, lo[1]:hi[1], None] - mu[1]
            Z = ctr[None, None, lo[2]:hi[2]] - mu[2]
            d2 = X * X + Y * Y + Z * Z
            grid[c, lo[0]:hi[0], lo[1]:hi[1], lo[2]:hi[2]] += (a * np.exp(-d2 / (2 * sg * sg)) * (d2 < r * r)).astype(np.float32)
        if unbounded:
            X = ctr[:, None, None]; Y = ctr[None, :, None]; Z = ctr[None, None, :]
            m = (np.abs(Z + 0.3) < max(0.02, cell / 2)) & (np.sqrt(X * X + Y * Y) < scale / 2)
            grid[c] += 50.0 * m
    # morton order
    ii = np.arange(G, dtype=np.uint32)
    mx = morton3d(ii[:, None, None], np.uint32(0), np.uint32(0))
    my = morton3d(np.uint32(0), ii[None, :, None], np.uint32(0))
    mz = morton3d(np.uint32(0), np.uint32(0), ii[None, None, :])
    midx = (mx | my | mz).reshape(-1).astype(np.int64)
    out = np.zeros((C, G ** 3), np.float32)
    for c in range(C):
        out[c, midx] = grid[c].reshape(-1)
    pos = out[out > 0]
    mean = float(pos.mean()) if len(pos) else 0.0
    return Scene(scale, C, exp_step_factor_for_scale(scale), mus, sig, amp, unbounded, out,
                 min(mean, DENSITY_THRESHOLD))


def packbits_np(grid: np.ndarray, thr: float) -> np.ndarray:
    """numpy statement of packbits (raymarching.cu:130-140): bit i of byte n = grid[8n+i] > thr."""
    g = (grid.reshape(-1, 8) > np.float32(thr)).astype(np.uint8)
    return (g << np.arange(8, dtype=np.uint8)).sum(1).astype(np.uint8)


@dataclasses.dataclass
class Cameras:
    H: int
    W: int
    fx: float
    fy: float
    cx: float
    cy: float
    poses: np.ndarray  # (n,3,4) c2w, OpenCV [right, down, front]


def make_cameras(n: int = 100, radius: float = 1.5, H: int = 800, W: int = 800,
                 fx: float = 1111.1, fy: float | None = None, seed: int = 5) -> Cameras:
    rng = np.random.RandomState(seed)
    poses = np.zeros((n, 3, 4), np.float32)
    for i in range(n):
        phi = rng.uniform(0, 2 * math.pi)
        elev = rng.uniform(math.radians(10), math.radians(80))
        t = radius * np.array([math.cos(elev) * math.cos(phi), math.cos(elev) * math.sin(phi), math.sin(elev)])
        front = -t / np.linalg.norm(t)
        up = np.array([0.0, 0.0, 1.0])
        right = np.cross(front, up); right /= np.linalg.norm(right)
        down = np.cross(front, right)
        poses[i, :, 0] = right; poses[i, :, 1] = down; poses[i, :, 2] = front; poses[i, :, 3] = t
    return Cameras(H, W, fx, fy if fy is not None else fx, W / 2.0, H / 2.0, poses)


def pixel_directions(cam: Cameras, pix_idxs: np.ndarray) -> np.ndarray:
    """Camera-space directions of flat pixel indices (ray_utils.py:34-35, pixel centre)."""
    v, u = np.divmod(np.asarray(pix_idxs, np.int64), cam.W)
    return np.stack([(u - cam.cx + 0.5) / cam.fx, (v - cam.cy + 0.5) / cam.fy, np.ones_like(u, np.float64)], -1).astype(np.float32)


def rays_from(cam: Cameras, img_idxs: np.ndarray, pix_idxs: np.ndarray):
    """World-space (rays_o, rays_d) -- rays_d = d_cam @ R^T, NOT normalised (ray_utils.py:62-68)."""
    d = pixel_directions(cam, pix_idxs)
    R = cam.poses[img_idxs, :, :3]
    rays_d = np.einsum('nc,nac->na', d, R).astype(np.float32)
    rays_o = np.ascontiguousarray(cam.poses[img_idxs, :, 3]).astype(np.float32)
    return rays_o, np.ascontiguousarray(rays_d)


def train_batch(cam: Cameras, batch: int = 8192, seed: int = 42):
    """iid (image, pixel) pairs as datasets/base.py:25-27."""
    rng = np.random.RandomState(seed)
    img = rng.choice(len(cam.poses), batch)
    pix = rng.choice(cam.H * cam.W, batch)
    o, d = rays_from(cam, img, pix)
    return o, d, img, pix


def full_image_rays(cam: Cameras, img_idx: int = 0, rows: slice | None = None):
    pix = np.arange(cam.H * cam.W, dtype=np.int64)
    if rows is not None:
        pix = pix.reshape(cam.H, cam.W)[rows].reshape(-1)
    return rays_from(cam, np.full(len(pix), img_idx, np.int64), pix)


# named workload shapes (BASELINE.json configs / SURVEY 8d)
WORKLOADS = {
    # BASELINE.json configs[0]: Instant-NGP volume-render FORWARD (AABB + march + composite), NSVF-style bounded scene, K = 1
    "ngp_nsvf_k1_cpu_fwd": dict(scale=0.5, K=1, H=800, W=800, fx=1111.1, radius=1.5, forward_only=True),
    "ngp_blender_k1": dict(scale=0.5, K=1, H=800, W=800, fx=1111.1, radius=1.5),
    "radnerf_tat_k2": dict(scale=0.5, K=2, H=1080, W=1920, fx=1166.0, radius=0.9),   # radius tuned: ~33 samples/ray
    "radnerf_360_k4": dict(scale=16.0, K=4, H=822, W=1237, fx=1050.0, radius=1.0),
    # test-time frames (BASELINE.json's second metric / configs[4]): one step = one full frame, rows sharded over the ranks
    "radnerf_tat_k8_infer": dict(scale=0.5, K=8, H=1080, W=1920, fx=1166.0, radius=0.9, infer=True),
    "ngp_blender_k1_infer": dict(scale=0.5, K=1, H=800, W=800, fx=1111.1, radius=1.5, infer=True),
}


def analytic_colors(scene: Scene, rays_o, rays_d, background: float = 1.0):
    """Ground-truth colours of `scene` for CUDA rays (B,3): the analytic density / colour integrated along the
    reference's own constant-step marching (dt = sqrt(3)/1024 through the scene's occupancy bitfield) and composited
    by the vren operators, blended on a uniform background.  Used to fit a model against the synthetic scene
    (scripts/fit_scene.py, tests/test_fit_gpu.py); bounded scenes only."""
    import torch
    from . import vren
    assert scene.cascades == 1, "bounded scenes only"
    dev = rays_o.device
    bf = torch.from_numpy(packbits_np(scene.density_grid, scene.threshold)).to(dev)
    center = torch.zeros(1, 3, device=dev); half = torch.full((1, 3), float(scene.scale), device=dev)
    out = torch.empty(rays_o.shape[0], 3, device=dev)
    mus = torch.from_numpy(scene.mus).to(dev); sg = torch.from_numpy(scene.sig).to(dev); amp = torch.from_numpy(scene.amp).to(dev)
    for s in range(0, rays_o.shape[0], 1 << 16):
        o, d = rays_o[s:s + (1 << 16)].contiguous(), rays_d[s:s + (1 << 16)].contiguous()
        _, hits_t, _ = vren.ray_aabb_intersect(o, d, center, half, 1)
        h = hits_t[:, 0].contiguous()
        h[(h[:, 0] >= 0) & (h[:, 0] < NEAR_DISTANCE), 0] = NEAR_DISTANCE
        noise = torch.full((o.shape[0],), 0.5, device=dev)
        rays_a, xyzs, _, deltas, ts, counter = vren.raymarching_train(o, d, h, bf, 1, float(scene.scale), 0.0, noise, GRID_SIZE, 1024)
        n = int(counter[0])
        x = xyzs[:n]
        sigma = torch.zeros(n, device=dev)
        for j in range(mus.shape[0]):
            d2 = ((x - mus[j]) ** 2).sum(1)
            sigma += amp[j] * torch.exp(-d2 / (2 * sg[j] * sg[j])) * (d2 < (4 * sg[j]) ** 2)
        rgb = (0.5 + 0.5 * torch.sin(8.0 * x)).contiguous()
        _, opacity, _, col, _ = vren.composite_train_fw(sigma, rgb, deltas[:n].contiguous(), ts[:n].contiguous(), rays_a, 1e-4)
        out[s:s + o.shape[0]] = col + background * (1 - opacity[:, None])
    return out
