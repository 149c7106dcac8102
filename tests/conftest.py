import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import radnerf_b200  # noqa: E402  (alias loader for the package in rad-nerf_b200/)
from radnerf_b200 import synthetic as syn  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:  # pragma: no cover
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


_scene_cache = {}


def get_scene(scale, seed=1234, jitter=None):
    key = (scale, seed, jitter)
    if key not in _scene_cache:
        sc = syn.make_scene(scale, seed, jitter)
        bf = syn.packbits_np(sc.density_grid, sc.threshold)
        _scene_cache[key] = (sc, bf)
    return _scene_cache[key]


def make_rays(scale, n_rays, seed=42, H=800, W=800, fx=1111.1):
    """(rays_o, rays_d, hits_t (n,2) with the NEAR_DISTANCE clamp of ml_rendering.py:50) via the CPU oracle."""
    from oracle import vren_cpu as O
    cam = syn.make_cameras(100, 1.5 if scale <= 0.5 else 1.0, H, W, fx)
    o, d, _, _ = syn.train_batch(cam, n_rays, seed)
    _, hits_t, _ = O.ray_aabb_intersect(o, d, np.zeros(3, np.float32), np.full(3, scale, np.float32))
    h = np.ascontiguousarray(hits_t[:, 0])
    m = (h[:, 0] >= 0) & (h[:, 0] < syn.NEAR_DISTANCE)
    h[m, 0] = syn.NEAR_DISTANCE
    return o, d, h


CONFIGS = {
    # name: (scale, exp_step_factor)
    "bounded": (0.5, 0.0),
    "unbounded": (16.0, 1.0 / 256),
}
