"""Single-field render entry (the K=1 Instant-NGP baseline path of train.py).
Same call convention and result keys as the reference's models/rendering.py:12-46."""
from __future__ import annotations

import torch

from ._render_core import MAX_SAMPLES, NEAR_DISTANCE, intersect_scene_box, render_rays_test, render_rays_train  # noqa: F401


def render(model, rays_o, rays_d, **kwargs):
    """model: NGP.  rays_o, rays_d (N,3).  kwargs: test_time, exp_step_factor, random_bg, T_threshold,
    max_samples, to_cpu, to_numpy.  -> dict with rgb, depth, opacity (+ marching bookkeeping)."""
    rays_o = rays_o.contiguous(); rays_d = rays_d.contiguous()
    with torch.autocast("cuda", enabled=rays_o.is_cuda):   # the reference decorates render with autocast
        hits_t = intersect_scene_box(model, rays_o, rays_d)
        fn = render_rays_test if kwargs.get("test_time", False) else render_rays_train
        results = fn(model, lambda x, d, **kw: model(x, d, **kw), model.density_bitfield, rays_o, rays_d, hits_t, **kwargs)
    for k, v in results.items():
        if kwargs.get("to_cpu", False) and isinstance(v, torch.Tensor):
            v = v.cpu()
            if kwargs.get("to_numpy", False):
                v = v.numpy()
        results[k] = v
    return results
