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


def moe_render(model_zoo, gating_net, rays_o, rays_d, warmup=False, **kwargs):
    """Gated blend of `model_zoo.size` independent NGPs (NGP_zoo): every member renders every ray through its own
    scene box / occupancy grid, rgb = sum_i g_i rgb_i, opacity = sum_i g_i opacity_i, depth[:, i] = depth_i.
    Reference: models/rendering.py:48-110.  -> dict rgb, depth (N,size), opacity, gating_code, gating_importance."""
    rays_o = rays_o.contiguous(); rays_d = rays_d.contiguous()
    kwargs["return_gate"] = False
    with torch.autocast("cuda", enabled=rays_o.is_cuda):
        gate_in = rays_o if gating_net.type == "position" else torch.cat((rays_o, rays_d), 1)
        gating_code, gating_importance, _ = gating_net(gate_in, warmup)
        n = rays_o.shape[0]
        depth = torch.zeros(n, model_zoo.size, device=rays_o.device)
        rgb = torch.zeros_like(rays_o)
        opacity = torch.zeros(n, device=rays_o.device)
        fn = render_rays_test if kwargs.get("test_time", False) else render_rays_train
        for i in range(model_zoo.size):
            model = getattr(model_zoo, f"model_{i}")
            hits_t = intersect_scene_box(model, rays_o, rays_d)
            res = fn(model, lambda x, d, _m=model, **kw: _m(x, d, **kw), model.density_bitfield, rays_o, rays_d, hits_t, **dict(kwargs))
            rgb = rgb + res["rgb"] * gating_code[:, i][:, None]
            depth[:, i] = res["depth"]
            opacity = opacity + res["opacity"] * gating_code[:, i]
    return {"rgb": rgb, "depth": depth, "opacity": opacity, "gating_code": gating_code, "gating_importance": gating_importance}
