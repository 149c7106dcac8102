"""Rad-NeRF render entry: gate -> K sub-NeRF passes -> gate-weighted blend.
Same call convention and result keys as the reference's models/ml_rendering.py:11-78:
  rgb (N,3), independent_rgbs [K x (N,3)], depth (N,K) UNGATED (Q19), opacity (N),
  gating_code (N,K), gating_importance (K).
The scene-box intersection is computed once and cloned per sub-NeRF (the reference recomputes
the identical intersection K times, Q22; each pass needs its own hits_t because the test-time
marcher advances it in place, Q12)."""
from __future__ import annotations

import torch

from ._render_core import MAX_SAMPLES, NEAR_DISTANCE, intersect_scene_box, render_rays_test, render_rays_train  # noqa: F401


def ml_render(model, gating_net, rays_o, rays_d, imgs_d, warmup=False, **kwargs):
    with torch.autocast("cuda", enabled=rays_o.is_cuda):   # the reference decorates ml_render with autocast
        device = rays_o.device
        rays_o = rays_o.contiguous(); rays_d = rays_d.contiguous(); imgs_d = imgs_d.contiguous()
        kwargs["return_gate"] = False
        # gate input: raw (unnormalised) ray data (Q18)
        second = imgs_d if gating_net.type == "image" else rays_d
        gating_code, gating_importance, _ = gating_net(torch.cat((rays_o, second), 1), warmup)

        depth_results = torch.zeros(rays_o.shape[0], model.size, device=device)
        rgb_results = torch.zeros(rays_o.shape, device=device)
        opacity_results = torch.zeros(rays_o.shape[0], device=device)
        independent_rgbs = []
        hits_shared = intersect_scene_box(model, rays_o, rays_d)
        render_func = render_rays_test if kwargs.get("test_time", False) else render_rays_train
        for i in range(model.size):
            hits_t = hits_shared.clone()
            bitfield = getattr(model, f"density_bitfield_{i}")
            results = render_func(model, lambda x, d, i=i, **kw: model(x, d, i, **kw), bitfield, rays_o, rays_d, hits_t, **kwargs)
            for k, v in results.items():
                if kwargs.get("to_cpu", False) and isinstance(v, torch.Tensor):
                    v = v.cpu()
                    if kwargs.get("to_numpy", False):
                        v = v.numpy()
                results[k] = v
            independent_rgbs.append(results["rgb"])
            rgb_results = rgb_results + results["rgb"] * gating_code[:, i][:, None]
            depth_results[:, i] = results["depth"]
            opacity_results = opacity_results + results["opacity"] * gating_code[:, i]
        return {"rgb": rgb_results, "independent_rgbs": independent_rgbs, "depth": depth_results,
                "opacity": opacity_results, "gating_code": gating_code, "gating_importance": gating_importance}
