"""Shared per-field render routines used by rendering.render (K=1) and ml_rendering.ml_render.

`field(xyzs, dirs) -> (sigmas (T) f32, rgbs (T,3))` evaluates one (sub-)NeRF; `bitfield` is its
occupancy bitfield.  Semantics follow the reference's __render_rays_train / __render_rays_test
(models/ml_rendering.py:81-202, models/rendering.py:113-239), including hazards Q11-Q14 and Q17.
"""
from __future__ import annotations

import torch

from .. import vren
from .custom_functions import RayAABBIntersector, RayMarcher, VolumeRenderer

MAX_SAMPLES = 1024
NEAR_DISTANCE = 0.01


def intersect_scene_box(model, rays_o, rays_d):
    """AABB hit interval per ray with the NEAR_DISTANCE clamp (ml_rendering.py:48-50). -> hits_t (N,1,2)"""
    _, hits_t, _ = RayAABBIntersector.apply(rays_o, rays_d, model.center, model.half_size, 1)
    near = (hits_t[:, 0, 0] >= 0) & (hits_t[:, 0, 0] < NEAR_DISTANCE)
    hits_t[near, 0, 0] = NEAR_DISTANCE
    return hits_t


def background(exp_step_factor, device, random_bg=False):
    if exp_step_factor == 0:          # synthetic scenes: white
        return torch.ones(3, device=device)
    if random_bg:                     # real scenes, training only: one draw per call (Q17)
        return torch.rand(3, device=device)
    return torch.zeros(3, device=device)


def render_rays_train(model, field, bitfield, rays_o, rays_d, hits_t, **kwargs):
    exp_step_factor = kwargs.get("exp_step_factor", 0.)
    results = {}
    (rays_a, xyzs, dirs, results["deltas"], results["ts"], results["rm_samples"]) = RayMarcher.apply(
        rays_o, rays_d, hits_t[:, 0], bitfield, model.cascades, model.scale, exp_step_factor,
        model.grid_size, MAX_SAMPLES)
    for k, v in kwargs.items():       # per-ray tensors are repeated per sample, as the reference does
        if isinstance(v, torch.Tensor):
            kwargs[k] = torch.repeat_interleave(v[rays_a[:, 0]], rays_a[:, 2], 0)
    sigmas, rgbs = field(xyzs, dirs, **kwargs)
    (results["vr_samples"], results["opacity"], results["depth"], results["rgb"], results["ws"]) = \
        VolumeRenderer.apply(sigmas, rgbs.contiguous(), results["deltas"], results["ts"], rays_a,
                             kwargs.get("T_threshold", 1e-4))
    results["rays_a"] = rays_a
    rgb_bg = background(exp_step_factor, rays_o.device, kwargs.get("random_bg", False))
    results["rgb"] = results["rgb"] + rgb_bg * (1 - results["opacity"])[:, None]
    return results


@torch.no_grad()
def render_rays_test(model, field, bitfield, rays_o, rays_d, hits_t, **kwargs):
    """Alive-ray loop: march N_samples more occupied steps per alive ray, evaluate, composite in
    place, drop converged rays; N_samples grows as rays die (ml_rendering.py:108-141)."""
    exp_step_factor = kwargs.get("exp_step_factor", 0.)
    N_rays = len(rays_o)
    device = rays_o.device
    opacity = torch.zeros(N_rays, device=device)
    depth = torch.zeros(N_rays, device=device)
    rgb = torch.zeros(N_rays, 3, device=device)
    hits = hits_t[:, 0].contiguous()   # (N,2); advanced in place by the marcher (Q12)
    samples = total_samples = 0
    alive_indices = torch.arange(N_rays, device=device)
    min_samples = 1 if exp_step_factor == 0 else 4
    deltas = None
    while samples < kwargs.get("max_samples", MAX_SAMPLES):
        N_alive = len(alive_indices)
        if N_alive == 0:
            break
        N_samples = max(min(N_rays // N_alive, 64), min_samples)
        samples += N_samples
        xyzs, dirs, deltas, ts, N_eff_samples = vren.raymarching_test(
            rays_o, rays_d, hits, alive_indices, bitfield, model.cascades, model.scale, exp_step_factor,
            model.grid_size, MAX_SAMPLES, N_samples)
        total_samples += N_eff_samples.sum()
        # valid slots = the first N_eff of each row (equivalent to the reference's "dir row != 0", Q14)
        valid_mask = (torch.arange(N_samples, device=device)[None, :] < N_eff_samples[:, None]).reshape(-1)
        if valid_mask.sum() == 0:
            break
        xyzs = xyzs.reshape(-1, 3); dirs = dirs.reshape(-1, 3)
        sigmas = torch.zeros(len(xyzs), device=device)
        rgbs = torch.zeros(len(xyzs), 3, device=device)
        _sigmas, _rgbs = field(xyzs[valid_mask], dirs[valid_mask], **kwargs)
        sigmas[valid_mask] = _sigmas.float(); rgbs[valid_mask] = _rgbs.float()
        vren.composite_test_fw(sigmas.view(-1, N_samples), rgbs.view(-1, N_samples, 3), deltas, ts, hits,
                               alive_indices, kwargs.get("T_threshold", 1e-4), N_eff_samples, opacity, depth, rgb)
        alive_indices = alive_indices[alive_indices >= 0]
    results = {"opacity": opacity, "depth": depth, "rgb": rgb, "total_samples": total_samples, "deltas": deltas}
    rgb_bg = background(exp_step_factor, device)
    results["rgb"] += rgb_bg * (1 - opacity)[:, None]
    return results
