"""Autograd operator surface of the Rad-NeRF hot path -- same class names, call signatures and
return conventions as the reference's models/custom_functions.py:8-173, computed by the
sm_100a kernels behind the C ABI (radnerf_b200.vren).

Behavioural notes:
  * RayMarcher emits canonical rays_a (ascending ray index, exclusive-scan offsets, Q2) and
    consumes the torch RNG exactly like the reference: one rand_like((N_rays,)) per call (Q5);
  * RayMarcher.backward (only used with --optimize_ext) uses the CSR layout directly instead
    of torch_scatter.segment_csr; with canonical ordering the reference's latent bug Q3 vanishes;
  * VolumeRenderer.backward skips the dL_dws path when no gradient reaches ws.
"""
from __future__ import annotations

import torch
from torch.amp import custom_bwd, custom_fwd

from .. import vren


class RayAABBIntersector(torch.autograd.Function):
    """(rays_o (N,3), rays_d (N,3), center (V,3), half_size (V,3), max_hits)
    -> hits_cnt (N) i32, hits_t (N,max_hits,2) [-1 = miss], hits_voxel_idx (N,max_hits) i64.
    Reference: custom_functions.py:8-29."""

    @staticmethod
    @custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, rays_o, rays_d, center, half_size, max_hits):
        out = vren.ray_aabb_intersect(rays_o.contiguous(), rays_d.contiguous(), center.contiguous(),
                                      half_size.contiguous(), max_hits)
        ctx.mark_non_differentiable(*out)
        return tuple(out)


class RaySphereIntersector(torch.autograd.Function):
    """Reference: custom_functions.py:32-52 (no caller on the Rad-NeRF path)."""

    @staticmethod
    @custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, rays_o, rays_d, center, radii, max_hits):
        out = vren.ray_sphere_intersect(rays_o.contiguous(), rays_d.contiguous(), center.contiguous(),
                                        radii.contiguous(), max_hits)
        ctx.mark_non_differentiable(*out)
        return tuple(out)


class RayMarcher(torch.autograd.Function):
    """Occupancy-grid marching with a perturbed first sample (training).
    (rays_o, rays_d, hits_t (N,2), density_bitfield, cascades, scale, exp_step_factor, grid_size,
     max_samples) -> rays_a (N,3) i64 [ray_idx, start_idx, N_samples], xyzs (T,3), dirs (T,3),
    deltas (T), ts (T), total_samples (0-d i32 tensor).  Reference: custom_functions.py:55-112."""

    @staticmethod
    @custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, rays_o, rays_d, hits_t, density_bitfield, cascades, scale, exp_step_factor,
                grid_size, max_samples):
        noise = torch.rand_like(rays_o[:, 0])   # one draw per call, as the reference (Q5)
        rays_a, xyzs, dirs, deltas, ts, counter = vren.raymarching_train(
            rays_o.contiguous(), rays_d.contiguous(), hits_t.contiguous(), density_bitfield, cascades, scale,
            exp_step_factor, noise, grid_size, max_samples)
        total_samples = counter[0]
        ctx.save_for_backward(rays_a, ts)
        ctx.mark_non_differentiable(rays_a, total_samples)
        return rays_a, xyzs, dirs, deltas, ts, total_samples

    @staticmethod
    @custom_bwd(device_type="cuda")
    def backward(ctx, dL_drays_a, dL_dxyzs, dL_ddirs, dL_ddeltas, dL_dts, dL_dtotal_samples):
        rays_a, ts = ctx.saved_tensors
        n_rays = rays_a.shape[0]
        if dL_dxyzs is None and dL_ddirs is None:
            return (None,) * 9
        seg = torch.repeat_interleave(rays_a[:, 0], rays_a[:, 2])      # sample -> ray index (rows are CSR-ordered)
        dL_drays_o = dL_drays_d = None
        if dL_dxyzs is not None:
            dL_drays_o = torch.zeros(n_rays, 3, dtype=dL_dxyzs.dtype, device=ts.device).index_add_(0, seg, dL_dxyzs)
            per_sample = dL_dxyzs * ts[:, None]
        else:
            per_sample = torch.zeros_like(dL_ddirs)
        if dL_ddirs is not None:
            per_sample = per_sample + dL_ddirs
        dL_drays_d = torch.zeros(n_rays, 3, dtype=per_sample.dtype, device=ts.device).index_add_(0, seg, per_sample)
        return dL_drays_o, dL_drays_d, None, None, None, None, None, None, None


class VolumeRenderer(torch.autograd.Function):
    """Front-to-back compositing over CSR ray segments with early termination (training).
    (sigmas (T), rgbs (T,3), deltas (T), ts (T), rays_a (N,3), T_threshold)
    -> total_samples (0-d i64), opacity (N), depth (N), rgb (N,3), ws (T).
    Reference: custom_functions.py:115-159."""

    @staticmethod
    @custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, sigmas, rgbs, deltas, ts, rays_a, T_threshold):
        sigmas, rgbs, deltas, ts = sigmas.contiguous(), rgbs.contiguous(), deltas.contiguous(), ts.contiguous()
        total_samples, opacity, depth, rgb, ws = vren.composite_train_fw(sigmas, rgbs, deltas, ts, rays_a, T_threshold)
        ctx.save_for_backward(sigmas, rgbs, deltas, ts, rays_a, opacity, depth, rgb, ws)
        ctx.T_threshold = T_threshold
        ctx.set_materialize_grads(False)
        total = total_samples.sum()
        ctx.mark_non_differentiable(total)
        return total, opacity, depth, rgb, ws

    @staticmethod
    @custom_bwd(device_type="cuda")
    def backward(ctx, dL_dtotal_samples, dL_dopacity, dL_ddepth, dL_drgb, dL_dws):
        sigmas, rgbs, deltas, ts, rays_a, opacity, depth, rgb, ws = ctx.saved_tensors
        z = lambda ref: torch.zeros_like(ref)  # noqa: E731
        dL_dopacity = z(opacity) if dL_dopacity is None else dL_dopacity.float().contiguous()
        dL_ddepth = z(depth) if dL_ddepth is None else dL_ddepth.float().contiguous()
        dL_drgb = z(rgb) if dL_drgb is None else dL_drgb.float().contiguous()
        if dL_dws is not None:
            dL_dws = dL_dws.float().contiguous()
        dL_dsigmas, dL_drgbs = vren.composite_train_bw(dL_dopacity, dL_ddepth, dL_drgb, dL_dws, sigmas, rgbs, ws,
                                                       deltas, ts, rays_a, opacity, depth, rgb, ctx.T_threshold)
        return dL_dsigmas, dL_drgbs, None, None, None, None


class TruncExp(torch.autograd.Function):
    """exp in fp32; the backward clamps the exponent to [-15, 15] (custom_functions.py:162-173)."""

    @staticmethod
    @custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, x):
        x = x.float()            # fp32 also outside autocast (e.g. update_density_grid called directly)
        ctx.save_for_backward(x)
        return torch.exp(x)

    @staticmethod
    @custom_bwd(device_type="cuda")
    def backward(ctx, dL_dout):
        (x,) = ctx.saved_tensors
        return dL_dout * torch.exp(x.clamp(-15, 15))
