"""Drop-in for the reference's pybind11 module ``vren`` (models/csrc/binding.cpp:234-251).

Same function names, positional signatures, return conventions and in-place contracts;
the work is done by hand-written sm_100a kernels behind the C ABI (include/radnerf_b200.h).
Differences a caller can observe (all legal outcomes of the reference's own races,
SURVEY.md section 9):
  * raymarching_train emits rays_a rows in ascending ray order with start_idx = exclusive
    prefix sum of N_samples (Q2), and returns exactly-sized (not N_rays*max_samples) outputs
    -- `xyzs[:counter[0]]` etc. slice to the same thing as with the reference;
  * kernels run on torch's CURRENT stream (the reference uses the legacy default stream).
"""
from __future__ import annotations

import torch

from . import _lib
from ._lib import call, check_input, f32, i32, i64, stream_arg


def ray_aabb_intersect(rays_o, rays_d, centers, half_sizes, max_hits):
    """binding.cpp:4-16.  -> [hit_cnt (N) i32, hits_t (N,max_hits,2) f32, hits_voxel_idx (N,max_hits) i64]"""
    for t, n in ((rays_o, "rays_o"), (rays_d, "rays_d"), (centers, "centers"), (half_sizes, "half_sizes")):
        check_input(t, n, torch.float32)
    n_rays, n_vox = rays_o.shape[0], centers.shape[0]
    dev = rays_o.device
    hit_cnt = torch.empty(n_rays, dtype=torch.int32, device=dev)
    hits_t = torch.empty(n_rays, max_hits, 2, dtype=torch.float32, device=dev)
    hits_idx = torch.empty(n_rays, max_hits, dtype=torch.int64, device=dev)
    call("rnb_ray_aabb_intersect", rays_o, rays_d, centers, half_sizes, i64(n_rays), i64(n_vox),
         i32(max_hits), hit_cnt, hits_t, hits_idx, stream_arg(dev))
    return [hit_cnt, hits_t, hits_idx]


def ray_sphere_intersect(rays_o, rays_d, centers, radii, max_hits):
    """binding.cpp:19-31."""
    for t, n in ((rays_o, "rays_o"), (rays_d, "rays_d"), (centers, "centers"), (radii, "radii")):
        check_input(t, n, torch.float32)
    n_rays, n_sph = rays_o.shape[0], centers.shape[0]
    dev = rays_o.device
    hit_cnt = torch.empty(n_rays, dtype=torch.int32, device=dev)
    hits_t = torch.empty(n_rays, max_hits, 2, dtype=torch.float32, device=dev)
    hits_idx = torch.empty(n_rays, max_hits, dtype=torch.int64, device=dev)
    call("rnb_ray_sphere_intersect", rays_o, rays_d, centers, radii, i64(n_rays), i64(n_sph),
         i32(max_hits), hit_cnt, hits_t, hits_idx, stream_arg(dev))
    return [hit_cnt, hits_t, hits_idx]


def morton3D(coords):
    """binding.cpp:46-50.  (N,3) i32 -> (N) i32"""
    check_input(coords, "coords", torch.int32)
    out = torch.empty(coords.shape[0], dtype=torch.int32, device=coords.device)
    call("rnb_morton3D", coords, i64(coords.shape[0]), out, stream_arg(coords.device))
    return out


def morton3D_invert(indices):
    """binding.cpp:53-57.  (N) i32 -> (N,3) i32"""
    check_input(indices, "indices", torch.int32)
    out = torch.empty(indices.shape[0], 3, dtype=torch.int32, device=indices.device)
    call("rnb_morton3D_invert", indices, i64(indices.shape[0]), out, stream_arg(indices.device))
    return out


def packbits(density_grid, density_threshold, density_bitfield):
    """binding.cpp:34-43.  In place into density_bitfield (u8, numel = grid.numel()/8)."""
    check_input(density_grid, "density_grid")
    check_input(density_bitfield, "density_bitfield", torch.uint8)
    if density_grid.dtype not in (torch.float32, torch.float16):
        raise RuntimeError(f"packbits: unsupported grid dtype {density_grid.dtype}")
    call("rnb_packbits", density_grid, i32(density_grid.dtype == torch.float16),
         i64(density_bitfield.shape[0]), f32(density_threshold), density_bitfield,
         stream_arg(density_grid.device))


_scratch_cache = {}


def _scratch(dev, numel):
    """Reusable fp32 scratch (the marcher's per-ray sample-t rows), one per (device, stream): callers on different streams
    of one device (a side-stream evaluation during training) must not share it."""
    key = (dev, torch.cuda.current_stream(dev).cuda_stream)
    buf = _scratch_cache.get(key)
    if buf is None or buf.numel() < numel:
        buf = torch.empty(max(numel, 1), dtype=torch.float32, device=dev)
        _scratch_cache[key] = buf
    return buf


def raymarching_train(rays_o, rays_d, hits_t, density_bitfield, cascades, scale, exp_step_factor,
                      noise, grid_size, max_samples):
    """binding.cpp:60-81.  -> [rays_a (N,3) i64, xyzs (T,3), dirs (T,3), deltas (T), ts (T),
    counter (2) i32 = [total_samples T, N_rays]]."""
    for t, n in ((rays_o, "rays_o"), (rays_d, "rays_d"), (hits_t, "hits_t"), (noise, "noise")):
        check_input(t, n, torch.float32)
    check_input(density_bitfield, "density_bitfield", torch.uint8)
    n_rays = rays_o.shape[0]
    dev = rays_o.device
    rays_a = torch.empty(n_rays, 3, dtype=torch.int64, device=dev)
    counter = torch.empty(2, dtype=torch.int32, device=dev)
    st = stream_arg(dev)
    t_scratch = _scratch(dev, n_rays * max_samples)   # un-initialised; only sample slots are written
    call("rnb_raymarching_train_count", rays_o, rays_d, hits_t, density_bitfield, i32(cascades), f32(scale),
         f32(exp_step_factor), noise, i32(grid_size), i32(max_samples), i64(n_rays), rays_a, counter,
         t_scratch, st)
    total = int(counter[0].item())  # the reference syncs here too (custom_functions.py:91-96)
    xyzs = torch.empty(total, 3, dtype=torch.float32, device=dev)
    dirs = torch.empty(total, 3, dtype=torch.float32, device=dev)
    deltas = torch.empty(total, dtype=torch.float32, device=dev)
    ts = torch.empty(total, dtype=torch.float32, device=dev)
    call("rnb_raymarching_train_write", rays_o, rays_d, f32(scale), f32(exp_step_factor), i32(grid_size),
         i32(max_samples), i64(n_rays), rays_a, t_scratch, i64(total), xyzs, dirs, deltas, ts, st)
    return [rays_a, xyzs, dirs, deltas, ts, counter]


def raymarching_test(rays_o, rays_d, hits_t, alive_indices, density_bitfield, cascades, scale,
                     exp_step_factor, grid_size, max_samples, N_samples):
    """binding.cpp:84-106.  hits_t (N_rays,2) is mutated.  -> [xyzs (A,Ns,3), dirs (A,Ns,3),
    deltas (A,Ns), ts (A,Ns), N_eff_samples (A) i32]."""
    for t, n in ((rays_o, "rays_o"), (rays_d, "rays_d"), (hits_t, "hits_t")):
        check_input(t, n, torch.float32)
    check_input(alive_indices, "alive_indices", torch.int64)
    check_input(density_bitfield, "density_bitfield", torch.uint8)
    A = alive_indices.shape[0]
    dev = rays_o.device
    xyzs = torch.empty(A, N_samples, 3, dtype=torch.float32, device=dev)
    dirs = torch.empty(A, N_samples, 3, dtype=torch.float32, device=dev)
    deltas = torch.empty(A, N_samples, dtype=torch.float32, device=dev)
    ts = torch.empty(A, N_samples, dtype=torch.float32, device=dev)
    n_eff = torch.empty(A, dtype=torch.int32, device=dev)
    call("rnb_raymarching_test", rays_o, rays_d, hits_t, alive_indices, i64(A), density_bitfield,
         i32(cascades), f32(scale), f32(exp_step_factor), i32(grid_size), i32(max_samples), i32(N_samples),
         xyzs, dirs, deltas, ts, n_eff, stream_arg(dev))
    return [xyzs, dirs, deltas, ts, n_eff]


def composite_train_fw(sigmas, rgbs, deltas, ts, rays_a, T_threshold):
    """binding.cpp:109-126.  -> [total_samples (N_rays) i64, opacity, depth, rgb (N_rays,3), ws (N)]"""
    for t, n in ((sigmas, "sigmas"), (rgbs, "rgbs"), (deltas, "deltas"), (ts, "ts")):
        check_input(t, n, torch.float32)
    check_input(rays_a, "rays_a", torch.int64)
    n_rays, N = rays_a.shape[0], sigmas.shape[0]
    dev = sigmas.device
    total_samples = torch.zeros(n_rays, dtype=torch.int64, device=dev)
    opacity = torch.zeros(n_rays, dtype=torch.float32, device=dev)
    depth = torch.zeros(n_rays, dtype=torch.float32, device=dev)
    rgb = torch.zeros(n_rays, 3, dtype=torch.float32, device=dev)
    ws = torch.empty(N, dtype=torch.float32, device=dev)
    call("rnb_composite_train_fw", sigmas, rgbs, deltas, ts, rays_a, f32(T_threshold), i64(n_rays), i64(N),
         total_samples, opacity, depth, rgb, ws, stream_arg(dev))
    return [total_samples, opacity, depth, rgb, ws]


def composite_train_bw(dL_dopacity, dL_ddepth, dL_drgb, dL_dws, sigmas, rgbs, ws, deltas, ts, rays_a,
                       opacity, depth, rgb, T_threshold):
    """binding.cpp:129-163.  dL_dws may be None (no gradient flows into ws).
    -> [dL_dsigmas (N), dL_drgbs (N,3)]"""
    for t, n in ((dL_dopacity, "dL_dopacity"), (dL_ddepth, "dL_ddepth"), (dL_drgb, "dL_drgb"),
                 (sigmas, "sigmas"), (rgbs, "rgbs"), (deltas, "deltas"), (ts, "ts"),
                 (opacity, "opacity"), (depth, "depth"), (rgb, "rgb")):
        check_input(t, n, torch.float32)
    if dL_dws is not None:
        check_input(dL_dws, "dL_dws", torch.float32)
        check_input(ws, "ws", torch.float32)
    check_input(rays_a, "rays_a", torch.int64)
    n_rays, N = rays_a.shape[0], sigmas.shape[0]
    dev = sigmas.device
    dL_dsigmas = torch.empty(N, dtype=torch.float32, device=dev)
    dL_drgbs = torch.empty(N, 3, dtype=torch.float32, device=dev)
    call("rnb_composite_train_bw", dL_dopacity, dL_ddepth, dL_drgb, dL_dws, sigmas, rgbs, ws, deltas, ts,
         rays_a, opacity, depth, rgb, f32(T_threshold), i64(n_rays), i64(N), dL_dsigmas, dL_drgbs,
         stream_arg(dev))
    return [dL_dsigmas, dL_drgbs]


def composite_test_fw(sigmas, rgbs, deltas, ts, hits_t, alive_indices, T_threshold, N_eff_samples,
                      opacity, depth, rgb):
    """binding.cpp:166-194.  In place: alive_indices (-1 = converged), opacity, depth, rgb."""
    for t, n in ((sigmas, "sigmas"), (rgbs, "rgbs"), (deltas, "deltas"), (ts, "ts"),
                 (opacity, "opacity"), (depth, "depth"), (rgb, "rgb")):
        check_input(t, n, torch.float32)
    check_input(alive_indices, "alive_indices", torch.int64)
    check_input(N_eff_samples, "N_eff_samples", torch.int32)
    A = alive_indices.shape[0]
    Ns = sigmas.shape[1] if sigmas.dim() == 2 else 1
    call("rnb_composite_test_fw", sigmas, rgbs, deltas, ts, hits_t, alive_indices, f32(T_threshold),
         N_eff_samples, i64(A), i32(Ns), opacity, depth, rgb, stream_arg(sigmas.device))


def distortion_loss_fw(ws, deltas, ts, rays_a):
    """binding.cpp:197-210 -> [loss (N_rays), ws_inclusive_scan (N), wts_inclusive_scan (N)]."""
    for t, n in ((ws, "ws"), (deltas, "deltas"), (ts, "ts")):
        check_input(t, n, torch.float32)
    check_input(rays_a, "rays_a", torch.int64)
    n_rays, N = rays_a.shape[0], ws.shape[0]
    loss = torch.zeros(n_rays, device=ws.device)
    wsi = torch.zeros(N, device=ws.device)
    wtsi = torch.zeros(N, device=ws.device)
    call("rnb_distortion_loss_fw", ws, deltas, ts, rays_a, i64(n_rays), i64(N), loss, wsi, wtsi, stream_arg(ws.device))
    return [loss, wsi, wtsi]


def distortion_loss_bw(dL_dloss, ws_inclusive_scan, wts_inclusive_scan, ws, deltas, ts, rays_a):
    """binding.cpp:212-231 -> dL_dws (N)."""
    for t, n in ((dL_dloss, "dL_dloss"), (ws_inclusive_scan, "ws_inclusive_scan"), (wts_inclusive_scan, "wts_inclusive_scan"),
                 (ws, "ws"), (deltas, "deltas"), (ts, "ts")):
        check_input(t, n, torch.float32)
    check_input(rays_a, "rays_a", torch.int64)
    n_rays, N = rays_a.shape[0], ws.shape[0]
    dL_dws = torch.zeros(N, device=ws.device)
    call("rnb_distortion_loss_bw", dL_dloss, ws_inclusive_scan, wts_inclusive_scan, ws, deltas, ts, rays_a, i64(n_rays), i64(N),
         dL_dws, stream_arg(ws.device))
    return dL_dws
