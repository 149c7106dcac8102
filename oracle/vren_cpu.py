"""TEST INFRASTRUCTURE ONLY -- numpy/ctypes front-end of oracle/libvren_oracle.so, the CPU
restatement (oracle/vren_oracle.c) of the reference's `vren` kernels on the hot path.

Function names/argument order follow the reference's pybind module
(/root/reference/models/csrc/binding.cpp:234-251) so parity tests read like calls into it.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module; the product package never does.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libvren_oracle.so")
_lib = None


def build():
    subprocess.check_call(["make", "-C", _HERE, "libvren_oracle.so"], stdout=subprocess.DEVNULL)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(os.path.join(_HERE, "vren_oracle.c")):
            build()
        _lib = ctypes.CDLL(_SO)
        _lib.orc_march_train_write.restype = ctypes.c_int64
    return _lib


def _p(a):
    return ctypes.c_void_p(0) if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def ray_aabb_intersect(rays_o, rays_d, center, half_size, max_hits=1):
    assert max_hits == 1 and np.asarray(center).reshape(-1, 3).shape[0] == 1
    rays_o, rays_d = _f32(rays_o), _f32(rays_d)
    center, half_size = _f32(center).reshape(3), _f32(half_size).reshape(3)
    n = len(rays_o)
    cnt = np.empty(n, np.int32); hits_t = np.empty((n, 1, 2), np.float32); idx = np.empty((n, 1), np.int64)
    lib().orc_ray_aabb_intersect(_p(rays_o), _p(rays_d), _p(center), _p(half_size), ctypes.c_int64(n),
                                 _p(cnt), _p(hits_t), _p(idx))
    return cnt, hits_t, idx


def morton3D(coords):
    coords = np.ascontiguousarray(coords, np.int32)
    out = np.empty(len(coords), np.int32)
    lib().orc_morton3D(_p(coords), ctypes.c_int64(len(coords)), _p(out))
    return out


def morton3D_invert(indices):
    indices = np.ascontiguousarray(indices, np.int32)
    out = np.empty((len(indices), 3), np.int32)
    lib().orc_morton3D_invert(_p(indices), ctypes.c_int64(len(indices)), _p(out))
    return out


def packbits(density_grid, thr, bitfield=None):
    g = _f32(density_grid).reshape(-1)
    n_bytes = g.size // 8
    if bitfield is None:
        bitfield = np.empty(n_bytes, np.uint8)
    lib().orc_packbits(_p(g), ctypes.c_int64(n_bytes), ctypes.c_float(thr), _p(bitfield))
    return bitfield


def raymarching_train(rays_o, rays_d, hits_t, bitfield, cascades, scale, exp_step_factor, noise,
                      grid_size=128, max_samples=1024, count_only=False):
    """-> rays_a (n,3) i64 canonical (ascending ray, exclusive-scan offsets), xyzs, dirs, deltas, ts, counter"""
    rays_o, rays_d, hits_t, noise = _f32(rays_o), _f32(rays_d), _f32(hits_t).reshape(-1, 2), _f32(noise)
    bitfield = np.ascontiguousarray(bitfield, np.uint8)
    n = len(rays_o)
    ns = np.empty(n, np.int32)
    args = (_p(rays_o), _p(rays_d), _p(hits_t), _p(bitfield), ctypes.c_int(cascades), ctypes.c_float(scale),
            ctypes.c_float(exp_step_factor), _p(noise), ctypes.c_int(grid_size), ctypes.c_int(max_samples),
            ctypes.c_int64(n))
    lib().orc_march_train_count(*args, _p(ns))
    rays_a = np.empty((n, 3), np.int64)
    total = int(ns.sum())
    if count_only:
        lib().orc_march_train_write(*args, _p(ns), _p(rays_a), None, None, None, None)
        return rays_a, np.array([total, n], np.int32)
    xyzs = np.zeros((total, 3), np.float32); dirs = np.zeros((total, 3), np.float32)
    deltas = np.zeros(total, np.float32); ts = np.zeros(total, np.float32)
    lib().orc_march_train_write(*args, _p(ns), _p(rays_a), _p(xyzs), _p(dirs), _p(deltas), _p(ts))
    return rays_a, xyzs, dirs, deltas, ts, np.array([total, n], np.int32)


def raymarching_test(rays_o, rays_d, hits_t, alive, bitfield, cascades, scale, exp_step_factor,
                     grid_size, max_samples, N_samples):
    """hits_t (n_rays,2) f32 contiguous numpy array, mutated in place."""
    rays_o, rays_d = _f32(rays_o), _f32(rays_d)
    assert hits_t.dtype == np.float32 and hits_t.flags.c_contiguous
    alive = np.ascontiguousarray(alive, np.int64)
    bitfield = np.ascontiguousarray(bitfield, np.uint8)
    A = len(alive)
    xyzs = np.zeros((A, N_samples, 3), np.float32); dirs = np.zeros((A, N_samples, 3), np.float32)
    deltas = np.zeros((A, N_samples), np.float32); ts = np.zeros((A, N_samples), np.float32)
    n_eff = np.zeros(A, np.int32)
    lib().orc_march_test(_p(rays_o), _p(rays_d), _p(hits_t), _p(alive), _p(bitfield), ctypes.c_int(cascades),
                         ctypes.c_float(scale), ctypes.c_float(exp_step_factor), ctypes.c_int(grid_size),
                         ctypes.c_int(max_samples), ctypes.c_int(N_samples), ctypes.c_int64(A),
                         _p(xyzs), _p(dirs), _p(deltas), _p(ts), _p(n_eff))
    return xyzs, dirs, deltas, ts, n_eff


def composite_train_fw(sigmas, rgbs, deltas, ts, rays_a, T_threshold):
    sigmas, rgbs, deltas, ts = _f32(sigmas), _f32(rgbs), _f32(deltas), _f32(ts)
    rays_a = np.ascontiguousarray(rays_a, np.int64)
    n, N = len(rays_a), len(sigmas)
    total = np.zeros(n, np.int64); opacity = np.zeros(n, np.float32); depth = np.zeros(n, np.float32)
    rgb = np.zeros((n, 3), np.float32); ws = np.zeros(N, np.float32)
    lib().orc_composite_train_fw(_p(sigmas), _p(rgbs), _p(deltas), _p(ts), _p(rays_a), ctypes.c_float(T_threshold),
                                 ctypes.c_int64(n), _p(total), _p(opacity), _p(depth), _p(rgb), _p(ws))
    return total, opacity, depth, rgb, ws


def composite_train_bw(dL_dopacity, dL_ddepth, dL_drgb, dL_dws, sigmas, rgbs, ws, deltas, ts, rays_a,
                       opacity, depth, rgb, T_threshold):
    a = [_f32(x) for x in (dL_dopacity, dL_ddepth, dL_drgb, dL_dws, sigmas, rgbs, ws, deltas, ts)]
    rays_a = np.ascontiguousarray(rays_a, np.int64)
    opacity, depth, rgb = _f32(opacity), _f32(depth), _f32(rgb)
    n, N = len(rays_a), len(a[4])
    gs = np.zeros(N, np.float32); gc = np.zeros((N, 3), np.float32)
    lib().orc_composite_train_bw(*[_p(x) for x in a], _p(rays_a), _p(opacity), _p(depth), _p(rgb),
                                 ctypes.c_float(T_threshold), ctypes.c_int64(n), _p(gs), _p(gc))
    return gs, gc


def composite_test_fw(sigmas, rgbs, deltas, ts, hits_t, alive, T_threshold, n_eff, opacity, depth, rgb):
    """alive (i64), opacity, depth, rgb (f32) contiguous numpy arrays mutated in place."""
    sigmas, rgbs, deltas, ts = _f32(sigmas), _f32(rgbs), _f32(deltas), _f32(ts)
    n_eff = np.ascontiguousarray(n_eff, np.int32)
    for x in (alive, opacity, depth, rgb):
        assert x.flags.c_contiguous
    A = len(alive)
    Ns = sigmas.shape[1]
    lib().orc_composite_test_fw(_p(sigmas), _p(rgbs), _p(deltas), _p(ts), _p(alive), ctypes.c_float(T_threshold),
                                _p(n_eff), ctypes.c_int(Ns), ctypes.c_int64(A), _p(opacity), _p(depth), _p(rgb))


def distortion_loss_fw(ws, deltas, ts, rays_a):
    """losses.cu:9-60 (prefix_sums_kernel: serial thrust scans per ray), :92-93 (per-sample term as an ATen
    expression: 2*(wts_incl*ws_excl - ws_incl*wts_excl) + ws*ws*deltas/3) and :44-60 (serial per-ray sum).
    fp32, sequential accumulation like the reference's one-thread-per-ray loops."""
    ws, deltas, ts = _f32(ws), _f32(deltas), _f32(ts)
    rays_a = np.ascontiguousarray(rays_a, np.int64)
    N = len(ws)
    wts = (ws * ts).astype(np.float32)
    wi = np.zeros(N, np.float32); we = np.zeros(N, np.float32); ti = np.zeros(N, np.float32); te = np.zeros(N, np.float32)
    for _, s, n in rays_a:
        if n <= 0:
            continue
        wi[s:s + n] = np.cumsum(ws[s:s + n], dtype=np.float32)
        ti[s:s + n] = np.cumsum(wts[s:s + n], dtype=np.float32)
        we[s + 1:s + n] = wi[s:s + n - 1]
        te[s + 1:s + n] = ti[s:s + n - 1]
    per = (np.float32(2) * (ti * we - wi * te) + np.float32(1.0 / 3) * ws * ws * deltas).astype(np.float32)
    loss = np.zeros(len(rays_a), np.float32)
    for r, s, n in rays_a:
        acc = np.float32(0)
        for k in range(s, s + n):
            acc = np.float32(acc + per[k])
        loss[r] = acc
    return loss, wi, ti


def distortion_loss_bw(dL_dloss, ws_inclusive_scan, wts_inclusive_scan, ws, deltas, ts, rays_a):
    """losses.cu:113-136."""
    g, wi, ti, ws, deltas, ts = (_f32(x) for x in (dL_dloss, ws_inclusive_scan, wts_inclusive_scan, ws, deltas, ts))
    rays_a = np.ascontiguousarray(rays_a, np.int64)
    out = np.zeros(len(ws), np.float32)
    two3 = np.float32(2) / np.float32(3)
    for r, s, n in rays_a:
        if n <= 0:
            continue
        e = s + n - 1
        ws_sum, wts_sum = wi[e], ti[e]
        sl = slice(s, e + 1)
        before = np.zeros(n, np.float32)
        before[1:] = ts[s + 1:e + 1] * wi[s:e] - ti[s:e]
        d = g[r] * np.float32(2) * (before + (wts_sum - ti[sl] - ts[sl] * (ws_sum - wi[sl])))
        out[sl] = (d + g[r] * two3 * ws[sl] * deltas[sl]).astype(np.float32)
    return out
