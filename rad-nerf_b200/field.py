"""ctypes front-end of the fused field kernels (rnb_field_fwd / rnb_field_bwd, include/radnerf_b200.h):
hash-grid encode -> geo MLP -> TruncExp -> SH4 -> rgb MLP for all K sub-NeRFs on one packed sample
list, forward and backward in one kernel each.  Replaces the chain of MNGP.forward
(reference models/networks.py:291-328) + tinycudann's backward.  No CPU path.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib
from ._lib import call, f32, stream_arg
from .tcnn import _GridMeta

MAX_K = 16
TILE = 128


class _FieldParams(ctypes.Structure):
    _fields_ = [("K", ctypes.c_int32), ("grid", _GridMeta),
                ("xyz_min", ctypes.c_float * 3), ("xyz_max", ctypes.c_float * 3),
                ("table_f16", ctypes.c_void_p),
                ("geo_w_f16", ctypes.c_void_p * MAX_K), ("rgb_w_f16", ctypes.c_void_p * MAX_K)]


class _Samples(ctypes.Structure):
    _fields_ = [("rays_o", ctypes.c_void_p), ("rays_d", ctypes.c_void_p), ("ray_id", ctypes.c_void_p),
                ("ts", ctypes.c_void_p), ("seg", ctypes.c_void_p), ("capacity", ctypes.c_int64)]


class _Saved(ctypes.Structure):
    _fields_ = [("enc", ctypes.c_void_p), ("rgb_in", ctypes.c_void_p), ("geo_h", ctypes.c_void_p),
                ("rgb_h1", ctypes.c_void_p), ("rgb_h2", ctypes.c_void_p), ("sraw", ctypes.c_void_p)]


def field_params(model, table_f16, geo_w_f16, rgb_w_f16):
    """model: NGP / MNGP (for the grid meta and the box); *_f16: fp16 compute copies (tensors / lists)."""
    P = _FieldParams()
    K = len(geo_w_f16)
    assert 1 <= K <= MAX_K
    P.K = K
    ctypes.memmove(ctypes.byref(P.grid), ctypes.byref(model.xyz_encoder._meta), ctypes.sizeof(_GridMeta))
    mn = model.xyz_min.reshape(-1).tolist(); mx = model.xyz_max.reshape(-1).tolist()
    for c in range(3):
        P.xyz_min[c] = mn[c]; P.xyz_max[c] = mx[c]
    P.table_f16 = table_f16.data_ptr()
    for i in range(K):
        P.geo_w_f16[i] = geo_w_f16[i].data_ptr()
        P.rgb_w_f16[i] = rgb_w_f16[i].data_ptr()
    P._keep = (table_f16, list(geo_w_f16), list(rgb_w_f16))
    return P


def samples(rays_o, rays_d, ray_id, ts, seg):
    S = _Samples()
    for t, n, dt in ((rays_o, "rays_o", torch.float32), (rays_d, "rays_d", torch.float32), (ray_id, "ray_id", torch.int32),
                     (ts, "ts", torch.float32), (seg, "seg", torch.int32)):
        _lib.check_input(t, n, dt)
    assert ray_id.numel() % TILE == 0 and ts.numel() == ray_id.numel()
    S.rays_o, S.rays_d, S.ray_id, S.ts, S.seg = (rays_o.data_ptr(), rays_d.data_ptr(), ray_id.data_ptr(),
                                                 ts.data_ptr(), seg.data_ptr())
    S.capacity = ray_id.numel()
    S._keep = (rays_o, rays_d, ray_id, ts, seg)
    return S


class Saved:
    """Workspace for the activations the backward pass needs (546 B / sample)."""

    def __init__(self, capacity, device):
        h = dict(dtype=torch.float16, device=device)
        self.enc = torch.empty(capacity, 32, **h)
        self.rgb_in = torch.empty(capacity, 32, **h)
        self.geo_h = torch.empty(capacity, 64, **h)
        self.rgb_h1 = torch.empty(capacity, 64, **h)
        self.rgb_h2 = torch.empty(capacity, 64, **h)
        self.sraw = torch.empty(capacity, **h)
        self.c = _Saved(self.enc.data_ptr(), self.rgb_in.data_ptr(), self.geo_h.data_ptr(), self.rgb_h1.data_ptr(),
                        self.rgb_h2.data_ptr(), self.sraw.data_ptr())

    def rows(self, name):
        """A saved activation as (capacity, W) rows.  The kernels keep it as per-128-sample operand tiles
        [W/8][128][8] (include/radnerf_b200.h, rnb_field_saved); this un-permutes a copy for inspection."""
        t = getattr(self, name)
        cap, W = t.shape
        return t.view(cap // 128, W // 8, 128, 8).permute(0, 2, 1, 3).reshape(cap, W)


def field_fwd(P, S, saved, sigmas, rgb4):
    call("rnb_field_fwd", ctypes.byref(P), ctypes.byref(S), ctypes.byref(saved.c) if saved is not None else None,
         sigmas, rgb4, stream_arg(sigmas.device))


def field_bwd(P, S, saved, rgb4, dsig, drgb, loss_scale, grad_table, grad_geo, grad_rgb):
    K = P.K
    gg = (ctypes.c_void_p * K)(*[g.data_ptr() for g in grad_geo])
    gr = (ctypes.c_void_p * K)(*[g.data_ptr() for g in grad_rgb])
    call("rnb_field_bwd", ctypes.byref(P), ctypes.byref(S), ctypes.byref(saved.c), rgb4, dsig, drgb, f32(loss_scale),
         grad_table, gg, gr, stream_arg(dsig.device))


def pack_passes(per_pass_ray_id, per_pass_ts, device):
    """Host-side packing helper (tests): lists of per-pass (ray_id i32, ts f32) -> ray_id, ts, seg."""
    K = len(per_pass_ray_id)
    bases, counts, off = [], [], 0
    for r in per_pass_ray_id:
        bases.append(off); counts.append(int(r.numel()))
        off = (off + int(r.numel()) + TILE - 1) // TILE * TILE
    cap = max(off, TILE)
    ray_id = torch.zeros(cap, dtype=torch.int32, device=device)
    ts = torch.zeros(cap, dtype=torch.float32, device=device)
    for b, r, t in zip(bases, per_pass_ray_id, per_pass_ts):
        ray_id[b:b + r.numel()] = r.int(); ts[b:b + r.numel()] = t
    seg = torch.tensor(bases + [off] + counts, dtype=torch.int32, device=device)
    return ray_id, ts, seg, bases, counts
