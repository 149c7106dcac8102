"""Drop-in for the slice of the ``tinycudann`` PyTorch API the reference uses
(models/networks.py:42-89, 234-289, 1075-1085):

    tcnn.Encoding(n_input_dims, encoding_config)      Grid/Hash or SphericalHarmonics
    tcnn.Network(n_input_dims, n_output_dims, network_config)   FullyFusedMLP / CutlassMLP

Each is an ``nn.Module`` with ONE flat fp32 ``params`` Parameter in tinycudann's layout (so
reference state-dicts load), ``n_output_dims``, and ``__call__((N,in)) -> (N,out) fp16``.
The math runs in hand-written sm_100a kernels behind the C ABI (rnb_hashgrid_*, rnb_sh4_fwd,
rnb_mlp_*); there is no CPU path.  Like tinycudann, backward runs with an internal loss
scale (fp16 activations gradients) and accumulates parameter gradients in fp32.
"""
from __future__ import annotations

import ctypes
import math
import os

import numpy as np
import torch
from torch import nn

from . import _lib
from ._lib import call, f32, i64, stream_arg

LOSS_SCALE = 128.0   # tinycudann's default for fp16 networks


class _GridMeta(ctypes.Structure):
    _fields_ = [("n_levels", ctypes.c_int32), ("n_features", ctypes.c_int32),
                ("scale", ctypes.c_float * 32), ("res", ctypes.c_uint32 * 32), ("offset", ctypes.c_uint32 * 33)]


class _MlpDesc(ctypes.Structure):
    _fields_ = [("in_width", ctypes.c_int32), ("out_width", ctypes.c_int32),
                ("n_hidden", ctypes.c_int32), ("out_activation", ctypes.c_int32)]


_LIBM = None


def _libm():
    global _LIBM
    if _LIBM is None:
        _LIBM = ctypes.CDLL("libm.so.6")
        for f in ("exp2f", "log2f", "ceilf"):
            getattr(_LIBM, f).restype = ctypes.c_float
            getattr(_LIBM, f).argtypes = [ctypes.c_float]
    return _LIBM


def level_scale_res(level, per_level_scale, base_resolution, rule="tcnn_fp32"):
    """(scale_l, res_l) of one level.

    rule "tcnn_fp32" (default) mirrors tinycudann's host arithmetic (grid.h grid_scale / grid_resolution, called from
    the GridEncodingTemplated constructor with std::log2(float)): everything in fp32 through the C library --
    exp2f(l * log2f(b)) * N_min - 1.0f, then (uint32)ceilf(scale) + 1.  This is the table a reference checkpoint's flat
    `params` vector was laid out with.  rule "fp64_snap" is round 1's rule (fp64, snapped to 6 decimals): it differs
    exactly where the real-valued scale is an integer -- at scene scale 0.5 (b = 64^(1/15)) levels 5, 10, 15 come out as
    64 / 256 / 1024 instead of tinycudann's 65 / 257 / 1025, because fp32(b) rounds up and log2f returns 0.40000004."""
    if rule == "fp64_snap":
        s = round(2.0 ** (level * math.log2(per_level_scale)) * base_resolution - 1.0, 6)
        return s, int(math.ceil(s)) + 1
    assert rule == "tcnn_fp32", rule
    m = _libm()
    f = np.float32
    log2b = f(m.log2f(ctypes.c_float(f(per_level_scale))))
    e = f(f(level) * log2b)
    s = f(f(f(m.exp2f(ctypes.c_float(e))) * f(base_resolution)) - f(1.0))
    return float(s), int(m.ceilf(ctypes.c_float(s))) + 1


def default_level_rule():
    return os.environ.get("RNB_GRID_LEVEL_RULE", "tcnn_fp32")


def grid_meta(n_levels, n_features, log2_hashmap_size, base_resolution, per_level_scale, rule=None):
    """Level table of the multiresolution grid (tinycudann grid.h): scale_l = N_min*b^l - 1,
    res_l = ceil(scale_l)+1, entries_l = min(round_up(res_l^3, 8), 2^log2_T); scale_l / res_l as
    level_scale_res() evaluates them (tinycudann's fp32 host arithmetic unless RNB_GRID_LEVEL_RULE=fp64_snap)."""
    rule = rule or default_level_rule()
    m = _GridMeta()
    m.n_levels, m.n_features = n_levels, n_features
    off = 0
    for l in range(n_levels):
        s, r = level_scale_res(l, per_level_scale, base_resolution, rule)
        n = min(r ** 3, 2 ** 31 - 1)
        n = (n + 7) // 8 * 8
        n = min(n, 1 << log2_hashmap_size)
        m.scale[l], m.res[l], m.offset[l] = s, r, off
        off += n
    m.offset[n_levels] = off
    return m, off


class _F16Shadow:
    """fp16 compute copy of an fp32 master parameter, refreshed when the master changes."""

    def __init__(self):
        self.buf = None
        self.version = None
        self.ptr = None

    def get(self, p: torch.Tensor) -> torch.Tensor:
        key = (p._version, p.data_ptr(), p.device)
        if self.buf is None or self.buf.device != p.device or self.buf.numel() != p.numel():
            self.buf = torch.empty(p.numel(), dtype=torch.float16, device=p.device)
            self.version = None
        if self.version != key:
            _lib.check_input(p.detach(), "params", torch.float32)
            call("rnb_cast_f32_to_f16", p.detach(), i64(p.numel()), self.buf, stream_arg(p.device))
            self.version = key
        return self.buf


class _HashGridFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, params, module):
        x = x.detach().float().contiguous()
        _lib.check_input(x, "x", torch.float32)
        table = module._shadow.get(params)
        n = x.shape[0]
        out = torch.empty(n, module.n_output_dims, dtype=torch.float16, device=x.device)
        call("rnb_hashgrid_fwd", x, i64(n), table, ctypes.byref(module._meta), out, stream_arg(x.device))
        ctx.save_for_backward(x)
        ctx.module = module
        return out

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        m = ctx.module
        grad = torch.zeros(m.params.numel(), dtype=torch.float32, device=x.device)
        dy = dy.contiguous()
        if dy.dtype != torch.float16:
            dy = dy.half()
        call("rnb_hashgrid_bwd", x, i64(x.shape[0]), dy, ctypes.byref(m._meta), f32(1.0), grad, stream_arg(x.device))
        return None, grad, None


class Encoding(nn.Module):
    def __init__(self, n_input_dims, encoding_config, seed=1337, dtype=None):
        super().__init__()
        assert n_input_dims == 3, "only 3-D encodings are on the Rad-NeRF path"
        self.n_input_dims = n_input_dims
        self.encoding_config = dict(encoding_config)
        otype = encoding_config["otype"]
        if otype in ("Grid", "HashGrid"):
            assert encoding_config.get("type", "Hash") == "Hash"
            assert encoding_config.get("interpolation", "Linear") == "Linear"
            L = int(encoding_config.get("n_levels", 16)); F = int(encoding_config.get("n_features_per_level", 2))
            assert F == 2, "n_features_per_level must be 2"
            self._meta, n_entries = grid_meta(L, F, int(encoding_config.get("log2_hashmap_size", 19)),
                                              int(encoding_config.get("base_resolution", 16)),
                                              float(encoding_config.get("per_level_scale", 2.0)))
            self.kind = "grid"
            self.n_output_dims = L * F
            g = torch.Generator().manual_seed(seed)
            # tinycudann initialises grid parameters U(-1e-4, 1e-4)
            self.params = nn.Parameter((torch.rand(n_entries * F, generator=g) * 2 - 1) * 1e-4)
            self._shadow = _F16Shadow()
        elif otype == "SphericalHarmonics":
            assert int(encoding_config.get("degree", 4)) == 4
            self.kind = "sh"
            self.n_output_dims = 16
            self.params = nn.Parameter(torch.zeros(0))   # keeps the reference's state-dict key
        else:
            raise ValueError(f"unsupported encoding {otype}")

    def forward(self, x):
        if x.requires_grad and torch.is_grad_enabled():
            # tinycudann returns dL/dx for its encodings; this build does not (the Rad-NeRF path never differentiates through the
            # sample positions unless --optimize_ext is set, train_ml.py:90-147).  Fail loudly instead of returning zero gradients.
            raise RuntimeError("radnerf_b200.tcnn.Encoding does not propagate gradients to its input (pose / extrinsics "
                               "optimisation, --optimize_ext, is not supported); detach the input")
        if self.kind == "grid":
            return _HashGridFn.apply(x, self.params, self)
        x = x.detach().float().contiguous()
        _lib.check_input(x, "x", torch.float32)
        out = torch.empty(x.shape[0], 16, dtype=torch.float16, device=x.device)
        call("rnb_sh4_fwd", x, i64(x.shape[0]), out, stream_arg(x.device))
        return out


class _MlpFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, params, module):
        # x: (N, in_pad) fp16 contiguous
        w = module._shadow.get(params)
        n = x.shape[0]
        dev = x.device
        out = torch.empty(n, module._desc.out_width, dtype=torch.float16, device=dev)
        need_bw = ctx.needs_input_grad[0] or ctx.needs_input_grad[1]   # (grad mode is off inside forward)
        hidden = torch.empty(module._desc.n_hidden, n, 64, dtype=torch.float16, device=dev) if need_bw else None
        call("rnb_mlp_fwd", ctypes.byref(module._desc), w, x, i64(n), out, hidden, stream_arg(dev))
        if need_bw:
            ctx.save_for_backward(x, hidden, out, w)
            ctx.module = module
            ctx.need_dx = ctx.needs_input_grad[0]
        return out

    @staticmethod
    def backward(ctx, dy):
        x, hidden, out, w = ctx.saved_tensors
        m = ctx.module
        n = x.shape[0]
        dev = x.device
        dy = (dy.float() * LOSS_SCALE).half().contiguous()
        dx = torch.empty_like(x) if ctx.need_dx else None
        gw = torch.zeros(m.params.numel(), dtype=torch.float32, device=dev)
        call("rnb_mlp_bwd", ctypes.byref(m._desc), w, x, hidden, out, dy, i64(n), dx, gw, f32(1.0 / LOSS_SCALE),
             stream_arg(dev))
        if dx is not None:
            dx = (dx.float() / LOSS_SCALE).half()
        return dx, gw, None


class Network(nn.Module):
    def __init__(self, n_input_dims, n_output_dims, network_config, seed=1337):
        super().__init__()
        otype = network_config.get("otype", "FullyFusedMLP")
        assert otype in ("FullyFusedMLP", "CutlassMLP", "CUTLASSMLP")
        assert int(network_config.get("n_neurons", 64)) == 64
        assert network_config.get("activation", "ReLU") == "ReLU"
        out_act = network_config.get("output_activation", "None")
        assert out_act in ("None", "Sigmoid")
        self.n_input_dims, self.n_output_dims = n_input_dims, n_output_dims
        self.network_config = dict(network_config)
        self.in_pad = (n_input_dims + 15) // 16 * 16
        self.out_pad = (n_output_dims + 15) // 16 * 16
        n_hidden = int(network_config.get("n_hidden_layers", 1))
        self._desc = _MlpDesc(self.in_pad, self.out_pad, n_hidden, 1 if out_act == "Sigmoid" else 0)
        self.shapes = [(64, self.in_pad)] + [(64, 64)] * (n_hidden - 1) + [(self.out_pad, 64)]
        g = torch.Generator().manual_seed(seed)
        parts = []
        for a, b in self.shapes:   # tinycudann: xavier_uniform per matrix
            s = math.sqrt(6.0 / (a + b))
            parts.append((torch.rand(a * b, generator=g) * 2 - 1) * s)
        self.params = nn.Parameter(torch.cat(parts))
        self._shadow = _F16Shadow()

    def forward(self, x):
        if x.shape[1] != self.n_input_dims:
            raise RuntimeError(f"expected {self.n_input_dims} input dims, got {x.shape[1]}")
        if x.dtype != torch.float16:
            x = x.half()
        if self.in_pad != self.n_input_dims:   # tinycudann pads the input with ones
            x = torch.cat([x, torch.ones(x.shape[0], self.in_pad - self.n_input_dims, dtype=x.dtype, device=x.device)], 1)
        x = x.contiguous()
        if not x.is_cuda:
            raise RuntimeError("x must be a CUDA tensor")
        y = _MlpFn.apply(x, self.params, self)
        return y[:, :self.n_output_dims]
