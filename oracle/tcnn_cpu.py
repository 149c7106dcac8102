"""TEST INFRASTRUCTURE ONLY -- plain PyTorch (CPU, fp32 + explicit fp16 rounding points)
restatement of the tinycudann objects the reference instantiates
(/root/reference/models/networks.py:42-89, 234-289, 1075-1085).

PARITY UNPINNED: tinycudann is an external, un-vendored, un-pinned dependency of the
reference (requirements.txt:14-15 holds only an install comment) and is absent from
/root/reference, this container and the GPU box, so nothing here could be checked against
it.  What follows restates its PUBLISHED algorithm (NVlabs/tiny-cuda-nn, grid.h /
spherical_harmonics.h / fully_fused_mlp.cu) as SURVEY.md section 2.4 records it:

  Grid/Hash encoding (L levels, F=2 features, T=2^log2_T, N_min, per_level_scale b):
    scale_l = exp2(l*log2(b))*N_min - 1 ; res_l = ceil(scale_l)+1
    params_l = min(round_up(res_l^3, 8), T)  (level-major flat table, F features per entry)
    pos = x*scale_l + 0.5 ; cell = floor(pos) ; w = pos - cell
    corner index = dense (x + y*res + z*res^2) when res^3 <= params_l, else
                   (x*1 ^ y*2654435761 ^ z*805459861), always  % params_l
    out = trilinear blend of the 8 corners (accumulated in fp32 here, rounded to fp16 once)
  SphericalHarmonics degree 4 on d = 2u-1.
  FullyFusedMLP: bias-free, fp16 weights/activations (fp32 accumulate here), row-major
    (out,in) matrices concatenated [first, hidden..., last]; input padded to a multiple of
    16 WITH ONES, output padded to a multiple of 16; ReLU hidden, None/Sigmoid output.

Rounding points (the "fp16-MLP mode" of BASELINE.json): parameters are rounded to fp16
before use, every layer output is rounded to fp16.  Autograd gives the backward oracle.
"""
from __future__ import annotations

import math

import numpy as np
import torch

PRIMES = (1, 2654435761, 805459861)


def h16(x: torch.Tensor) -> torch.Tensor:
    """Round to fp16 and back (straight-through for autograd)."""
    return x + (x.half().float() - x).detach()


_LIBM = None


def _libm():
    import ctypes
    global _LIBM
    if _LIBM is None:
        _LIBM = ctypes.CDLL("libm.so.6")
        for f in ("exp2f", "log2f", "ceilf"):
            getattr(_LIBM, f).restype = ctypes.c_float
            getattr(_LIBM, f).argtypes = [ctypes.c_float]
    return _LIBM


def level_scale_res(level, b, base_res, rule="tcnn_fp32"):
    """tinycudann grid.h: grid_scale() = exp2f(level * log2_per_level_scale) * base_resolution - 1.0f and
    grid_resolution() = (uint32_t)ceilf(scale) + 1, evaluated on the host in fp32 by the encoding's constructor with
    log2_per_level_scale = std::log2(float per_level_scale).  `rule="fp64_snap"` is the round-1 rule (fp64, 6-decimal
    snap), kept for comparison: it differs where the real-valued scale is an integer (scene scale 0.5: levels 5/10/15)."""
    import ctypes
    if rule == "fp64_snap":
        s = round(float(np.exp2(level * np.log2(b)) * base_res - 1.0), 6)
        return s, int(np.ceil(s)) + 1
    m = _libm()
    f = np.float32
    log2b = f(m.log2f(ctypes.c_float(f(b))))
    e = f(f(level) * log2b)
    s = f(f(f(m.exp2f(ctypes.c_float(e))) * f(base_res)) - f(1.0))
    return float(s), int(m.ceilf(ctypes.c_float(s))) + 1


class GridMeta:
    def __init__(self, n_levels=16, n_features=2, log2_T=19, base_res=16, per_level_scale=1.3195079565048218, rule=None):
        import os
        rule = rule or os.environ.get("RNB_GRID_LEVEL_RULE", "tcnn_fp32")
        self.L, self.F, self.log2_T, self.N_min = n_levels, n_features, log2_T, base_res
        self.b = float(per_level_scale)
        self.scales, self.res, self.offsets, self.sizes = [], [], [0], []
        for l in range(n_levels):
            s, r = level_scale_res(l, self.b, base_res, rule)
            n = min(r ** 3, 2 ** 31 - 1)
            n = (n + 7) // 8 * 8
            n = min(n, 1 << log2_T)
            self.scales.append(np.float32(s)); self.res.append(r); self.sizes.append(n)
            self.offsets.append(self.offsets[-1] + n)
        self.n_entries = self.offsets[-1]
        self.n_params = self.n_entries * n_features


def grid_indices(meta: GridMeta, x: torch.Tensor, level: int):
    """x (N,3) in [0,1] -> (idx (N,8) int64 into the level's slice, w (N,8) fp32)."""
    s = float(meta.scales[level]); res = meta.res[level]; size = meta.sizes[level]
    pos = torch.addcmul(torch.full_like(x, 0.5), x, torch.tensor(s, dtype=x.dtype))   # fma(scale, x, 0.5) up to 1 rounding
    cell = torch.floor(pos)
    w = pos - cell
    cell = cell.long()
    idxs, ws = [], []
    for c in range(8):
        wt = torch.ones(x.shape[0], dtype=x.dtype)
        p = []
        for dim in range(3):
            if (c >> dim) & 1:
                wt = wt * w[:, dim]; p.append(cell[:, dim] + 1)
            else:
                wt = wt * (1 - w[:, dim]); p.append(cell[:, dim])
        p = [(q & 0xFFFFFFFF) for q in p]
        # tcnn grid_index: accumulate dense index while stride <= hashmap_size; hash if the grid does not fit
        stride, index = 1, torch.zeros_like(p[0])
        for dim in range(3):
            if stride <= size:
                index = (index + p[dim] * stride) & 0xFFFFFFFF
                stride *= res
        if size < stride:
            index = ((p[0] * PRIMES[0]) ^ ((p[1] * PRIMES[1]) & 0xFFFFFFFF) ^ ((p[2] * PRIMES[2]) & 0xFFFFFFFF)) & 0xFFFFFFFF
        idxs.append(index % size); ws.append(wt)
    return torch.stack(idxs, 1), torch.stack(ws, 1)


def hashgrid_encode(meta: GridMeta, params: torch.Tensor, x: torch.Tensor, round_out=True) -> torch.Tensor:
    """params flat fp32 (n_entries*F,), x (N,3) fp32 in [0,1] -> (N, L*F)."""
    table = h16(params).view(-1, meta.F)
    outs = []
    for l in range(meta.L):
        idx, w = grid_indices(meta, x.detach(), l)
        v = table[meta.offsets[l] + idx]                  # (N,8,F)
        outs.append((v * w[..., None]).sum(1))
    out = torch.cat(outs, 1)
    return h16(out) if round_out else out


_SH = None


def sh4(u: torch.Tensor, round_out=True) -> torch.Tensor:
    """u (N,3) in [0,1] (the reference feeds (d/|d|+1)/2, networks.py:322-323) -> (N,16)."""
    x, y, z = (u[:, 0] * 2 - 1), (u[:, 1] * 2 - 1), (u[:, 2] * 2 - 1)
    xy, xz, yz, x2, y2, z2 = x * y, x * z, y * z, x * x, y * y, z * z
    o = [
        torch.full_like(x, 0.28209479177387814),
        -0.48860251190291987 * y,
        0.48860251190291987 * z,
        -0.48860251190291987 * x,
        1.0925484305920792 * xy,
        -1.0925484305920792 * yz,
        0.94617469575755997 * z2 - 0.31539156525251999,
        -1.0925484305920792 * xz,
        0.54627421529603959 * x2 - 0.54627421529603959 * y2,
        0.59004358992664352 * y * (-3.0 * x2 + y2),
        2.8906114426405538 * xy * z,
        0.45704579946446572 * y * (1.0 - 5.0 * z2),
        0.3731763325901154 * z * (5.0 * z2 - 3.0),
        0.45704579946446572 * x * (1.0 - 5.0 * z2),
        1.4453057213202769 * z * (x2 - y2),
        0.59004358992664352 * x * (-x2 + 3.0 * y2),
    ]
    out = torch.stack(o, 1)
    return h16(out) if round_out else out


class MLPMeta:
    def __init__(self, n_in, n_out, n_hidden_layers, n_neurons=64, out_act="None"):
        assert n_neurons == 64
        self.n_in, self.n_out, self.n_hidden, self.width, self.out_act = n_in, n_out, n_hidden_layers, n_neurons, out_act
        self.in_pad = (n_in + 15) // 16 * 16
        self.out_pad = (n_out + 15) // 16 * 16
        self.shapes = [(64, self.in_pad)] + [(64, 64)] * (n_hidden_layers - 1) + [(self.out_pad, 64)]
        self.n_params = sum(a * b for a, b in self.shapes)

    def split(self, params):
        ws, o = [], 0
        for a, b in self.shapes:
            ws.append(params[o:o + a * b].view(a, b)); o += a * b
        return ws

    def init(self, seed=0):
        """Xavier-uniform per matrix (tcnn default), flat fp32."""
        g = torch.Generator().manual_seed(seed)
        out = []
        for a, b in self.shapes:
            s = math.sqrt(6.0 / (a + b))
            out.append((torch.rand(a * b, generator=g) * 2 - 1) * s)
        return torch.cat(out)


def mlp_forward(meta: MLPMeta, params: torch.Tensor, x: torch.Tensor, return_padded=False) -> torch.Tensor:
    """x (N,n_in) -> (N,n_out) with fp16 rounding after every layer."""
    N = x.shape[0]
    if meta.in_pad != meta.n_in:
        x = torch.cat([x, torch.ones(N, meta.in_pad - meta.n_in, dtype=x.dtype)], 1)
    h = h16(x)
    ws = [h16(w) for w in meta.split(params)]
    for w in ws[:-1]:
        h = h16(torch.relu(h @ w.t()))
    y = h @ ws[-1].t()
    if meta.out_act == "Sigmoid":
        y = torch.sigmoid(y)
    y = h16(y)
    return y if return_padded else y[:, :meta.n_out]


def trunc_exp(x):
    """custom_functions.py:162-173: exp fw; bw multiplies by exp(clamp(x,-15,15))."""
    class _F(torch.autograd.Function):
        @staticmethod
        def forward(ctx, x):
            ctx.save_for_backward(x)
            return torch.exp(x)

        @staticmethod
        def backward(ctx, g):
            (x,) = ctx.saved_tensors
            return g * torch.exp(x.clamp(-15, 15))
    return _F.apply(x.float())
