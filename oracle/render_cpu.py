"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the Rad-NeRF render path end to end:
gate -> K x (AABB, march, field, composite, background) -> gated blend -> NeRFLoss.

Follows /root/reference/models/ml_rendering.py:11-78, :158-202 (train) and :81-155 (test),
models/networks.py:291-328 (MNGP.density/forward), :1087-1093 (Ray_Gate), losses.py:52-74.
Marching / AABB use the C restatement (oracle/vren_cpu.py, bit-exact vs the reference kernels);
compositing is restated in plain PyTorch so that autograd provides the backward oracle; the
tinycudann objects are restated in oracle/tcnn_cpu.py (parity unpinned, see there).
This is also bench.py's `cpu_baseline` / `--impl reference` workload ("the reference's
volume-rendering math in plain PyTorch on the host cores").
"""
from __future__ import annotations

import numpy as np
import torch

from . import tcnn_cpu as T
from . import vren_cpu as O

MAX_SAMPLES = 1024
NEAR_DISTANCE = 0.01


class FieldParams:
    """Flat fp32 parameter vectors in tinycudann layout, keyed like the reference state-dict."""

    def __init__(self, scale, K, log2_T=19, seed=0, grid_init=1e-4):
        self.scale, self.K = scale, K
        b = float(np.exp(np.log(2048 * scale / 16) / 15))
        self.grid = T.GridMeta(16, 2, log2_T, 16, b)
        self.geo = T.MLPMeta(32, 17, 1, out_act="None")
        self.rgb = T.MLPMeta(32, 3, 2, out_act="Sigmoid")
        self.gate = T.MLPMeta(6, K, 4, out_act="None")
        g = torch.Generator().manual_seed(seed)
        self.p = {"xyz_encoder.params": (torch.rand(self.grid.n_params, generator=g) * 2 - 1) * grid_init}
        for i in range(K):
            self.p[f"geo_net_{i}.params"] = self.geo.init(seed=seed * 1000 + 10 + i)
            self.p[f"rgb_net_{i}.params"] = self.rgb.init(seed=seed * 1000 + 50 + i)
        self.p["gate.encoder.params"] = self.gate.init(seed=seed * 1000 + 99)

    def requires_grad_(self, flag=True):
        for v in self.p.values():
            v.requires_grad_(flag)
        return self


def field(P: FieldParams, i, xyzs, dirs):
    """MNGP.forward (networks.py:311-328): -> sigmas (N) f32, rgbs (N,3) (fp16-rounded values)."""
    lo, hi = -P.scale, P.scale
    x = ((xyzs - lo) / (hi - lo)).clip(0.0, 1.0)
    feat = T.hashgrid_encode(P.grid, P.p["xyz_encoder.params"], x)
    h = T.mlp_forward(P.geo, P.p[f"geo_net_{i}.params"], feat)
    sigmas = T.trunc_exp(h[:, 0])
    d = dirs / torch.norm(dirs, dim=1, keepdim=True)
    sh = T.sh4((d + 1) / 2)
    rgbs = T.mlp_forward(P.rgb, P.p[f"rgb_net_{i}.params"], torch.cat([sh, h[:, 1:]], 1))
    return sigmas, rgbs


def density(P: FieldParams, i, xyzs):
    lo, hi = -P.scale, P.scale
    x = ((xyzs - lo) / (hi - lo)).clip(0.0, 1.0)
    h = T.mlp_forward(P.geo, P.p[f"geo_net_{i}.params"], T.hashgrid_encode(P.grid, P.p["xyz_encoder.params"], x))
    return T.trunc_exp(h[:, 0])


def gate(P: FieldParams, rays_o, rays_d):
    """Ray_Gate.forward (networks.py:1087-1093)."""
    logits = T.mlp_forward(P.gate, P.p["gate.encoder.params"], torch.cat([rays_o, rays_d], 1))
    g = torch.softmax(logits.float(), 1)
    return g, g.sum(0)


def composite_train(sigmas, rgbs, deltas, ts, rays_a, T_threshold=1e-4):
    """Differentiable plain-PyTorch statement of volumerendering.cu:23-44 on padded segments.
    -> opacity (R), depth (R), rgb (R,3), ws (N), total_samples (R) i64 (all indexed by ray_idx)."""
    rays_a = torch.as_tensor(rays_a)
    R = rays_a.shape[0]
    lens = rays_a[:, 2]
    Lmax = int(lens.max()) if R else 0
    N = sigmas.shape[0]
    if Lmax == 0:
        z = torch.zeros(R)
        return z, z.clone(), torch.zeros(R, 3), torch.zeros(N), torch.zeros(R, dtype=torch.int64)
    pos = torch.arange(Lmax)[None, :]
    valid = pos < lens[:, None]
    idx = (rays_a[:, 1:2] + pos).clamp(max=max(N - 1, 0))
    a = torch.where(valid, 1.0 - torch.exp(-sigmas[idx] * deltas[idx]), torch.zeros(()))
    T_incl = torch.cumprod(1.0 - a, 1)
    T_excl = torch.cat([torch.ones(R, 1), T_incl[:, :-1]], 1)
    term = (T_incl <= T_threshold) & valid                         # the sample on which the ray stops
    stopped_before = torch.cumsum(term.long(), 1) - term.long() > 0
    live = valid & ~stopped_before
    w = torch.where(live, a * T_excl, torch.zeros(()))
    opacity = w.sum(1)
    depth = (w * ts[idx]).sum(1)
    rgb = (w[..., None] * rgbs[idx]).sum(1)
    ws = torch.zeros(N, dtype=w.dtype).index_put((idx[valid],), w[valid])
    first_term = torch.where(term.any(1), term.float().argmax(1), lens)
    out_o = torch.zeros(R).index_put((rays_a[:, 0],), opacity)
    out_d = torch.zeros(R).index_put((rays_a[:, 0],), depth)
    out_c = torch.zeros(R, 3).index_put((rays_a[:, 0],), rgb)
    out_t = torch.zeros(R, dtype=torch.int64).index_put((rays_a[:, 0],), first_term)
    return out_o, out_d, out_c, ws, out_t


def hits_for(rays_o, rays_d, scale):
    _, hits_t, _ = O.ray_aabb_intersect(rays_o, rays_d, np.zeros(3, np.float32), np.full(3, scale, np.float32))
    h = np.ascontiguousarray(hits_t[:, 0])
    m = (h[:, 0] >= 0) & (h[:, 0] < NEAR_DISTANCE)
    h[m, 0] = NEAR_DISTANCE
    return h


def ml_render_train(P: FieldParams, rays_o, rays_d, bitfields, noises, cascades, exp_step_factor=0.0,
                    T_threshold=1e-4, field_fn=None, bgs=None):
    """ml_render(test_time=False).  rays_* numpy (N,3); bitfields[i] numpy u8; noises[i] numpy (N).
    field_fn(i, xyzs, dirs) overrides the learned field (e.g. the analytic scene).
    -> dict like the reference's (rgb, independent_rgbs, depth (N,K), opacity, gating_code,
       gating_importance) + per-i bookkeeping under 'per_i'."""
    K = P.K
    to = torch.from_numpy(np.ascontiguousarray(rays_o)); td = torch.from_numpy(np.ascontiguousarray(rays_d))
    code, importance = gate(P, to, td)
    n = len(rays_o)
    h = hits_for(rays_o, rays_d, P.scale)
    rgb_res = torch.zeros(n, 3); opa_res = torch.zeros(n); depth_cols = []; ind = []; per_i = []
    for i in range(K):
        rays_a, xyzs, dirs, deltas, ts, counter = O.raymarching_train(rays_o, rays_d, h, bitfields[i], cascades, P.scale,
                                                                      exp_step_factor, noises[i])
        xyzs_t, dirs_t = torch.from_numpy(xyzs), torch.from_numpy(dirs)
        sig, col = field_fn(i, xyzs_t, dirs_t) if field_fn else field(P, i, xyzs_t, dirs_t)
        opacity, depth, rgb, ws, tot = composite_train(sig.float(), col.float(), torch.from_numpy(deltas), torch.from_numpy(ts),
                                                       rays_a, T_threshold)
        bg = bgs[i] if bgs is not None else (torch.ones(3) if exp_step_factor == 0 else torch.zeros(3))
        rgb = rgb + bg * (1 - opacity)[:, None]
        ind.append(rgb); depth_cols.append(depth)
        rgb_res = rgb_res + rgb * code[:, i:i + 1]
        opa_res = opa_res + opacity * code[:, i]
        per_i.append({"rays_a": rays_a, "ws": ws, "vr_samples": int(tot.sum()), "rm_samples": int(counter[0]),
                      "sigmas": sig, "rgbs": col, "xyzs": xyzs, "dirs": dirs, "deltas": deltas, "ts": ts})
    return {"rgb": rgb_res, "independent_rgbs": ind, "depth": torch.stack(depth_cols, 1), "opacity": opa_res,
            "gating_code": code, "gating_importance": importance, "per_i": per_i}


def nerf_loss(results, target_rgb, lambda_opacity=1e-3, lambda_cv=0.0, lambda_dm=0.0):
    """losses.py:52-74 + the reduction of train_ml.py:192 (sum of term means)."""
    loss = {"rgb": (results["rgb"] - target_rgb) ** 2}
    o = results["opacity"] + 1e-10
    loss["opacity"] = lambda_opacity * (-o * torch.log(o))
    K = results["gating_code"].shape[-1]
    if lambda_cv > 0 and K > 1:
        imp = results["gating_importance"].float()
        loss["cv_importance"] = lambda_cv * (imp.var() / (imp.mean() ** 2 + 1e-10))
    if lambda_dm > 0 and K > 1:
        fused = torch.sum(results["depth"] * results["gating_code"], 1, keepdim=True).detach()
        loss["depth_mutual"] = lambda_dm * ((results["depth"] - fused) ** 2)
    return sum(v.mean() for v in loss.values()), loss


def render_test(P: FieldParams, rays_o, rays_d, bitfield, i, cascades, exp_step_factor=0.0, T_threshold=1e-4,
                field_fn=None, max_samples=MAX_SAMPLES):
    """__render_rays_test (ml_rendering.py:81-155) for sub-NeRF i."""
    n = len(rays_o)
    hits = hits_for(rays_o, rays_d, P.scale)
    opacity = np.zeros(n, np.float32); depth = np.zeros(n, np.float32); rgb = np.zeros((n, 3), np.float32)
    alive = np.arange(n, dtype=np.int64)
    samples = 0
    min_samples = 1 if exp_step_factor == 0 else 4
    total = 0
    while samples < max_samples:
        if len(alive) == 0:
            break
        Ns = max(min(n // len(alive), 64), min_samples)
        samples += Ns
        x, d, dl, ts, ne = O.raymarching_test(rays_o, rays_d, hits, alive, bitfield, cascades, P.scale, exp_step_factor,
                                              128, MAX_SAMPLES, Ns)
        total += int(ne.sum())
        valid = (np.arange(Ns)[None, :] < ne[:, None]).reshape(-1)
        if valid.sum() == 0:
            break
        sig = np.zeros(len(valid), np.float32); col = np.zeros((len(valid), 3), np.float32)
        xv = torch.from_numpy(x.reshape(-1, 3)[valid]); dv = torch.from_numpy(d.reshape(-1, 3)[valid])
        with torch.no_grad():
            s_, c_ = field_fn(i, xv, dv) if field_fn else field(P, i, xv, dv)
        sig[valid] = s_.float().numpy(); col[valid] = c_.float().numpy()
        O.composite_test_fw(sig.reshape(-1, Ns), col.reshape(-1, Ns, 3), dl, ts, hits, alive, T_threshold, ne, opacity, depth, rgb)
        alive = np.ascontiguousarray(alive[alive >= 0])
    bg = np.ones(3, np.float32) if exp_step_factor == 0 else np.zeros(3, np.float32)
    rgb = rgb + bg * (1 - opacity)[:, None]
    return {"opacity": opacity, "depth": depth, "rgb": rgb, "total_samples": total}


# ---------------------------------------------------------------------------------------------------------------
# Occupancy-grid update: MNGP.update_density_grid (networks.py:374-409) with its helpers get_all_cells (:330-342)
# and sample_uniform_and_occupied_cells (:344-372), for ONE sub-NeRF i.  Every random draw of the reference
# (`torch.randint` coords1 :359, `torch.randint(len(indices2))` :365, `torch.rand_like` jitter :391) is an INPUT
# here so that the CUDA path can be fed the same numbers:
#   coords1 (C, M, 3) int32 in [0,G)                              -- the uniform cells (M = G^3 // 4, :381)
#   pick_u  (C, M)    f32  in [0,1)   rand_idx = min(int(u*n), n-1) -- stands for randint(n) over the occupied list
#   noise   (C, M_c, 3) f32 in [0,1)                               -- the jitter, M_c = G^3 (warm-up) or 2M
# Returns a dict:
#   cells[c]  int64 indices evaluated in cascade c, in the reference's order (uniform cells, then occupied picks)
#   sigma[c]  density at each of them
#   tmp_last  density_grid_tmp as numpy's last-write-wins scatter leaves it (ONE legal outcome of the reference's
#             index_put with duplicate indices, whose winner CUDA leaves undefined), with tmp_lo / tmp_hi the
#             smallest / largest candidate per cell (every legal outcome lies between them)
#   grid, grid_lo, grid_hi  where(grid<0, grid, max(grid*decay, tmp)) for the three tmp variants (:397-400)
#   mean      grid[grid>0].mean() (:402);  thr = min(mean, density_threshold);  bitfield = packbits(grid, thr) (:403-404)
def update_density_grid(P: FieldParams, i, density_grid, density_threshold, warmup, noise, coords1=None, pick_u=None,
                        decay=0.95, grid_size=128):
    G = grid_size
    G3 = G ** 3
    grid = np.ascontiguousarray(density_grid, dtype=np.float32)
    C = grid.shape[0]
    cells, coords = [], []
    if warmup:                                                     # get_all_cells: every cell, morton order = index order
        all_idx = np.arange(G3, dtype=np.int32)
        all_coords = O.morton3D_invert(all_idx)
        for c in range(C):
            cells.append(all_idx.astype(np.int64)); coords.append(all_coords)
    else:
        for c in range(C):
            c1 = np.ascontiguousarray(coords1[c], dtype=np.int32)
            idx1 = O.morton3D(c1).astype(np.int64)
            occ = np.nonzero(grid[c] > np.float32(density_threshold))[0]
            if len(occ) > 0:
                n = len(occ)
                ridx = np.minimum((np.asarray(pick_u[c], np.float32) * np.float32(n)).astype(np.int64), n - 1)
                idx2 = occ[ridx]
                c2 = O.morton3D_invert(idx2.astype(np.int32))
                cells.append(np.concatenate([idx1, idx2])); coords.append(np.concatenate([c1, c2]))
            else:
                cells.append(idx1); coords.append(c1)
    tmp_last = np.zeros_like(grid); tmp_lo = np.full_like(grid, np.inf); tmp_hi = np.zeros_like(grid)
    touched = np.zeros(grid.shape, bool)
    sigmas = []
    for c in range(C):
        s = min(2 ** (c - 1), P.scale)
        half = s / G
        cc = torch.from_numpy(coords[c].astype(np.float32))
        nz = torch.from_numpy(np.ascontiguousarray(noise[c][:len(cells[c])], dtype=np.float32))
        xyzs_w = (cc / (G - 1) * 2 - 1) * (s - half)
        xyzs_w = xyzs_w + (nz * 2 - 1) * half
        with torch.no_grad():
            sig = density(P, i, xyzs_w).numpy().astype(np.float32)
        sigmas.append(sig)
        tmp_last[c][cells[c]] = sig
        np.minimum.at(tmp_lo[c], cells[c], sig); np.maximum.at(tmp_hi[c], cells[c], sig)
        touched[c][cells[c]] = True
    tmp_lo[~touched] = 0.0

    def ema(t):
        return np.where(grid < 0, grid, np.maximum(grid * np.float32(decay), t)).astype(np.float32)
    new = ema(tmp_last)
    mean = float(new[new > 0].astype(np.float64).mean()) if (new > 0).any() else 0.0
    thr = min(mean, float(density_threshold))
    return {"cells": cells, "sigma": sigmas, "touched": touched, "tmp_last": tmp_last, "tmp_lo": tmp_lo, "tmp_hi": tmp_hi,
            "grid": new, "grid_lo": ema(tmp_lo), "grid_hi": ema(tmp_hi), "mean": mean, "thr": thr,
            "bitfield": O.packbits(new, thr)}
