"""Fused occupancy-grid update (rnb_grid_update, csrc/grid.cu + density_cells_tc_kernel) against the drop-in
Python body of MNGP.update_density_grid (networks.py:374-409) evaluated with the SAME random inputs.

Bars: the cell lists (ordered compaction = torch.nonzero, morton codes, picks) are integer work and must match
exactly -- checked through which cells change; densities go through fp16 MLPs: 1e-2 relative; the bitfield must be
bit-exact packbits(grid, min(mean(grid > 0), threshold)) of the grid the kernel produced; the mean within 1e-5.
"""
import numpy as np
import pytest
import torch

from radnerf_b200 import synthetic as syn

pytestmark = pytest.mark.gpu
G = 128
G3 = G ** 3


def _model(scale, K):
    from radnerf_b200.engine import Engine
    from radnerf_b200.models.networks import MNGP, Ray_Gate
    torch.manual_seed(0)
    model = MNGP(scale, size=K).cuda()
    gate = Ray_Gate(K).cuda()
    with torch.no_grad():
        model.xyz_encoder.params.uniform_(-0.5, 0.5)
    eng = Engine(model, gate, n_rays=1024, exp_step_factor=syn.exp_step_factor_for_scale(scale))
    return model, eng


def _points(model, coords, c, noise):
    s = min(2 ** (c - 1), model.scale)
    half = s / G
    return (coords / (G - 1) * 2 - 1) * (s - half) + (noise * 2 - 1) * half


def _check_bitfield_and_mean(model, eng, K, thr):
    mean = eng.grid_mean().cpu()
    for i in range(K):
        g = getattr(model, f"density_grid_{i}")
        ref_mean = float(g[g > 0].double().mean())
        assert abs(float(mean[i]) - ref_mean) <= 1e-5 * ref_mean
        exp = syn.packbits_np(g.cpu().numpy(), min(float(mean[i]), thr))
        assert np.array_equal(getattr(model, f"density_bitfield_{i}").cpu().numpy(), exp)


@pytest.mark.parametrize("scale", [0.5, 2.0])
def test_grid_update_warmup(scale):
    from radnerf_b200 import vren
    K = 2
    model, eng = _model(scale, K)
    C = model.cascades
    torch.manual_seed(1)
    noise = torch.rand(K, C, G3, 3, device="cuda")
    thr = syn.DENSITY_THRESHOLD
    # reference: the drop-in Python path, cell m = morton index m
    coords = vren.morton3D_invert(torch.arange(G3, dtype=torch.int32, device="cuda")).float()
    expect = []
    with torch.no_grad():
        for i in range(K):
            tmp = torch.zeros(C, G3, device="cuda")
            for c in range(C):
                tmp[c] = model.density(_points(model, coords, c, noise[i, c]), i)
            g = getattr(model, f"density_grid_{i}")
            expect.append(torch.where(g < 0, g, torch.maximum(g * 0.95, tmp)))
    eng.update_grid(thr, warmup=True, noise=noise)
    for i in range(K):
        g = getattr(model, f"density_grid_{i}")
        rel = (g - expect[i]).abs() / expect[i].abs().clamp_min(1e-6)
        assert float(rel.max()) < 1e-2, float(rel.max())
    _check_bitfield_and_mean(model, eng, K, thr)
    # built-in generator: every cell of every cascade is touched, values are plausible densities
    for i in range(K):
        getattr(model, f"density_grid_{i}").zero_()
    eng.update_grid(thr, warmup=True)
    for i in range(K):
        g = getattr(model, f"density_grid_{i}")
        assert bool((g > 0).all())
        assert abs(float(g.mean()) - float(expect[i].mean())) < 0.05 * float(expect[i].mean())
    _check_bitfield_and_mean(model, eng, K, thr)


def test_grid_update_steady_state():
    from radnerf_b200 import vren
    K, scale = 2, 0.5
    model, eng = _model(scale, K)
    C = model.cascades
    eng.update_grid(1e9, warmup=True)                    # fill the grids
    thr = float(torch.quantile(model.density_grid_0[0, ::64], 0.9))    # ~10 % of the cells count as occupied
    M1 = G3 // 4
    torch.manual_seed(2)
    coords1 = torch.randint(G, (K, C, M1, 3), dtype=torch.int32, device="cuda")
    pick_u = torch.rand(K, C, M1, device="cuda")
    noise = torch.rand(K, C, 2 * M1, 3, device="cuda")
    before = [getattr(model, f"density_grid_{i}").clone() for i in range(K)]
    lo, hi, touched = [], [], []
    with torch.no_grad():
        for i in range(K):
            tl = torch.full((C, G3), float("inf"), device="cuda"); th = torch.zeros(C, G3, device="cuda")
            tm = torch.zeros(C, G3, dtype=torch.bool, device="cuda")
            for c in range(C):
                idx1 = vren.morton3D(coords1[i, c]).long()
                occ = torch.nonzero(before[i][c] > thr)[:, 0]
                assert len(occ) > 1000
                idx2 = occ[(pick_u[i, c] * len(occ)).long().clamp_max(len(occ) - 1)]
                idx = torch.cat([idx1, idx2])
                cc = torch.cat([coords1[i, c], vren.morton3D_invert(idx2.int())]).float()
                sig = model.density(_points(model, cc, c, noise[i, c]), i)
                tl[c].scatter_reduce_(0, idx, sig, "amin"); th[c].scatter_reduce_(0, idx, sig, "amax")
                tm[c, idx] = True
            lo.append(torch.where(tm, tl, torch.zeros_like(tl))); hi.append(th); touched.append(tm)
    eng.update_grid(thr, warmup=False, noise=noise, coords1=coords1, pick_u=pick_u)
    for i in range(K):
        g = getattr(model, f"density_grid_{i}")
        # untouched cells only decay; touched cells take max(decayed, one of the candidate densities)
        un = ~touched[i]
        assert torch.equal(g[un], (before[i] * 0.95)[un])
        low = torch.maximum(before[i] * 0.95, lo[i]); high = torch.maximum(before[i] * 0.95, hi[i])
        t = touched[i]
        assert bool((g[t] >= low[t] * (1 - 1e-2)).all()) and bool((g[t] <= high[t] * (1 + 1e-2)).all())
    _check_bitfield_and_mean(model, eng, K, thr)
    # the scratch grid is left at zero for the next update
    L = eng._grid_layout
    assert int(eng._grid_ws[L.tmp:L.tmp + K * C * G3 * 4].view(torch.float32).count_nonzero()) == 0
    eng.update_grid(thr, warmup=False)                        # built-in generator path runs
    _check_bitfield_and_mean(model, eng, K, thr)


def test_trainer_uses_fused_grid_update():
    from radnerf_b200.trainer import Trainer
    tr = Trainer(scale=0.5, K=2, freeze_occupancy=False, device="cuda", engine=True, batch_rays=1024)
    tr._grid_update()
    for i in range(2):
        assert int(getattr(tr.model, f"density_bitfield_{i}").count_nonzero()) > 0
        assert bool((getattr(tr.model, f"density_grid_{i}") > 0).all())
