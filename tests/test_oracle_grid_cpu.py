"""CPU checks of the occupancy-update oracle (oracle/render_cpu.update_density_grid, restating
/root/reference/models/networks.py:330-409) at a small grid size, against an independent statement of the same steps in
numpy written from the reference text: cell selection (get_all_cells :330-342; sample_uniform_and_occupied_cells :344-372),
cell-centre + jitter positions (:385-391), the EMA-max update (:397-400), the masked mean (:402) and packbits (:403-404).
The density network itself is the (unpinned) tinycudann restatement; here it is evaluated by the oracle's own `density`, so what
these tests pin is the bookkeeping around it.  Also: bench.py prints the identical `config` object in both arms."""
import numpy as np
import pytest
import torch

from oracle import render_cpu as R
from oracle import vren_cpu as O

G = 16
G3 = G ** 3


def _params(scale=0.5, K=2):
    return R.FieldParams(scale, K, log2_T=12, seed=3, grid_init=0.5)


def _expected_positions(coords, c, noise, scale):
    s = min(2 ** (c - 1), scale)
    half = s / G
    return (coords.astype(np.float32) / (G - 1) * 2 - 1) * (s - half) + (noise * 2 - 1) * half


def test_warmup_visits_every_cell_and_applies_the_ema_max_rule():
    P = _params()
    rng = np.random.RandomState(0)
    grid = (rng.rand(1, G3) * 2).astype(np.float32)
    grid[0, rng.rand(G3) < 0.2] = 0.0
    grid[0, rng.rand(G3) < 0.1] = -1.0                      # invisible cells keep their value (networks.py:397-399)
    noise = rng.rand(1, G3, 3).astype(np.float32)
    thr = 0.7
    o = R.update_density_grid(P, 1, grid, thr, True, noise, grid_size=G)
    assert np.array_equal(o["cells"][0], np.arange(G3))     # get_all_cells: indices = morton3D(grid_coords) of all cells, in order
    assert o["touched"].all()
    coords = O.morton3D_invert(np.arange(G3, dtype=np.int32))
    assert np.array_equal(O.morton3D(coords), np.arange(G3))
    x = _expected_positions(coords, 0, noise[0], P.scale)
    with torch.no_grad():
        sig = R.density(P, 1, torch.from_numpy(x)).numpy()
    assert np.allclose(o["sigma"][0], sig, rtol=1e-6, atol=0)
    exp = np.where(grid < 0, grid, np.maximum(grid * np.float32(0.95), sig[None, :])).astype(np.float32)
    assert np.array_equal(o["grid"], exp)
    assert np.array_equal(o["grid_lo"], exp) and np.array_equal(o["grid_hi"], exp)     # no duplicates in warm-up
    mean = exp[exp > 0].astype(np.float64).mean()
    assert abs(o["mean"] - mean) < 1e-12 and o["thr"] == min(mean, thr)
    assert np.array_equal(o["bitfield"], O.packbits(exp, min(mean, thr)))


def test_steady_state_cell_lists_and_duplicate_bounds():
    P = _params()
    rng = np.random.RandomState(1)
    grid = (rng.rand(1, G3) * 2).astype(np.float32)
    M = G3 // 4
    coords1 = rng.randint(0, G, (1, M, 3)).astype(np.int32)
    pick_u = rng.rand(1, M).astype(np.float32)
    noise = rng.rand(1, 2 * M, 3).astype(np.float32)
    thr = 1.5
    o = R.update_density_grid(P, 0, grid, thr, False, noise, coords1, pick_u, grid_size=G)
    occ = np.nonzero(grid[0] > np.float32(thr))[0]          # torch.nonzero order = ascending index (:362)
    assert len(occ) > 10
    idx1 = O.morton3D(coords1[0]).astype(np.int64)
    idx2 = occ[np.minimum((pick_u[0] * np.float32(len(occ))).astype(np.int64), len(occ) - 1)]
    assert np.array_equal(o["cells"][0], np.concatenate([idx1, idx2]))      # uniform cells first, then the picks (:370)
    touched = np.zeros(G3, bool); touched[o["cells"][0]] = True
    assert np.array_equal(o["touched"][0], touched)
    # untouched cells only decay; touched ones lie between the smallest and the largest candidate written to them
    un = ~touched
    assert np.array_equal(o["grid"][0][un], (grid[0] * np.float32(0.95))[un])
    assert np.all(o["grid_lo"] <= o["grid"]) and np.all(o["grid"] <= o["grid_hi"])
    cells, sig = o["cells"][0], o["sigma"][0]
    dup = np.bincount(cells, minlength=G3) > 1
    assert dup.any()                                         # sampling with replacement does produce duplicates
    c = int(np.nonzero(dup)[0][0])
    cand = sig[cells == c]
    assert o["tmp_lo"][0, c] == cand.min() and o["tmp_hi"][0, c] == cand.max() and o["tmp_last"][0, c] == cand[-1]


def test_steady_state_without_occupied_cells_keeps_the_uniform_half():
    P = _params()
    rng = np.random.RandomState(2)
    grid = np.zeros((1, G3), np.float32)
    M = G3 // 4
    o = R.update_density_grid(P, 0, grid, 1e9, False, rng.rand(1, 2 * M, 3).astype(np.float32),
                              rng.randint(0, G, (1, M, 3)).astype(np.int32), rng.rand(1, M).astype(np.float32), grid_size=G)
    assert len(o["cells"][0]) == M                           # networks.py:363-371: `else: cells += [(indices1, coords1)]`


def test_cascade_extents_follow_the_reference():
    """s = min(2^(c-1), scale), half_grid_size = s / G (networks.py:385-388): cascade c's points stay inside [-s, s]."""
    P = R.FieldParams(4.0, 1, log2_T=12, seed=0, grid_init=0.5)
    C = 3
    rng = np.random.RandomState(4)
    noise = rng.rand(C, G3, 3).astype(np.float32)
    o = R.update_density_grid(P, 0, np.zeros((C, G3), np.float32), 1.0, True, noise, grid_size=G)
    coords = O.morton3D_invert(np.arange(G3, dtype=np.int32))
    for c in range(C):
        x = _expected_positions(coords, c, noise[c], P.scale)
        s = min(2 ** (c - 1), P.scale)
        assert np.abs(x).max() <= s + 1e-6 and np.abs(x).max() > 0.8 * s
        with torch.no_grad():
            assert np.allclose(o["sigma"][c], R.density(P, 0, torch.from_numpy(x)).numpy(), rtol=1e-6, atol=0)


def test_bench_arms_print_the_same_config():
    import argparse
    import bench
    for name in ("radnerf_tat_k2", "radnerf_360_k4", "ngp_nsvf_k1_cpu_fwd"):
        syn, w = bench.workload(name)
        a = argparse.Namespace(workload=name)
        assert bench.arm_config(a, w, 1) == bench.arm_config(a, w, 1)
        cfg = bench.arm_config(a, w, 8)
        assert cfg["workload"] == name and cfg["K"] == w["K"] and cfg["batch_rays_per_gpu"] == bench.BATCH
        assert set(cfg) >= {"workload", "K", "scale", "batch_rays_per_gpu", "step", "params", "occupancy", "parallelism", "l2"}
