"""CPU tests of the oracle itself (oracle/vren_oracle.c): analytic known answers and
cross-checks against independent numpy / torch-autograd statements.  The oracle is pinned
against the REAL reference by tests/test_golden.py (fixtures produced by the reference's own
kernels, tests/golden/make_golden_vren.py)."""
import numpy as np
import pytest
import torch

from conftest import CONFIGS, get_scene, make_rays
from oracle import vren_cpu as O
from radnerf_b200 import synthetic as syn


def test_morton_known_answers():
    c = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1], [127, 127, 127], [5, 9, 100]], np.int32)
    idx = O.morton3D(c)
    assert idx[:5].tolist() == [0, 1, 2, 4, 128 ** 3 - 1]
    assert np.array_equal(O.morton3D_invert(idx), c)
    assert np.array_equal(idx.astype(np.uint32), syn.morton3d(c[:, 0], c[:, 1], c[:, 2]))


def test_packbits_known_answers():
    g = np.zeros(64, np.float32); g[[0, 9, 63]] = 1.0; g[10] = 0.5
    assert O.packbits(g, 0.5).tolist() == [1, 2, 0, 0, 0, 0, 0, 128]   # strict '>' : 0.5 is not set
    sc, bf = get_scene(0.5)
    assert np.array_equal(O.packbits(sc.density_grid, sc.threshold), bf)


def test_aabb_known_answers():
    o = np.array([[0, 0, -2], [0, 0, -2], [0, 0, 0], [3, 3, 3]], np.float32)
    d = np.array([[0, 0, 1], [0, 0, -1], [1, 0, 0], [1, 0, 0]], np.float32)
    cnt, ht, idx = O.ray_aabb_intersect(o, d, np.zeros(3), np.full(3, 0.5))
    assert cnt.tolist() == [1, 0, 1, 0]
    assert ht[0, 0].tolist() == [1.5, 2.5]
    assert ht[1, 0].tolist() == [-1, -1] and idx[1, 0] == -1
    assert ht[2, 0].tolist() == [0.0, 0.5]           # origin inside: near clamped to 0


@pytest.mark.parametrize("cfg", list(CONFIGS))
def test_march_train_invariants(cfg):
    scale, esf = CONFIGS[cfg]
    sc, bf = get_scene(scale)
    n = 2048
    o, d, h = make_rays(scale, n)
    noise = np.random.RandomState(7).rand(n).astype(np.float32)
    rays_a, xyzs, dirs, deltas, ts, counter = O.raymarching_train(o, d, h, bf, sc.cascades, scale, esf, noise)
    assert counter[0] == rays_a[:, 2].sum() == len(ts) and counter[1] == n
    assert np.array_equal(rays_a[:, 0], np.arange(n))
    assert np.array_equal(rays_a[:, 1], np.concatenate([[0], np.cumsum(rays_a[:, 2])[:-1]]))
    assert rays_a[:, 2].max() <= 1024 and rays_a[h[:, 0] < 0, 2].sum() == 0
    rid = np.repeat(np.arange(n), rays_a[:, 2])
    # samples lie on their ray, inside [t1, t2), in increasing t, with dirs = rays_d
    assert np.allclose(xyzs, o[rid] + ts[:, None] * d[rid], atol=1e-5 * scale * 4)
    assert np.array_equal(dirs, d[rid])
    assert (ts >= h[rid, 0]).all() and (ts < h[rid, 1]).all()
    same = rid[1:] == rid[:-1]
    assert (np.diff(ts)[same] > 0).all()
    if esf == 0:
        assert np.all(deltas == np.float32(1.73205080757) / np.float32(1024))
    else:
        assert deltas.min() >= np.float32(1.73205080757) / 1024 and deltas.max() <= np.float32(1.73205080757 * 2 * scale / 128) * 1.0001
    # every sample sits in an occupied cell of the cascade the reference selects (raymarching.cu:208-224)
    mx = np.abs(xyzs).max(1)
    mip_pos = np.clip(np.frexp(mx)[1] + 1, 0, sc.cascades - 1)
    mip_dt = np.clip(np.frexp(deltas * np.float32(128))[1], 0, sc.cascades - 1)
    mip = np.maximum(mip_pos, mip_dt)
    bound = np.minimum(np.ldexp(1.0, mip - 1), scale).astype(np.float32)
    cell = np.clip(0.5 * (xyzs / bound[:, None] + 1) * 128, 0, 127).astype(np.uint32)
    idx = mip.astype(np.int64) * 128 ** 3 + syn.morton3d(cell[:, 0], cell[:, 1], cell[:, 2]).astype(np.int64)
    occ = (bf[idx // 8] >> (idx % 8).astype(np.uint8)) & 1
    assert occ.mean() > 0.999   # (numpy recomputation of the cell may round differently on cell faces)


def test_march_test_equals_train_without_noise():
    """Bounded scene: the test-time marcher with a large budget visits the same samples as the
    training marcher with zero noise (dt is constant, so hazard Q1 does not bite)."""
    sc, bf = get_scene(0.5)
    n = 512
    o, d, h = make_rays(0.5, n)
    rays_a, xyzs, _, deltas, ts, _ = O.raymarching_train(o, d, h, bf, 1, 0.5, 0.0, np.zeros(n, np.float32))
    alive = np.arange(n, dtype=np.int64)
    hh = h.copy()
    got_ts = [[] for _ in range(n)]
    for _ in range(1024 // 64):
        x, dd, dl, t, ne = O.raymarching_test(o, d, hh, alive, bf, 1, 0.5, 0.0, 128, 1024, 64)
        for r in range(n):
            got_ts[r].extend(t[r, :ne[r]].tolist())
    for r in range(n):
        s, c = rays_a[r, 1], rays_a[r, 2]
        assert np.array_equal(np.array(got_ts[r], np.float32), ts[s:s + c])


def _torch_composite(sigmas, rgbs, deltas, ts, rays_a, thr):
    """Independent float64 autograd statement of volumerendering.cu:23-44 (per ray, serial)."""
    outs = []
    for ray_idx, start, n in rays_a.tolist():
        s = slice(start, start + n)
        a = 1 - torch.exp(-sigmas[s] * deltas[s])
        T_incl = torch.cumprod(1 - a, 0)
        T_excl = torch.cat([torch.ones(1, dtype=a.dtype), T_incl[:-1]])
        # sample k contributes iff no earlier sample terminated the ray
        alive = torch.cat([torch.ones(1, dtype=torch.bool), (T_incl > thr)[:-1]]) if n else torch.ones(0, dtype=torch.bool)
        alive = torch.cumprod(alive.to(a.dtype), 0)
        w = a * T_excl * alive
        outs.append((ray_idx, w.sum(), (w * ts[s]).sum(), (w[:, None] * rgbs[s]).sum(0), w))
    return outs


def test_composite_fw_bw_against_autograd():
    rng = np.random.RandomState(0)
    lens = [0, 1, 5, 33, 64, 200, 31, 32]
    n = len(lens); N = sum(lens)
    starts = np.concatenate([[0], np.cumsum(lens)[:-1]])
    rays_a = np.stack([np.arange(n)[::-1], starts, lens], 1).astype(np.int64)   # ray_idx != row on purpose
    sig = np.exp(rng.normal(1.0, 2.0, N)).astype(np.float32)
    rgbs = rng.uniform(0, 1, (N, 3)).astype(np.float32)
    dl = rng.uniform(0.002, 0.02, N).astype(np.float32)
    ts = np.concatenate([np.sort(rng.uniform(0.2, 3, l)) for l in lens]).astype(np.float32)
    total, op, dp, col, ws = O.composite_train_fw(sig, rgbs, dl, ts, rays_a, 1e-4)
    tsig = torch.tensor(sig, dtype=torch.float64, requires_grad=True)
    trgb = torch.tensor(rgbs, dtype=torch.float64, requires_grad=True)
    outs = _torch_composite(tsig, trgb, torch.tensor(dl, dtype=torch.float64), torch.tensor(ts, dtype=torch.float64), rays_a, 1e-4)
    gO = rng.normal(size=n).astype(np.float32); gD = rng.normal(size=n).astype(np.float32)
    gC = rng.normal(size=(n, 3)).astype(np.float32); gW = rng.normal(size=N).astype(np.float32)
    loss = 0
    for ray_idx, o_, d_, c_, w_ in outs:
        assert abs(float(o_.detach()) - op[ray_idx]) < 1e-5 and abs(float(d_.detach()) - dp[ray_idx]) < 1e-5
        assert np.abs(c_.detach().numpy() - col[ray_idx]).max() < 1e-5
        row = [r for r in rays_a.tolist() if r[0] == ray_idx][0]
        if row[2]:
            assert np.abs(w_.detach().numpy() - ws[row[1]:row[1] + row[2]]).max() < 1e-6
        loss = loss + gO[ray_idx] * o_ + gD[ray_idx] * d_ + (torch.tensor(gC[ray_idx], dtype=torch.float64) * c_).sum() \
            + (torch.tensor(gW[row[1]:row[1] + row[2]], dtype=torch.float64) * w_).sum()
    loss.backward()
    gs, gc = O.composite_train_bw(gO, gD, gC, gW, sig, rgbs, ws, dl, ts, rays_a, op, dp, col, 1e-4)
    assert np.abs(gc - trgb.grad.numpy()).max() < 1e-5
    # the reference's closed form ignores d(alive)/d(sigma) (piecewise constant) -- equal a.e.
    ref_gs = tsig.grad.numpy()
    assert np.abs(gs - ref_gs).max() < 2e-4 * max(1.0, np.abs(ref_gs).max())
    # early termination bookkeeping (Q8): total_samples excludes the terminating sample; ws = 0 behind it
    for ray_idx, start, ln in rays_a.tolist():
        t = total[ray_idx]
        assert 0 <= t <= ln
        if t < ln:
            assert np.all(ws[start + t + 1:start + ln] == 0) and ws[start + t] > 0
