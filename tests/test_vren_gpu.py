"""GPU parity tests of the vren-equivalent operators (a1-a3, a5, a9-a12).

Every test calls the product through the C ABI (radnerf_b200.vren -> libradnerf_b200.so)
and compares with (1) the CPU oracle oracle/vren_oracle.c and, when it was built,
(2) the reference's own kernels compiled for sm_100a (oracle/_ref/vren_ref*.so).
Integer / index / byte outputs: bit-exact.  fp32 outputs: tolerances stated per test.
"""
import os
import sys

import numpy as np
import pytest
import torch

from conftest import CONFIGS, ROOT, get_scene, make_rays
from radnerf_b200 import synthetic as syn

pytestmark = pytest.mark.gpu


def _vren():
    from radnerf_b200 import vren
    return vren


def _vren_ref():
    p = os.path.join(ROOT, "oracle", "_ref")
    if p not in sys.path:
        sys.path.insert(0, p)
    try:
        import vren_ref
        return vren_ref
    except Exception:
        return None


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def test_device_is_blackwell():
    from radnerf_b200 import _lib
    assert _lib.lib().rnb_device_arch() == 100


# ---- morton / packbits -------------------------------------------------------------------
def test_morton_roundtrip_and_oracle():
    from oracle import vren_cpu as O
    vren = _vren()
    rng = np.random.RandomState(0)
    coords = rng.randint(0, 128, (100003, 3)).astype(np.int32)
    idx = vren.morton3D(cu(coords))
    assert np.array_equal(idx.cpu().numpy(), O.morton3D(coords))
    back = vren.morton3D_invert(idx)
    assert np.array_equal(back.cpu().numpy(), coords)
    assert np.array_equal(back.cpu().numpy(), O.morton3D_invert(idx.cpu().numpy()))
    ref = _vren_ref()
    if ref is not None:
        assert torch.equal(ref.morton3D(cu(coords)), idx)
        assert torch.equal(ref.morton3D_invert(idx), back)
    # full grid: a permutation of 0..128^3-1 (idempotence / bijection property at full size)
    g = np.stack(np.meshgrid(*[np.arange(128, dtype=np.int32)] * 3, indexing="ij"), -1).reshape(-1, 3)
    full = vren.morton3D(cu(g))
    assert torch.equal(torch.sort(full.long()).values, torch.arange(128 ** 3, device="cuda"))
    assert vren.morton3D(torch.empty(0, 3, dtype=torch.int32, device="cuda")).numel() == 0


@pytest.mark.parametrize("cfg", list(CONFIGS))
def test_packbits_bit_exact(cfg):
    from oracle import vren_cpu as O
    vren = _vren()
    scale, _ = CONFIGS[cfg]
    sc, bf = get_scene(scale)
    grid = cu(sc.density_grid)
    out = torch.zeros(sc.cascades * 128 ** 3 // 8, dtype=torch.uint8, device="cuda")
    vren.packbits(grid, sc.threshold, out)
    assert np.array_equal(out.cpu().numpy(), bf)
    assert np.array_equal(out.cpu().numpy(), O.packbits(sc.density_grid, sc.threshold))
    ref = _vren_ref()
    if ref is not None:
        out2 = torch.zeros_like(out)
        ref.packbits(grid, sc.threshold, out2)
        assert torch.equal(out, out2)
    # half grid + ragged size (tail path) + threshold on equality boundary
    g16 = torch.rand(8 * 77, device="cuda").half()
    thr = float(g16[5])
    o16 = torch.zeros(77, dtype=torch.uint8, device="cuda")
    vren.packbits(g16, thr, o16)
    exp = syn.packbits_np(g16.float().cpu().numpy(), thr)
    assert np.array_equal(o16.cpu().numpy(), exp)


def test_packbits_device_threshold():
    from radnerf_b200 import _lib
    sc, bf = get_scene(0.5)
    grid = cu(sc.density_grid)
    pos = grid[grid > 0]
    mean = pos.mean().reshape(1)
    out = torch.zeros(128 ** 3 // 8, dtype=torch.uint8, device="cuda")
    _lib.call("rnb_packbits_devthr", grid, _lib.i64(out.numel()), mean, _lib.f32(syn.DENSITY_THRESHOLD), out,
              _lib.stream_arg())
    thr = min(float(mean), syn.DENSITY_THRESHOLD)
    assert np.array_equal(out.cpu().numpy(), syn.packbits_np(sc.density_grid, thr))


# ---- ray / box -------------------------------------------------------------------------------
@pytest.mark.parametrize("cfg", list(CONFIGS))
def test_ray_aabb_bit_exact(cfg):
    from oracle import vren_cpu as O
    vren = _vren()
    scale, _ = CONFIGS[cfg]
    cam = syn.make_cameras(100, 1.5 if scale <= 0.5 else 1.0)
    o, d, _, _ = syn.train_batch(cam, 8192, 3)
    d[:17, 0] = 0.0            # axis-parallel rays: 1/0 = inf must behave like the reference
    o[17:40] *= 10             # misses
    center = np.zeros((1, 3), np.float32); half = np.full((1, 3), scale, np.float32)
    cnt, ht, hi = vren.ray_aabb_intersect(cu(o), cu(d), cu(center), cu(half), 1)
    ocnt, oht, ohi = O.ray_aabb_intersect(o, d, center, half)
    assert np.array_equal(cnt.cpu().numpy(), ocnt)
    assert np.array_equal(ht.cpu().numpy().view(np.uint32), oht.view(np.uint32))
    assert np.array_equal(hi.cpu().numpy(), ohi)
    ref = _vren_ref()
    if ref is not None:
        rc, rt, ri = ref.ray_aabb_intersect(cu(o), cu(d), cu(center), cu(half), 1)
        assert torch.equal(rc, cnt) and torch.equal(ri, hi)
        assert torch.equal(rt.view(torch.int32), ht.view(torch.int32))


def test_ray_aabb_multi_voxel_sorted():
    vren = _vren()
    rng = np.random.RandomState(1)
    o = rng.uniform(-2, 2, (1000, 3)).astype(np.float32); d = rng.normal(size=(1000, 3)).astype(np.float32)
    centers = rng.uniform(-1, 1, (6, 3)).astype(np.float32); half = rng.uniform(0.1, 0.5, (6, 3)).astype(np.float32)
    cnt, ht, hi = vren.ray_aabb_intersect(cu(o), cu(d), cu(centers), cu(half), 6)
    t1 = ht[..., 0]
    assert bool((t1[:, 1:] >= t1[:, :-1]).all())          # ascending, -1 rows first (torch::sort)
    assert torch.equal((hi >= 0).sum(1).int(), cnt)


# ---- marching (training) -----------------------------------------------------------------------
def _canon(rays_a, *arrs):
    """Reference rows are in race order: canonicalise to per-ray lists sorted by ray index."""
    rays_a = rays_a.cpu().numpy()
    order = np.argsort(rays_a[:, 0])
    return rays_a[order], [a.cpu().numpy() for a in arrs]


@pytest.mark.parametrize("cfg", list(CONFIGS))
@pytest.mark.parametrize("n_rays", [8192, 1000])
def test_raymarching_train_bit_exact(cfg, n_rays):
    from oracle import vren_cpu as O
    vren = _vren()
    scale, esf = CONFIGS[cfg]
    sc, bf = get_scene(scale)
    o, d, h = make_rays(scale, n_rays)
    noise = np.random.RandomState(7).rand(n_rays).astype(np.float32)
    rays_a, xyzs, dirs, deltas, ts, counter = vren.raymarching_train(
        cu(o), cu(d), cu(h), cu(bf), sc.cascades, scale, esf, cu(noise), 128, 1024)
    ora, oxyz, odir, odl, ots, ocnt = O.raymarching_train(o, d, h, bf, sc.cascades, scale, esf, noise)
    assert np.array_equal(counter.cpu().numpy(), ocnt)
    assert np.array_equal(rays_a.cpu().numpy(), ora)                    # counts, offsets, indices
    for a, b in ((xyzs, oxyz), (dirs, odir), (deltas, odl), (ts, ots)):
        assert np.array_equal(a.cpu().numpy().view(np.uint32), b.view(np.uint32))
    assert int(counter[0]) > n_rays  # the scene is not empty
    ref = _vren_ref()
    if ref is not None:
        r_a, rx, rd, rdl, rts, rc = ref.raymarching_train(cu(o), cu(d), cu(h), cu(bf), sc.cascades, scale, esf,
                                                          cu(noise), 128, 1024)
        assert int(rc[0]) == int(counter[0])
        ra = r_a.cpu().numpy(); order = np.argsort(ra[:, 0]); ra = ra[order]
        mine = rays_a.cpu().numpy()
        assert np.array_equal(ra[:, 0], mine[:, 0]) and np.array_equal(ra[:, 2], mine[:, 2])   # per-ray counts
        # payload per ray via its own [start, start+n)
        rx, rd, rdl, rts = (t.cpu().numpy() for t in (rx, rd, rdl, rts))
        mx, md, mdl, mts = (t.cpu().numpy() for t in (xyzs, dirs, deltas, ts))
        gather = np.concatenate([np.arange(s, s + n) for s, n in zip(ra[:, 1], ra[:, 2])]) if len(ra) else np.zeros(0, int)
        assert np.array_equal(rx[gather].view(np.uint32), mx.view(np.uint32))
        assert np.array_equal(rd[gather].view(np.uint32), md.view(np.uint32))
        assert np.array_equal(rdl[gather].view(np.uint32), mdl.view(np.uint32))
        assert np.array_equal(rts[gather].view(np.uint32), mts.view(np.uint32))


def test_raymarching_train_edge_cases():
    vren = _vren()
    sc, bf = get_scene(0.5)
    z3 = torch.empty(0, 3, device="cuda"); z2 = torch.empty(0, 2, device="cuda"); z1 = torch.empty(0, device="cuda")
    rays_a, xyzs, dirs, deltas, ts, counter = vren.raymarching_train(z3, z3, z2, cu(bf), 1, 0.5, 0.0, z1, 128, 1024)
    assert rays_a.shape == (0, 3) and xyzs.shape == (0, 3) and counter.tolist() == [0, 0]
    # all rays miss (t1 = -1): every ray still gets a row with 0 samples (Q6)
    o, d, h = make_rays(0.5, 256)
    h[:] = -1
    rays_a, xyzs, _, _, _, counter = vren.raymarching_train(cu(o), cu(d), cu(h), cu(bf), 1, 0.5, 0.0,
                                                            torch.rand(256, device="cuda"), 128, 1024)
    assert counter.tolist() == [0, 256] and xyzs.shape[0] == 0
    assert rays_a[:, 0].tolist() == list(range(256)) and int(rays_a[:, 1:].abs().sum()) == 0
    # fully occupied unbounded grid: the max_samples cap is honoured and matches the oracle
    full = np.full(6 * 128 ** 3 // 8, 255, np.uint8)
    o, d, h = make_rays(16.0, 512)
    rays_a, *_ = vren.raymarching_train(cu(o), cu(d), cu(h), cu(full), 6, 16.0, 1 / 256, torch.zeros(512, device="cuda"), 128, 8)
    assert int(rays_a[:, 2].max()) == 8
    from oracle import vren_cpu as O
    ora, _ = O.raymarching_train(o, d, h, full, 6, 16.0, 1 / 256, np.zeros(512, np.float32), 128, 8, count_only=True)
    assert np.array_equal(rays_a.cpu().numpy(), ora)


# ---- marching (test time) ---------------------------------------------------------------------
@pytest.mark.parametrize("cfg", list(CONFIGS))
@pytest.mark.parametrize("N_samples", [1, 4, 64])
def test_raymarching_test_bit_exact(cfg, N_samples):
    from oracle import vren_cpu as O
    vren = _vren()
    scale, esf = CONFIGS[cfg]
    sc, bf = get_scene(scale)
    n_rays = 4096
    o, d, h = make_rays(scale, n_rays, seed=11)
    alive = np.random.RandomState(3).permutation(n_rays)[:3000].astype(np.int64)
    ref = _vren_ref()
    hg = cu(h); ho = h.copy(); hr = cu(h)
    for rnd in range(3):  # resume semantics: hits_t[:,0] is advanced between rounds
        xyzs, dirs, deltas, ts, n_eff = vren.raymarching_test(cu(o), cu(d), hg, cu(alive), cu(bf), sc.cascades, scale,
                                                              esf, 128, 1024, N_samples)
        oo = O.raymarching_test(o, d, ho, alive, bf, sc.cascades, scale, esf, 128, 1024, N_samples)
        assert np.array_equal(n_eff.cpu().numpy(), oo[4])
        for a, b in zip((xyzs, dirs, deltas, ts), oo[:4]):
            assert np.array_equal(a.cpu().numpy().view(np.uint32), b.view(np.uint32))
        assert np.array_equal(hg.cpu().numpy().view(np.uint32), ho.view(np.uint32))
        if ref is not None:
            rr = ref.raymarching_test(cu(o), cu(d), hr, cu(alive), cu(bf), sc.cascades, scale, esf, 128, 1024, N_samples)
            for a, b in zip((xyzs, dirs, deltas, ts, n_eff), rr):
                assert torch.equal(a.view(torch.int32), b.view(torch.int32))
            assert torch.equal(hg.view(torch.int32), hr.view(torch.int32))


# ---- compositing ----------------------------------------------------------------------------------
def _composite_inputs(cfg, n_rays=8192, seed=0):
    from oracle import vren_cpu as O
    scale, esf = CONFIGS[cfg]
    sc, bf = get_scene(scale)
    o, d, h = make_rays(scale, n_rays)
    noise = np.random.RandomState(7).rand(n_rays).astype(np.float32)
    rays_a, xyzs, dirs, deltas, ts, _ = O.raymarching_train(o, d, h, bf, sc.cascades, scale, esf, noise)
    rng = np.random.RandomState(seed)
    N = len(ts)
    sigmas = np.minimum(np.exp(rng.normal(1.5, 2.0, N)), np.exp(15)).astype(np.float32)   # SURVEY 8d operator inputs
    rgbs = rng.uniform(0, 1, (N, 3)).astype(np.float32)
    return rays_a, sigmas, rgbs, deltas, ts


@pytest.mark.parametrize("cfg", list(CONFIGS))
def test_composite_train_fw(cfg):
    from oracle import vren_cpu as O
    vren = _vren()
    rays_a, sigmas, rgbs, deltas, ts = _composite_inputs(cfg)
    # shuffle the rows: rays_a row order must not matter (reference rows are in race order)
    perm = np.random.RandomState(5).permutation(len(rays_a))
    ra = rays_a[perm]
    total, opacity, depth, rgb, ws = vren.composite_train_fw(cu(sigmas), cu(rgbs), cu(deltas), cu(ts), cu(ra), 1e-4)
    ot, oo, od, oc, ow = O.composite_train_fw(sigmas, rgbs, deltas, ts, ra, 1e-4)
    # CPU oracle uses expf, the GPU __expf (Q10): tolerance 2e-6 abs on weights, 1e-5 on ray sums
    assert np.abs(ws.cpu().numpy() - ow).max() < 2e-6
    assert np.abs(opacity.cpu().numpy() - oo).max() < 1e-5
    assert np.abs(rgb.cpu().numpy() - oc).max() < 1e-5
    assert np.abs(depth.cpu().numpy() - od).max() < 1e-5 * max(1.0, float(ts.max()))
    # termination index may differ from the CPU only where T straddles the threshold within fp error
    assert (total.cpu().numpy() != ot).mean() < 1e-3
    ref = _vren_ref()
    if ref is not None:
        rt, ro, rd, rc, rw = ref.composite_train_fw(cu(sigmas), cu(rgbs), cu(deltas), cu(ts), cu(ra), 1e-4)
        assert torch.equal(rt, total)                                   # same __expf, same serial T chain
        assert torch.equal(rw.view(torch.int32), ws.view(torch.int32))  # weights bit-exact
        assert (ro - opacity).abs().max() < 2e-6 and (rc - rgb).abs().max() < 2e-6
        assert (rd - depth).abs().max() < 2e-6 * max(1.0, float(ts.max()))


@pytest.mark.parametrize("cfg", list(CONFIGS))
@pytest.mark.parametrize("with_dws", [False, True])
def test_composite_train_bw(cfg, with_dws):
    from oracle import vren_cpu as O
    vren = _vren()
    rays_a, sigmas, rgbs, deltas, ts = _composite_inputs(cfg, n_rays=4096)
    n = len(rays_a); N = len(ts)
    rng = np.random.RandomState(9)
    gO = rng.normal(size=n).astype(np.float32); gD = rng.normal(size=n).astype(np.float32)
    gC = rng.normal(size=(n, 3)).astype(np.float32)
    gW = rng.normal(size=N).astype(np.float32) if with_dws else np.zeros(N, np.float32)
    total, opacity, depth, rgb, ws = vren.composite_train_fw(cu(sigmas), cu(rgbs), cu(deltas), cu(ts), cu(rays_a), 1e-4)
    gs, gc = vren.composite_train_bw(cu(gO), cu(gD), cu(gC), cu(gW) if with_dws else None, cu(sigmas), cu(rgbs), ws,
                                     cu(deltas), cu(ts), cu(rays_a), opacity, depth, rgb, 1e-4)
    ogs, ogc = O.composite_train_bw(gO, gD, gC, gW, sigmas, rgbs, ws.cpu().numpy(), deltas, ts, rays_a,
                                    opacity.cpu().numpy(), depth.cpu().numpy(), rgb.cpu().numpy(), 1e-4)
    scale_s = np.abs(ogs).max()
    assert np.abs(gc.cpu().numpy() - ogc).max() < 1e-5
    assert np.abs(gs.cpu().numpy() - ogs).max() < 1e-4 * max(1.0, scale_s)   # gradients: 1e-2 rel allowed, we hold 1e-4
    ref = _vren_ref()
    if ref is not None:
        rgs, rgc = ref.composite_train_bw(cu(gO), cu(gD), cu(gC), cu(gW), cu(sigmas), cu(rgbs), ws, cu(deltas), cu(ts),
                                          cu(rays_a), opacity, depth, rgb, 1e-4)
        assert (rgc - gc).abs().max() < 1e-6
        assert (rgs - gs).abs().max() < 1e-4 * max(1.0, float(rgs.abs().max()))


@pytest.mark.parametrize("N_samples", [1, 4, 64])
def test_composite_test_fw(N_samples):
    from oracle import vren_cpu as O
    vren = _vren()
    rng = np.random.RandomState(2)
    n_rays, A = 5000, 3000
    alive = rng.permutation(n_rays)[:A].astype(np.int64)
    n_eff = rng.randint(0, N_samples + 1, A).astype(np.int32)
    sig = np.exp(rng.normal(1.5, 2.0, (A, N_samples))).astype(np.float32)
    rgbs = rng.uniform(0, 1, (A, N_samples, 3)).astype(np.float32)
    dl = np.full((A, N_samples), 1.7320508 / 1024, np.float32) * 20
    ts = np.sort(rng.uniform(0.5, 2, (A, N_samples)).astype(np.float32), 1)
    op = rng.uniform(0, 0.9, n_rays).astype(np.float32); dp = rng.uniform(0, 1, n_rays).astype(np.float32)
    col = rng.uniform(0, 0.9, (n_rays, 3)).astype(np.float32)
    hits = np.zeros((n_rays, 2), np.float32)
    g_alive, g_op, g_dp, g_col = cu(alive), cu(op), cu(dp), cu(col)
    vren.composite_test_fw(cu(sig), cu(rgbs), cu(dl), cu(ts), cu(hits), g_alive, 1e-4, cu(n_eff), g_op, g_dp, g_col)
    o_alive, o_op, o_dp, o_col = alive.copy(), op.copy(), dp.copy(), col.copy()
    O.composite_test_fw(sig, rgbs, dl, ts, hits, o_alive, 1e-4, n_eff, o_op, o_dp, o_col)
    assert (g_alive.cpu().numpy() != o_alive).mean() < 1e-3
    assert np.abs(g_op.cpu().numpy() - o_op).max() < 1e-5
    assert np.abs(g_dp.cpu().numpy() - o_dp).max() < 2e-5
    assert np.abs(g_col.cpu().numpy() - o_col).max() < 1e-5
    ref = _vren_ref()
    if ref is not None:
        r_alive, r_op, r_dp, r_col = cu(alive), cu(op), cu(dp), cu(col)
        ref.composite_test_fw(cu(sig), cu(rgbs), cu(dl), cu(ts), cu(hits), r_alive, 1e-4, cu(n_eff), r_op, r_dp, r_col)
        assert torch.equal(r_alive, g_alive)
        assert (r_op - g_op).abs().max() < 2e-6 and (r_col - g_col).abs().max() < 2e-6
