"""GPU parity of the fused training step (rnb_engine_*, radnerf_b200.engine.Engine) against the CPU
oracle of the whole path (oracle/render_cpu.py: gate -> K x (AABB, march, field, composite, background)
-> gated blend -> NeRFLoss -> autograd backward) and against torch.optim.Adam.

Tolerances (BASELINE.json north_star, fp16-MLP mode): per-ray sample counts bit-exact; composited rgb /
depth / opacity |delta| <= 1e-3; loss within 1e-3 relative; gradients within 1e-2 relative (L2 per
parameter vector).
"""
import numpy as np
import pytest
import torch

from conftest import CONFIGS
from radnerf_b200 import synthetic as syn
from test_e2e_gpu import build

pytestmark = pytest.mark.gpu


def _names(K):
    out = [("table", "xyz_encoder.params")]
    for i in range(K):
        out += [(f"geo{i}", f"geo_net_{i}.params"), (f"rgb{i}", f"rgb_net_{i}.params")]
    out.append(("gate", "gate.encoder.params"))
    return out


@pytest.mark.parametrize("cfg,K,grid_init,n", [("bounded", 2, 0.2, 1024), ("bounded", 2, 1e-4, 777), ("unbounded", 4, 0.2, 1024)])
def test_engine_step_matches_oracle(cfg, K, grid_init, n):
    scale, esf = CONFIGS[cfg]
    cam = syn.make_cameras(100, 1.5 if scale <= 0.5 else 1.0)
    _engine_vs_oracle(scale, esf, K, grid_init, n, cam, 256)


@pytest.mark.parametrize("workload", ["radnerf_tat_k2", "radnerf_360_k4"])
def test_engine_step_matches_oracle_at_bench_shape(workload):
    """The very batch shape bench.py times (BASELINE.json configs 3 and 4): the workload's cameras (1920x1080 fx 1166
    radius 0.9 / 1237x822 fx 1050 radius 1.0), 8192 rays, K = 2 / 4, against the CPU oracle of the whole step."""
    w = syn.WORKLOADS[workload]
    cam = syn.make_cameras(100, w["radius"], w["H"], w["W"], w["fx"])
    _engine_vs_oracle(w["scale"], syn.exp_step_factor_for_scale(w["scale"]), w["K"], 0.2, 8192, cam, 128)


def _engine_vs_oracle(scale, esf, K, grid_init, n, cam, samples_per_ray):
    from oracle import render_cpu as R
    from radnerf_b200.engine import Engine
    P, model, gate, bitfields = build(scale, K, grid_init)
    eng = Engine(model, gate, n_rays=n, exp_step_factor=esf, samples_per_ray=samples_per_ray, loss_scale=128.0 * 1024.0)
    o, d, _, _ = syn.train_batch(cam, n, 42)
    target = np.random.RandomState(1).rand(n, 3).astype(np.float32)
    torch.manual_seed(5)
    noise = torch.rand(K, n, device="cuda")
    bg = eng.background()
    eng.forward_backward(torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda(), torch.from_numpy(target).cuda(), noise, bg, True)
    torch.cuda.synchronize()
    # ---- oracle ----
    P.requires_grad_(True)
    noises = [noise[i].cpu().numpy() for i in range(K)]
    ores = R.ml_render_train(P, o, d, bitfields, noises, syn.cascades_for_scale(scale), esf)
    oloss, oterms = R.nerf_loss(ores, torch.from_numpy(target), 1e-3, 1e-2, 5e-3)
    oloss.backward()
    # integer outputs: per-ray sample counts and packed offsets, bit-exact
    counts, overflow = eng.sample_stats()
    assert not overflow
    for i in range(K):
        ra = ores["per_i"][i]["rays_a"]
        assert np.array_equal(eng.counts[i].cpu().numpy(), ra[:, 2].astype(np.int32)), i
        base = int(eng.seg[i])
        assert base % 128 == 0
        assert np.array_equal(eng.starts[i].cpu().numpy() - base, ra[:, 1].astype(np.int32))
        nn = counts[i]
        assert np.array_equal(eng.ts[base:base + nn].cpu().numpy().view(np.uint32), ores["per_i"][i]["ts"].view(np.uint32))
        assert np.array_equal(eng.deltas[base:base + nn].cpu().numpy().view(np.uint32), ores["per_i"][i]["deltas"].view(np.uint32))
    # forward
    assert float((eng.gate_prob.cpu() - ores["gating_code"].detach()).abs().max()) < 2e-3
    for key, a in (("rgb", eng.rgb), ("opacity", eng.opacity), ("depth", eng.depth)):
        b = ores[key].detach()
        scale_k = max(1.0, float(b.abs().max())) if key == "depth" else 1.0
        assert float((a.cpu() - b).abs().max()) < 1e-3 * scale_k, key
    loss = eng.loss.cpu()
    assert abs(float(loss[0]) - float(oloss)) < 1e-3 * max(1.0, abs(float(oloss)))
    assert abs(float(loss[1]) - float(oterms["rgb"].mean())) < 1e-3
    assert abs(float(loss[3]) - float(oterms["cv_importance"])) < 1e-3 * max(1.0, float(oterms["cv_importance"]))
    assert abs(float(loss[4]) - float(oterms["depth_mutual"].mean())) < 1e-3
    # gradients
    for key, name in _names(K):
        off = eng.offsets[key]
        go = P.p[name].grad
        g = eng.grads[off:off + go.numel()].cpu()
        rel = float((g - go).norm() / go.norm().clamp_min(1e-20))
        assert rel < 1e-2, (name, rel, float(go.norm()))


def test_engine_matches_python_surface_and_adam():
    """Same weights, same noise: the fused step and the drop-in Python surface (ml_render + NeRFLoss + autograd
    + torch.optim.Adam) produce the same loss and the same updated parameters."""
    from radnerf_b200.engine import Engine
    from radnerf_b200.losses import NeRFLoss
    from radnerf_b200.models.ml_rendering import ml_render
    scale, esf, K, n = 0.5, 0.0, 2, 1024
    P, model, gate, bitfields = build(scale, K, 0.2)
    eng = Engine(model, gate, n_rays=n, exp_step_factor=esf, lr=1e-2)
    cam = syn.make_cameras(100, 1.5)
    o, d, _, _ = syn.train_batch(cam, n, 7)
    o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    target = torch.rand(n, 3, device="cuda")
    before = eng.flat.clone()
    # python surface on the very same (re-homed) parameters
    params = [m.params for _, m in eng._mods]
    opt = torch.optim.Adam(params, 1e-2, eps=1e-15)
    torch.manual_seed(11)
    res = ml_render(model, gate, o, d, d, test_time=False, exp_step_factor=esf)
    loss_d = NeRFLoss()(res, {"rgb": target}, lambda_opacity=1e-3, lambda_cv_importance=1e-2, lambda_depth_mutual=5e-3)
    loss = sum(v.mean() for v in loss_d.values())
    (loss * 1024.0).backward()
    for p in params:
        p.grad.mul_(1.0 / 1024.0)
    grads_py = torch.zeros_like(eng.grads)
    for (name, m) in eng._mods:
        grads_py[eng.offsets[name]:eng.offsets[name] + m.params.numel()] = m.params.grad
    opt.step()
    after_py = eng.flat.clone()
    # restore and run the engine with the noise the python path consumed
    with torch.no_grad():
        eng.flat.copy_(before)
    for _, m in eng._mods:
        m.params.grad = None
    eng.sync_params()
    for _, m in eng._mods:          # force the fp16 copies to be recast from the restored values
        m._shadow.version = None
    eng.sync_params()
    torch.manual_seed(11)
    noise = torch.stack([torch.rand(n, device="cuda") for _ in range(K)])
    out = eng.step(o, d, target, noise=noise)
    torch.cuda.synchronize()
    assert abs(float(out[0]) - float(loss)) < 1e-3 * max(1.0, abs(float(loss)))
    assert float((eng.rgb - res["rgb"].float()).abs().max()) < 1e-3
    # the engine zeroes its gradient vector inside Adam
    assert float(eng.grads.abs().max()) == 0.0
    delta_py = after_py - before
    delta_en = eng.flat - before
    # Adam's first step moves every touched parameter by ~lr*sign(g): compare where the gradient is not marginal
    big = grads_py.abs() > 1e-7 * grads_py.abs().max()
    agree = (torch.sign(delta_py[big]) == torch.sign(delta_en[big])).float().mean()
    assert float(agree) > 0.98
    assert float((delta_py - delta_en).abs().mean()) < 2e-4
    # fp16 copy refreshed by the optimizer kernel
    assert torch.equal(eng.flat16, eng.flat.half())


def test_march_ahead_matches_sequential():
    """Marching batch n+1 on the side stream while batch n trains (Engine.prefetch / rnb_engine_march) must give the
    steps of the sequential order: same samples (bit-exact counts), same losses, same parameters after 4 steps."""
    from radnerf_b200.engine import Engine
    scale, esf, K, n = 0.5, 0.0, 2, 1024
    cam = syn.make_cameras(100, 1.5)
    batches, noises = [], []
    for b in range(4):
        o, d, _, _ = syn.train_batch(cam, n, 100 + b)
        t = np.random.RandomState(b).rand(n, 3).astype(np.float32)
        batches.append(tuple(torch.from_numpy(a).cuda() for a in (o, d, t)))
        torch.manual_seed(50 + b)
        noises.append(torch.rand(K, n, device="cuda"))
    runs = []
    for ahead in (False, True):
        P, model, gate, bitfields = build(scale, K, 0.2)
        eng = Engine(model, gate, n_rays=n, exp_step_factor=esf, lr=1e-2)
        losses, totals = [], []
        for b in range(4):
            o, d, t = batches[b]
            if ahead and b + 1 < 4:
                eng.prefetch(batches[b + 1][0], batches[b + 1][1], noise=noises[b + 1])
            loss = eng.step(o, d, t, noise=None if (ahead and b > 0) else noises[b])
            losses.append(float(loss[0]))
            totals.append(eng.sample_stats()[0])
        torch.cuda.synchronize()
        runs.append((losses, totals, eng.flat.clone()))
    assert runs[0][1] == runs[1][1]                       # per-pass sample counts, every step
    for a, b in zip(runs[0][0], runs[1][0]):
        assert abs(a - b) < 1e-4 * max(1.0, abs(a))
    # Adam's early steps are ~lr*sign(g): entries whose gradient is at rounding level may flip with the order of the
    # atomic reductions; everything else must agree
    differ = ((runs[0][2] - runs[1][2]).abs() > 1e-3).float().mean()
    assert float(differ) < 0.01


@pytest.mark.parametrize("scale,esf", [(0.5, 0.0), (0.5, 1.0 / 256.0), (2.0, 1.0 / 256.0), (16.0, 1.0 / 256.0)])
def test_empty_space_test_is_exact_on_sparse_grids(scale, esf):
    """The marcher's conservative empty-space test (csrc/coarse.cuh) may only skip rays -- and cut marching intervals --
    where no sample can arise.  Adversarial occupancy: isolated random cells in every cascade (most rays miss everything,
    the others graze single cells), two densities.  Every ray's sample count and sample parameters from rnb_engine_march
    must be bit-identical to the operator-level marcher (vren.raymarching_train), which has no such test."""
    from radnerf_b200 import vren
    from radnerf_b200.engine import Engine
    K, n = 2, 16384
    C = syn.cascades_for_scale(scale)
    P, model, gate, _ = build(scale, K, 0.2)
    rs = np.random.RandomState(7)
    for i, p in enumerate((3e-5, 2e-3)):
        cells = rs.rand(C * 128 ** 3) < p
        getattr(model, f"density_bitfield_{i}").copy_(torch.from_numpy(np.packbits(cells, bitorder="little")).cuda())
    eng = Engine(model, gate, n_rays=n, exp_step_factor=esf, lr=1e-2)
    cam = syn.make_cameras(100, 1.5 if scale <= 0.5 else 1.0)
    o, d, _, _ = syn.train_batch(cam, n, 77)
    o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    torch.manual_seed(3)
    noise = torch.rand(K, n, device="cuda")
    eng.prefetch(o, d, noise=noise)
    torch.cuda.synchronize()
    slot = next(iter(eng._ahead.values()))[0]
    M = eng.cfg.max_samples
    counts = eng._slot_view("counts", torch.int32, (K, n), slot)
    t_scr = eng._slot_view("t_scratch", torch.float32, (K, n, M), slot)
    center = torch.zeros(1, 3, device="cuda"); half = torch.full((1, 3), scale, device="cuda")
    _, hits_t, _ = vren.ray_aabb_intersect(o, d, center, half, 1)
    h = hits_t[:, 0].contiguous()
    h[(h[:, 0] >= 0) & (h[:, 0] < syn.NEAR_DISTANCE), 0] = syn.NEAR_DISTANCE
    hit_any = 0
    for i in range(K):
        bf = getattr(model, f"density_bitfield_{i}")
        rays_a, _, _, _, ts, _ = vren.raymarching_train(o, d, h.clone(), bf, C, scale, esf, noise[i].contiguous(), 128, M)
        ra = rays_a.cpu().numpy(); tsn = ts.cpu().numpy()
        ref_cnt = np.zeros(n, np.int32); ref_cnt[ra[:, 0]] = ra[:, 2]
        got_cnt = counts[i].cpu().numpy()
        assert np.array_equal(got_cnt, ref_cnt), (i, int((got_cnt != ref_cnt).sum()))
        tsc = t_scr[i].cpu().numpy()
        for ray, start, cnt in ra[ra[:, 2] > 0]:
            assert np.array_equal(tsc[ray, :cnt], tsn[start:start + cnt])
        hit_any += int((ref_cnt > 0).sum())
        assert 0 < (ref_cnt > 0).sum()
    assert hit_any > 50


def test_gate_on_side_stream_matches_in_line():
    """The gate forward / backward issued on a second stream next to the compositing kernels (rnb_engine_fwd_bwd_ex) must
    give the step of the in-line order: same outputs bit for bit, same gate gradients up to reduction order."""
    from radnerf_b200.engine import Engine
    scale, esf, K, n = 0.5, 0.0, 2, 2048
    cam = syn.make_cameras(100, 1.5)
    o, d, _, _ = syn.train_batch(cam, n, 21)
    o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    target = torch.rand(n, 3, device="cuda")
    torch.manual_seed(9)
    noise = torch.rand(K, n, device="cuda")
    out = []
    for overlap in (False, True):
        P, model, gate, bitfields = build(scale, K, 0.2)
        eng = Engine(model, gate, n_rays=n, exp_step_factor=esf, lr=1e-2)
        eng.overlap_gate = overlap
        for _ in range(3):
            loss = eng.step(o, d, target, noise=noise).clone()
        torch.cuda.synchronize()
        assert (eng._gate_stream is not None) == overlap
        out.append((loss, eng.rgb.clone(), eng.gate_prob.clone(), eng.flat.clone()))
    a, b = out
    assert torch.allclose(a[0], b[0], rtol=1e-4, atol=1e-7)
    assert float((a[1] - b[1]).abs().max()) < 1e-3 and float((a[2] - b[2]).abs().max()) < 1e-3
    assert float(((a[3] - b[3]).abs() > 1e-3).float().mean()) < 0.01
