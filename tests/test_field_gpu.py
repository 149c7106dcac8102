"""GPU parity of the fused field kernels (rnb_field_fwd / rnb_field_bwd) against the operator-level
path (hash grid -> geo MLP -> TruncExp -> SH4 -> rgb MLP through radnerf_b200.tcnn + autograd), which
is itself checked against the CPU oracle in test_tcnn_gpu.py / test_e2e_gpu.py.

Tolerances: both paths round at the same points (fp16 features / activations / outputs, fp32
accumulation), so outputs agree to a few fp16 ulps: rgb |d| <= 2e-3, sigma relative 2e-2 (exp of an
fp16 value); parameter gradients within 1e-2 relative (L2), the bar of BASELINE.json's north_star.
"""
import numpy as np
import pytest
import torch

from conftest import get_scene, make_rays

pytestmark = pytest.mark.gpu


def _setup(K, n_rays, scale=0.5, esf=0.0, seed=0):
    from radnerf_b200 import vren
    from radnerf_b200.models.networks import MNGP
    torch.manual_seed(seed)
    dev = torch.device("cuda")
    model = MNGP(scale, size=K).to(dev)
    with torch.no_grad():
        model.xyz_encoder.params.mul_(2000.0)      # U(-0.2, 0.2): non-trivial features
    o, d, h = make_rays(scale, n_rays)
    o, d, h = (torch.from_numpy(a).to(dev) for a in (o, d, h))
    passes = []
    for i in range(K):
        _, bf = get_scene(scale, jitter=i if i else None)
        bf = torch.from_numpy(bf).to(dev)
        noise = torch.rand(n_rays, device=dev)
        rays_a, xyzs, dirs, deltas, ts, _ = vren.raymarching_train(o, d, h, bf, model.cascades, scale, esf, noise, 128, 1024)
        ray_id = torch.repeat_interleave(rays_a[:, 0], rays_a[:, 2]).int()
        passes.append((ray_id, ts, xyzs, dirs))
    return model, o, d, passes


def _fused_inputs(model, o, d, passes):
    from radnerf_b200 import field as F
    K = len(passes)
    table = model.xyz_encoder._shadow.get(model.xyz_encoder.params)
    geo = [getattr(model, f"geo_net_{i}")._shadow.get(getattr(model, f"geo_net_{i}").params) for i in range(K)]
    rgb = [getattr(model, f"rgb_net_{i}")._shadow.get(getattr(model, f"rgb_net_{i}").params) for i in range(K)]
    P = F.field_params(model, table, geo, rgb)
    ray_id, ts, seg, bases, counts = F.pack_passes([p[0] for p in passes], [p[1] for p in passes], o.device)
    S = F.samples(o, d, ray_id, ts, seg)
    return F, P, S, bases, counts


@pytest.mark.parametrize("K,n_rays", [(1, 300), (2, 1024), (3, 77)])
def test_field_fwd_matches_operator_path(K, n_rays):
    model, o, d, passes = _setup(K, n_rays)
    F, P, S, bases, counts = _fused_inputs(model, o, d, passes)
    cap = S.capacity
    sig = torch.full((cap,), -7.0, device=o.device)
    rgb4 = torch.zeros(cap, 4, dtype=torch.float16, device=o.device)
    saved = F.Saved(cap, o.device)
    F.field_fwd(P, S, saved, sig, rgb4)
    torch.cuda.synchronize()
    for i, (ray_id, ts, xyzs, dirs) in enumerate(passes):
        n = counts[i]
        assert n > 0
        with torch.no_grad(), torch.autocast("cuda"):
            s_ref, c_ref = model(xyzs, dirs, i)
        s_f = sig[bases[i]:bases[i] + n]; c_f = rgb4[bases[i]:bases[i] + n, :3].float()
        rel = ((s_f - s_ref.float()).abs() / (s_ref.float().abs() + 1e-6)).max().item()
        assert rel < 2e-2, (i, rel)
        assert (c_f - c_ref.float()).abs().max().item() < 2e-3
        # saved activations equal the operator path's encoder output bit for bit
        with torch.no_grad():
            x = ((xyzs - model.xyz_min) / (model.xyz_max - model.xyz_min)).clip(0.0, 1.0)
            enc_ref = model.xyz_encoder(x)
        assert torch.equal(saved.rows("enc")[bases[i]:bases[i] + n], enc_ref)
    # padding slots of sigma were not touched
    if K > 1 and bases[1] > counts[0]:
        assert float(sig[counts[0]:bases[1]].max()) == -7.0


@pytest.mark.parametrize("K,n_rays", [(1, 256), (2, 1024)])
def test_field_bwd_matches_autograd(K, n_rays):
    model, o, d, passes = _setup(K, n_rays, seed=1)
    F, P, S, bases, counts = _fused_inputs(model, o, d, passes)
    cap = S.capacity
    dev = o.device
    sig = torch.zeros(cap, device=dev); rgb4 = torch.zeros(cap, 4, dtype=torch.float16, device=dev)
    saved = F.Saved(cap, dev)
    F.field_fwd(P, S, saved, sig, rgb4)
    g = torch.Generator(device="cpu").manual_seed(5)
    dsig = torch.zeros(cap, device=dev); drgb = torch.zeros(cap, 3, device=dev)
    for i in range(K):
        n = counts[i]
        dsig[bases[i]:bases[i] + n] = (torch.randn(n, generator=g) * 1e-3).to(dev)
        drgb[bases[i]:bases[i] + n] = (torch.randn(n, 3, generator=g) * 1e-2).to(dev)
    # reference: autograd through the operator path
    model.zero_grad()
    for i, (ray_id, ts, xyzs, dirs) in enumerate(passes):
        n = counts[i]
        with torch.autocast("cuda"):
            s_ref, c_ref = model(xyzs, dirs, i)
        loss = (s_ref.float() * dsig[bases[i]:bases[i] + n]).sum() + (c_ref.float() * drgb[bases[i]:bases[i] + n]).sum()
        loss.backward()
    gt = torch.zeros_like(model.xyz_encoder.params)
    gg = [torch.zeros_like(getattr(model, f"geo_net_{i}").params) for i in range(K)]
    gr = [torch.zeros_like(getattr(model, f"rgb_net_{i}").params) for i in range(K)]
    F.field_bwd(P, S, saved, rgb4, dsig, drgb, 128.0, gt, gg, gr)
    torch.cuda.synchronize()

    def rel(a, b):
        return float((a - b).norm() / (b.norm() + 1e-20))
    assert rel(gt, model.xyz_encoder.params.grad) < 1e-2
    for i in range(K):
        assert rel(gg[i], getattr(model, f"geo_net_{i}").params.grad) < 1e-2, i
        assert rel(gr[i], getattr(model, f"rgb_net_{i}").params.grad) < 1e-2, i
    # padded geo rows 17..31 and rgb rows 3..15 of the last layers get exactly zero gradient
    assert float(gg[0][64 * 32 + 17 * 64:].abs().max()) == 0.0
