"""Baseline model variants of the comparison matrix on the same operators:
NGP_zoo + moe_render (networks.py:1100-1163, rendering.py:48-110) and unshared_MNGP (networks.py:424-628).

Cross-check: an unshared_MNGP whose sub-NeRF i carries the parameters and occupancy of NGP_zoo.model_i computes the
same function, so ml_render(unshared) and moe_render(zoo) must agree output for output and gradient for gradient
(same kernels, same noise draws) -- and both must agree with blending K independent render() calls by hand."""
import numpy as np
import pytest
import torch

from radnerf_b200 import synthetic as syn

pytestmark = pytest.mark.gpu
SCALE, K, N = 0.5, 3, 2048


def _pair():
    from radnerf_b200.models.networks import NGP_zoo, Ray_Gate, unshared_MNGP
    torch.manual_seed(0)
    zoo = NGP_zoo(SCALE, size=K, t=15).cuda()
    un = unshared_MNGP(SCALE, size=K, t=15).cuda()
    gate = Ray_Gate(K).cuda()
    with torch.no_grad():
        for i, m in enumerate(zoo.models()):
            m.xyz_encoder.params.uniform_(-0.3, 0.3)
            sc = syn.make_scene(SCALE, 1234, jitter_seed=i if i else None)
            m.density_bitfield.copy_(torch.from_numpy(syn.packbits_np(sc.density_grid, sc.threshold)).cuda())
            getattr(un, f"xyz_encoder_{i}").params.copy_(m.xyz_encoder.params)
            getattr(un, f"geo_net_{i}").params.copy_(m.geo_net.params)
            getattr(un, f"rgb_net_{i}").params.copy_(m.rgb_net.params)
            getattr(un, f"density_bitfield_{i}").copy_(m.density_bitfield)
    return zoo, un, gate


def _rays():
    cam = syn.make_cameras(100, 1.5)
    o, d, _, _ = syn.train_batch(cam, N, 11)
    return torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()


def test_state_dict_keys_follow_the_reference():
    zoo, un, _ = _pair()
    kz, ku = set(zoo.state_dict()), set(un.state_dict())
    for i in range(K):
        assert {f"model_{i}.xyz_encoder.params", f"model_{i}.geo_net.params", f"model_{i}.rgb_net.params",
                f"model_{i}.density_bitfield", f"model_{i}.density_grid", f"model_{i}.grid_coords"} <= kz
        assert {f"xyz_encoder_{i}.params", f"geo_net_{i}.params", f"rgb_net_{i}.params", f"density_bitfield_{i}",
                f"density_grid_{i}"} <= ku
    assert {"center", "xyz_min", "xyz_max", "half_size"} <= kz and {"center", "xyz_min", "xyz_max", "half_size", "grid_coords"} <= ku


def test_moe_render_equals_unshared_ml_render_and_manual_blend():
    from radnerf_b200.models.ml_rendering import ml_render
    from radnerf_b200.models.rendering import moe_render, render
    zoo, un, gate = _pair()
    o, d = _rays()
    target = torch.rand(N, 3, device="cuda")
    torch.manual_seed(3)
    a = moe_render(zoo, gate, o, d)
    torch.manual_seed(3)
    b = ml_render(un, gate, o, d, d)
    for k in ("rgb", "opacity", "depth", "gating_code", "gating_importance"):
        assert torch.allclose(a[k].float(), b[k].float(), atol=1e-6), k
    assert float(a["opacity"].detach().max()) > 0.05             # the scene is actually hit (random-init field: thin)
    # manual blend of K independent single-field renders (same noise sequence: one draw per member, in order)
    torch.manual_seed(3)
    g = a["gating_code"].detach()
    rgb = torch.zeros(N, 3, device="cuda")
    for i, m in enumerate(zoo.models()):
        r = render(m, o, d)
        rgb += r["rgb"].detach().float() * g[:, i:i + 1]
        assert torch.allclose(r["depth"].detach().float(), a["depth"][:, i].detach(), atol=1e-6)
    assert torch.allclose(rgb, a["rgb"].detach().float(), atol=1e-5)
    # gradients: same loss on both formulations -> same parameter gradients member by member
    ((a["rgb"].float() - target) ** 2).mean().backward()
    ga = [(m.xyz_encoder.params.grad.clone(), m.geo_net.params.grad.clone(), m.rgb_net.params.grad.clone()) for m in zoo.models()]
    gg = gate.encoder.params.grad.clone()
    gate.encoder.params.grad = None
    ((b["rgb"].float() - target) ** 2).mean().backward()
    for i in range(K):
        for x, name in zip(ga[i], ("xyz_encoder", "geo_net", "rgb_net")):
            y = getattr(un, f"{name}_{i}").params.grad
            assert float(x.norm()) > 0
            assert float((x - y).norm() / x.norm()) < 1e-3, (i, name)      # atomics order only
    assert float((gg - gate.encoder.params.grad).norm() / gg.norm()) < 1e-3


def test_variants_update_density_grid_and_test_time():
    from radnerf_b200.models.ml_rendering import ml_render
    from radnerf_b200.models.rendering import moe_render
    zoo, un, gate = _pair()
    o, d = _rays()
    with torch.no_grad():
        a = moe_render(zoo, gate, o, d, test_time=True)
        b = ml_render(un, gate, o, d, d, test_time=True)
    assert torch.allclose(a["rgb"].float(), b["rgb"].float(), atol=1e-5)
    zoo.update_density_grid(0.01 * 1024 / 3 ** 0.5, warmup=True)
    un.update_density_grid(0.01 * 1024 / 3 ** 0.5, warmup=True)
    for i, m in enumerate(zoo.models()):
        assert bool((m.density_grid > 0).all()) and int(m.density_bitfield.count_nonzero()) > 0
        assert bool((getattr(un, f"density_grid_{i}") > 0).all())
