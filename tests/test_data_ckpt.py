"""Callers of the path: device-side batch assembly (f1), LR schedule, checkpoint compatibility (f2)."""
import math
import os

import numpy as np
import pytest
import torch


@pytest.mark.gpu
def test_batch_rays_matches_reference_formulas():
    """rnb_batch_rays vs datasets/ray_utils.py:34-35 (directions), :62-68 (get_rays) and the index gathers of
    train_ml.py:183-185, evaluated in plain torch on the same indices."""
    from radnerf_b200.data import RaySampler, get_ray_directions
    from radnerf_b200 import synthetic as syn
    H, W, fx = 60, 80, 100.0
    cam = syn.make_cameras(12, 1.5, H, W, fx)
    dirs = get_ray_directions(H, W, fx)
    d_np = syn.pixel_directions(cam, np.arange(H * W))
    assert np.allclose(dirs.cpu().numpy(), d_np, atol=1e-6)
    poses = torch.from_numpy(np.ascontiguousarray(cam.poses).astype(np.float32)).cuda()
    images = torch.rand(12, H * W, 3, device="cuda")
    sm = RaySampler(dirs, poses, images, batch_size=4096, seed=3)
    o, d, rgb, ii, pi = sm.sample(return_indices=True)
    assert int(ii.min()) >= 0 and int(ii.max()) < 12 and int(pi.min()) >= 0 and int(pi.max()) < H * W
    rd = (dirs[pi].unsqueeze(1) @ poses[ii][..., :3].transpose(1, 2)).squeeze(1)     # ray_utils.py:62-66
    assert float((d - rd).abs().max()) < 1e-5
    assert torch.equal(o, poses[ii][..., 3]) and torch.equal(rgb, images[ii, pi])
    # uniformity of the built-in draws: every image is hit, pixel indices spread over the frame
    assert len(torch.unique(ii)) == 12 and float(pi.float().mean()) == pytest.approx(H * W / 2, rel=0.05)
    o2, d2, rgb2 = sm.sample()
    assert not torch.equal(d2, d)                                                     # a new batch every call
    # supplied indices
    gi = torch.randint(12, (777,), device="cuda"); gp = torch.randint(H * W, (777,), device="cuda")
    o3, d3, rgb3 = sm.sample(gi, gp)
    assert torch.equal(rgb3, images[gi, gp]) and o3.shape == (777, 3)


def test_cosine_lr_matches_torch_scheduler():
    from radnerf_b200.trainer import cosine_lr
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.Adam([p], 1e-2)
    sch = torch.optim.lr_scheduler.CosineAnnealingLR(opt, 200, 1e-2 / 30)
    for step in range(200):
        assert math.isclose(opt.param_groups[0]["lr"], cosine_lr(1e-2, step, 200), rel_tol=1e-9, abs_tol=1e-12)
        opt.step(); sch.step()


@pytest.mark.gpu
def test_checkpoint_round_trip_with_reference_key_layout(tmp_path):
    """A Lightning-style checkpoint with the reference's keys (model.xyz_encoder.params, model.geo_net_i.params,
    model.density_bitfield_i, ray_gate.encoder.params ...) loads into the drop-in modules; save -> slim -> load is
    the identity on the parameters."""
    from radnerf_b200 import checkpoint as C
    from radnerf_b200.models.networks import MNGP, Ray_Gate
    torch.manual_seed(1)
    model = MNGP(0.5, size=2).cuda(); gate = Ray_Gate(2).cuda()
    with torch.no_grad():
        model.xyz_encoder.params.normal_()
        model.density_bitfield_1.fill_(0x5A)
    path = os.path.join(tmp_path, "epoch=0.ckpt")
    C.save_ckpt(path, model, gate, extra={"epoch": 0})
    sd = torch.load(path, weights_only=False)["state_dict"]
    for k in ("model.xyz_encoder.params", "model.geo_net_0.params", "model.rgb_net_1.params", "model.density_bitfield_0",
              "model.density_grid_1", "ray_gate.encoder.params"):
        assert k in sd, k
    assert sd["model.xyz_encoder.params"].dtype == torch.float32 and sd["model.xyz_encoder.params"].dim() == 1   # tcnn's flat layout
    m2 = MNGP(0.5, size=2).cuda(); g2 = Ray_Gate(2).cuda()
    C.load_ckpt(m2, path, "model"); C.load_ckpt(g2, path, "ray_gate")
    assert torch.equal(m2.xyz_encoder.params, model.xyz_encoder.params)
    assert torch.equal(m2.density_bitfield_1, model.density_bitfield_1)
    assert torch.equal(g2.encoder.params, gate.encoder.params)
    slim = C.slim_ckpt(path)
    assert "model.grid_coords" not in slim and "model.xyz_encoder.params" in slim
    # the loaded model renders the same frame
    from radnerf_b200.engine import Engine
    from radnerf_b200 import synthetic as syn
    for i in range(2):
        sc = syn.make_scene(0.5, 1234, jitter_seed=i if i else None)
        bf = torch.from_numpy(syn.packbits_np(sc.density_grid, sc.threshold)).cuda()
        getattr(model, f"density_bitfield_{i}").copy_(bf); getattr(m2, f"density_bitfield_{i}").copy_(bf)
    cam = syn.make_cameras(4, 1.5, 24, 32, 40.0)
    o, d = syn.full_image_rays(cam, 1)
    o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    a = Engine(model, gate, n_rays=256).render(o, d)["rgb"]
    b = Engine(m2, g2, n_rays=256).render(o, d)["rgb"]
    assert torch.equal(a, b)
