"""Fits a Rad-NeRF (K sub-NeRFs + ray gate) to the analytic synthetic scene with the fused training engine -- live
occupancy updates, cosine schedule, device-side batch sampling -- and reports PSNR on held-out views rendered by the
fused test-time renderer (the "fitted" parameter state; 8f f1: the end-to-end driver).
Usage: python scripts/fit_scene.py [--steps 2000] [--K 2] [--res 200] [--batch 8192]  -> one JSON line."""
import argparse
import json
import math
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def fit(steps=2000, K=2, res=200, batch=8192, n_train=60, n_test=4, log_every=0, seed=0, log2_T=19, return_trainer=False):
    from radnerf_b200 import synthetic as syn
    from radnerf_b200.data import RaySampler, get_ray_directions
    from radnerf_b200.trainer import Trainer
    dev = torch.device("cuda:0")
    scene = syn.make_scene(0.5, 1234)
    fx = 1111.1 * res / 800.0
    cam = syn.make_cameras(n_train + n_test, 1.5, res, res, fx)
    poses = torch.from_numpy(np.ascontiguousarray(cam.poses[:, :3, :4])).float().to(dev)
    dirs = get_ray_directions(res, res, fx, device=dev)
    every = RaySampler(dirs, poses, None, batch_size=res * res)
    images = torch.empty(n_train + n_test, res * res, 3, device=dev)
    pix = torch.arange(res * res, device=dev)
    for i in range(n_train + n_test):
        o, d, _ = every.sample(torch.full((res * res,), i, device=dev, dtype=torch.int64), pix)
        images[i] = syn.analytic_colors(scene, o, d)
    test_ids = list(range(n_train, n_train + n_test))
    sampler = RaySampler(dirs, poses[:n_train].contiguous(), images[:n_train].contiguous(), batch_size=batch, seed=seed)
    tr = Trainer(scale=0.5, K=K, device="cuda:0", engine=True, batch_rays=batch, samples_per_ray=1024, total_steps=steps,
                 seed=seed, log2_T=log2_T)
    torch.cuda.synchronize()
    t0 = time.time()
    nxt = sampler.sample()
    curve = []
    for s in range(steps):
        o, d, rgb = nxt
        nxt = sampler.sample()
        loss = tr.step(o, d, rgb, next_rays=(nxt[0], nxt[1]))
        if log_every and (s + 1) % log_every == 0:
            curve.append((s + 1, float(loss)))
    torch.cuda.synchronize()
    train_s = time.time() - t0

    def psnr_of(ids):
        out = []
        for i in ids:
            o, d, _ = every.sample(torch.full((res * res,), i, device=dev, dtype=torch.int64), pix)
            r = tr.engine.render(o, d)
            mse = float(((r["rgb"] - images[i]) ** 2).mean())
            out.append(-10.0 * math.log10(max(mse, 1e-12)))
        return out
    test_psnr = psnr_of(test_ids)
    train_psnr = psnr_of([0, 1])
    occ = [float((b != 0).float().mean()) for b in tr.engine.bitfields()]
    info = {"steps": steps, "K": K, "res": res, "batch": batch, "train_views": n_train, "train_seconds": round(train_s, 2),
            "rays_per_s": round(steps * batch / train_s), "final_loss": float(loss), "psnr_test": [round(p, 2) for p in test_psnr],
            "psnr_test_mean": round(sum(test_psnr) / len(test_psnr), 2), "psnr_train_views": [round(p, 2) for p in train_psnr],
            "occupied_bitfield_bytes_frac": [round(x, 4) for x in occ], "loss_curve": curve}
    return (info, tr) if return_trainer else info


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--K", type=int, default=2)
    ap.add_argument("--res", type=int, default=200)
    ap.add_argument("--batch", type=int, default=8192)
    ap.add_argument("--log-every", type=int, default=250)
    ap.add_argument("--save", default=None, help="write the fitted flat parameters, grids and bitfields (bench_render_dist.py --fitted=FILE)")
    a = ap.parse_args()
    info, tr = fit(a.steps, a.K, a.res, a.batch, log_every=a.log_every, return_trainer=True)
    if a.save:
        torch.save({"flat": tr.engine.flat, "grids": tr.engine.grids(), "bitfields": tr.engine.bitfields(), "info": info}, a.save)
    print(json.dumps(info))
