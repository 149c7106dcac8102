"""Operator-level roofline table (metric 3: "HBM % of peak for encode and composite, tensor-pipe use for
the MLPs"): every drop-in operator on inputs of the training shape, CUDA-event timed, L2 flushed between repetitions.
Achieved = ALGORITHMIC bytes (or FLOP) per call / time; peaks from MEASURED_PEAKS.json.
Usage: python scripts/bench_ops.py  -> one JSON line per operator (tee into profiles/)."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import measured_peaks, scene_bitfields, workload  # noqa: E402


def main():
    from radnerf_b200 import synthetic as syn, vren, tcnn
    from radnerf_b200.models.networks import MNGP
    hbm, tflops, which = measured_peaks()
    flush = torch.empty(256 * 1024 * 1024 // 4, device="cuda")

    def timed(fn, reps=10):
        fn(); torch.cuda.synchronize()
        tot = 0.0
        for _ in range(reps):
            flush.add_(1.0)
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record()
            torch.cuda.synchronize()
            tot += e0.elapsed_time(e1)
        return tot / reps

    def line(op, ms, nbytes=0, flops=0, note=""):
        gbs = nbytes / (ms * 1e-3) / 1e9
        tf = flops / (ms * 1e-3) / 1e12
        print(json.dumps({"op": op, "ms": round(ms, 4), "algorithmic_GB": round(nbytes / 1e9, 4), "GB/s": round(gbs, 1),
                          "hbm_frac": round(gbs / hbm, 4), "TFLOP/s": round(tf, 2), "tensor_frac": round(tf / tflops, 4),
                          "peaks": f"{hbm} GB/s, {tflops} TF/s ({which})", "note": note}), flush=True)

    # ---- inputs: 64 batches' worth of rays so that streaming operators run long enough to measure ----
    _, w = workload("radnerf_tat_k2")
    cam = syn.make_cameras(100, w["radius"], w["H"], w["W"], w["fx"])
    R = 8192 * 16
    o, d, _, _ = syn.train_batch(cam, R, 42)
    o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    sc = syn.make_scene(0.5, 1234)
    grid = torch.from_numpy(sc.density_grid).cuda()
    bf = torch.zeros(128 ** 3 // 8, dtype=torch.uint8, device="cuda")
    ms = timed(lambda: vren.packbits(grid, float(sc.threshold), bf))
    line("packbits", ms, 4.125 * grid.numel(), note="1 cascade of 128^3 cells (8.7 MB: launch-latency sized)")
    big = torch.rand(64, 128 ** 3, device="cuda")
    bbf = torch.zeros(64 * 128 ** 3 // 8, dtype=torch.uint8, device="cuda")
    ms = timed(lambda: vren.packbits(big, 0.5, bbf))
    line("packbits x64", ms, 4.125 * big.numel(), note="64 cascades (554 MB): streaming rate")
    center = torch.zeros(1, 3, device="cuda"); half = torch.full((1, 3), 0.5, device="cuda")
    ms = timed(lambda: vren.ray_aabb_intersect(o, d, center, half, 1))
    line("ray_aabb_intersect", ms, 44 * R, note=f"{R} rays")
    _, hits_t, _ = vren.ray_aabb_intersect(o, d, center, half, 1)
    h = hits_t[:, 0].contiguous()
    h[(h[:, 0] >= 0) & (h[:, 0] < syn.NEAR_DISTANCE), 0] = syn.NEAR_DISTANCE
    noise = torch.rand(R, device="cuda")
    out = vren.raymarching_train(o, d, h, bf, 1, 0.5, 0.0, noise, 128, 1024)
    rays_a, xyzs, dirs, deltas, ts, counter = out
    N = int(counter[0])
    ms = timed(lambda: vren.raymarching_train(o, d, h, bf, 1, 0.5, 0.0, noise, 128, 1024))
    line("raymarching_train", ms, 60 * R + 32 * N, note=f"{R} rays, {N} samples (incl. the wrapper's counter read-back)")
    sig = torch.rand(N, device="cuda") * 20; rgb = torch.rand(N, 3, device="cuda")
    ms = timed(lambda: vren.composite_train_fw(sig, rgb, deltas, ts, rays_a, 1e-4))
    line("composite_train_fw", ms, 28 * N + 52 * R, note=f"{N} samples")
    tot, op, dp, col, ws = vren.composite_train_fw(sig, rgb, deltas, ts, rays_a, 1e-4)
    g1 = torch.rand(R, device="cuda"); g3 = torch.rand(R, 3, device="cuda"); gw = torch.zeros(N, device="cuda")
    ms = timed(lambda: vren.composite_train_bw(g1, g1, g3, gw, sig, rgb, ws, deltas, ts, rays_a, op, dp, col, 1e-4))
    line("composite_train_bw", ms, 48 * N + 64 * R, note=f"{N} samples")
    ms = timed(lambda: vren.distortion_loss_fw(ws, deltas, ts, rays_a))
    line("distortion_loss_fw", ms, 20 * N + 28 * R, note="incl. the wrapper's three zero-fills")
    # ---- field operators (operator level: the tinycudann-shaped modules) ----
    model = MNGP(0.5, size=2).cuda()
    x = ((xyzs + 0.5).clamp(0, 1)).contiguous()
    enc = model.xyz_encoder
    with torch.no_grad():
        ms = timed(lambda: enc(x))
    line("hashgrid_fwd (operator)", ms, 76 * N, note="12 B in + 64 B out per sample; gathers hit the L2-resident table (512 B/sample of L2 traffic)")
    feat = enc(x)
    gy = torch.randn_like(feat) * 1e-3
    xr = x.clone()
    def enc_bw():
        enc.params.grad = None
        y = enc(xr)
        y.backward(gy)
    ms_fb = timed(enc_bw)
    line("hashgrid fwd+bwd (operator, autograd)", ms_fb, 2 * 76 * N, note="scatter = 512 B/sample of L2 reductions")
    geo = model.geo_net_0
    with torch.no_grad():
        ms = timed(lambda: geo(feat))
    line("geo MLP fwd (32-64-17)", ms, (64 + 64) * N, flops=2 * (32 * 64 + 64 * 17) * N, note="operator-level (mma.sync) kernel")
    rgbn = model.rgb_net_0
    xin = torch.randn(N, 32, device="cuda", dtype=torch.float16)
    with torch.no_grad():
        ms = timed(lambda: rgbn(xin))
    line("rgb MLP fwd (32-64-64-3)", ms, (64 + 32) * N, flops=2 * (32 * 64 + 64 * 64 + 64 * 3) * N, note="operator-level (mma.sync) kernel")


if __name__ == "__main__":
    main()
