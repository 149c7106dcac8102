"""GPU timing anchor: the reference's OWN kernels (oracle/_ref/vren_ref*.so = /root/reference/models/csrc
compiled unmodified for sm_100a by oracle/build_ref.sh) next to this repository's operators on identical inputs of the
training shape.  Test infrastructure, not product: it lives under tests/ because it loads oracle/_ref.  Not collected by
pytest; run on a GPU box:   python tests/anchor_vren_ref.py | tee profiles/rNN_anchor_vren_ref.jsonl
One JSON line per operator: ms of the reference kernel path, ms of ours, ratio; outputs are compared as they go
(integer outputs bit-exact, floats to 1e-5)."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
from bench import workload  # noqa: E402


def main():
    import vren_ref
    from radnerf_b200 import synthetic as syn, vren
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

    def line(op, ms_ref, ms_ours, note=""):
        print(json.dumps({"op": op, "reference_kernels_ms": round(ms_ref, 4), "ours_ms": round(ms_ours, 4),
                          "speedup": round(ms_ref / ms_ours, 2), "note": note}), flush=True)

    _, w = workload("radnerf_tat_k2")
    cam = syn.make_cameras(100, w["radius"], w["H"], w["W"], w["fx"])
    for R in ((8192,) if "--training-shape-only" in sys.argv else (8192, 8192 * 16)):
        o, d, _, _ = syn.train_batch(cam, R, 42)
        o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
        sc = syn.make_scene(0.5, 1234)
        grid = torch.from_numpy(sc.density_grid).cuda()
        bf = torch.zeros(128 ** 3 // 8, dtype=torch.uint8, device="cuda")
        bf2 = torch.zeros_like(bf)
        if R == 8192:
            a = timed(lambda: vren_ref.packbits(grid, float(sc.threshold), bf2))
            b = timed(lambda: vren.packbits(grid, float(sc.threshold), bf))
            assert torch.equal(bf, bf2)
            line("packbits (1 cascade, 128^3)", a, b)
        else:
            vren.packbits(grid, float(sc.threshold), bf)
        center = torch.zeros(1, 3, device="cuda"); half = torch.full((1, 3), 0.5, device="cuda")
        a = timed(lambda: vren_ref.ray_aabb_intersect(o, d, center, half, 1))
        b = timed(lambda: vren.ray_aabb_intersect(o, d, center, half, 1))
        _, hits_t, _ = vren.ray_aabb_intersect(o, d, center, half, 1)
        assert torch.equal(hits_t, vren_ref.ray_aabb_intersect(o, d, center, half, 1)[1])
        line(f"ray_aabb_intersect ({R} rays)", a, b)
        h = hits_t[:, 0].contiguous()
        h[(h[:, 0] >= 0) & (h[:, 0] < syn.NEAR_DISTANCE), 0] = syn.NEAR_DISTANCE
        noise = torch.rand(R, device="cuda")
        args = (o, d, h, bf, 1, 0.5, 0.0, noise, 128, 1024)
        a = timed(lambda: vren_ref.raymarching_train(*args))
        b = timed(lambda: vren.raymarching_train(*args))
        ref = vren_ref.raymarching_train(*args)
        rays_a, xyzs, dirs, deltas, ts, counter = vren.raymarching_train(*args)
        N = int(counter[0])
        # the reference orders segments by atomic arrival: compare per-ray counts and each ray's parameters
        rc = torch.zeros(R, dtype=torch.int64, device="cuda"); rc[ref[0][:, 0]] = ref[0][:, 2]
        oc = torch.zeros(R, dtype=torch.int64, device="cuda"); oc[rays_a[:, 0]] = rays_a[:, 2]
        assert torch.equal(rc, oc) and int(ref[5][0]) == N
        assert torch.equal(torch.sort(ref[4][:N])[0], torch.sort(ts[:N])[0])
        line(f"raymarching_train ({R} rays, {N} samples)", a, b, "operator level, incl. the counter read-back both sides")
        sig = torch.rand(N, device="cuda") * 20; rgb = torch.rand(N, 3, device="cuda")
        dl, tt = deltas[:N].contiguous(), ts[:N].contiguous()
        a = timed(lambda: vren_ref.composite_train_fw(sig, rgb, dl, tt, rays_a, 1e-4))
        b = timed(lambda: vren.composite_train_fw(sig, rgb, dl, tt, rays_a, 1e-4))
        r1 = vren_ref.composite_train_fw(sig, rgb, dl, tt, rays_a, 1e-4)
        tot, op, dp, col, ws = vren.composite_train_fw(sig, rgb, dl, tt, rays_a, 1e-4)
        assert torch.allclose(r1[3], col, atol=1e-5) and torch.allclose(r1[4], ws, atol=1e-6)
        line(f"composite_train_fw ({N} samples)", a, b)
        g1 = torch.rand(R, device="cuda"); g3 = torch.rand(R, 3, device="cuda"); gw = torch.zeros(N, device="cuda")
        bw = (g1, g1, g3, gw, sig, rgb, ws, dl, tt, rays_a, op, dp, col, 1e-4)
        a = timed(lambda: vren_ref.composite_train_bw(*bw))
        b = timed(lambda: vren.composite_train_bw(*bw))
        ga, gb = vren_ref.composite_train_bw(*bw), vren.composite_train_bw(*bw)
        assert torch.allclose(ga[0], gb[0], rtol=1e-3, atol=1e-5)
        line(f"composite_train_bw ({N} samples)", a, b)
        a = timed(lambda: vren_ref.distortion_loss_fw(ws, dl, tt, rays_a))
        b = timed(lambda: vren.distortion_loss_fw(ws, dl, tt, rays_a))
        line(f"distortion_loss_fw ({N} samples)", a, b)


if __name__ == "__main__":
    main()
