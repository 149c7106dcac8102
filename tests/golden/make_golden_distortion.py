"""Generates tests/golden/distortion.npz from the REFERENCE's own CUDA kernels (losses.cu) through
oracle/_ref/vren_ref*.so (oracle/build_ref.sh).  Run on a GPU box:
    gpurun -- python tests/golden/make_golden_distortion.py
Output: gpurun_out/golden/distortion.npz (copied into tests/golden/ and committed)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
import vren_ref  # noqa: E402


def main():
    rs = np.random.RandomState(7)
    counts = rs.randint(0, 90, 160).astype(np.int64)
    counts[[3, 17, 40]] = 0                      # rays without samples
    counts[5] = 1; counts[6] = 33; counts[7] = 64
    starts = np.concatenate([[0], np.cumsum(counts)[:-1]])
    perm = rs.permutation(len(counts))           # rays_a rows in arbitrary order, as the reference's atomics leave them
    rays_a = np.stack([perm, starts[perm], counts[perm]], 1).astype(np.int64)
    N = int(counts.sum())
    ws = (rs.rand(N) ** 3 * 0.2).astype(np.float32)
    deltas = (rs.rand(N) * 0.01 + 0.001).astype(np.float32)
    ts = np.zeros(N, np.float32)
    for s, n in zip(starts, counts):
        ts[s:s + n] = 0.3 + np.cumsum(deltas[s:s + n])
    dL = rs.randn(len(counts)).astype(np.float32)
    cu = lambda a: torch.from_numpy(a).cuda()
    loss, wi, ti = vren_ref.distortion_loss_fw(cu(ws), cu(deltas), cu(ts), cu(rays_a))
    dws = vren_ref.distortion_loss_bw(cu(dL), wi, ti, cu(ws), cu(deltas), cu(ts), cu(rays_a))
    out = os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(out, exist_ok=True)
    np.savez_compressed(os.path.join(out, "distortion.npz"), ws=ws, deltas=deltas, ts=ts, rays_a=rays_a, dL_dloss=dL,
                        loss=loss.cpu().numpy(), ws_inclusive_scan=wi.cpu().numpy(), wts_inclusive_scan=ti.cpu().numpy(),
                        dL_dws=dws.cpu().numpy())
    print("wrote", out, "N =", N)


if __name__ == "__main__":
    main()
