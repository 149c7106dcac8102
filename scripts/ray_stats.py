"""Distribution of samples per (sub-NeRF, ray) on the bench workload: the longest ray sets the duration of every
warp-per-ray kernel (march, expand, compositing).  Usage: python scripts/ray_stats.py [workload]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import BATCH, make_batches, scene_bitfields, workload  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "radnerf_tat_k2"
    syn, w = workload(name)
    from radnerf_b200.trainer import Trainer
    tr = Trainer(scale=w["scale"], K=w["K"], freeze_occupancy=True, device="cuda:0", engine=True, batch_rays=BATCH)
    tr.set_occupancy(scene_bitfields(syn, w))
    b = [torch.from_numpy(a).cuda() for a in make_batches(syn, w, 0)[0]]
    tr.step(*b)
    torch.cuda.synchronize()
    c = tr.engine.counts.flatten().float()
    nz = c[c > 0]
    q = torch.tensor([0.5, 0.9, 0.99, 0.999])
    print(f"{name}: rows {c.numel()}, non-empty {nz.numel()}, mean {c.mean():.1f}, mean(non-empty) {nz.mean():.1f}, "
          f"quantiles(non-empty) {torch.quantile(nz, q.cuda()).tolist()}, max {int(c.max())}")
    hist = torch.histc(c, bins=16, min=0, max=1024).int().tolist()
    print("histogram (64-sample bins):", hist)


if __name__ == "__main__":
    main()
