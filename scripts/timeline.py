"""Device timeline of the overlapped training step (march-ahead and gate streams running next to the step's own
stream): start/end of every stage from CUDA events, for the last of a few steps.  Usage: python scripts/timeline.py"""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import BATCH, STAGES, make_batches, scene_bitfields, workload  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "radnerf_tat_k2"
    syn, w = workload(name)
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:       # under torchrun: the multi-rank step (peer-memory optimizer, stream barriers) on rank 0's clock
        import torch.distributed as dist
        opts = dist.ProcessGroupNCCL.Options(); opts.is_high_priority_stream = True
        dist.init_process_group("nccl", device_id=torch.device("cuda", local), pg_options=opts)
    from radnerf_b200 import _lib
    from radnerf_b200.trainer import Trainer
    lib = _lib.lib()
    tr = Trainer(scale=w["scale"], K=w["K"], freeze_occupancy=True, device=f"cuda:{local}", engine=True, batch_rays=BATCH)
    tr.set_occupancy(scene_bitfields(syn, w))
    dev = [tuple(torch.from_numpy(a).cuda() for a in b) for b in make_batches(syn, w, rank)]
    flush = torch.empty(256 * 1024 * 1024 // 4, device="cuda")
    torch.cuda.set_stream(torch.cuda.Stream(priority=-1))

    def step(s):
        torch.cuda.current_stream().wait_stream(tr.engine._march_stream) if tr.engine._march_stream is not None else None
        flush.add_(1.0)
        tr.step(*dev[s % len(dev)], next_rays=dev[(s + 1) % len(dev)][:2])
    for s in range(1, 9):       # steps 1..8: no occupancy update inside
        step(s)
    torch.cuda.synchronize()
    lib.rnb_engine_profile(1)
    for s in range(9, 15):
        step(s)
    n_max = 256
    st = (ctypes.c_int32 * n_max)(); t0 = (ctypes.c_double * n_max)(); t1 = (ctypes.c_double * n_max)(); n = ctypes.c_int32(0)
    lib.rnb_engine_profile_timeline(st, t0, t1, n_max, ctypes.byref(n))
    lib.rnb_engine_profile(0)
    if rank != 0:
        dist.destroy_process_group()
        return
    recs = sorted((t0[i], t1[i], STAGES[st[i]]) for i in range(n.value))
    # print the last full step: from the last field_fwd's predecessor gate to its adam
    adams = [r for r in recs if r[2] == "adam"]
    print("adam end -> adam end (us, includes the 256 MiB flush):", [round(1e3 * (adams[i + 1][1] - adams[i][1]), 1) for i in range(len(adams) - 1)])
    lo = adams[-2][1] if len(adams) > 1 else 0.0
    print(f"{'stage':14s} {'start us':>9s} {'end us':>9s} {'dur us':>8s}")
    for a, b, nm in recs:
        if a >= lo - 1e-6:
            print(f"{nm:14s} {1e3 * (a - lo):9.1f} {1e3 * (b - lo):9.1f} {1e3 * (b - a):8.1f}")
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
