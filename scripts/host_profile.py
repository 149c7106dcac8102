"""Host-side cost of the training loop: wall time per step of enqueueing (no device wait inside the loop) and a cProfile of
where the Python time goes.  Usage: python scripts/host_profile.py [--graph]"""
import cProfile
import os
import pstats
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import BATCH, make_batches, scene_bitfields, workload  # noqa: E402


def main():
    syn, w = workload("radnerf_tat_k2")
    from radnerf_b200.data import HostFeeder
    from radnerf_b200.trainer import Trainer
    graph = "--graph" in sys.argv
    tr = Trainer(scale=w["scale"], K=w["K"], freeze_occupancy=True, device="cuda:0", engine=True, batch_rays=BATCH, graph=graph)
    tr.set_occupancy(scene_bitfields(syn, w))
    host = [tuple(torch.from_numpy(a).pin_memory() for a in b) for b in make_batches(syn, w, 0)]
    feeder = HostFeeder(host, "cuda:0")
    torch.cuda.set_stream(torch.cuda.Stream(priority=-1))

    def loop(n, s0):
        for s in range(s0, s0 + n):
            o, d, t = feeder.get(s)
            nxt = feeder.peek(s + 1)
            tr.step(o, d, t, next_rays=nxt if graph else nxt[:2])
            feeder.done(s)
    loop(20, 0)
    torch.cuda.synchronize()
    t0 = time.perf_counter(); loop(200, 20); t1 = time.perf_counter()
    torch.cuda.synchronize(); t2 = time.perf_counter()
    print(f"{'graph replay' if graph else 'eager'}: host enqueue {1e3 * (t1 - t0) / 200:.3f} ms/step; with device drain {1e3 * (t2 - t0) / 200:.3f} ms/step")
    pr = cProfile.Profile(); pr.enable(); loop(200, 220); pr.disable()
    torch.cuda.synchronize()
    pstats.Stats(pr).sort_stats("cumulative").print_stats(28)


if __name__ == "__main__":
    main()
