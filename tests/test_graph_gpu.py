"""CUDA-graph replay of the training step (radnerf_b200.engine.GraphStepper, Trainer(graph=True)) against the
eager step.  Both draw their marching noise from the library's counter generator (keyed by the marching-call count), so
two trainers with the same seed see the same rays, the same noise and the same occupancy updates: per-step sample counts
must be IDENTICAL (integer work), losses equal up to the order of the fp32 gradient reductions."""
import numpy as np
import pytest
import torch

from conftest import get_scene
from radnerf_b200 import synthetic as syn

pytestmark = pytest.mark.gpu


def _trainer(graph, K=2, n=2048):
    from radnerf_b200.trainer import Trainer
    tr = Trainer(scale=0.5, K=K, freeze_occupancy=True, device="cuda", engine=True, batch_rays=n, graph=graph, seed=0)
    tr.set_occupancy([get_scene(0.5, jitter=i if i else None)[1] for i in range(K)])
    return tr


def _batches(n, count):
    cam = syn.make_cameras(100, 1.5)
    out = []
    for b in range(count):
        o, d, _, _ = syn.train_batch(cam, n, 500 + b)
        t = np.random.RandomState(b).rand(n, 3).astype(np.float32)
        out.append(tuple(torch.from_numpy(a).cuda() for a in (o, d, t)))
    return out


@pytest.mark.parametrize("K", [2, 1])
def test_graph_replay_matches_eager(K):
    n, steps = 2048, 40                      # crosses two occupancy updates (every 16 steps): kinds first / steady / last
    bs = _batches(n, 8)
    runs = {}
    for graph in (False, True):
        tr = _trainer(graph, K, n)
        losses, counts = [], []
        for s in range(steps):
            o, d, t = bs[s % len(bs)]
            nxt = bs[(s + 1) % len(bs)]
            loss = tr.step(o, d, t, next_rays=nxt if graph else nxt[:2])
            losses.append(loss.clone())
            counts.append(tr.engine._slot_view("seg", torch.int32, (2 * 8 + 4,), tr.engine._slot)[:2 * K + 3].clone())
        torch.cuda.synchronize()
        st = tr.engine.guard_status()
        runs[graph] = ([float(x) for x in losses], [c.tolist() for c in counts], st, tr)
    le, ce, se, _ = runs[False]
    lg, cg, sg, trg = runs[True]
    assert trg.stepper is not None and trg.stepper.replays >= steps - 4 and len(trg.stepper.graphs) >= (2 if trg.stepper.hybrid else 3)
    assert se["applied"] == sg["applied"] == steps and sg["skipped_nonfinite"] == 0
    assert ce == cg                                                       # same samples per pass in every step
    for a, b in zip(le, lg):
        assert abs(a - b) < 2e-2 * max(abs(a), 1e-3), (a, b)
    assert lg[-1] < lg[0]                                                 # and it trains


def test_graph_replay_follows_the_learning_rate_schedule():
    """lr is read from device memory at replay time: a graph captured at one lr must apply the lr written before each replay."""
    n = 1024
    bs = _batches(n, 4)
    tr = _trainer(True, 2, n)
    tr.update_interval = 10 ** 9
    for s in range(6):
        tr.step(*bs[s % 4], next_rays=bs[(s + 1) % 4])
    torch.cuda.synchronize()
    before = tr.engine.flat.clone()
    tr.lr = 0.0                                  # (Trainer.step passes self.lr)
    tr.step(*bs[2], next_rays=bs[3])
    torch.cuda.synchronize()
    assert torch.equal(tr.engine.flat, before)   # lr = 0: the replayed Adam moved nothing
    tr.lr = 1e-2
    tr.step(*bs[3], next_rays=bs[0])
    torch.cuda.synchronize()
    assert not torch.equal(tr.engine.flat, before)
