"""Two-GPU test of the data-parallel step (skipped on a single-GPU box; run with `gpurun --gpus 2`).

Two NCCL ranks train on halves of a 2048-ray batch with the sharded optimizer (csrc/peer.cu: reduce-scatter + Adam +
all-gather in one kernel over NVLink peer memory); a single process trains on the whole batch.  After three steps:
  * both ranks hold bit-identical fp16 parameters (every shard is written by its one owner);
  * they agree with the single-process parameters except where the gradient is at rounding level (Adam's early steps
    are ~lr * sign(g));
  * the reported losses agree (every term is normalised by the global ray count);
  * the sharded path and the plain all-reduce path (sharded_optimizer=False) produce the same parameters.
"""
import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

SCALE, K, N, STEPS = 0.5, 2, 2048, 3


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _make(n_rays, sharded=True):
    from radnerf_b200 import synthetic as syn
    from radnerf_b200.engine import Engine
    from radnerf_b200.models.networks import MNGP, Ray_Gate
    torch.manual_seed(0)
    model = MNGP(SCALE, size=K).cuda()
    gate = Ray_Gate(K).cuda()
    with torch.no_grad():
        model.xyz_encoder.params.uniform_(-0.2, 0.2)
    for i in range(K):
        sc = syn.make_scene(SCALE, 1234, jitter_seed=i if i else None)
        getattr(model, f"density_bitfield_{i}").copy_(torch.from_numpy(syn.packbits_np(sc.density_grid, sc.threshold)).cuda())
    return Engine(model, gate, n_rays=n_rays, exp_step_factor=0.0, lr=1e-2, sharded_optimizer=sharded)


def _batches():
    from radnerf_b200 import synthetic as syn
    cam = syn.make_cameras(100, 1.5)
    out = []
    for b in range(STEPS):
        o, d, _, _ = syn.train_batch(cam, N, 300 + b)
        t = np.random.RandomState(b).rand(N, 3).astype(np.float32)
        nz = np.random.RandomState(100 + b).rand(K, N).astype(np.float32)
        out.append((o, d, t, nz))
    return out


def _worker(rank, world, port, out, sharded):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        eng = _make(N // world, sharded)
        assert eng.sharded == sharded
        per = N // world
        sl = slice(rank * per, (rank + 1) * per)
        losses = []
        for o, d, t, nz in _batches():
            cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
            loss = eng.step(cu(o[sl]), cu(d[sl]), cu(t[sl]), noise=cu(nz[:, sl]))
            tot = loss[:1].clone()
            dist.all_reduce(tot)              # per-ray terms add up; the shared CV term is counted `world` times
            losses.append((float(tot[0]), float(loss[3])))
        torch.cuda.synchronize()
        eng.gather_master()
        out[rank] = (eng.flat16.cpu().clone(), eng.flat.cpu().clone(), losses)
    finally:
        dist.destroy_process_group()


def _run(world, sharded):
    import torch.multiprocessing as mp
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out, sharded), nprocs=world, join=True)
    return out


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_rank_step_matches_single_process():
    res = _run(2, True)
    p16_0, p32_0, losses0 = res[0]
    p16_1, p32_1, losses1 = res[1]
    assert torch.equal(p16_0, p16_1)                      # one owner per shard: replicas cannot drift
    assert torch.equal(p32_0, p32_1)                      # after gather_master
    # single process on the whole batch
    eng = _make(N, False)
    ref_losses = []
    for o, d, t, nz in _batches():
        cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
        loss = eng.step(cu(o), cu(d), cu(t), noise=cu(nz))
        ref_losses.append((float(loss[0]), float(loss[3])))
    torch.cuda.synchronize()
    for (tot, cv), (rtot, rcv) in zip(losses0, ref_losses):
        assert abs(cv - rcv) < 1e-4 * max(1.0, abs(rcv))
        assert abs((tot - cv) - rtot) < 1e-3 * max(1.0, abs(rtot))     # tot counts the shared CV term twice
    differ = ((p32_0 - eng.flat.cpu()).abs() > 1e-3).float().mean()
    assert float(differ) < 0.01
    # plain NCCL all-reduce + replicated Adam gives the same parameters as the fused peer-memory kernel
    res2 = _run(2, False)
    differ2 = ((res2[0][1] - p32_0).abs() > 1e-3).float().mean()
    assert float(differ2) < 0.01
    assert abs(res2[0][2][-1][0] - losses0[-1][0]) < 1e-3 * max(1.0, abs(losses0[-1][0]))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("width", [4, 8])
def test_wider_peer_kernel_instantiations(width, monkeypatch):
    """The 4- and 8-peer instantiations of the peer-memory optimizer kernel (used at N = 3..8) run here on two ranks
    through the RNB_PEER_MIN_WIDTH hook.  Same reduction order as the 2-peer kernel, but two runs never see the same
    gradient bits (the table scatter is fp32 atomics), so the comparison is the one used against the single process:
    replicas bit-identical within the run, parameters equal except where the gradient is at rounding level, same loss."""
    monkeypatch.delenv("RNB_PEER_MIN_WIDTH", raising=False)
    base = _run(2, True)
    monkeypatch.setenv("RNB_PEER_MIN_WIDTH", str(width))
    wide = _run(2, True)
    assert torch.equal(wide[0][0], wide[1][0]) and torch.equal(wide[0][1], wide[1][1])
    diff = (base[0][1] - wide[0][1]).abs()
    print("width", width, "max |dp|", float(diff.max()), "fraction > 1e-3", float((diff > 1e-3).float().mean()))
    assert float((diff > 1e-3).float().mean()) < 0.01
    for (a, _), (b, _) in zip(base[0][2], wide[0][2]):
        assert abs(a - b) < 1e-3 * max(1.0, abs(a))
