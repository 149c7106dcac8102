"""world_size-2 `gloo` tests of the data-parallel host logic (radnerf_b200/parallel.py), on CPU.

The claim under test: with rays sharded by rank, the gate importance all-reduced
before the CV loss and every loss term normalised by the GLOBAL element count, the SUM all-reduce of the per-rank
gradients equals the gradient of the single-process step on the concatenated batch.  The per-rank computation is the
CPU oracle (oracle/render_cpu.py -- test infrastructure); the exchange goes through the product's own helpers.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import get_scene
from radnerf_b200 import parallel
from radnerf_b200 import synthetic as syn

SCALE, K, N, LOG2_T = 0.5, 2, 96, 12


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _inputs():
    cam = syn.make_cameras(100, 1.5)
    o, d, _, _ = syn.train_batch(cam, N, 42)
    target = np.random.RandomState(1).rand(N, 3).astype(np.float32)
    noise = np.random.RandomState(2).rand(K, N).astype(np.float32)
    bitfields = [get_scene(SCALE, jitter=i if i else None)[1] for i in range(K)]
    return o, d, target, noise, bitfields


def _grads(P):
    return torch.cat([P.p[n].grad.reshape(-1) for n in sorted(P.p)])


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import render_cpu as R
        torch.set_num_threads(2)
        o, d, target, noise, bitfields = _inputs()
        cas = syn.cascades_for_scale(SCALE)
        # ---- this rank's shard through the product's exchange helpers ----
        sl = parallel.shard_slice(N)
        assert parallel.world_size() == world and parallel.rank() == rank
        P = R.FieldParams(SCALE, K, log2_T=LOG2_T, seed=0, grid_init=0.2); P.requires_grad_(True)
        res = R.ml_render_train(P, o[sl], d[sl], bitfields, [noise[i][sl] for i in range(K)], cas, 0.0)
        res["gating_importance"] = parallel.global_importance(res["gating_importance"])
        _, terms = R.nerf_loss(res, torch.from_numpy(target[sl]), 1e-3, 1e-2, 5e-3)
        loss = parallel.total_loss(terms)
        loss.backward()
        flat = _grads(P)
        parallel.allreduce_sum_(flat)
        reported = parallel.reported_loss(terms)
        # ---- single-process reference on the whole batch ----
        Pr = R.FieldParams(SCALE, K, log2_T=LOG2_T, seed=0, grid_init=0.2); Pr.requires_grad_(True)
        rres = R.ml_render_train(Pr, o, d, bitfields, [noise[i] for i in range(K)], cas, 0.0)
        rloss, _ = R.nerf_loss(rres, torch.from_numpy(target), 1e-3, 1e-2, 5e-3)
        rloss.backward()
        ref = _grads(Pr)
        rel = float((flat - ref).norm() / ref.norm())
        out[rank] = (rel, float(reported), float(rloss), float(res["gating_importance"].sum()), float(rres["gating_importance"].sum()))
        # frame sharding: interleaved rows reassemble to the full frame
        H, W = 7, 5
        rows = parallel.frame_rows(H)
        full = torch.arange(H * W * 3, dtype=torch.float32).reshape(H, W, 3)
        got = parallel.gather_frame(full[rows].reshape(-1, 3), H, W)
        assert torch.equal(got, full.reshape(-1, 3))
    finally:
        dist.destroy_process_group()


def test_sharded_gradient_equals_single_process():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    assert len(out) == world
    for r in range(world):
        rel, reported, rloss, imp, rimp = out[r]
        assert rel < 1e-4, rel                              # fp32 summation order only
        assert abs(reported - rloss) < 1e-5 * max(1.0, abs(rloss))
        assert abs(imp - rimp) < 1e-3


def test_single_process_helpers_are_identity():
    t = torch.arange(4.0)
    assert parallel.world_size() == 1 and parallel.rank() == 0
    assert parallel.allreduce_sum_(t) is t and torch.equal(t, torch.arange(4.0))
    assert parallel.global_importance(t) is t
    assert parallel.shard_slice(10, 1, 4) == slice(3, 6) and parallel.shard_slice(10, 3, 4) == slice(9, 10)
    assert parallel.frame_rows(7, 1, 3).tolist() == [1, 4]
    terms = {"rgb": torch.ones(8, 3) * 2.0, "cv": torch.tensor(0.5)}
    assert abs(float(parallel.total_loss(terms)) - 2.5) < 1e-6
    assert abs(float(parallel.reported_loss(terms)) - 2.5) < 1e-6


def test_adam_shards_tile_the_parameter_vector():
    """parallel.adam_shard mirrors rnb_adam_shard_bounds (csrc/peer.cu): shards are multiples of 4 parameters, disjoint,
    ordered, and cover [0, n) for every world size -- checked against the C function through the C ABI (no GPU needed)."""
    import ctypes
    from radnerf_b200 import _lib
    lib = _lib.lib()
    for n in (8, 11_400_000, 4 * 7 + 4, 1024):
        for world in (1, 2, 3, 4, 8):
            end = 0
            for r in range(world):
                sl = parallel.adam_shard(n, r, world)
                lo, hi = ctypes.c_int64(), ctypes.c_int64()
                assert lib.rnb_adam_shard_bounds(ctypes.c_int64(n), r, world, ctypes.byref(lo), ctypes.byref(hi)) == 0
                assert (sl.start, sl.stop) == (lo.value, hi.value)
                assert sl.start == end and sl.start % 4 == 0 and sl.stop >= sl.start
                end = sl.stop
            assert end == n
