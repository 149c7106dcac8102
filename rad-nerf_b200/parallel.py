"""Host-side data-parallel plumbing of the path: one process per GPU, rays shard by rank,
parameters are replicated.  Two exchanges per training step, both plain SUM all-reduces over torch.distributed
(NCCL on GPUs, gloo in the CPU tests):

  1. the K gate-importance sums, before the CV loss (losses.py:64-68 is non-linear in the batch-wide importance);
  2. the flat gradient vector -- every loss term is normalised by the GLOBAL element count, so the sum over ranks
     IS the gradient of the single-process step on the concatenated batch; no rescale afterwards.

Test-time frames shard by interleaved image rows (row % world == rank: bands differ in occupancy, interleaving
balances them); no communication until the final gather.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def world_size() -> int:
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def rank() -> int:
    return dist.get_rank() if dist.is_available() and dist.is_initialized() else 0


def allreduce_sum_(t: torch.Tensor) -> torch.Tensor:
    """In-place SUM over ranks (no-op for a single process)."""
    if world_size() > 1:
        dist.all_reduce(t)
    return t


def global_importance(local: torch.Tensor) -> torch.Tensor:
    """Batch-wide gate importance whose gradient still flows to this rank's rays only: value = sum over ranks,
    d(value)/d(local) = 1.  (The fused engine all-reduces its importance buffer in place instead.)"""
    if world_size() == 1:
        return local
    tot = local.detach().clone()
    dist.all_reduce(tot)
    return local + (tot - local.detach())


def shard_slice(n_global: int, rk: int | None = None, world: int | None = None) -> slice:
    """Rows of a global ray batch owned by a rank (strong scaling: the global batch is split evenly)."""
    rk = rank() if rk is None else rk
    world = world_size() if world is None else world
    per = (n_global + world - 1) // world
    return slice(min(rk * per, n_global), min((rk + 1) * per, n_global))


def total_loss(loss_terms: dict, world: int | None = None) -> torch.Tensor:
    """Sum of term means over the GLOBAL batch, as seen from one rank: per-ray terms (any tensor with a leading ray
    dimension) contribute sum/(numel*world); the CV term is one global scalar that every rank computes identically from
    the all-reduced importance -- its gradient pieces add up over ranks, its VALUE is shared, so it enters each rank's
    backward in full and the reported loss is de-duplicated by `reported_loss`."""
    world = world_size() if world is None else world
    tot = 0.0
    for v in loss_terms.values():
        tot = tot + (v.sum() / (v.numel() * world) if v.dim() > 0 else v)
    return tot


def reported_loss(loss_terms: dict, world: int | None = None) -> torch.Tensor:
    """The global loss value: all-reduced per-ray terms plus the (shared) scalar terms counted once."""
    world = world_size() if world is None else world
    per_ray = sum((v.detach().sum() / (v.numel() * world) for v in loss_terms.values() if v.dim() > 0), torch.zeros(()))
    per_ray = per_ray.clone()
    allreduce_sum_(per_ray)
    return per_ray + sum((v.detach() for v in loss_terms.values() if v.dim() == 0), torch.zeros((), device=per_ray.device))


def frame_rows(H: int, rk: int | None = None, world: int | None = None) -> torch.Tensor:
    """Image rows rendered by a rank: interleaved assignment."""
    rk = rank() if rk is None else rk
    world = world_size() if world is None else world
    return torch.arange(rk, H, world)


def gather_frame(local: torch.Tensor, H: int, W: int) -> torch.Tensor:
    """local: (rows_of_this_rank * W, C) results of frame_rows(H) -> the full (H*W, C) frame on every rank."""
    world = world_size()
    C = local.shape[-1] if local.dim() > 1 else 1
    loc = local.reshape(-1, W, C)
    if world == 1:
        return loc.reshape(H * W, C)
    max_rows = (H + world - 1) // world
    pad = torch.zeros(max_rows, W, C, dtype=local.dtype, device=local.device)
    pad[:loc.shape[0]] = loc
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad)
    out = torch.empty(H, W, C, dtype=local.dtype, device=local.device)
    for r in range(world):
        rows = torch.arange(r, H, world)
        out[rows] = parts[r][:len(rows)]
    return out.reshape(H * W, C)


# ---- peer memory (NVLink) for the fused reduce-scatter + Adam + all-gather kernel (csrc/peer.cu) -----------------------
class _RawCuda:
    """Exposes a raw device allocation through __cuda_array_interface__ so that torch can view it."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class PeerBuffer:
    """A device buffer of `nbytes` that every rank of the node can address: allocated here with rnb_peer_alloc, its CUDA
    IPC handle exchanged over torch.distributed, the peers' allocations mapped with rnb_peer_open.
    .local: uint8 torch view of this rank's allocation; .ptrs: this buffer on every rank, as mapped in this process."""

    def __init__(self, nbytes, device):
        import ctypes
        from . import _lib
        self._lib = _lib
        lib = _lib.lib()
        ptr = ctypes.c_void_p()
        handle = (ctypes.c_ubyte * 64)()
        _lib.call("rnb_peer_alloc", ctypes.c_int64(nbytes), ctypes.byref(ptr), handle)
        self.ptr = ptr.value
        self.local = torch.as_tensor(_RawCuda(self.ptr, nbytes), device=device)
        world, rk = world_size(), rank()
        handles = [None] * world
        dist.all_gather_object(handles, bytes(handle))
        self.ptrs = []
        self._opened = []
        for r in range(world):
            if r == rk:
                self.ptrs.append(self.ptr)
                continue
            q = ctypes.c_void_p()
            buf = (ctypes.c_ubyte * 64).from_buffer_copy(handles[r])
            _lib.call("rnb_peer_open", buf, ctypes.byref(q))
            self.ptrs.append(q.value)
            self._opened.append(q.value)
        del lib

    def pointer_array(self):
        import ctypes
        return (ctypes.c_void_p * len(self.ptrs))(*self.ptrs)


def adam_shard(n: int, rk: int | None = None, world: int | None = None) -> slice:
    """Slice of the flat parameter vector a rank owns in the sharded optimizer (multiples of 4 parameters; mirrors
    rnb_adam_shard_bounds)."""
    rk = rank() if rk is None else rk
    world = world_size() if world is None else world
    n4 = n // 4
    per = (n4 + world - 1) // world
    return slice(min(per * rk, n4) * 4, min(per * (rk + 1), n4) * 4)


_token = {}


def stream_barrier(device) -> None:
    """A collective that orders the ranks' CURRENT STREAMS without blocking the host: everything enqueued before it on
    every rank completes before anything enqueued after it on any rank starts."""
    if world_size() == 1:
        return
    t = _token.get(device)
    if t is None:
        t = _token[device] = torch.zeros(1, device=device)
    dist.all_reduce(t)
