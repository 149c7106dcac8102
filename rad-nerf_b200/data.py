"""Device-resident training data for the path: camera-space pixel directions, poses and
images live in HBM; one kernel per step (rnb_batch_rays) draws the pixel batch, forms world-space rays and gathers the
target colours -- what the reference does on the host in BaseDataset.__getitem__ (datasets/base.py:22-29) followed by
train_ml.py:183-186 (index gathers + get_rays, datasets/ray_utils.py:46-70).  There is no CPU path."""
from __future__ import annotations

import ctypes

import torch

from . import _lib
from ._lib import call, i64, stream_arg


def get_ray_directions(H, W, fx, fy=None, cx=None, cy=None, device="cuda"):
    """datasets/ray_utils.py:7-43 with random=False: ((u - cx + 0.5)/fx, (v - cy + 0.5)/fy, 1), flattened row-major."""
    fy = fx if fy is None else fy
    cx = W / 2 if cx is None else cx
    cy = H / 2 if cy is None else cy
    v, u = torch.meshgrid(torch.arange(H, device=device, dtype=torch.float32), torch.arange(W, device=device, dtype=torch.float32),
                          indexing="ij")
    return torch.stack([(u - cx + 0.5) / fx, (v - cy + 0.5) / fy, torch.ones_like(u)], -1).reshape(-1, 3).contiguous()


class RaySampler:
    """directions (H*W,3), poses (n_img,3,4), images (n_img,H*W,3) or None -- all fp32 CUDA tensors."""

    def __init__(self, directions, poses, images=None, batch_size=8192, seed=0):
        _lib.lib()
        for t, n in ((directions, "directions"), (poses, "poses")) + (((images, "images"),) if images is not None else ()):
            _lib.check_input(t, n, torch.float32)
        assert poses.shape[1:] == (3, 4) and directions.shape[1] == 3
        assert images is None or images.shape == (poses.shape[0], directions.shape[0], 3)
        self.directions, self.poses, self.images = directions, poses, images
        self.batch_size, self.seed, self.calls = batch_size, seed, 0

    def sample(self, img_idxs=None, pix_idxs=None, return_indices=False):
        """-> rays_o, rays_d (B,3), rgb (B,3) or None [, img_idxs, pix_idxs].  Indices default to iid uniform draws."""
        dev = self.directions.device
        B = self.batch_size if img_idxs is None else img_idxs.shape[0]
        o = torch.empty(B, 3, device=dev); d = torch.empty(B, 3, device=dev)
        rgb = torch.empty(B, 3, device=dev) if self.images is not None else None
        io = torch.empty(B, dtype=torch.int64, device=dev) if return_indices else None
        po = torch.empty(B, dtype=torch.int64, device=dev) if return_indices else None
        self.calls += 1
        call("rnb_batch_rays", self.directions, self.poses, self.images, img_idxs, pix_idxs, i64(self.poses.shape[0]),
             i64(self.directions.shape[0]), i64(B), ctypes.c_uint64((self.seed << 32) + self.calls), o, d, rgb, io, po, stream_arg(dev))
        return (o, d, rgb, io, po) if return_indices else (o, d, rgb)


class HostFeeder:
    """Host -> device staging of training batches that live in pinned host memory (a DataLoader's output): batch n+1 is
    copied on a dedicated copy stream into one of `depth` preallocated device slots while step n computes, so the upload
    never sits between two kernels of the step.  `get(n)` returns the device tensors of batch n (the current stream is
    made to wait for their copy) and keeps the copies of the following `depth - 2` batches in flight."""

    def __init__(self, host_batches, device, depth=4, cycle=True):
        self.host = host_batches
        for b in host_batches:
            for t in b:
                assert t.is_pinned(), "HostFeeder wants pinned host tensors"
        self.device = torch.device(device)
        self.cycle = cycle
        self.stream = torch.cuda.Stream(self.device)
        self.slots = [tuple(torch.empty_like(t, device=self.device) for t in host_batches[0]) for _ in range(depth)]
        self.ready = {}              # batch index -> (slot, event)
        self.free_evt = [None] * depth   # last consumer of every slot
        self.bytes_per_batch = sum(t.numel() * t.element_size() for t in host_batches[0])

    def _issue(self, n):
        if n in self.ready or (not self.cycle and n >= len(self.host)):
            return
        k = n % len(self.slots)
        with torch.cuda.stream(self.stream):
            if self.free_evt[k] is not None:
                self.stream.wait_event(self.free_evt[k])       # the step that last read this slot has finished
            for dst, src in zip(self.slots[k], self.host[n % len(self.host)]):
                dst.copy_(src, non_blocking=True)
            ev = torch.cuda.Event(); ev.record(self.stream)
        self.ready[n] = (k, ev)

    def get(self, n):
        for j in range(len(self.slots) - 1):     # batch n for this step, n+1 for the march-ahead of this step, n+2 in flight
            self._issue(n + j)
        k, ev = self.ready.pop(n)
        torch.cuda.current_stream(self.device).wait_event(ev)
        return self.slots[k]

    def peek(self, n):
        """Device tensors of batch n (already issued by get(n - 1)) WITHOUT consuming them: for Trainer.step(next_rays=...).
        The caller's side stream must wait for `event_of(n)` -- Engine.prefetch records its own dependency on the current
        stream, so make the current stream wait first."""
        self._issue(n)
        k, ev = self.ready[n]
        torch.cuda.current_stream(self.device).wait_event(ev)
        return self.slots[k]

    def done(self, n):
        """Call after the step that consumed batch n has been enqueued: its slot may be overwritten once that work is done."""
        k = n % len(self.slots)
        ev = torch.cuda.Event(); ev.record(torch.cuda.current_stream(self.device))
        self.free_evt[k] = ev
