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
