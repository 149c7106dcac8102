"""NGP / MNGP / Ray_Gate with the reference's public interface (models/networks.py:17-211,
214-421, 1070-1097): density(), forward(), update_density_grid(), register_bbox(), the same
buffer / parameter names (state-dict keys: xyz_encoder.params, dir_encoder.params,
geo_net[_i].params, rgb_net[_i].params, encoder.params; buffers center, xyz_min, xyz_max,
half_size, grid_coords, density_bitfield[_i], density_grid[_i]) -- on top of the sm_100a
operators (radnerf_b200.tcnn / radnerf_b200.vren).
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F
from torch import nn

from .. import tcnn, vren
from .custom_functions import TruncExp

GRID_SIZE = 128


def _grid_encoder(scale, log2_T):
    L, Fdim, N_min = 16, 2, 16
    b = float(np.exp(np.log(2048 * scale / N_min) / (L - 1)))   # networks.py:39-40
    return tcnn.Encoding(3, {"otype": "Grid", "type": "Hash", "n_levels": L, "n_features_per_level": Fdim,
                             "log2_hashmap_size": log2_T, "base_resolution": N_min, "per_level_scale": b,
                             "interpolation": "Linear"})


def _mlp(n_in, n_out, hidden, out_act, otype="FullyFusedMLP", seed=1337):
    return tcnn.Network(n_in, n_out, {"otype": otype, "activation": "ReLU", "output_activation": out_act,
                                      "n_neurons": 64, "n_hidden_layers": hidden}, seed=seed)


def _all_grid_coords(G):
    ii = torch.arange(G, dtype=torch.int32)
    return torch.stack(torch.meshgrid(ii, ii, ii, indexing="ij"), -1).reshape(-1, 3).contiguous()


class _OccupancyMixin:
    """Occupancy-grid maintenance shared by NGP and MNGP (networks.py:128-201 / 330-409)."""

    def _cascade_points(self, coords, c):
        s = min(2 ** (c - 1), self.scale)
        half_grid_size = s / self.grid_size
        xyzs_w = (coords / (self.grid_size - 1) * 2 - 1) * (s - half_grid_size)
        xyzs_w += (torch.rand_like(xyzs_w) * 2 - 1) * half_grid_size   # random position inside the cell
        return xyzs_w

    @torch.no_grad()
    def get_all_cells(self):
        indices = vren.morton3D(self.grid_coords).long()
        return [(indices, self.grid_coords)] * self.cascades

    @torch.no_grad()
    def _sample_cells(self, density_grid, M, density_threshold):
        cells = []
        for c in range(self.cascades):
            coords1 = torch.randint(self.grid_size, (M, 3), dtype=torch.int32, device=density_grid.device)
            indices1 = vren.morton3D(coords1).long()
            indices2 = torch.nonzero(density_grid[c] > density_threshold)[:, 0]
            if len(indices2) > 0:
                rand_idx = torch.randint(len(indices2), (M,), device=density_grid.device)
                indices2 = indices2[rand_idx]
                coords2 = vren.morton3D_invert(indices2.int())
                cells += [(torch.cat([indices1, indices2]), torch.cat([coords1, coords2]))]
            else:
                cells += [(indices1, coords1)]
        return cells

    @torch.no_grad()
    def _update_one_grid(self, density_grid, density_bitfield, density_fn, density_threshold, warmup, decay):
        density_grid_tmp = torch.zeros_like(density_grid)
        cells = self.get_all_cells() if warmup else \
            self._sample_cells(density_grid, self.grid_size ** 3 // 4, density_threshold)
        for c in range(self.cascades):
            indices, coords = cells[c]
            density_grid_tmp[c, indices] = density_fn(self._cascade_points(coords, c))
        density_grid = torch.where(density_grid < 0, density_grid,
                                   torch.maximum(density_grid * decay, density_grid_tmp))
        mean_density = density_grid[density_grid > 0].mean().item()
        vren.packbits(density_grid.contiguous(), min(mean_density, density_threshold), density_bitfield)
        return density_grid


class NGP(nn.Module, _OccupancyMixin):
    """Single Instant-NGP field (the K=1 baseline of train.py; reference networks.py:17-211)."""

    def __init__(self, scale, rgb_act="Sigmoid", t=19):
        super().__init__()
        self.rgb_act = rgb_act
        self.scale = scale
        self.register_buffer("center", torch.zeros(1, 3))
        self.register_buffer("xyz_min", -torch.ones(1, 3) * scale)
        self.register_buffer("xyz_max", torch.ones(1, 3) * scale)
        self.register_buffer("half_size", (self.xyz_max - self.xyz_min) / 2)
        self.cascades = max(1 + int(np.ceil(np.log2(2 * scale))), 1)
        self.grid_size = GRID_SIZE
        self.register_buffer("density_bitfield", torch.zeros(self.cascades * self.grid_size ** 3 // 8, dtype=torch.uint8))
        self.register_buffer("density_grid", torch.zeros(self.cascades, self.grid_size ** 3))
        self.register_buffer("grid_coords", _all_grid_coords(self.grid_size))
        self.xyz_encoder = _grid_encoder(scale, t)
        self.geo_net = _mlp(self.xyz_encoder.n_output_dims, 16 + 1, 1, "None", "CutlassMLP", seed=11)
        self.dir_encoder = tcnn.Encoding(3, {"otype": "SphericalHarmonics", "degree": 4})
        self.rgb_net = _mlp(32, 3, 2, rgb_act, "CutlassMLP", seed=12)

    def density(self, x, return_feat=False):
        x = (x - self.xyz_min) / (self.xyz_max - self.xyz_min)
        x = x.clip(0.0, 1.0)
        h = self.geo_net(self.xyz_encoder(x))
        sigmas = TruncExp.apply(h[:, 0])
        if return_feat:
            return sigmas, h[..., 1:]
        return sigmas

    def forward(self, x, d, **kwargs):
        sigmas, h = self.density(x, return_feat=True)
        d = d / torch.norm(d, dim=1, keepdim=True)
        d = self.dir_encoder((d + 1) / 2)
        rgbs = self.rgb_net(torch.cat([d, h], 1))
        return sigmas, rgbs

    @torch.no_grad()
    def sample_uniform_and_occupied_cells(self, M, density_threshold):
        return self._sample_cells(self.density_grid, M, density_threshold)

    @torch.no_grad()
    def update_density_grid(self, density_threshold, warmup=False, decay=0.95, erode=False):
        self.density_grid = self._update_one_grid(self.density_grid, self.density_bitfield, self.density,
                                                  density_threshold, warmup, decay)

    @torch.no_grad()
    def register_bbox(self, bbox):
        dev = self.xyz_min.device
        self.xyz_min = torch.from_numpy(bbox[0][None, :]).float().to(dev)
        self.xyz_max = torch.from_numpy(bbox[1][None, :]).float().to(dev)
        self.half_size = torch.from_numpy((bbox[1] - bbox[0]) / 2)[None, :].float().to(dev)
        self.center = torch.from_numpy(np.mean(bbox, axis=0))[None, :].float().to(dev)
        self.cascades = max(1 + int(np.ceil(np.log2(2 * float(self.half_size.max())))), 1)
        self.density_bitfield = torch.zeros(self.cascades * self.grid_size ** 3 // 8, dtype=torch.uint8, device=dev)
        self.density_grid = torch.zeros(self.cascades, self.grid_size ** 3, device=dev)


class MNGP(nn.Module, _OccupancyMixin):
    """Rad-NeRF field zoo: ONE shared hash grid + SH encoder, `size` x (geo_net_i, rgb_net_i,
    density_grid_i, density_bitfield_i).  Reference: networks.py:214-421."""

    def __init__(self, scale, rgb_act="Sigmoid", size=2, t=19):
        super().__init__()
        self.rgb_act = rgb_act
        self.scale = scale
        self.size = size
        self.register_buffer("center", torch.zeros(1, 3))
        self.register_buffer("xyz_min", -torch.ones(1, 3) * scale)
        self.register_buffer("xyz_max", torch.ones(1, 3) * scale)
        self.register_buffer("half_size", (self.xyz_max - self.xyz_min) / 2)
        self.xyz_encoder = _grid_encoder(scale, t)
        self.dir_encoder = tcnn.Encoding(3, {"otype": "SphericalHarmonics", "degree": 4})
        self.cascades = max(1 + int(np.ceil(np.log2(2 * scale))), 1)
        self.grid_size = GRID_SIZE
        self.register_buffer("grid_coords", _all_grid_coords(self.grid_size))
        for i in range(size):
            self.register_buffer(f"density_bitfield_{i}", torch.zeros(self.cascades * self.grid_size ** 3 // 8, dtype=torch.uint8))
            self.register_buffer(f"density_grid_{i}", torch.zeros(self.cascades, self.grid_size ** 3))
            setattr(self, f"geo_net_{i}", _mlp(self.xyz_encoder.n_output_dims, 16 + 1, 1, "None", seed=100 + i))
            setattr(self, f"rgb_net_{i}", _mlp(32, 3, 2, rgb_act, seed=200 + i))

    def density(self, x, ind, return_feat=False):
        x = (x - self.xyz_min) / (self.xyz_max - self.xyz_min)
        x = x.clip(0.0, 1.0)
        h = getattr(self, f"geo_net_{ind}")(self.xyz_encoder(x))
        sigmas = TruncExp.apply(h[:, 0])
        if return_feat:
            return sigmas, h[..., 1:]
        return sigmas

    def forward(self, x, d, ind, **kwargs):
        sigmas, h = self.density(x, ind, return_feat=True)
        d = d / torch.norm(d, dim=1, keepdim=True)
        d = self.dir_encoder((d + 1) / 2)
        rgbs = getattr(self, f"rgb_net_{ind}")(torch.cat([d, h], 1))
        return sigmas, rgbs

    @torch.no_grad()
    def sample_uniform_and_occupied_cells(self, M, density_threshold, ind):
        return self._sample_cells(getattr(self, f"density_grid_{ind}"), M, density_threshold)

    @torch.no_grad()
    def update_density_grid(self, density_threshold, warmup=False, decay=0.95, erode=False):
        for i in range(self.size):
            grid = self._update_one_grid(getattr(self, f"density_grid_{i}"), getattr(self, f"density_bitfield_{i}"),
                                         lambda p, i=i: self.density(p, i), density_threshold, warmup, decay)
            setattr(self, f"density_grid_{i}", grid)

    @torch.no_grad()
    def register_bbox(self, bbox):
        dev = self.xyz_min.device
        self.xyz_min = torch.from_numpy(bbox[0][None, :]).float().to(dev)
        self.xyz_max = torch.from_numpy(bbox[1][None, :]).float().to(dev)
        self.half_size = torch.from_numpy((bbox[1] - bbox[0]) / 2)[None, :].float().to(dev)
        self.center = torch.from_numpy(np.mean(bbox, axis=0))[None, :].float().to(dev)
        self.cascades = max(1 + int(np.ceil(np.log2(2 * float(self.half_size.max())))), 1)
        for i in range(self.size):
            setattr(self, f"density_bitfield_{i}", torch.zeros(self.cascades * self.grid_size ** 3 // 8, dtype=torch.uint8, device=dev))
            setattr(self, f"density_grid_{i}", torch.zeros(self.cascades, self.grid_size ** 3, device=dev))


class Ray_Gate(nn.Module):
    """Per-ray soft gate: 6 -> 64x4 -> K MLP, softmax over K, importance = column sums.
    Reference: networks.py:1070-1097."""

    def __init__(self, out_dim, type="ray"):
        super().__init__()
        self.out_dim = out_dim
        self.type = type
        self.encoder = _mlp(6, out_dim, 4, "None", seed=300)

    def forward(self, x, warmup=False):
        x = self.encoder(x)
        gate = F.softmax(x.float(), 1)   # fp32 softmax (the reference runs it under autocast)
        importance = gate.sum(0)
        return gate, importance, None

    def freeze_dict(self):
        for _, param in self.named_parameters():
            param.requires_grad_(False)


class unshared_MNGP(nn.Module, _OccupancyMixin):
    """Ablation of MNGP without parameter sharing: every sub-NeRF owns its hash grid, SH encoder, geo / rgb MLPs and
    occupancy grid (xyz_encoder_i, dir_encoder_i, geo_net_i, rgb_net_i, density_grid_i, density_bitfield_i); same call
    interface as MNGP, so ml_render drives it unchanged.  Reference: networks.py:424-628."""

    def __init__(self, scale, rgb_act="Sigmoid", size=2, t=19):
        super().__init__()
        self.rgb_act = rgb_act
        self.scale = scale
        self.size = size
        self.register_buffer("center", torch.zeros(1, 3))
        self.register_buffer("xyz_min", -torch.ones(1, 3) * scale)
        self.register_buffer("xyz_max", torch.ones(1, 3) * scale)
        self.register_buffer("half_size", (self.xyz_max - self.xyz_min) / 2)
        self.cascades = max(1 + int(np.ceil(np.log2(2 * scale))), 1)
        self.grid_size = GRID_SIZE
        self.register_buffer("grid_coords", _all_grid_coords(self.grid_size))
        for i in range(size):
            self.register_buffer(f"density_bitfield_{i}", torch.zeros(self.cascades * self.grid_size ** 3 // 8, dtype=torch.uint8))
            self.register_buffer(f"density_grid_{i}", torch.zeros(self.cascades, self.grid_size ** 3))
            setattr(self, f"xyz_encoder_{i}", _grid_encoder(scale, t))
            setattr(self, f"dir_encoder_{i}", tcnn.Encoding(3, {"otype": "SphericalHarmonics", "degree": 4}))
            setattr(self, f"geo_net_{i}", _mlp(32, 16 + 1, 1, "None", seed=100 + i))
            setattr(self, f"rgb_net_{i}", _mlp(32, 3, 2, rgb_act, seed=200 + i))

    def density(self, x, ind, return_feat=False):
        x = ((x - self.xyz_min) / (self.xyz_max - self.xyz_min)).clip(0.0, 1.0)
        h = getattr(self, f"geo_net_{ind}")(getattr(self, f"xyz_encoder_{ind}")(x))
        sigmas = TruncExp.apply(h[:, 0])
        return (sigmas, h[..., 1:]) if return_feat else sigmas

    def forward(self, x, d, ind, **kwargs):
        sigmas, h = self.density(x, ind, return_feat=True)
        d = getattr(self, f"dir_encoder_{ind}")((d / torch.norm(d, dim=1, keepdim=True) + 1) / 2)
        return sigmas, getattr(self, f"rgb_net_{ind}")(torch.cat([d, h], 1))

    @torch.no_grad()
    def sample_uniform_and_occupied_cells(self, M, density_threshold, ind):
        return self._sample_cells(getattr(self, f"density_grid_{ind}"), M, density_threshold)

    update_density_grid = MNGP.update_density_grid
    register_bbox = MNGP.register_bbox


class NGP_zoo(nn.Module):
    """`size` independent Instant-NGP fields (model_0 .. model_{size-1}) blended by a gate in moe_render -- the
    "mixture of whole NGPs" baseline.  Reference: networks.py:1100-1163."""

    def __init__(self, scale, rgb_act="Sigmoid", size=10, t=19):
        super().__init__()
        self.rgb_act = rgb_act
        self.scale = scale
        self.size = size
        self.register_buffer("center", torch.zeros(1, 3))
        self.register_buffer("xyz_min", -torch.ones(1, 3) * scale)
        self.register_buffer("xyz_max", torch.ones(1, 3) * scale)
        self.register_buffer("half_size", (self.xyz_max - self.xyz_min) / 2)
        for i in range(size):
            setattr(self, f"model_{i}", NGP(scale=scale, rgb_act=rgb_act, t=t))

    def models(self):
        return [getattr(self, f"model_{i}") for i in range(self.size)]

    @torch.no_grad()
    def update_density_grid(self, density_threshold, warmup=False, decay=0.95, erode=False):
        for m in self.models():
            m.update_density_grid(density_threshold, warmup=warmup, decay=decay, erode=erode)

    @torch.no_grad()
    def register_bbox(self, bbox):
        for m in self.models():
            m.register_bbox(bbox)
