"""Lightning-free training driver for the Rad-NeRF hot path: one
process per GPU, the step of train_ml.py:172-200 --

    every 16 steps: update_density_grid            (train_ml.py:173-177)
    ml_render(model, gate, rays_o, rays_d)         (ml_rendering.py:11-78)
    NeRFLoss: rgb + opacity + cv_importance + depth_mutual, loss = sum of term means
    backward -> gradient exchange when world_size > 1 -> Adam(eps=1e-15)

Rays shard naturally: with world_size W each rank renders its own slice of the global batch.  Exchanges per step: the
K-float gating importance before the (non-linear) CV loss (an NCCL all-reduce), and the gradients -- with the fused engine
ONE kernel over NVLink peer memory (csrc/peer.cu: reduce-scatter + Adam + fp16 all-gather, flag barriers in peer memory);
the Python operator surface (engine=False) all-reduces one flat gradient bucket over NCCL instead.
graph=True replays the single-process step from captured CUDA graphs (engine.GraphStepper).
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from . import parallel
from . import synthetic as syn
from .losses import NeRFLoss
from .models.ml_rendering import ml_render
from .models.networks import MNGP, NGP, Ray_Gate
from .models.rendering import render


def cosine_lr(base_lr, step, total_steps, eta_min_ratio=1.0 / 30):
    """torch.optim.lr_scheduler.CosineAnnealingLR closed form with eta_min = base_lr / 30 (train_ml.py:150-153)."""
    import math
    eta_min = base_lr * eta_min_ratio
    return eta_min + (base_lr - eta_min) * (1 + math.cos(math.pi * min(step, total_steps) / total_steps)) / 2


class Trainer:
    def __init__(self, scale=0.5, K=2, lr=1e-2, log2_T=19, lambda_opacity=1e-3, lambda_cv=1e-2, lambda_dm=5e-3,
                 loss_scale=1024.0, device="cuda", update_interval=16, warmup_steps=256, freeze_occupancy=False,
                 random_bg=False, seed=0, engine=True, batch_rays=8192, samples_per_ray=128, total_steps=None,
                 on_overflow="grow", graph=False):
        self.device = torch.device(device)
        self.K, self.scale = K, scale
        self.exp_step_factor = syn.exp_step_factor_for_scale(scale)
        torch.manual_seed(seed)
        if K == 1:
            self.model = NGP(scale, t=log2_T).to(self.device)
            self.gate = None
        else:
            self.model = MNGP(scale, size=K, t=log2_T).to(self.device)
            self.gate = Ray_Gate(K).to(self.device)
        self.params = [p for p in list(self.model.parameters()) + (list(self.gate.parameters()) if self.gate else [])
                       if p.numel() > 0]
        self.opt = torch.optim.Adam(self.params, lr, eps=1e-15, fused=True)   # FusedAdam(lr, eps=1e-15), train_ml.py:143
        self.loss_fn = NeRFLoss()
        self.lam = dict(lambda_opacity=lambda_opacity, lambda_cv_importance=lambda_cv if K > 1 else 0,
                        lambda_depth_mutual=lambda_dm if K > 1 else 0)
        self.loss_scale = loss_scale
        self.update_interval, self.warmup_steps = update_interval, warmup_steps
        self.freeze_occupancy = freeze_occupancy
        self.random_bg = random_bg
        self.global_step = 0
        self.world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
        self._shadow = None
        self.lr = self.base_lr = lr
        # CosineAnnealingLR(opt, num_epochs * 1000, lr / 30) stepped once per iteration (train_ml.py:143-153); None = constant
        self.total_steps = total_steps
        self.engine = None
        if engine:
            # the fused step (csrc/engine.cu): same modules, their parameters re-homed into one flat vector
            from .engine import Engine
            self.engine = Engine(self.model, self.gate, n_rays=batch_rays, exp_step_factor=self.exp_step_factor,
                                 samples_per_ray=samples_per_ray, lambda_opacity=lambda_opacity, lambda_cv=lambda_cv,
                                 lambda_dm=lambda_dm, lr=lr, on_overflow=on_overflow)
        # graph=True: the step is replayed from captured CUDA graphs (engine.GraphStepper); batches pass through two static slots
        self.stepper = self.engine.graph_stepper() if (graph and self.engine is not None and self.world == 1 and not random_bg) else None
        self._slot, self._staged = 0, None

    # -- occupancy ---------------------------------------------------------------------------
    def set_occupancy(self, bitfields):
        """Install externally computed bitfields (e.g. the synthetic scenes of bench.py)."""
        for i, bf in enumerate(bitfields):
            name = "density_bitfield" if self.K == 1 else f"density_bitfield_{i}"
            getattr(self.model, name).copy_(torch.as_tensor(bf, device=self.device))

    def _grid_update(self):
        thr = syn.DENSITY_THRESHOLD
        warm = self.global_step < self.warmup_steps
        if self.engine is not None:
            # fused update (csrc/grid.cu).  Benchmark mode runs the SAME work against shadow buffers so that the
            # occupancy used for marching -- and therefore the samples per ray -- stays that of the synthetic scene.
            if not self.freeze_occupancy:
                self.engine.update_grid(thr, warmup=warm)
                return
            if self._shadow is None:
                self._shadow = ([torch.zeros_like(g) for g in self.engine.grids()],
                                [torch.zeros_like(b) for b in self.engine.bitfields()])
            self.engine.update_grid(thr, warmup=warm, grids=self._shadow[0], bitfields=self._shadow[1])
            return
        if not self.freeze_occupancy:
            self.model.update_density_grid(thr, warmup=warm)
            return
        # Benchmark mode: the SAME work (cell sampling, density evaluation, EMA-max, mean, packbits)
        # runs against shadow buffers so that the occupancy used for marching -- and therefore the
        # number of samples per ray -- stays that of the synthetic scene.
        names = ["density_grid", "density_bitfield"] if self.K == 1 else \
            [f"density_{k}_{i}" for i in range(self.K) for k in ("grid", "bitfield")]
        keep = {n: getattr(self.model, n) for n in names}
        if self._shadow is None:
            self._shadow = {n: torch.zeros_like(v) for n, v in keep.items()}
        for n, v in self._shadow.items():
            setattr(self.model, n, v)
        self.model.update_density_grid(thr, warmup=warm)
        for n in names:
            self._shadow[n] = getattr(self.model, n)
            setattr(self.model, n, keep[n])

    def _graph_step(self, rays_o, rays_d, target_rgb, next_rays):
        """One step through the captured graphs.  next_rays = (rays_o, rays_d[, target_rgb]) of the FOLLOWING step: staged
        into the other static slot so that this step's graph can march it ahead.  `_staged` = (slot, has_target)."""
        st, s = self.stepper, self._slot
        if self._staged is not None and self._staged[0] == s:      # the rays are in place (marched by the previous graph)
            if not self._staged[1]:
                st.load(s, None, None, target_rgb)
            first = False
        else:
            st.load(s, rays_o, rays_d, target_rgb)
            first = True
        after_update = self.global_step % self.update_interval == 0
        before_update = (self.global_step + 1) % self.update_interval == 0
        look_ahead = next_rays is not None and not before_update
        kind = ("first" if (first or after_update) else "steady") if look_ahead else "last"
        if kind == "last" and (first or after_update):
            # nothing marched ahead for this batch and nothing to march ahead: the plain eager step
            self.engine._ahead.clear()
            loss = self.engine.step(self.stepper.slots[s][0], self.stepper.slots[s][1], self.stepper.slots[s][2], lr=self.lr)
        else:
            if look_ahead:
                st.load(s ^ 1, *next_rays)
            loss = st.step(s, lr=self.lr, kind=kind)
        self._staged = (s ^ 1, len(next_rays) > 2) if look_ahead else None
        self._slot = s ^ 1
        self.global_step += 1
        self._last_results = None
        return loss[0]

    @property
    def last_results(self):
        """The reference's result dict of the last step (ml_rendering.py:72-77)."""
        if getattr(self, "_last_results", None) is None and self.engine is not None:
            self._last_results = self.engine.results()
        return self._last_results

    # -- one step ----------------------------------------------------------------------------
    def step(self, rays_o, rays_d, target_rgb, next_rays=None):
        """rays_o, rays_d, target_rgb: (B_local,3) fp32 CUDA tensors.  Returns the (unscaled) loss tensor.
        next_rays = (rays_o, rays_d) of the FOLLOWING batch, when it is already on the device: its marching then
        overlaps this step's field kernels (Engine.prefetch).  Skipped when the following step starts with an
        occupancy update, so that every batch is marched against the same bitfields as in the sequential order."""
        if self.global_step % self.update_interval == 0:
            self._grid_update()
        if self.total_steps:
            self.lr = cosine_lr(self.base_lr, self.global_step, self.total_steps)
            for g in self.opt.param_groups:
                g["lr"] = self.lr
        if self.stepper is not None:
            return self._graph_step(rays_o, rays_d, target_rgb, next_rays)
        if self.engine is not None:
            if next_rays is not None and (self.global_step + 1) % self.update_interval == 0:
                next_rays = None
            loss = self.engine.step(rays_o, rays_d, target_rgb, lr=self.lr, random_bg=self.random_bg, next_rays=next_rays)
            self.global_step += 1
            self._last_results = None             # formed on demand (Trainer.last_results): nothing per step on the host
            return loss[0]
        kw = dict(test_time=False, random_bg=self.random_bg)
        if self.exp_step_factor:
            kw["exp_step_factor"] = self.exp_step_factor
        if self.K == 1:
            results = render(self.model, rays_o, rays_d, **kw)
        else:
            results = ml_render(self.model, self.gate, rays_o, rays_d, rays_d, **kw)
            results["gating_importance"] = parallel.global_importance(results["gating_importance"])   # 8e: CV loss is non-linear
        loss_d = self.loss_fn(results, {"rgb": target_rgb}, **self.lam)
        # every term is normalised by the GLOBAL element count: the gradient exchange is then a plain sum
        loss = parallel.total_loss(loss_d, self.world)
        self.opt.zero_grad(set_to_none=True)
        (loss * self.loss_scale).backward()
        flat = torch.cat([p.grad.reshape(-1) for p in self.params])
        parallel.allreduce_sum_(flat)          # one flat bucket over NCCL / NVLink
        flat.mul_(1.0 / self.loss_scale)
        o = 0
        for p in self.params:
            p.grad.copy_(flat[o:o + p.numel()].view_as(p)); o += p.numel()
        self.opt.step()
        self.global_step += 1
        self._last_results = results
        self.last_loss = loss.detach()
        return self.last_loss
