"""Fused Rad-NeRF training step (host side of rnb_engine_*, include/radnerf_b200.h).

`Engine` takes the drop-in modules (`MNGP`/`NGP` + `Ray_Gate`, same classes the reference builds in
train_ml.py:74-80), re-homes their flat tinycudann-layout parameters into ONE flat fp32 vector (the
modules keep working: their `.params` become views), and runs the step of train_ml.py:172-200 --
gate, K x (AABB, march, field, composite, background), gated blend, NeRFLoss, backward, Adam -- as
~12 kernel launches without a host synchronisation.  Data parallel: rays shard by rank; the only
exchanges are the K-float gate importance (before the non-linear CV loss) and one all-reduce of the
flat gradient vector (NCCL over NVLink), both issued from here through torch.distributed.
There is no CPU path.
"""
from __future__ import annotations

import ctypes
import os

import torch
import torch.distributed as dist

from . import _lib, parallel
from ._lib import call, f32, i32, i64, stream_arg
from .field import MAX_K
from .tcnn import _GridMeta


class _Config(ctypes.Structure):
    _fields_ = [("K", ctypes.c_int32), ("n_rays", ctypes.c_int32), ("max_samples", ctypes.c_int32),
                ("grid_size", ctypes.c_int32), ("cascades", ctypes.c_int32), ("scale", ctypes.c_float),
                ("exp_step_factor", ctypes.c_float), ("T_threshold", ctypes.c_float), ("near_distance", ctypes.c_float),
                ("sample_capacity", ctypes.c_int64), ("grid", _GridMeta),
                ("xyz_min", ctypes.c_float * 3), ("xyz_max", ctypes.c_float * 3), ("center", ctypes.c_float * 3),
                ("half_size", ctypes.c_float * 3),
                ("lambda_opacity", ctypes.c_float), ("lambda_cv", ctypes.c_float), ("lambda_dm", ctypes.c_float),
                ("loss_scale", ctypes.c_float),
                ("n_params", ctypes.c_int64), ("table_off", ctypes.c_int64),
                ("geo_off", ctypes.c_int64 * MAX_K), ("rgb_off", ctypes.c_int64 * MAX_K), ("gate_off", ctypes.c_int64),
                ("noise_seed", ctypes.c_uint64)]


_LAYOUT_FIELDS = ["total_bytes", "gate_in", "gate_hidden", "gate_logits", "gate_dlogits", "gate_prob", "dgate", "importance",
                  "t_scratch", "counts", "starts", "seg", "ray_id", "ts", "deltas", "enc", "rgb_in", "geo_h", "rgb_h1", "rgb_h2",
                  "sraw", "sigmas", "rgb4", "ws", "dsigmas", "drgbs", "pass_opacity", "pass_depth", "pass_rgb", "pass_total",
                  "d_pass_opacity", "d_pass_depth", "d_pass_rgb", "rgb", "opacity", "depth", "loss", "march_slot1", "march_slot_bytes", "coarse",
                  "status"]

GRAD_TAIL = 64      # RNB_GRAD_TAIL: floats behind the gradients; word [n_params] is the step guard's flag word


class SampleCapacityError(_lib.RadNerfB200Error):
    """A step's samples did not fit the workspace (`samples_per_ray`); the step was skipped on the device."""


class _Layout(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int64) for n in _LAYOUT_FIELDS]


_RENDER_LAYOUT_FIELDS = ["total_bytes", "gate_in", "gate_logits", "gate_prob", "importance", "t_cur", "t2", "acc_opacity", "acc_depth",
                         "acc_rgb", "item_start", "item_n", "alive0", "alive1", "counters", "seg", "ray_id", "ts", "sigmas", "rgb4", "coarse"]


class _RenderLayout(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int64) for n in _RENDER_LAYOUT_FIELDS]


_GRID_LAYOUT_FIELDS = ["total_bytes", "tmp", "cell_idx", "occ_list", "block_counts", "occ_count", "part_sum", "part_cnt", "mean"]


class _GridLayout(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int64) for n in _GRID_LAYOUT_FIELDS]


class _Span:
    """Caller-recorded profile span (rnb_engine_profile_span_*): a no-op unless rnb_engine_profile(1) is on."""
    ADAM, BARRIER = 10, 11

    def __init__(self, stage, device):
        self.device = device
        self.h = _lib.lib().rnb_engine_profile_span_begin(i32(stage), stream_arg(device))

    def end(self):
        if self.h >= 0:
            _lib.lib().rnb_engine_profile_span_end(i32(self.h), stream_arg(self.device))


class Engine:
    def __init__(self, model, gate=None, n_rays=8192, exp_step_factor=0.0, samples_per_ray=128, max_samples=1024,
                 T_threshold=1e-4, near_distance=0.01, lambda_opacity=1e-3, lambda_cv=1e-2, lambda_dm=5e-3,
                 loss_scale=128.0 * 1024.0, lr=1e-2, betas=(0.9, 0.999), eps=1e-15, sharded_optimizer=True,
                 on_overflow="raise", dynamic_loss_scale=True, growth_interval=2000):
        """on_overflow: what happens when a step's samples exceed the workspace (n_rays * K * samples_per_ray): the
        device skips that optimizer step either way (its tail samples were dropped, so its gradient is biased);
        "raise" -> SampleCapacityError at the next step boundary, "grow" -> the workspace is re-allocated with headroom
        and training continues.  dynamic_loss_scale: GradScaler policy of the reference's precision=16 run
        (train_ml.py:304): halve on a skipped non-finite step, double after `growth_interval` applied steps."""
        _lib.lib()
        self.model, self.gate = model, gate
        self.K = K = getattr(model, "size", 1)
        assert (gate is not None) == (K > 1), "a gate is needed iff model_zoo_size > 1"
        dev = model.xyz_encoder.params.device
        assert dev.type == "cuda", "the engine needs the model on a CUDA device (no CPU path)"
        self.device, self.n_rays = dev, int(n_rays)
        self.exp_step_factor = float(exp_step_factor)
        # ---- flat parameter vector: [table | geo_i, rgb_i ... | gate] -------------------------------
        mods = [("table", model.xyz_encoder)]
        for i in range(K):
            sfx = f"_{i}" if hasattr(model, "size") else ""
            mods += [(f"geo{i}", getattr(model, "geo_net" + sfx)), (f"rgb{i}", getattr(model, "rgb_net" + sfx))]
        if gate is not None:
            mods.append(("gate", gate.encoder))
        offs, off = {}, 0
        for name, m in mods:
            offs[name] = off
            off += (m.params.numel() + 7) // 8 * 8      # keep every block 16-byte aligned in fp16
        self.n_params = off
        self.offsets = offs
        self.flat = torch.zeros(off, dtype=torch.float32, device=dev)
        self.world = parallel.world_size()
        self.sharded = self.world > 1 and sharded_optimizer
        if self.sharded:
            # buffers the other ranks address over NVLink (csrc/peer.cu): gradients are read, fp16 parameters written
            # (two alternating gradient buffers: the one peers read in step n is cleared by this rank's kernel of step n+1)
            # (+ GRAD_TAIL floats: the step guard's flag word, which the peers' optimizer kernels read too)
            self._peer_g = [parallel.PeerBuffer((off + GRAD_TAIL) * 4, dev), parallel.PeerBuffer((off + GRAD_TAIL) * 4, dev)]
            self._peer_p = parallel.PeerBuffer(off * 2, dev)
            self._g = 0
            # barrier words of the flag barrier (csrc/peer.cu): words 16.. of buffer 0's tail on every rank, as mapped here
            self._flag_ptrs = (ctypes.c_void_p * self.world)(*[p + (off + 16) * 4 for p in self._peer_g[0].ptrs])
            self._grad_buf = self._peer_g[0].local.view(torch.float32)
            self.flat16 = self._peer_p.local.view(torch.float16)
        else:
            self.flat16 = torch.zeros(off, dtype=torch.float16, device=dev)
            self._grad_buf = torch.zeros(off + GRAD_TAIL, dtype=torch.float32, device=dev)
        self.grads = self._grad_buf[:off]
        self.exp_avg = torch.zeros(off, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(off, dtype=torch.float32, device=dev)
        self._mods = mods
        with torch.no_grad():
            for name, m in mods:
                n = m.params.numel()
                self.flat[offs[name]:offs[name] + n].copy_(m.params.data)
                m.params.data = self.flat[offs[name]:offs[name] + n]          # the module now views the flat vector
                m._shadow.buf = self.flat16[offs[name]:offs[name] + n]        # ... and its fp16 copy
                m._shadow.version = None
        self.sync_params()
        # ---- config / workspace -------------------------------------------------------------------------
        c = _Config()
        c.K, c.n_rays, c.max_samples, c.grid_size, c.cascades = K, self.n_rays, max_samples, model.grid_size, model.cascades
        c.scale, c.exp_step_factor, c.T_threshold, c.near_distance = model.scale, exp_step_factor, T_threshold, near_distance
        c.sample_capacity = (self.n_rays * K * int(samples_per_ray) + 127) // 128 * 128
        ctypes.memmove(ctypes.byref(c.grid), ctypes.byref(model.xyz_encoder._meta), ctypes.sizeof(_GridMeta))
        for name in ("xyz_min", "xyz_max", "center", "half_size"):
            v = getattr(model, name).reshape(-1).tolist()
            for k in range(3):
                getattr(c, name)[k] = v[k]
        c.lambda_opacity, c.lambda_cv, c.lambda_dm = lambda_opacity, (lambda_cv if K > 1 else 0.0), (lambda_dm if K > 1 else 0.0)
        c.loss_scale = loss_scale
        c.n_params, c.table_off = off, offs["table"]
        for i in range(K):
            c.geo_off[i], c.rgb_off[i] = offs[f"geo{i}"], offs[f"rgb{i}"]
        c.gate_off = offs.get("gate", 0)
        c.noise_seed = 0x5EED5EED + 7919 * parallel.rank()
        self.cfg = c
        self.samples_per_ray = int(samples_per_ray)
        self.on_overflow = on_overflow
        assert on_overflow in ("raise", "grow")
        self.dynamic_loss_scale, self.growth_interval = bool(dynamic_loss_scale), int(growth_interval)
        self._status_seen = [0, 0, 0, 0]          # host copy of the status counters last acted upon
        self._status_ring = None                  # pinned read-back slots + events (no host synchronisation per step)
        self._good_since_growth = 0
        self.status_interval = 4                  # optimizer steps between two read-backs of the (cumulative) status counters
        self._alloc_workspace()
        self.lr, self.betas, self.eps = lr, betas, eps
        self.opt_step = 0
        # marching ahead (rnb_engine_march): the next batch's samples are produced on a side stream into the other slot
        self._march_stream = None
        self._gate_stream = None        # gate forward / backward next to the compositing kernels (forward_backward)
        self.overlap_gate = os.environ.get("RNB_NO_GATE_OVERLAP", "0") != "1"
        self._ahead = {}                # (rays_o ptr, rays_d ptr) -> (slot, rays_o, rays_d, noise, done_event)
        self._slot = 0                  # slot holding the samples of the batch last trained on

    def _alloc_workspace(self, carry_status=None):
        """Carve the workspace for cfg.sample_capacity and (re)create the tensor views of its regions."""
        c, K = self.cfg, self.K
        L = _Layout()
        call("rnb_engine_layout_of", ctypes.byref(c), ctypes.byref(L))
        self.layout = L
        self.ws = torch.empty(L.total_bytes, dtype=torch.uint8, device=self.device)
        R, S = self.n_rays, c.sample_capacity
        v = self._view
        self.gate_prob = v("gate_prob", torch.float32, (R, K))
        self.importance = v("importance", torch.float32, (MAX_K,))[:K]
        self.seg = v("seg", torch.int32, (2 * MAX_K + 4,))
        self.counts = v("counts", torch.int32, (K, R)); self.starts = v("starts", torch.int32, (K, R))
        self.ts = v("ts", torch.float32, (S,)); self.deltas = v("deltas", torch.float32, (S,)); self.ray_id = v("ray_id", torch.int32, (S,))
        self.sigmas = v("sigmas", torch.float32, (S,)); self.rgb4 = v("rgb4", torch.float16, (S, 4)); self.sample_ws = v("ws", torch.float32, (S,))
        self.pass_rgb = v("pass_rgb", torch.float32, (K, R, 3)); self.pass_opacity = v("pass_opacity", torch.float32, (K, R))
        self.pass_depth = v("pass_depth", torch.float32, (K, R)); self.pass_total = v("pass_total", torch.int32, (K, R))
        self.rgb = v("rgb", torch.float32, (R, 3)); self.opacity = v("opacity", torch.float32, (R,)); self.depth = v("depth", torch.float32, (R, K))
        self.loss = v("loss", torch.float32, (8,))
        self.status = v("status", torch.int32, (16,))
        self.status.zero_()
        if carry_status is not None:      # re-allocation (capacity growth): counters, noise-stream positions and lr carry over
            self.status.copy_(carry_status)

    def _view(self, name, dtype, shape):
        off = getattr(self.layout, name)
        n = 1
        for s in shape:
            n *= s
        nbytes = n * torch.empty(0, dtype=dtype).element_size()
        return self.ws[off:off + nbytes].view(dtype).view(*shape)

    # -- parameters -------------------------------------------------------------------------------------
    def sync_params(self):
        """Refresh the fp16 copies of modules whose fp32 parameters were changed through torch
        (load_state_dict, manual edits); a no-op otherwise."""
        for _, m in self._mods:
            m._shadow.get(m.params)

    def bitfields(self):
        if self.K == 1 and not hasattr(self.model, "size"):
            return [self.model.density_bitfield]
        return [getattr(self.model, f"density_bitfield_{i}") for i in range(self.K)]

    def _slot_view(self, name, dtype, shape, slot):
        v = self._view(name, dtype, shape)
        if slot == 0:
            return v
        L = self.layout
        off = L.march_slot1 + (getattr(L, name) - L.t_scratch)
        return self.ws[off:off + v.numel() * v.element_size()].view(dtype).view(*shape)

    # -- phases -------------------------------------------------------------------------------------------
    def prefetch(self, rays_o, rays_d, noise=None, slot=None):
        """March the NEXT batch (AABB, march, scan, expand: everything that depends on rays and occupancy only) on a
        side stream while the kernels of the current step run.  Call it BEFORE step() of the current batch; the
        following step()/forward_backward() on these same tensors picks the samples up.  The bitfields must not
        change in between (the trainer does not prefetch across an occupancy update)."""
        _lib.check_input(rays_o, "rays_o", torch.float32)
        _lib.check_input(rays_d, "rays_d", torch.float32)
        assert rays_o.shape[0] == self.n_rays
        while len(self._ahead) > 1:             # never picked up: let it finish, then forget it
            torch.cuda.current_stream(self.device).wait_event(self._ahead.pop(next(iter(self._ahead)))[4])
        if noise is not None:       # (None: the library draws the marching noise itself -- no ATen launch in the step)
            _lib.check_input(noise, "noise", torch.float32)
        if self._march_stream is None:
            self._march_stream = torch.cuda.Stream(self.device)
        # the batch being (or about to be) trained sits in the slot of its own prefetch if that is still pending, else in
        # self._slot (forward_backward has already claimed it)
        if slot is None and getattr(self, "_in_step", False):
            # called from forward_backward: the batch being trained sits in self._slot whatever the bookkeeping of older
            # prefetches says; an entry still parked in the other slot was never picked up (stale): let it finish, forget it
            slot = self._slot ^ 1
            for k in [k for k, v in self._ahead.items() if v[0] == slot]:
                torch.cuda.current_stream(self.device).wait_event(self._ahead.pop(k)[4])
        busy = next(iter(self._ahead.values()))[0] if self._ahead else self._slot
        slot = (busy ^ 1) if slot is None else slot
        cur = torch.cuda.current_stream(self.device)
        ready = torch.cuda.Event()
        ready.record(cur)                       # inputs produced / the slot's previous reader (two steps back) finished
        bfs = self.bitfields()
        bf_arr = (ctypes.c_void_p * self.K)(*[b.data_ptr() for b in bfs])
        with torch.cuda.stream(self._march_stream):
            self._march_stream.wait_event(ready)
            call("rnb_engine_march", ctypes.byref(self.cfg), self.ws, rays_o, rays_d, noise, bf_arr, i32(slot),
                 stream_arg(self.device))
            done = torch.cuda.Event()
            done.record(self._march_stream)
        self._ahead[(rays_o.data_ptr(), rays_d.data_ptr())] = (slot, rays_o, rays_d, noise, done)

    def forward_backward(self, rays_o, rays_d, target_rgb, noise, bg, with_backward=True, next_rays=None):
        """noise (K,R) f32 in [0,1); bg: K x 3 python floats.  Fills the result views and, with
        `with_backward`, accumulates into self.grads.  Returns nothing; no host synchronisation.
        next_rays = (rays_o, rays_d) of the following batch: marched ahead on the side stream (prefetch), issued after
        the gate's importance all-reduce so that the small collective does not queue behind 2048 marching CTAs."""
        c = self.cfg
        for t, n in ((rays_o, "rays_o"), (rays_d, "rays_d"), (target_rgb, "target_rgb")):
            _lib.check_input(t, n, torch.float32)
        assert rays_o.shape[0] == self.n_rays
        if not torch.cuda.is_current_stream_capturing():
            self.sync_params()      # cheap version check: fp32 parameters changed through torch (load_state_dict) -> fp16 copies
            self._poll_guard()      # act on the status the device reported for earlier steps (loss scale, capacity)
        cur = torch.cuda.current_stream(self.device)
        hit = self._ahead.pop((rays_o.data_ptr(), rays_d.data_ptr()), None)
        if hit is not None:
            march_slot, _, _, noise, done = hit
            cur.wait_event(done)
            _lib.launch_count -= 3      # the march / scan / expand launches were counted with rnb_engine_march
        else:
            # not marched ahead: march here, into the slot no prefetch is using.  Entries older than the newest one
            # were never picked up (the caller changed its mind): let them finish, then forget them.
            keys = list(self._ahead)
            for k in (keys if next_rays is not None else keys[:-1]):      # (next_rays given: this step prefetches itself, all are stale)
                cur.wait_event(self._ahead.pop(k)[4])
            keys = list(self._ahead)
            march_slot = (self._ahead[keys[-1]][0] ^ 1) if keys else 0
            if getattr(self, "_force_slot", None) is not None:      # (graph replay: static batch slot == march slot)
                march_slot = self._force_slot
            if noise is not None:
                _lib.check_input(noise, "noise", torch.float32)
                assert noise.numel() == self.K * self.n_rays
        self._slot = march_slot
        inline = hit is None
        st = stream_arg(self.device)
        # The gate network is off the critical path where the data flow allows it (rnb_engine_fwd_bwd_ex): its forward is
        # needed from the blend on and its backward needs only the blend, so both run on a second stream next to the
        # compositing kernels.  (With several ranks and the in-line order the importance all-reduce doubles as the barrier
        # that keeps this rank's field kernels behind every peer's parameter writes; with the gate on the side stream
        # adam_step issues that barrier itself.)
        side_gate = self.K > 1 and self.overlap_gate
        gate_evt = blend_evt = None
        if side_gate and self._gate_stream is None:
            self._gate_stream = torch.cuda.Stream(self.device, priority=-1)      # ahead of the marcher's stream: the blend waits for it
            self._blend_evt = torch.cuda.Event(); self._blend_evt.record(cur)      # materialise the handle once
        if self.K > 1:
            if side_gate:
                ready = torch.cuda.Event(); ready.record(cur)
                with torch.cuda.stream(self._gate_stream):
                    self._gate_stream.wait_event(ready)
                    call("rnb_engine_gate", ctypes.byref(c), self.ws, self.flat16, rays_o, rays_d, stream_arg(self.device))
                    if self.world > 1:
                        sp = _Span(_Span.BARRIER, self.device); parallel.allreduce_sum_(self.importance); sp.end()
                    gate_evt = torch.cuda.Event(); gate_evt.record(self._gate_stream)
            else:
                call("rnb_engine_gate", ctypes.byref(c), self.ws, self.flat16, rays_o, rays_d, st)
                parallel.allreduce_sum_(self.importance)     # CV loss is non-linear in the batch: the importance must be global (8e)
        if next_rays is not None:
            self._in_step = True
            try:
                self.prefetch(next_rays[0], next_rays[1])
            finally:
                self._in_step = False
        bfs = self.bitfields()
        bf_arr = (ctypes.c_void_p * self.K)(*[b.data_ptr() for b in bfs])
        bg_arr = (ctypes.c_float * (3 * self.K))(*[float(x) for row in bg for x in row])
        self._last_bg = [[float(x) for x in row] for row in bg]
        if inline:
            call("rnb_engine_march", ctypes.byref(c), self.ws, rays_o, rays_d, noise, bf_arr, i32(march_slot), st)
            _lib.launch_count -= 3
        side_bwd = side_gate and with_backward
        if side_bwd:
            blend_evt = self._blend_evt
        call("rnb_engine_fwd_bwd_ex", ctypes.byref(c), self.ws, self.flat16, self.grads if with_backward else None,
             rays_o, rays_d, target_rgb, noise, bf_arr, bg_arr, i64(self.n_rays * self.world),
             self.importance if self.K > 1 else None, i32(1 if with_backward else 0), i32(march_slot),
             ctypes.c_void_p(gate_evt.cuda_event) if gate_evt is not None else None,
             ctypes.c_void_p(blend_evt.cuda_event) if blend_evt is not None else None, i32(1 if side_bwd else 0), st)
        _lib.launch_count += 1 if (self.K > 1 and with_backward and not side_bwd) else 0
        if side_bwd:
            with torch.cuda.stream(self._gate_stream):
                self._gate_stream.wait_event(blend_evt)
                call("rnb_engine_gate_bwd", ctypes.byref(c), self.ws, self.flat16, self._grad_buf, stream_arg(self.device))
                joined = torch.cuda.Event(); joined.record(self._gate_stream)
            cur.wait_event(joined)            # whatever follows on the step's stream (the optimizer) sees the gate gradients

    def adam_step(self, lr=None, lr_from_device=False):
        """Optimizer step behind the step guard: an inf / nan in the dense networks' gradients is looked for here
        (rnb_engine_grad_check; the table gradient and the fp16 seeds are checked where they are produced), then the
        Adam kernel applies the step -- or, when the flag word is set on ANY rank, clears the gradients and leaves
        parameters and moments alone.  The count of applied steps (for the bias corrections) lives on the device."""
        self.opt_step += 1
        n = self.n_params
        flags = self._grad_buf[n:]
        mlp_lo = self.offsets["geo0"]
        call("rnb_engine_grad_check", self._grad_buf, i64(mlp_lo), i64(n), flags, stream_arg(self.device))
        if self.sharded:
            # one kernel: this rank's shard of sum_p grads_p (peer loads) -> Adam -> fp16 parameters into every rank's buffer
            # (peer stores); the gate block is updated redundantly on every rank; the previous step's gradient buffer is
            # cleared.  The stream barrier in front orders it after every rank's backward (and gradient check).  Behind it a
            # second barrier keeps every rank's next step (field kernels, occupancy update, rendering: all readers of the
            # fp16 parameters) behind every peer's parameter stores.
            cur, prev = self._peer_g[self._g], self._peer_g[self._g ^ 1]
            n_sharded = self.offsets["gate"] if self.K > 1 else self.n_params
            flag_barrier = os.environ.get("RNB_NCCL_BARRIER", "0") != "1"
            if not flag_barrier:      # A/B: order the ranks with two 4-byte NCCL all-reduces instead of the flag barrier
                sp = _Span(_Span.BARRIER, self.device); parallel.stream_barrier(self.device); sp.end()
            sp = _Span(_Span.ADAM, self.device)
            call("rnb_engine_adam_sharded_guarded", i32(parallel.rank()), i32(self.world), cur.pointer_array(),
                 self._peer_p.pointer_array(), self.flat, self.exp_avg, self.exp_avg_sq, i64(self.n_params), i64(n_sharded),
                 prev.local, f32(-1.0 if lr_from_device else (self.lr if lr is None else lr)), f32(self.betas[0]), f32(self.betas[1]),
                 f32(self.eps), f32(1.0), self.status, self._flag_ptrs if flag_barrier else None, i32(self.opt_step), stream_arg(self.device))
            sp.end()
            if not flag_barrier:
                sp = _Span(_Span.BARRIER, self.device); parallel.stream_barrier(self.device); sp.end()
            self._g ^= 1
            self._grad_buf = self._peer_g[self._g].local.view(torch.float32)      # (cleared by the kernel of the previous step)
            self.grads = self._grad_buf[:n]
            self._master_stale = True
        else:
            parallel.allreduce_sum_(self.grads)      # loss terms are normalised by the GLOBAL ray count: a plain sum is the gradient
            if self.world > 1:                       # ... and every rank must take the same skip decision
                dist.all_reduce(flags[:1].view(torch.int32), op=dist.ReduceOp.MAX)
            call("rnb_engine_adam_guarded", self.flat, self.flat16, self._grad_buf, self.exp_avg, self.exp_avg_sq, i64(n),
                 f32(-1.0 if lr_from_device else (self.lr if lr is None else lr)), f32(self.betas[0]), f32(self.betas[1]),
                 f32(self.eps), f32(1.0), flags, self.status, stream_arg(self.device))
        if self.opt_step % self.status_interval == 0:
            self._post_status()

    # -- step guard: host side ---------------------------------------------------------------------------------
    def _post_status(self):
        """Asynchronous read-back of the status counters into a pinned ring; consumed by _poll_guard one or more steps
        later (the host never waits for the device here)."""
        if torch.cuda.is_current_stream_capturing():
            return          # (inside a CUDA-graph capture: the read-back is issued by the replaying side)
        if self._status_ring is None:
            self._status_ring = {"pin": torch.zeros(4, 16, dtype=torch.int32).pin_memory(), "evt": [None] * 4, "k": 0,
                                 "stream": torch.cuda.Stream(self.device)}
        r = self._status_ring
        k = r["k"] % 4
        done = torch.cuda.Event(); done.record(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(r["stream"]):
            r["stream"].wait_event(done)
            r["pin"][k].copy_(self.status, non_blocking=True)
            ev = torch.cuda.Event(); ev.record(r["stream"]); r["evt"][k] = ev
        r["k"] += 1

    def _poll_guard(self, wait=False):
        if wait:
            self._post_status()
        r = self._status_ring
        if r is None:
            return
        newest = None
        for j in range(r["k"] - 1, max(r["k"] - 5, -1), -1):       # newest finished slot
            ev = r["evt"][j % 4]
            if ev is None:
                continue
            if wait:
                ev.synchronize()
            if ev.query():
                newest = r["pin"][j % 4].tolist()
                break
        if newest is None:
            return
        skipped_inf, skipped_cap, needed, applied = newest[:4]
        seen = self._status_seen
        d_inf, d_cap, d_app = skipped_inf - seen[0], skipped_cap - seen[1], applied - seen[3]
        self._status_seen = [skipped_inf, skipped_cap, needed, applied]
        if self.dynamic_loss_scale:
            if d_inf > 0:        # GradScaler: backoff_factor 0.5 per skipped step
                self.cfg.loss_scale = max(self.cfg.loss_scale * 0.5 ** d_inf, 1.0)
                self._good_since_growth = 0
            elif d_app > 0:
                self._good_since_growth += d_app
                if self._good_since_growth >= self.growth_interval:      # growth_factor 2 every growth_interval clean steps
                    self.cfg.loss_scale = min(self.cfg.loss_scale * 2.0, 2.0 ** 24)
                    self._good_since_growth = 0
        if d_cap > 0:
            cap = self.cfg.sample_capacity
            if self.on_overflow == "raise":
                raise SampleCapacityError(
                    f"{d_cap} training step(s) needed up to {needed} samples but the workspace holds {cap} "
                    f"(n_rays {self.n_rays} x K {self.K} x samples_per_ray {self.samples_per_ray}); those optimizer steps "
                    f"were skipped on the device.  Re-create the Engine with samples_per_ray >= "
                    f"{-(-needed // (self.n_rays * self.K)) + 1} or on_overflow='grow'.")
            self.grow(int(needed * 1.25))

    def grow(self, min_capacity):
        """Re-allocate the workspace for at least `min_capacity` samples (on_overflow='grow')."""
        torch.cuda.synchronize(self.device)
        self._ahead.clear()
        old = self.status.clone()          # (the device is idle: this is the exact state, not the host's lagging copy)
        self._status_seen = old[:4].tolist()
        self.cfg.sample_capacity = (max(int(min_capacity), self.cfg.sample_capacity) + 127) // 128 * 128
        self.samples_per_ray = self.cfg.sample_capacity // (self.n_rays * self.K)
        self._alloc_workspace(carry_status=old)
        self._status_ring = None           # pending read-backs refer to the old block
        self._render_cache = None

    def guard_status(self):
        """{'skipped_nonfinite', 'skipped_capacity', 'max_samples_needed', 'applied', 'loss_scale'} -- synchronises."""
        torch.cuda.synchronize(self.device)
        s = self.status.tolist()
        return {"skipped_nonfinite": s[0], "skipped_capacity": s[1], "max_samples_needed": s[2], "applied": s[3],
                "loss_scale": float(self.cfg.loss_scale)}

    def gather_master(self):
        """Sharded optimizer: fp32 master parameters and Adam moments are current only inside each rank's own shard.
        Gathers them (for a checkpoint, or before using the Python operator surface on these modules)."""
        if not self.sharded or not getattr(self, "_master_stale", False):
            return
        n_sh = self.offsets["gate"] if self.K > 1 else self.n_params      # the gate block is replicated (and current) everywhere
        sl = parallel.adam_shard(n_sh)
        for t in (self.flat, self.exp_avg, self.exp_avg_sq):
            own, rep = t[sl].clone(), t[n_sh:].clone()
            t.zero_()
            t[sl] = own
            if parallel.rank() == 0:
                t[n_sh:] = rep
            parallel.allreduce_sum_(t)
        self._master_stale = False

    def background(self, random_bg=False):
        if self.exp_step_factor == 0:
            return [[1.0, 1.0, 1.0]] * self.K
        if random_bg:
            return torch.rand(self.K, 3).tolist()
        return [[0.0, 0.0, 0.0]] * self.K

    def step(self, rays_o, rays_d, target_rgb, lr=None, random_bg=False, noise=None, next_rays=None):
        """One training step.  Returns the loss vector view [total, rgb, opacity, cv, depth_mutual, ...] (device)."""
        self.forward_backward(rays_o, rays_d, target_rgb, noise, self.background(random_bg), True, next_rays)
        self.adam_step(lr)
        return self.loss

    def graph_stepper(self):
        """CUDA-graph replay of the step (GraphStepper): single-process engines only."""
        return GraphStepper(self)

    def results(self):
        """The reference's ml_render result dict (ml_rendering.py:72-77) as views of the workspace ("independent_rgbs",
        the K per-sub-NeRF images after the background blend, is formed when this is called: K small elementwise ops)."""
        out = {"rgb": self.rgb, "depth": self.depth, "opacity": self.opacity,
               "independent_rgbs": self.independent_rgbs(), "pass_rgb": self.pass_rgb, "pass_opacity": self.pass_opacity}
        if self.K > 1:
            out["gating_code"] = self.gate_prob
            out["gating_importance"] = self.importance
        return out

    def independent_rgbs(self):
        """The K per-sub-NeRF images AFTER the background blend (ml_rendering.py:46, 65, 73: `independent_rgbs`).  The
        workspace keeps the composited colour and opacity of every pass before the blend, so this is
        rgb_i + bg_i (1 - opacity_i), formed lazily (K small elementwise ops; not part of the timed step)."""
        last = getattr(self, "_last_bg", None)
        if last is None:
            return None
        key = tuple(tuple(r) for r in last)
        cache = getattr(self, "_bg_cache", None)
        if cache is None or cache[0] != key:      # (the background only changes with --random_bg)
            cache = (key, torch.tensor(last, dtype=torch.float32, device=self.device))
            self._bg_cache = cache
        bg = cache[1]
        return [self.pass_rgb[i] + bg[i][None, :] * (1.0 - self.pass_opacity[i])[:, None] for i in range(self.K)]

    # -- occupancy grids ------------------------------------------------------------------------------------
    def grids(self):
        if self.K == 1 and not hasattr(self.model, "size"):
            return [self.model.density_grid]
        return [getattr(self.model, f"density_grid_{i}") for i in range(self.K)]

    @torch.no_grad()
    def update_grid(self, density_threshold, warmup=False, decay=0.95, grids=None, bitfields=None, noise=None, coords1=None,
                    pick_u=None, seed=None):
        """MNGP.update_density_grid (networks.py:374-409) for all K sub-NeRFs through rnb_grid_update: no host
        synchronisation, 6-10 launches.  `grids` / `bitfields` default to the model's buffers (updated in place);
        noise / coords1 / pick_u are optional explicit random inputs (see include/radnerf_b200.h)."""
        self.sync_params()
        if getattr(self, "_grid_ws", None) is None:
            L = _GridLayout()
            call("rnb_grid_update_layout_of", ctypes.byref(self.cfg), ctypes.byref(L))
            self._grid_layout = L
            self._grid_ws = torch.zeros(L.total_bytes, dtype=torch.uint8, device=self.device)
            self._grid_calls = 0
        grids = self.grids() if grids is None else grids
        bitfields = self.bitfields() if bitfields is None else bitfields
        for g in grids:
            _lib.check_input(g, "density_grid", torch.float32)
        g_arr = (ctypes.c_void_p * self.K)(*[g.data_ptr() for g in grids])
        b_arr = (ctypes.c_void_p * self.K)(*[b.data_ptr() for b in bitfields])
        self._grid_calls += 1
        if seed is None:
            seed = 0x5EED0000 + self._grid_calls
        call("rnb_grid_update", ctypes.byref(self.cfg), self._grid_ws, self.flat16, g_arr, b_arr, i32(1 if warmup else 0),
             f32(decay), f32(density_threshold), noise, coords1, pick_u, ctypes.c_uint64(seed), i32(0), stream_arg(self.device))

    def grid_mean(self):
        """mean density over grid > 0 of the last update, per sub-NeRF (device tensor view)."""
        L = self._grid_layout
        return self._grid_ws[L.mean:L.mean + 4 * self.K].view(torch.float32)

    # -- test-time rendering ----------------------------------------------------------------------------------
    def _render_plan(self, n_rays, samples_per_ray):
        key = (n_rays, samples_per_ray)
        plan = getattr(self, "_render_cache", None)
        if plan is None or plan[0] != key:
            c = _Config()
            ctypes.memmove(ctypes.byref(c), ctypes.byref(self.cfg), ctypes.sizeof(_Config))
            c.n_rays = n_rays
            c.sample_capacity = (n_rays * self.K * max(int(samples_per_ray), 8) + 127) // 128 * 128 + 128 * self.K
            L = _RenderLayout()
            call("rnb_render_layout_of", ctypes.byref(c), ctypes.byref(L))
            ws = torch.empty(L.total_bytes, dtype=torch.uint8, device=self.device)
            plan = (key, c, L, ws)
            self._render_cache = plan
        return plan[1:]

    @torch.no_grad()
    def render(self, rays_o, rays_d, chunk=1 << 19, samples_per_ray=16, with_pass_rgb=False):
        """ml_render(test_time=True) (ml_rendering.py:18-77, 81-155) for a whole frame: all K sub-NeRFs advance
        together through rnb_render, `chunk` rays per call.  Returns the reference's result dict (+ "total_samples")."""
        _lib.check_input(rays_o, "rays_o", torch.float32)
        _lib.check_input(rays_d, "rays_d", torch.float32)
        self.sync_params()
        N, K, dev = rays_o.shape[0], self.K, self.device
        rgb = torch.empty(N, 3, device=dev); depth = torch.empty(N, K, device=dev); opacity = torch.empty(N, device=dev)
        pass_rgb = torch.empty(K, N, 3, device=dev) if with_pass_rgb else None
        gating = torch.empty(N, K, device=dev) if K > 1 else None
        bfs = self.bitfields()
        bf_arr = (ctypes.c_void_p * K)(*[b.data_ptr() for b in bfs])
        bg_arr = (ctypes.c_float * (3 * K))(*[float(x) for row in self.background(False) for x in row])
        stats = (ctypes.c_int64 * 2)()
        total, rounds = 0, 0
        for a in range(0, N, chunk):
            n = min(chunk, N - a)
            c, L, ws = self._render_plan(n, samples_per_ray)
            tmp_pass = torch.empty(K, n, 3, device=dev) if with_pass_rgb else None
            call("rnb_render", ctypes.byref(c), ws, self.flat16, rays_o[a:a + n], rays_d[a:a + n], bf_arr, bg_arr,
                 rgb[a:a + n], depth[a:a + n], opacity[a:a + n], tmp_pass, stats, stream_arg(dev))
            if with_pass_rgb:
                pass_rgb[:, a:a + n] = tmp_pass
            if K > 1:
                gating[a:a + n] = ws[L.gate_prob:L.gate_prob + n * K * 4].view(torch.float32).view(n, K)
            total += stats[0]; rounds += stats[1]
        out = {"rgb": rgb, "depth": depth, "opacity": opacity, "total_samples": total, "rounds": rounds,
               "independent_rgbs": [pass_rgb[i] for i in range(K)] if with_pass_rgb else None}     # ml_rendering.py:46, 65, 73
        if with_pass_rgb:
            out["pass_rgb"] = pass_rgb
        if K > 1:
            out["gating_code"] = gating
        return out

    def sample_stats(self):
        """(samples per pass list, overflow flag) -- synchronises."""
        seg = self._slot_view("seg", torch.int32, (2 * MAX_K + 4,), self._slot).tolist()
        K = self.K
        return seg[K + 1:2 * K + 1], bool(seg[2 * K + 1])


class GraphStepper:
    """The training step as CUDA graphs ("CUDA-graph-captured step").

    The eager step costs ~0.45 ms of host time (a dozen ctypes calls, stream and event bookkeeping) against ~0.63 ms on the
    device; replaying a captured graph costs a few tens of microseconds.  What makes the step capturable: sample counts never
    leave the device, the marching noise comes from the library's counter generator (no ATen launch), the count of applied
    optimizer steps and the learning rate are read from device memory (status[3], status[8]), and the step guard skips a bad
    step on the device.  Outside the graphs: the occupancy update (every 16 steps), the upload of the next batch, the
    4-byte learning-rate write and the status read-back.

    Two static batch slots alternate: the graph of slot s trains on the batch in slot s and (kind "steady" / "first")
    marches the batch in slot s ^ 1 ahead on the side stream, joined before the graph ends.  Kinds: "first" marches its own
    batch in line (first step, step after an occupancy update), "last" marches nothing ahead (step before an update),
    "steady" finds its samples marched by the previous replay.  A change of the loss scale (step guard) or of the workspace
    (capacity growth) drops the captured graphs; they are re-captured on the next step.
    """

    def __init__(self, engine):
        assert engine.world == 1, "graph replay is implemented for single-process engines"
        self.e = engine
        n, dev = engine.n_rays, engine.device
        self.slots = [tuple(torch.zeros(n, 3, device=dev) for _ in range(3)) for _ in range(2)]
        self.graphs, self.launches = {}, {}
        self._key = None
        self._pool = None
        self._warm = False
        self.replays = 0
        # hybrid: the graphs hold the parameter-dependent chain only and the marching stays an eager side-stream launch (host
        # 0.18 instead of 0.12 ms/step; same device time unless something -- the benchmark's L2 flush -- sits between replays)
        self.hybrid = os.environ.get("RNB_GRAPH_HYBRID", "0") == "1"

    def load(self, slot, rays_o=None, rays_d=None, target_rgb=None):
        """Copy (parts of) a batch into static slot `slot` on the current stream (device or pinned-host sources)."""
        for dst, src in zip(self.slots[slot], (rays_o, rays_d, target_rgb)):
            if src is not None:
                dst.copy_(src, non_blocking=True)

    def _body(self, slot, kind):
        e = self.e
        o, d, t = self.slots[slot]
        if kind == "first":
            e._ahead.clear()
        nxt = None if kind == "last" else self.slots[slot ^ 1][:2]
        e._force_slot = slot
        try:
            e.forward_backward(o, d, t, None, e.background(False), True, nxt)
        finally:
            e._force_slot = None
        e.adam_step(lr_from_device=True)
        if nxt is not None and torch.cuda.is_current_stream_capturing():
            # a capture must end with every forked stream joined: the march of the next batch ends inside this graph
            torch.cuda.current_stream(e.device).wait_event(e._ahead[(nxt[0].data_ptr(), nxt[1].data_ptr())][4])

    def _set_lr(self, lr, force=False):
        e = self.e
        v = float(e.lr if lr is None else lr)
        if force or v != getattr(self, "_lr_dev", None):       # (a constant rate is written once)
            if getattr(self, "_lr_ring", None) is None:
                self._lr_ring, self._lr_k = torch.zeros(64, dtype=torch.float32).pin_memory(), 0
            k = self._lr_k % 64                                  # a ring of pinned words: an earlier copy may still be in flight
            self._lr_k += 1
            self._lr_ring[k] = v
            e.status[8:9].view(torch.float32).copy_(self._lr_ring[k:k + 1], non_blocking=True)
            self._lr_dev = v

    def _capture(self, slot, kind):
        e = self.e
        g = torch.cuda.CUDAGraph()
        saved, n_before, l_before = dict(e._ahead), e.opt_step, _lib.launch_count
        o, d, _ = self.slots[slot]
        ev = torch.cuda.Event()
        # the bookkeeping of marched-ahead batches must look, at capture time, like it will at replay time
        e._ahead = {} if kind == "first" else {(o.data_ptr(), d.data_ptr()): (slot, o, d, None, ev)}
        torch.cuda.synchronize(e.device)
        if getattr(self, "_cap_stream", None) is None:
            # kernel nodes inherit the priority of the stream they were captured on: the step's own chain is captured at high
            # priority, so the side-stream work (next batch's march) only fills what the chain leaves free -- as in the eager
            # step, which the caller runs on a high-priority stream
            self._cap_stream = torch.cuda.Stream(e.device, priority=-1)
        with torch.cuda.graph(g, pool=self._pool, stream=self._cap_stream, capture_error_mode="thread_local"):
            if kind != "first":
                ev.record()        # stands for "the previous replay's march is complete" (it is: replays are stream-ordered)
            self._body(slot, kind)
        self._pool = self._pool or g.pool()
        e.opt_step, e._ahead = n_before, saved          # capturing enqueues nothing
        self.launches[(slot, kind)] = _lib.launch_count - l_before
        _lib.launch_count = l_before
        return g

    def step(self, slot, lr=None, kind="steady"):
        """Train on the batch in static slot `slot`; unless kind == "last", march the batch in slot ^ 1 ahead (it must have
        been loaded).  Returns the engine's loss vector view."""
        e = self.e
        e.sync_params()
        e._poll_guard()
        key = (e.cfg.loss_scale, e.ws.data_ptr())       # both are baked into the graphs
        if key != self._key:
            self.graphs.clear(); self._key = key; self._lr_dev = None      # (a new workspace has a fresh status block)
        cur = torch.cuda.current_stream(e.device)
        self._set_lr(lr)
        if not self._warm:
            # the library sets kernel attributes lazily and the engine creates its side streams on first use: one eager step
            self._body(slot, kind); self._warm = True; self.replays += 1
            if e.opt_step % e.status_interval == 0:
                e._post_status()
            return e.loss
        if self.hybrid:
            # graphs hold the parameter-dependent chain only (field, compositing, losses, backward, Adam: kind "last"); the
            # marching of this batch (kind "first") and of the next one stays an eager launch on the side stream
            if kind == "first":
                for rec in e._ahead.values():
                    cur.wait_event(rec[4])
                e._ahead.clear()
                e.prefetch(self.slots[slot][0], self.slots[slot][1], slot=slot)
            mine = e._ahead.pop((self.slots[slot][0].data_ptr(), self.slots[slot][1].data_ptr()), None)
            if mine is not None:
                cur.wait_event(mine[4])
            if kind != "last":
                while len(e._ahead) > 0:
                    cur.wait_event(e._ahead.pop(next(iter(e._ahead)))[4])
                e.prefetch(self.slots[slot ^ 1][0], self.slots[slot ^ 1][1], slot=slot ^ 1)
            gk = (slot, "last")
            if gk not in self.graphs:
                keep = dict(e._ahead)
                self.graphs[gk] = self._capture(slot, "last")
                e._ahead = keep
            self.graphs[gk].replay()
            _lib.launch_count += self.launches[gk]
            e.opt_step += 1
            e._slot = slot
            self.replays += 1
            if e.opt_step % e.status_interval == 0:
                e._post_status()
            return e.loss
        gk = (slot, kind)
        if gk not in self.graphs:
            self.graphs[gk] = self._capture(slot, kind)
        for rec in e._ahead.values():          # a batch marched ahead by an EAGER step (the warm-up): order the replay behind it
            cur.wait_event(rec[4])
        e._ahead.clear()
        self.graphs[gk].replay()
        _lib.launch_count += self.launches[gk]
        e.opt_step += 1
        e._slot = slot
        self.replays += 1
        if e.opt_step % e.status_interval == 0:
            e._post_status()
        return e.loss
