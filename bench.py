#!/usr/bin/env python
"""bench.py -- Rad-NeRF hot-path benchmark on B200 (contract: see the task statement).

    python bench.py --gpus N --steps K --warmup W            # this repo's sm_100a path
    python bench.py --impl reference --steps K --warmup W    # the reference's math on host cores

Metric (BASELINE.json): train rays/s for Rad-NeRF K=2 (config "radnerf_tat_k2": TAT-shape
1920x1080 cameras, scale 0.5, batch 8192 rays per GPU, lambda_dm=5e-3, lambda_cv=1e-2,
lambda_opacity=1e-3), synthetic analytic scene, tinycudann-style random init.  One "step" =
the training step of train_ml.py:172-200 (grid update every 16 steps, gate, K x (march, field,
composite), NeRFLoss, backward, gradient all-reduce when N>1, Adam).  Weak scaling: every rank
gets its own 8192-ray batch.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BATCH = int(os.environ.get("RNB_BENCH_BATCH", "8192"))      # rays per step (BASELINE: 8192; the override is for debugging small shards)
N_BATCHES = 8          # distinct pre-generated batches cycled through
TRAFFIC_JSON = "r02_traffic.json"   # ncu --set full DRAM bytes per launch of the current build (scripts/profile_round.sh)


def workload(name):
    import radnerf_b200  # noqa: F401
    from radnerf_b200 import synthetic as syn
    w = dict(syn.WORKLOADS[name])
    return syn, w


def make_batches(syn, w, rank, n_batches=N_BATCHES, batch=BATCH):
    cam = syn.make_cameras(100, w["radius"], w["H"], w["W"], w["fx"])
    scene = syn.make_scene(w["scale"], 1234)
    out = []
    for b in range(n_batches):
        o, d, _, _ = syn.train_batch(cam, batch, seed=42 + 1000 * rank + b)
        # target colours: the analytic scene's colour at the ray's box midpoint (cheap stand-in for images)
        tgt = scene.color(o + d * 1.5)
        out.append((o, d, tgt.astype(np.float32)))
    return out


def scene_bitfields(syn, w):
    bfs = []
    for i in range(w["K"]):
        sc = syn.make_scene(w["scale"], 1234, jitter_seed=i if i else None)
        bfs.append(syn.packbits_np(sc.density_grid, sc.threshold))
    return bfs


class ClockSampler:
    """Samples SM clocks / throttle reasons of one GPU while the timed region runs: NVML from a thread every 5 ms (a timed
    region of a few tens of milliseconds still gets samples), nvidia-smi -lms as the fallback."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index
        self.nv, self.h, self.stop_flag, self.thread = None, None, False, None
        self.sm, self.bits, self.sm_max = [], 0, None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[self.index]) if vis and all(x.strip().isdigit() for x in vis.split(",")) else self.index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.sm_max = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.nv = pynvml
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.nv = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _poll(self):
        nv = self.nv
        while not self.stop_flag:
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                self.bits |= int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
            except Exception:
                try:
                    self.bits |= int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                except Exception:
                    pass
            time.sleep(0.005)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.nv is not None:
            self.stop_flag = True
            self.thread.join(timeout=1.0)
            nv = self.nv
            names = [("hw_slowdown", getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8)),
                     ("hw_thermal_slowdown", getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40)),
                     ("sw_thermal_slowdown", getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20)),
                     ("sw_power_cap", getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4))]
            if not self.sm:
                return {"sm_mhz": None, "sm_max_mhz": self.sm_max, "reasons": ["no samples"]}
            sm = sorted(self.sm)
            return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.sm_max, "reasons": [n for n, b in names if self.bits & b],
                    "samples": len(sm), "source": "nvml"}
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for r in self.rows if len(r) >= 7]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = sorted(float(r[0]) for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(r[3 + k].lower().startswith("active") for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][1]), "reasons": reasons, "samples": len(rows), "source": "nvidia-smi"}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return d.get("hbm_gbs", 6650.0), d.get("bf16_tflops_sustained", 1400.0), "measured"
    return 6650.0, 1400.0, "fallback"


# Algorithmic work of one launch of each stage of the fused step (SURVEY.md 8d per-unit figures), as a
# function of the samples N of all K passes and the K*R marched rays of that launch.
#   bytes: march 60 B/ray + 32 B/sample (incl. expand); field fwd (fused) 34 B/sample; composite fw 28 B/sample
#   + 52 B/ray; composite bw 48 B/sample + 64 B/ray; field bwd = encode bwd 76 B/sample; adam 28 B/param.
#   flops: MLP fwd 18 944 FLOP/sample (unpadded), bwd 37 888 FLOP/sample; gate 25 344 + 128 K FLOP/ray.
STAGES = ["gate", "march", "scan", "expand", "field_fwd", "composite_fw", "blend_loss", "composite_bw", "field_bwd",
          "gate_bwd", "adam", "barrier"]
STAGE_BYTES = {
    "march": lambda N, KR, P: 60 * KR + 20 * N,
    "expand": lambda N, KR, P: 12 * N,
    "field_fwd": lambda N, KR, P: 34 * N,
    "composite_fw": lambda N, KR, P: 28 * N + 52 * KR,
    "composite_bw": lambda N, KR, P: 48 * N + 64 * KR,
    "field_bwd": lambda N, KR, P: 76 * N,
    "adam": lambda N, KR, P: 28 * P,
}
STAGE_FLOPS = {
    "field_fwd": lambda N, KR, P: 18944 * N,
    "field_bwd": lambda N, KR, P: 37888 * N,
}


def arm_config(args, w, world):
    """The workload description, IDENTICAL in both arms (product and --impl reference)."""
    return {"workload": args.workload, "K": w["K"], "scale": w["scale"], "batch_rays_per_gpu": BATCH,
            "step": "forward only (AABB + march + composite)" if w.get("forward_only") else "fwd + loss + bwd + Adam",
            "params": "tcnn-style random init", "occupancy": "analytic scene, fixed", "parallelism": f"ray-sharded dp{world}",
            "l2": "GPU arm: flushed between steps (256 MiB write, its time subtracted; the un-flushed loop is reported beside it "
                  "as ms_per_step_noflush); reference arm: host run"}


def cpu_reference(args, steps, warmup):
    """The reference's math for the same step on the host cores: plain PyTorch + the C restatement of its kernels
    (the reference itself has no CPU implementation of this path).  FIXED batch of BATCH rays per step -- never sized
    from a timing -- `warmup` untimed steps, then `steps` timed ones; the value is BATCH / median step time."""
    import torch
    syn, w = workload(args.workload)
    from oracle import render_cpu as R
    from oracle import vren_cpu as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    os.environ.setdefault("OMP_NUM_THREADS", str(cores))
    K = w["K"]
    bfs = scene_bitfields(syn, w)
    esf = syn.exp_step_factor_for_scale(w["scale"])
    C = syn.cascades_for_scale(w["scale"])
    batches = make_batches(syn, w, 0, n_batches=2)
    rng = np.random.RandomState(7)
    n = BATCH
    if w.get("forward_only"):
        # BASELINE config 1: NGP K = 1 volume-render FORWARD: ray-AABB + marching in the C restatement
        # (OpenMP over rays; raymarching.cu:184-280), compositing in plain PyTorch CPU ops following volumerendering.cu:23-44,
        # white background (rendering.py:229-237).  sigma / rgb come from the analytic scene, computed per batch outside
        # the timed region, so the timed region is sampling + compositing only.
        scene = syn.make_scene(w["scale"], 1234)
        pre = []
        for (o, d, _) in batches:
            nz = rng.rand(n).astype(np.float32)
            h = R.hits_for(o, d, w["scale"])
            _, xyzs, _, _, _, _ = O.raymarching_train(o, d, h, bfs[0], C, w["scale"], esf, nz)
            pre.append((nz, torch.from_numpy(scene.density(xyzs)), torch.from_numpy(scene.color(xyzs))))

        def step(i):
            o, d, _ = batches[i % 2]
            nz, sig, col = pre[i % 2]
            h = R.hits_for(o, d, w["scale"])
            rays_a, xyzs, dirs, deltas, ts, _ = O.raymarching_train(o, d, h, bfs[0], C, w["scale"], esf, nz)
            with torch.no_grad():
                opacity, depth, rgb, ws, tot = R.composite_train(sig, col, torch.from_numpy(deltas), torch.from_numpy(ts), rays_a, 1e-4)
                rgb = rgb + (1 - opacity)[:, None]
            return float(rgb.sum())
        what = "forward only: C marcher (OpenMP) + plain-PyTorch composite"
    else:
        P = R.FieldParams(w["scale"], K, seed=0).requires_grad_(True)
        opt = torch.optim.Adam(list(P.p.values()), 1e-2, eps=1e-15)

        def step(i):
            o, d, t = batches[i % 2]
            noises = [rng.rand(n).astype(np.float32) for _ in range(K)]
            res = R.ml_render_train(P, o, d, bfs, noises, C, esf)
            loss, _ = R.nerf_loss(res, torch.from_numpy(t), 1e-3, 1e-2 if K > 1 else 0, 5e-3 if K > 1 else 0)
            opt.zero_grad(); loss.backward(); opt.step()
            return float(loss.detach())
        what = "fwd+loss+bwd+Adam"
    for i in range(warmup):
        step(i)
    times = []
    for i in range(steps):
        t0 = time.perf_counter(); step(warmup + i); times.append(time.perf_counter() - t0)
    med = float(np.median(times))
    return {"value": n / med, "unit": "rays/s", "cores": cores, "kind": "port",
            "sample": f"{n} of {BATCH} rays per step (fixed; same scene/cameras/K), {what}; {warmup} warm-up + median of {steps} timed steps",
            "ms_per_step": 1e3 * med, "ms_min": 1e3 * min(times), "ms_max": 1e3 * max(times)}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    syn, w = workload(args.workload)
    cb = cpu_reference(args, max(args.steps, 1), args.warmup)
    val = cb["value"]
    line = {"impl": "reference", "metric": "train_rays_per_s", "value": val, "unit": "rays/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": cb["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32 (fp16 rounding points)", "data": "synthetic",
            "config": arm_config(args, w, args.gpus),
            "cpu_baseline": cb,
            "e2e": {"value": val, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def run_product(args):
    import ctypes
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        opts = dist.ProcessGroupNCCL.Options()
        opts.is_high_priority_stream = True      # the step's two small collectives must not queue behind side-stream work
        dist.init_process_group("nccl", device_id=torch.device("cuda", local), pg_options=opts)
    syn, w = workload(args.workload)
    from radnerf_b200 import _lib
    from radnerf_b200.trainer import Trainer
    lib = _lib.lib()   # fail loudly if the CUDA library is missing: there is no fallback
    K = w["K"]
    # --graph (single process): the step is replayed from captured CUDA graphs (Trainer(graph=True))
    # (default: eager launches -- the host enqueues a step in 0.29 ms against 0.62 ms on the device, so the device never waits,
    # and a graph launch adds ~0.01 ms at every step boundary; --graph: replay, host 0.12 ms/step)
    use_graph = world == 1 and args.graph and not args.python_surface and not args.no_march_ahead
    # weak scaling (default): 8192 rays per GPU; strong: a global batch of 8192 rays split over the ranks
    batch = BATCH // world if args.scaling == "strong" else BATCH
    tr = Trainer(scale=w["scale"], K=K, freeze_occupancy=True, device=f"cuda:{local}", engine=not args.python_surface,
                 batch_rays=batch, graph=use_graph)
    tr.set_occupancy(scene_bitfields(syn, w))
    host = [tuple(torch.from_numpy(a).pin_memory() for a in b) for b in make_batches(syn, w, rank, batch=batch)]
    dev = [tuple(a.cuda(non_blocking=True) for a in b) for b in host]
    flush = torch.empty(256 * 1024 * 1024 // 4, device="cuda")      # 256 MiB > 126 MB L2

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(fn, steps):
        sync_all()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for s in range(steps):
            fn(s)
        e1.record()
        sync_all()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    ahead = tr.engine is not None and not args.no_march_ahead
    trace = os.environ.get("RNB_BENCH_TRACE") == "1"

    def mark(what):      # progress markers on stderr (RNB_BENCH_TRACE=1): where a multi-rank run stops, if it does
        if trace:
            print(f"[bench rank {rank}] {what}", file=sys.stderr, flush=True)

    def flush_l2():
        # evict L2 between steps (inputs are smaller than L2).  When the next batch is marched ahead on the side stream,
        # that work must be finished before the flush starts: otherwise it could hide behind the flush, whose time is
        # subtracted below, and the step would look faster than it is.
        if ahead and tr.engine._march_stream is not None:
            torch.cuda.current_stream().wait_stream(tr.engine._march_stream)
        flush.add_(1.0)

    def dev_step(s):
        flush_l2()
        nxt = dev[(s + 1) % len(dev)]
        tr.step(*dev[s % len(dev)], next_rays=(nxt if use_graph else nxt[:2]) if ahead else None)

    def plain_step(s):      # no marching ahead: clean per-stage times for the roofline leg
        flush.add_(1.0)
        tr.step(*dev[s % len(dev)])

    losses = []
    staged = {}
    loss_pin = torch.zeros(4, 8, dtype=torch.float32).pin_memory()      # ring of pinned read-back slots
    loss_evt = [None] * 4

    def read_back(s, loss_vec):
        # device->host read of every step's loss: an asynchronous copy into pinned memory, consumed one step later so that
        # the host keeps launching ahead (what a training loop that logs its loss does); the last ones are drained below
        k = s % 4
        if rb_stream is None:
            loss_pin[k].copy_(loss_vec, non_blocking=True)
            ev = torch.cuda.Event(); ev.record(); loss_evt[k] = ev
        else:
            # the copy engine's turn sits on its own stream: the step's stream never waits for a DMA between two kernels
            done = torch.cuda.Event(); done.record()
            with torch.cuda.stream(rb_stream):
                rb_stream.wait_event(done)
                loss_pin[k].copy_(loss_vec, non_blocking=True)
                ev = torch.cuda.Event(); ev.record(rb_stream); loss_evt[k] = ev
        j = (s - 1) % 4
        if s > 0 and loss_evt[j] is not None:
            loss_evt[j].synchronize()
            losses.append(float(loss_pin[j][0]))

    feeder = None
    rb_stream = torch.cuda.Stream() if os.environ.get("RNB_E2E_INLINE_READBACK") != "1" else None

    def e2e_step(s):
        # every step uploads ONE batch from pinned host memory and reads the step's loss back.  The upload goes through
        # the library's HostFeeder: batch s+1 is copied on a copy stream while step s computes (when marching ahead it is
        # also the batch whose rays are marched on the side stream), so the copy never sits between two kernels
        nonlocal feeder
        flush_l2()
        if rb_stream is not None and s > 0 and loss_evt[(s - 1) % 4] is not None:
            # the next step clears the loss vector: order it behind the previous step's read-back (long finished by now)
            torch.cuda.current_stream().wait_event(loss_evt[(s - 1) % 4])
        if feeder is None:
            from radnerf_b200.data import HostFeeder
            feeder = HostFeeder(host, f"cuda:{local}", depth=4)
        if os.environ.get("RNB_E2E_INLINE_COPY") == "1":     # A/B: copies on the step's own stream
            if s not in staged:
                staged[s] = tuple(a.cuda(non_blocking=True) for a in host[s % len(host)])
            staged[s + 1] = tuple(a.cuda(non_blocking=True) for a in host[(s + 1) % len(host)])
            o, d, t = staged.pop(s)
            tr.step(o, d, t, next_rays=staged[s + 1][:2] if ahead else None)
        else:
            o, d, t = feeder.get(s)
            if not ahead:
                tr.step(o, d, t)
            else:
                nxt = feeder.peek(s + 1)
                tr.step(o, d, t, next_rays=nxt if use_graph else nxt[:2])
            feeder.done(s)
        read_back(s, tr.engine.loss if tr.engine is not None else tr.last_loss.reshape(1).expand(8))

    if ahead and os.environ.get("RNB_BENCH_PRIO", "1") == "1":
        # the step runs on a high-priority stream: the side stream that marches the next batch then only fills what the
        # step's own kernels leave free instead of delaying them
        torch.cuda.synchronize()
        torch.cuda.set_stream(torch.cuda.Stream(priority=-1))
    mark("set-up done")
    for s in range(args.warmup):
        dev_step(s)
    torch.cuda.synchronize(); mark("warm-up done")
    # time the L2 flush alone so that it can be subtracted (it is not part of the step)
    flush_ms = timed(lambda s: flush.add_(1.0), args.steps) / args.steps
    clocks = ClockSampler(local); clocks.start()
    n0 = _lib.launch_count
    ms = timed(lambda s: dev_step(s + args.warmup), args.steps)      # (indices continue: the batch marched ahead is the next one trained)
    launches = _lib.launch_count - n0
    clk = clocks.stop()
    ms_step = ms / args.steps - flush_ms
    mark("timed steps done")
    # the same loop without the L2 flush (nothing subtracted): parameters and inputs of the previous step stay in L2
    def noflush_step(s):
        nxt = dev[(s + 1) % len(dev)]
        tr.step(*dev[s % len(dev)], next_rays=(nxt if use_graph else nxt[:2]) if ahead else None)
    noflush_step(args.warmup + args.steps)
    ms_noflush = timed(lambda s: noflush_step(s + 1 + args.warmup + args.steps), args.steps) / args.steps
    for s in range(args.warmup):         # untimed: feeder slots, copy / read-back streams and pinned buffers come up here
        e2e_step(s)
    e2e_ms = timed(lambda s: e2e_step(s + args.warmup), args.steps) / args.steps - flush_ms
    mark("noflush + e2e done")
    # per-stage device time of the same steps (separate pass: CUDA events around every stage of the fused step)
    stage_ms = {}
    if tr.engine is not None:
        tr.engine._ahead.clear()
        tr.stepper, tr._staged = None, None      # the per-stage events are recorded by the library's host code: eager launches here
        torch.cuda.synchronize()
        lib.rnb_engine_profile(1)
        timed(plain_step, args.steps)
        arr_ms = (ctypes.c_double * 16)(); arr_n = (ctypes.c_int32 * 16)()
        lib.rnb_engine_profile_read(arr_ms, arr_n, 16)
        lib.rnb_engine_profile(0)
        stage_ms = {STAGES[i]: (arr_ms[i] / max(arr_n[i], 1), arr_n[i]) for i in range(len(STAGES)) if arr_n[i]}
    mark("stage profile done")
    # multi-rank correctness the driver can see: every rank must hold bit-identical fp16 parameters after the timed steps
    # (one writer per shard in the peer-memory optimizer) and must have seen the same global gate importance
    replica = None
    if world > 1 and tr.engine is not None:
        p16 = tr.engine.flat16.view(torch.int16).to(torch.int64)
        idx = torch.arange(1, p16.numel() + 1, device=p16.device, dtype=torch.int64)
        chk = torch.stack([p16.sum(), (p16 * (idx % 65521)).sum(),
                           tr.engine.importance.double().sum().mul(1e6).round().to(torch.int64)])
        allc = [torch.zeros_like(chk) for _ in range(world)]
        dist.all_gather(allc, chk)
        same = all(bool(torch.equal(allc[0], c)) for c in allc)
        replica = {"status": "ok" if same else "MISMATCH", "ranks": world,
                   "flat16_checksum": [int(x) for x in allc[0][:2].tolist()],
                   "what": "fp16 parameter vector (two 64-bit checksums) and global gate importance equal on every rank after the timed steps"}
        assert same, f"replicas diverged: {[c.tolist() for c in allc]}"
    rays_total = batch * world
    value = rays_total / (ms_step * 1e-3)
    e2e_val = rays_total / (e2e_ms * 1e-3)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    hbm, tflops, which = measured_peaks()
    line = {"metric": "train_rays_per_s", "value": value, "unit": "rays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "f16 (fp32 accumulate; fp32 march/composite)", "data": "synthetic",
            "config": arm_config(args, w, world),
            "run": {"batch_rays_this_rank": batch,
                    "occupancy_update": "every 16 steps, on shadow grids (the marched occupancy stays the analytic scene's)",
                    "march_ahead": bool(ahead), "cuda_graph": bool(use_graph),
                    "path": "fused engine (rnb_engine_*)" if tr.engine is not None else "python operator surface"},
            "ms_per_step_noflush": ms_noflush, "value_noflush": rays_total / (ms_noflush * 1e-3),
            "clocks": clk,
            "e2e": {"value": e2e_val, "unit": "rays/s", "h2d_bytes_per_step": 3 * batch * 12, "d2h_bytes_per_step": 32,
                    "ms_per_step": e2e_ms},
            "gpu_launches": launches, "final_loss": losses[-1] if losses else None}
    if tr.engine is not None:
        counts, overflow = tr.engine.sample_stats()
        N = float(sum(counts)); KR = float(K * batch); Pn = float(tr.engine.n_params)
        line["run"]["samples_per_ray"] = N / KR
        line["run"]["sample_overflow"] = overflow
        top = max(stage_ms, key=lambda k: stage_ms[k][0] * stage_ms[k][1])
        t_ms = stage_ms[top][0]
        fb = STAGE_BYTES.get(top, lambda *a: 0)(N, KR, Pn) / (t_ms * 1e-3) / 1e9
        ff = STAGE_FLOPS.get(top, lambda *a: 0)(N, KR, Pn) / (t_ms * 1e-3) / 1e12
        if ff / tflops > fb / hbm:
            roof = {"kernel": top, "bound": "tensor", "achieved": ff, "peak": tflops, "unit": "TFLOP/s", "frac": ff / tflops}
        else:
            roof = {"kernel": top, "bound": "hbm", "achieved": fb, "peak": hbm, "unit": "GB/s", "frac": fb / hbm}
        # DRAM bytes per launch of that kernel from the committed ncu --set full capture (profiles/, same workload and batch)
        traffic = None
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", TRAFFIC_JSON)))["kernels"]
            kname = {"field_bwd": "field_bwd_tc_kernel<7>", "field_fwd": "field_fwd_tc_kernel<1>", "march": "eng_march_kernel<1>",
                     "adam": "eng_adam_kernel"}.get(top)
            if kname in tj and args.workload == "radnerf_tat_k2" and tr.engine is not None:
                traffic = tj[kname]["dram_bytes_per_launch"]
        except (OSError, KeyError, ValueError):
            pass
        roof.update({"traffic": traffic, "traffic_unit": "bytes/launch (ncu dram__bytes_read.sum + dram__bytes_write.sum)",
                     "algorithmic_bytes_per_launch": STAGE_BYTES.get(top, lambda *a: 0)(N, KR, Pn), "peak_source": which, "hbm_frac": fb / hbm, "tensor_frac": ff / tflops,
                     "launch_ms": t_ms, "share_of_step": t_ms * stage_ms[top][1] / args.steps / ms_step,
                     "stage_ms_per_launch": {k: round(v[0], 4) for k, v in stage_ms.items()},
                     "stages_share_of_step": sum(v[0] * v[1] for v in stage_ms.values()) / args.steps / ms_step})
        # what actually bounds the field kernels: scattered accesses to the L2-resident tables.  Reference rates from the
        # microbenchmark (scripts/microbench_l2.py), sector counts per sample from the ncu capture of this build (profiles/)
        try:
            ref = json.load(open(os.path.join(ROOT, "profiles", "r02_l2_roofline_refs.json")))
            if "field_bwd" in stage_ms and args.workload == ref.get("workload"):
                t_b = stage_ms["field_bwd"][0] * 1e-3
                ach = ref["bwd_red_sectors_per_sample"] * N / t_b / 1e9
                roof["l2_scatter"] = {"kernel": "field_bwd", "what": "32-byte reduction sectors/s into the fp32 gradient table (L2)",
                                      "achieved": ach, "peak": ref["red_peak_G_per_s"], "unit": "G sectors/s",
                                      "frac": ach / ref["red_peak_G_per_s"], "sectors_per_sample": ref["bwd_red_sectors_per_sample"],
                                      "source": ref["source"]}
        except (OSError, KeyError, ValueError):
            pass
        line["roofline"] = roof
    if replica is not None:
        line["replica_check"] = replica["status"]
        line["replica"] = replica
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(args)
        line["gpu_anchor"] = gpu_anchor()
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def run_product_fwd_only(args):
    """BASELINE config 1 on the GPU: the volume-render FORWARD (ray-AABB, training march, compositing, white background) of
    an Instant-NGP K = 1 model through this library's `vren` operator surface, sigma / rgb supplied by the analytic scene
    (computed per batch outside the timed region): the timed region is sampling + compositing, as in the CPU arm."""
    import torch
    torch.cuda.set_device(0)
    syn, w = workload(args.workload)
    from radnerf_b200 import _lib, vren
    _lib.lib()
    sc = syn.make_scene(w["scale"], 1234)
    bf = torch.from_numpy(scene_bitfields(syn, w)[0]).cuda()
    host = [tuple(torch.from_numpy(a).pin_memory() for a in b) for b in make_batches(syn, w, 0)]
    center = torch.zeros(1, 3, device="cuda"); half = torch.full((1, 3), float(w["scale"]), device="cuda")
    rng = np.random.RandomState(7)
    noise = [torch.from_numpy(rng.rand(BATCH).astype(np.float32)).cuda() for _ in host]
    flush = torch.empty(256 * 1024 * 1024 // 4, device="cuda")

    def fwd(o, d, nz, sig=None, col=None):
        _, hits_t, _ = vren.ray_aabb_intersect(o, d, center, half, 1)
        h = hits_t[:, 0].contiguous()
        h[(h[:, 0] >= 0) & (h[:, 0] < syn.NEAR_DISTANCE), 0] = syn.NEAR_DISTANCE
        rays_a, xyzs, dirs, deltas, ts, counter = vren.raymarching_train(o, d, h, bf, 1, w["scale"], 0.0, nz, 128, 1024)
        if sig is None:
            return xyzs
        _, opacity, depth, rgb, ws = vren.composite_train_fw(sig, col, deltas, ts, rays_a, 1e-4)
        return rgb + (1 - opacity)[:, None]

    pre = []
    for (o, d, _), nz in zip(host, noise):
        x = fwd(o.cuda(), d.cuda(), nz).cpu().numpy()
        pre.append((torch.from_numpy(sc.density(x)).cuda(), torch.from_numpy(sc.color(x)).cuda()))
    dev = [tuple(a.cuda() for a in b) for b in host]
    n0 = None

    def timed(fn, steps):
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for s in range(steps):
            fn(s)
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / steps

    def dev_step(s):
        flush.add_(1.0)
        k = s % len(dev)
        fwd(dev[k][0], dev[k][1], noise[k], *pre[k])

    out_pin = torch.zeros(BATCH, 3).pin_memory()

    def e2e_step(s):
        flush.add_(1.0)
        k = s % len(host)
        o, d = host[k][0].cuda(non_blocking=True), host[k][1].cuda(non_blocking=True)
        out_pin.copy_(fwd(o, d, noise[k], *pre[k]), non_blocking=True)

    for s in range(args.warmup):
        dev_step(s); e2e_step(s)
    flush_ms = timed(lambda s: flush.add_(1.0), args.steps)
    clocks = ClockSampler(0); clocks.start()
    n0 = _lib.launch_count
    ms = timed(dev_step, args.steps) - flush_ms
    launches = _lib.launch_count - n0
    clk = clocks.stop()
    e2e_ms = timed(e2e_step, args.steps) - flush_ms
    line = {"metric": "train_rays_per_s", "value": BATCH / (ms * 1e-3), "unit": "rays/s", "n_gpus": 1, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": arm_config(args, w, 1),
            "run": {"path": "vren operator surface (rnb_ray_aabb_intersect, rnb_march_train_*, rnb_composite_train_fw)"},
            "clocks": clk, "gpu_launches": launches,
            "e2e": {"value": BATCH / (e2e_ms * 1e-3), "unit": "rays/s", "h2d_bytes_per_step": 2 * BATCH * 12,
                    "d2h_bytes_per_step": BATCH * 12, "ms_per_step": e2e_ms}}
    if not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(args)
    print(json.dumps(line))


def run_render(args):
    """Test-time frames (ml_render(test_time=True), ml_rendering.py:81-155): one step = one full frame of the workload's shape,
    image rows interleaved over the ranks (parallel.frame_rows), random-init parameters (no early termination: worst case).
    metric render_rays_per_s = H * W / frame time (ms_per_step is the frame time: BASELINE.json's "800^2 render ms")."""
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    syn, w = workload(args.workload)
    K, H, W = w["K"], w["H"], w["W"]
    if args.impl == "reference":
        if rank != 0:
            return
        # the oracle's frame loop (oracle/render_cpu.render_test per sub-NeRF + gate blend) on a bounded sample of the frame's rays
        from oracle import render_cpu as R
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        n = max(512, 4096 // K)          # a frame's worth of K alive-ray loops on the CPU takes minutes: bounded sample
        cam = syn.make_cameras(100, w["radius"], H, W, w["fx"])
        o, d = syn.full_image_rays(cam, 3)
        sel = np.random.RandomState(0).choice(len(o), n, replace=False)
        o, d = np.ascontiguousarray(o[sel]), np.ascontiguousarray(d[sel])
        P = R.FieldParams(w["scale"], K, seed=0)
        bfs = scene_bitfields(syn, w)

        def frame():
            with torch.no_grad():
                g = R.gate(P, torch.from_numpy(o), torch.from_numpy(d))[0].numpy() if K > 1 else np.ones((n, 1), np.float32)
                rgb = np.zeros((n, 3), np.float32)
                for i in range(K):
                    rgb += g[:, i:i + 1] * R.render_test(P, o, d, bfs[i], i, 1, 0.0)["rgb"]
            return rgb
        for _ in range(args.warmup):
            frame()
        ts = []
        for _ in range(max(args.steps, 1)):
            t0 = time.perf_counter(); frame(); ts.append(time.perf_counter() - t0)
        med = float(np.median(ts))
        cb = {"value": n / med, "unit": "rays/s", "cores": cores, "kind": "port",
              "sample": f"{n} of {H * W} rays of the frame per step (fixed), K = {K} alive-ray loops + gate blend; median of {len(ts)}"}
        print(json.dumps({"impl": "reference", "metric": "render_rays_per_s", "value": n / med, "unit": "rays/s", "n_gpus": args.gpus,
                          "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * med, "higher_is_better": True, "scaling": "strong",
                          "vs_baseline": None, "dtype": "f32 (fp16 rounding points)", "data": "synthetic", "config": arm_config(args, w, args.gpus),
                          "cpu_baseline": cb, "e2e": {"value": n / med, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from radnerf_b200 import _lib, parallel
    from radnerf_b200.trainer import Trainer
    _lib.lib()
    tr = Trainer(scale=w["scale"], K=K, freeze_occupancy=True, device=f"cuda:{local}", engine=True, batch_rays=8192)
    tr.set_occupancy(scene_bitfields(syn, w))
    cam = syn.make_cameras(100, w["radius"], H, W, w["fx"])
    o, d = syn.full_image_rays(cam, 3)
    rows = parallel.frame_rows(H)
    sel = (rows[:, None] * W + torch.arange(W)[None, :]).reshape(-1)
    ho = torch.from_numpy(o)[sel].contiguous().pin_memory(); hd = torch.from_numpy(d)[sel].contiguous().pin_memory()
    do, dd = ho.cuda(), hd.cuda()
    out_pin = torch.zeros(len(sel), 3).pin_memory()
    spr = int(os.environ.get("RNB_RENDER_SPR", "32"))
    last = {}

    def frame(src_o, src_d, to_host):
        if to_host:
            src_o, src_d = src_o.cuda(non_blocking=True), src_d.cuda(non_blocking=True)
        res = tr.engine.render(src_o, src_d, samples_per_ray=spr)
        last.update(samples=res["total_samples"], rounds=res["rounds"])
        if world > 1:
            full = parallel.gather_frame(torch.cat([res["rgb"], res["opacity"][:, None], res["depth"]], 1), H, W)
            assert full.shape[0] == H * W
        if to_host:
            out_pin.copy_(res["rgb"], non_blocking=True)

    def timed(fn, steps):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier(); torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record(); torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / steps], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)
    for _ in range(args.warmup):
        frame(do, dd, False)
    clocks = ClockSampler(local); clocks.start()
    n0 = _lib.launch_count
    ms = timed(lambda: frame(do, dd, False), args.steps)
    launches = _lib.launch_count - n0
    clk = clocks.stop()
    e2e_ms = timed(lambda: frame(ho, hd, True), args.steps)
    smp = torch.tensor([float(last["samples"])], device="cuda")
    if world > 1:
        dist.all_reduce(smp)
    if rank == 0:
        hbm, tflops, which = measured_peaks()
        ns = float(smp)
        line = {"metric": "render_rays_per_s", "value": H * W / (ms * 1e-3), "unit": "rays/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f16 (fp32 accumulate; fp32 march/composite)", "data": "synthetic", "config": arm_config(args, w, world),
                "run": {"frame": f"{W}x{H}", "frame_ms": ms, "samples": int(ns), "rounds": int(last["rounds"]), "samples_per_round": spr,
                        "sharding": "interleaved rows; all_gather of (3+1+K) floats per ray" if world > 1 else "single GPU",
                        "l2": "working set (rays, samples of a round) larger than L2: no flush"},
                "clocks": clk, "gpu_launches": launches,
                "e2e": {"value": H * W / (e2e_ms * 1e-3), "unit": "rays/s", "h2d_bytes_per_step": int(2 * len(sel) * 12),
                        "d2h_bytes_per_step": int(len(sel) * 12), "ms_per_step": e2e_ms},
                "roofline": {"kernel": "frame (inf_march + field_fwd_tc_kernel<0> + inf_composite, all rounds)", "bound": "tensor",
                             "achieved": 18944 * ns / (ms * 1e-3) / 1e12, "peak": tflops, "unit": "TFLOP/s",
                             "frac": 18944 * ns / (ms * 1e-3) / 1e12 / tflops, "traffic": None, "peak_source": which,
                             "hbm_frac": (34 + 24 + 32) * ns / (ms * 1e-3) / 1e9 / hbm,
                             "note": "bounded by scattered L2 gathers (128 per sample): %.0f G gathers/s against 282 G/s from L2 "
                                     "(profiles/r02_microbench_l2.jsonl); coarse levels hit L1" % (128 * ns / (ms * 1e-3) / 1e9)}}
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args)
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def _json_lines(cmd, timeout):
    import subprocess as sp
    out = sp.run(cmd, capture_output=True, text=True, timeout=timeout)
    rows = []
    for ln in out.stdout.splitlines():
        if ln.startswith("{"):
            try:
                rows.append(json.loads(ln))
            except ValueError:
                pass
    return rows, out.stderr[-300:]


def cpu_baseline(args):
    """The reference arm (oracle port of the reference's math, all host cores) on the same fixed 8192-ray batch:
    3 warm-up steps + median of 5.  Runs in a subprocess so that this (product) process never loads oracle/."""
    rows, err = _json_lines([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", "5", "--warmup", "3",
                             "--workload", args.workload], 900)
    if rows:
        return rows[-1]["cpu_baseline"]
    return {"value": None, "unit": "rays/s", "cores": os.cpu_count(), "kind": "port", "sample": "failed: " + err}


def gpu_anchor():
    """The reference's OWN kernels (oracle/_ref/vren_ref, compiled unmodified for sm_100a) next to this library's operators
    on the training shape (tests/anchor_vren_ref.py; subprocess: the product process never loads the reference)."""
    import glob
    if not glob.glob(os.path.join(ROOT, "oracle", "_ref", "vren_ref*.so")):
        return {"unavailable": "oracle/_ref/vren_ref*.so not built (needs /root/reference at build time)"}
    try:
        rows, err = _json_lines([sys.executable, os.path.join(ROOT, "tests", "anchor_vren_ref.py"), "--training-shape-only"], 600)
    except Exception as e:  # noqa: BLE001
        return {"unavailable": str(e)[:200]}
    if not rows:
        return {"unavailable": "anchor failed: " + err}
    return {r["op"]: {"ref_ms": r["reference_kernels_ms"], "ours_ms": r["ours_ms"]} for r in rows}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=32)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="radnerf_tat_k2")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: 8192 rays per GPU (default); strong: a global batch of 8192 rays split over the ranks")
    ap.add_argument("--graph", action="store_true", help="replay the step from captured CUDA graphs (single process) instead of eager launches")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-march-ahead", action="store_true", help="march every batch inside its own step (single stream)")
    ap.add_argument("--python-surface", action="store_true", help="time the drop-in Python operator surface instead of the fused engine")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if workload(args.workload)[1].get("infer"):
        run_render(args)
    elif args.impl == "reference":
        run_reference(args)
    elif workload(args.workload)[1].get("forward_only"):
        run_product_fwd_only(args)
    else:
        run_product(args)


if __name__ == "__main__":
    main()
