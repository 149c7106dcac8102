"""Full-frame test-time rendering sharded over the ranks by interleaved image rows (BASELINE.json configs[4]:
1920x1080, K = 8, 8 GPUs).  Launch: python -m torch.distributed.run --nproc-per-node N scripts/bench_render_dist.py [workload]
Rank 0 prints one JSON line: frame ms = max over ranks of the device time (render + gather), 3 repetitions after a warm-up."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import scene_bitfields, workload  # noqa: E402


def main():
    fitted = any(a.startswith("--fitted") for a in sys.argv)
    names = [a for a in sys.argv[1:] if not a.startswith("--")]
    name = names[0] if names else "radnerf_tat_k8_infer"
    spr = int(os.environ.get("RNB_RENDER_SPR", "32"))
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from radnerf_b200 import parallel
    from radnerf_b200.trainer import Trainer
    syn, w = workload(name)
    tr = Trainer(scale=w["scale"], K=w["K"], freeze_occupancy=True, device=f"cuda:{local}", engine=True, batch_rays=8192)
    tr.set_occupancy(scene_bitfields(syn, w))
    if fitted:
        # SURVEY 8d "fitted" state: 2000 training steps on the analytic scene.  The fit is a SINGLE-PROCESS job (its engine must
        # not be a data-parallel one: a rank-0-only fit with a sharded engine waits for collectives nobody joins), so it has
        # to be produced beforehand: python scripts/fit_scene.py --K 8 --save fitted_k8.pt, then pass --fitted=fitted_k8.pt
        path = [a.split("=", 1)[1] for a in sys.argv[1:] if a.startswith("--fitted=")]
        assert path, "--fitted=<file written by scripts/fit_scene.py --save>"
        st = torch.load(path[0], map_location=f"cuda:{local}")
        eng = tr.engine
        with torch.no_grad():
            eng.flat.copy_(st["flat"])
            for a, b in zip(eng.grids() + eng.bitfields(), st["grids"] + st["bitfields"]):
                a.copy_(b)
        for _, m in eng._mods:
            m._shadow.version = None
        eng.sync_params()
    cam = syn.make_cameras(100, w["radius"], w["H"], w["W"], w["fx"])
    H, W = w["H"], w["W"]
    o, d = syn.full_image_rays(cam, 3)
    rows = parallel.frame_rows(H)
    sel = (rows[:, None] * W + torch.arange(W)[None, :]).reshape(-1)
    o = torch.from_numpy(o)[sel].contiguous().cuda(); d = torch.from_numpy(d)[sel].contiguous().cuda()

    def frame():
        res = tr.engine.render(o, d, samples_per_ray=spr)
        out = torch.cat([res["rgb"], res["opacity"][:, None], res["depth"]], 1)      # (3 + 1 + K) floats per ray
        return parallel.gather_frame(out, H, W), res

    frame()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        full, res = frame()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / 3], device="cuda")
    smp = torch.tensor([float(res["total_samples"])], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(smp)
    if rank == 0:
        assert full.shape == (H * W, 4 + w["K"]) and bool(torch.isfinite(full).all())
        print(json.dumps({"workload": name, "params": "fitted (2000 steps)" if fitted else "random init", "samples_per_round": spr, "n_gpus": world, "rays": H * W, "K": w["K"], "frame_ms": round(float(ms), 3),
                          "samples": int(smp), "Msamples_per_s": round(float(smp) / float(ms) / 1e3, 1),
                          "sharding": "interleaved rows, all_gather of (3+1+K) floats per ray"}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
