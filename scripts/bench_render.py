"""Test-time frame rendering: the fused renderer (Engine.render -> rnb_render) next to the reference-shaped
python loop (ml_render(test_time=True) over the vren/tcnn drop-ins) on one synthetic frame per workload.
Prints one JSON line per workload; device time by CUDA events, 3 timed repetitions after 1 warm-up.
Usage: python scripts/bench_render.py [--fitted] [workload ...]
--fitted: parameters and occupancy grids after 2000 training steps on the analytic scene (scripts/fit_scene.py) instead of the
random initialisation, so that rays terminate on surfaces (SURVEY 8d asks for both states); RNB_RENDER_SPR: samples per ray and round."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import scene_bitfields, workload  # noqa: E402


def timed(fn, reps):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out


def main():
    fitted = "--fitted" in sys.argv
    names = [a for a in sys.argv[1:] if not a.startswith("--")] or ["ngp_blender_k1", "radnerf_tat_k2", "radnerf_tat_k8_infer"]
    from radnerf_b200.trainer import Trainer
    from radnerf_b200.models.ml_rendering import ml_render
    from radnerf_b200.models.rendering import render
    for name in names:
        syn, w = workload(name)
        if fitted:
            sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
            from fit_scene import fit
            info, tr = fit(steps=2000, K=w["K"], return_trainer=True)
        else:
            tr = Trainer(scale=w["scale"], K=w["K"], freeze_occupancy=True, device="cuda:0", engine=True, batch_rays=8192)
            tr.set_occupancy(scene_bitfields(syn, w))
        cam = syn.make_cameras(100, w["radius"], w["H"], w["W"], w["fx"])
        o, d = syn.full_image_rays(cam, 3)
        o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
        eng = tr.engine
        spr = int(os.environ.get("RNB_RENDER_SPR", "32"))
        eng.render(o, d, samples_per_ray=spr)
        ms, res = timed(lambda: eng.render(o, d, samples_per_ray=spr), 3)
        line = {"workload": name, "params": f"fitted ({info['psnr_test_mean']} dB after 2000 steps)" if fitted else "random init",
                "rays": int(o.shape[0]), "K": w["K"], "samples_per_round": spr, "fused_ms": round(ms, 3),
                "samples": int(res["total_samples"]), "rounds": int(res["rounds"]),
                "Msamples_per_s": round(res["total_samples"] / ms / 1e3, 1)}
        if os.environ.get("RNB_SKIP_PY") != "1":
            esf = tr.exp_step_factor

            def py():
                with torch.no_grad():
                    if w["K"] == 1:
                        return render(tr.model, o, d, test_time=True, exp_step_factor=esf)
                    return ml_render(tr.model, tr.gate, o, d, d, test_time=True, exp_step_factor=esf)
            py()
            ms_py, ref = timed(py, 1)
            line["python_surface_ms"] = round(ms_py, 3)
            line["max_abs_rgb_diff"] = float((ref["rgb"].float() - res["rgb"]).abs().max())
        print(json.dumps(line), flush=True)
        del tr, eng
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
