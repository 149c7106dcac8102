"""bench.py's reference arm (`--impl reference`: the oracle port of the reference's math on the host cores) run end to end on a
tiny batch (RNB_BENCH_BATCH), checking the JSON contract of the task statement: metric / unit / impl / cpu_baseline / e2e keys,
a fixed batch named in `sample`, and the same `config` object the product arm prints (bench.arm_config)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(workload, batch):
    env = dict(os.environ, RNB_BENCH_BATCH=str(batch), OMP_NUM_THREADS="4")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "1",
                          "--workload", workload], capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-800:]
    lines = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1, out.stdout[-400:]
    return json.loads(lines[0])


@pytest.mark.parametrize("workload,metric", [("radnerf_tat_k2", "train_rays_per_s"), ("ngp_nsvf_k1_cpu_fwd", "train_rays_per_s"),
                                             ("ngp_blender_k1_infer", "render_rays_per_s")])
def test_reference_arm_contract(workload, metric):
    d = _run(workload, 256)
    assert d["impl"] == "reference" and d["metric"] == metric and d["unit"] == "rays/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["steps"] == 2 and d["warmup"] == 1 and d["vs_baseline"] is None
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and "rays" in cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["workload"] == workload and "model" not in d["config"]
    # the product arm prints the same config object for the same workload and rank count
    sys.path.insert(0, ROOT)
    import argparse
    import bench
    os.environ["RNB_BENCH_BATCH"] = "256"
    try:
        import importlib
        importlib.reload(bench)
        _, w = bench.workload(workload)
        assert bench.arm_config(argparse.Namespace(workload=workload), w, 1) == d["config"]
    finally:
        os.environ.pop("RNB_BENCH_BATCH", None)
        importlib.reload(bench)
