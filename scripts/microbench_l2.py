#!/usr/bin/env python
"""Roofline reference for the fused field kernels: the rate at which a B200 sustains SCATTERED L2 accesses of the kind they
issue (rnb_diag_l2_rate, csrc/microbench.cu): 4/8-byte gathers from an fp16 hash table of 5.7 M entries (22 MB) and
8/16-byte fp32 reductions into the gradient table (46 MB), both L2-resident.  One JSON line per variant:
accesses/s, per-SM cycles per access at the measured SM clock, and the 32-byte sectors/s that corresponds to.
Usage: python scripts/microbench_l2.py | tee profiles/r02_microbench_l2.jsonl"""
import ctypes
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import radnerf_b200  # noqa: F401,E402
from radnerf_b200 import _lib  # noqa: E402
from radnerf_b200._lib import call, i32, i64, stream_arg  # noqa: E402

ENTRIES = 5722520          # table entries at scene scale 0.5 (tests/test_level_table_cpu.py)


def run(kind, entries, iters, lanes, blocks, reps=5):
    nbytes = entries * (8 if kind <= 2 else 4)
    table = torch.zeros(nbytes // 4 + 16, dtype=torch.float32, device="cuda")
    ops = ctypes.c_int64(0)
    best = None
    for _ in range(reps + 1):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        call("rnb_diag_l2_rate", i32(kind), table, i64(entries), i32(iters), i32(lanes), i32(blocks), ctypes.byref(ops),
             stream_arg(table.device))
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        best = ms if best is None else min(best, ms)
    return ops.value, best


def main():
    _lib.lib()
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    try:
        import pynvml
        pynvml.nvmlInit()
        mhz = pynvml.nvmlDeviceGetMaxClockInfo(pynvml.nvmlDeviceGetHandleByIndex(0), pynvml.NVML_CLOCK_SM)
    except Exception:  # noqa: BLE001
        mhz = 1965
    names = {0: "red.global.add.f32", 1: "red.global.add.v2.f32", 2: "red.global.add.v4.f32 (aligned pair)",
             3: "ld.global.nc.b32", 4: "ld.global.nc.b64"}
    for kind in (1, 2, 0, 3, 4):
        for entries in (ENTRIES, 1 << 15):
            for lanes in ((32, 8) if kind <= 2 else (32,)):
                for bps in (8, 2):
                    ops, ms = run(kind, entries, 64, lanes, sms * bps)
                    rate = ops / (ms * 1e-3)
                    print(json.dumps({"access": names[kind], "table_entries": entries,
                                      "table_MB": round(entries * (8 if kind <= 2 else 4) / 1e6, 1), "active_lanes": lanes,
                                      "warps_per_sm": bps * 8, "accesses": ops, "ms": round(ms, 4), "G_accesses_per_s": round(rate / 1e9, 2),
                                      "sm_cycles_per_access": round(sms * mhz * 1e6 / rate, 3), "sm_mhz": mhz}), flush=True)


if __name__ == "__main__":
    main()
