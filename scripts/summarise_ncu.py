"""Turns the scratch ncu outputs of scripts/profile_round.sh into the tracked summaries under profiles/:
   <tag>_launches.md   per-kernel launch count, mean duration and share of the profiled steps (from the launch-list csv)
   <tag>_kernels.md    per-kernel metrics of the full capture (duration, DRAM bytes, unit throughputs, occupancy, stalls)
   <tag>_traffic.json  DRAM bytes per launch of each captured kernel (bench.py's roofline.traffic)
Usage: python scripts/summarise_ncu.py <tag-in-gpurun_out> <tag-in-profiles>"""
import collections
import csv
import io
import json
import subprocess
import sys

src, dst = sys.argv[1], sys.argv[2]

# ---- launch list ----
lines = [l for l in open(f"gpurun_out/{src}_launches.csv") if not l.startswith("==")]
agg = collections.OrderedDict()
for row in csv.DictReader(io.StringIO("".join(lines))):
    k = row["Kernel Name"].split("(")[0].replace("void ", "")
    v = float(row["Metric Value"].replace(",", ""))
    u = row["Metric Unit"]
    v = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v
    agg.setdefault(k, []).append(v)
tot = sum(sum(v) for v in agg.values())
with open(f"profiles/{dst}_launches.md", "w") as f:
    f.write(f"# ncu launch list ({src}): `ncu --metrics gpu__time_duration.sum --clock-control none` over `bench.py --steps 2 --warmup 3`\n\n")
    f.write("Per-launch times under ncu are cold-cache and serialised: the SHARE column is what is comparable with bench.py's stage table.\n")
    f.write("The first profiled step includes the every-16-steps occupancy update (density_cells_tc, ema, mean, packbits).\n\n")
    f.write(f"total profiled GPU time: {tot:.1f} us\n\n| kernel | launches | mean us | share |\n|---|---|---|---|\n")
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        f.write(f"| `{k[:90]}` | {len(v)} | {sum(v) / len(v):.1f} | {sum(v) / tot * 100:.1f}% |\n")

# ---- full capture ----
raw = subprocess.run(["ncu", "-i", f"gpurun_out/{src}_full.ncu-rep", "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
idx = {h: i for i, h in enumerate(hdr)}
want = [("gpu__time_duration.sum", "duration"), ("dram__bytes_read.sum", "DRAM read"), ("dram__bytes_write.sum", "DRAM write"),
        ("lts__t_sector_hit_rate.pct", "L2 hit %"), ("l1tex__t_sector_hit_rate.pct", "L1 hit %"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM %"), ("l1tex__throughput.avg.pct_of_peak_sustained_active", "L1TEX %"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 %"), ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %"),
        ("sm__pipe_tc_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe %"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue active %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %"), ("launch__registers_per_thread", "regs/thread"),
        ("smsp__inst_executed.sum", "warp instructions")]
seen = {}
traffic = {}


def to_bytes(v, u):
    v = float(v.replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)


with open(f"profiles/{dst}_kernels.md", "w") as f:
    f.write(f"# ncu --set full --clock-control none ({src}), one launch per kernel, `bench.py --steps 2 --warmup 3`\n\n")
    for r in data:
        name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "")
        if name in seen:
            continue
        seen[name] = 1
        f.write(f"## `{name}`\n\n| metric | value |\n|---|---|\n")
        for m, label in want:
            if m in idx:
                f.write(f"| {label} (`{m}`) | {r[idx[m]]} {units[idx[m]]} |\n")
        st = [(h[33:], float(r[i].replace(",", ""))) for h, i in idx.items()
              if h.startswith("smsp__pcsamp_warps_issue_stalled") and not h.endswith("_not_issued") and r[i] not in ("", "n/a")]
        t = sum(v for _, v in st) or 1.0
        f.write("| top stall reasons (pc sampling) | " + ", ".join(f"{h} {v / t * 100:.0f}%" for h, v in sorted(st, key=lambda x: -x[1])[:5]) + " |\n\n")
        if "dram__bytes_read.sum" in idx:
            traffic[name] = {"dram_bytes_per_launch": to_bytes(r[idx["dram__bytes_read.sum"]], units[idx["dram__bytes_read.sum"]]) +
                             to_bytes(r[idx["dram__bytes_write.sum"]], units[idx["dram__bytes_write.sum"]]),
                             "duration_us": float(r[idx["gpu__time_duration.sum"]].replace(",", ""))}
json.dump({"source": f"{src}_full.ncu-rep (ncu --set full, --clock-control none)", "kernels": traffic}, open(f"profiles/{dst}_traffic.json", "w"), indent=1)
print(open(f"profiles/{dst}_launches.md").read()[:3000])
