#!/usr/bin/env python
"""Per-source-line view of one kernel in an ncu report (captured with --import-source on, library built with -lineinfo):
stall samples and executed warp instructions per CUDA source line, from `ncu --page source --csv` joined with
`nvdisasm -g` of the library's cubin.  Usage: python scripts/ncu_lines.py gpurun_out/x.ncu-rep field_bwd [cubin-name] [top]"""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    rep, kern = sys.argv[1], sys.argv[2]
    cub = sys.argv[3] if len(sys.argv) > 3 else "field_tc"
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "rad-nerf_b200", "lib", "libradnerf_b200.so")], cwd=tmp,
                   capture_output=True)
    dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, f"{cub}.sm_100a.cubin")], capture_output=True, text=True).stdout
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kern}"], capture_output=True,
                         text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    k0 = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    kname = rows[k0 - 1][1] if k0 else ""
    hdr, data = rows[k0], []
    for r in rows[k0 + 1:]:
        if not r or r[0] in ("Kernel Name", "Address"):
            break
        data.append(r)
    ia, isamp, iex, isrc = hdr.index("Address"), hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Source")
    stalls = [(i, h[6:]) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    # the kernel's function in the disassembly: match by instruction count
    fns, fn, line = collections.OrderedDict(), None, None
    for l in dis.splitlines():
        m = re.match(r"\s*\.text\.(\S+):", l)
        if m:
            fn = m.group(1); fns[fn] = {}; continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            line = (m.group(1).split("/")[-1], int(m.group(2))); continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
        if m and fn:
            fns[fn][int(m.group(1), 16)] = line
    cand = [f for f in fns if kern.split("|")[0] in f and len(fns[f]) == len(data)]
    if not cand:
        print("no function with", len(data), "instructions matching", kern, [(f, len(v)) for f, v in fns.items() if kern in f]); return
    a2l = fns[cand[0]]
    base = int(data[0][ia], 16)
    samp, ex, why = collections.Counter(), collections.Counter(), collections.defaultdict(collections.Counter)
    for r in data:
        ln = a2l.get(int(r[ia], 16) - base)
        samp[ln] += int(r[isamp] or 0); ex[ln] += int(r[iex] or 0)
        for i, h in stalls:
            why[ln][h] += int(r[i] or 0)
    ts, te = sum(samp.values()), sum(ex.values())
    print(f"{kname[:80]}: {len(data)} SASS instructions, {ts} samples, {te} warp instructions executed")
    allwhy = collections.Counter()
    for c in why.values():
        allwhy.update(c)
    print("stalls:", ", ".join(f"{k} {100 * v / max(ts, 1):.0f}%" for k, v in allwhy.most_common(8)))
    print(f"{'line':28s} {'samples':>8s} {'%':>6s} {'instr':>10s} {'%':>6s}  top stalls")
    for ln, c in samp.most_common(top):
        w = ", ".join(f"{k} {v}" for k, v in why[ln].most_common(2))
        print(f"{str(ln):28s} {c:8d} {100 * c / max(ts, 1):6.1f} {ex[ln]:10d} {100 * ex[ln] / max(te, 1):6.1f}  {w}")


if __name__ == "__main__":
    main()
