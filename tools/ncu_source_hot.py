"""Aggregate the --page source CSV of an ncu report: instructions executed and shared-memory wavefronts per opcode,
and the hottest instructions by stall samples.  usage: ncu_source_hot.py report.ncu-rep [top]"""
import collections
import csv
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hdr_i]
col = {n: i for i, n in enumerate(hdr)}
data = rows[hdr_i + 1:]
tot_inst = sum(float(r[col["Instructions Executed"]] or 0) for r in data)
tot_wf = sum(float(r[col["L1 Wavefronts Shared"]] or 0) for r in data)
tot_smp = sum(float(r[col["# Samples"]] or 0) for r in data)
by = collections.defaultdict(lambda: [0.0, 0.0, 0.0])
for r in data:
    op = r[col["Source"]].split()[0] if r[col["Source"]].strip() else "?"
    if op.startswith("@"):
        op = r[col["Source"]].split()[1]
    a = by[op]
    a[0] += float(r[col["Instructions Executed"]] or 0)
    a[1] += float(r[col["L1 Wavefronts Shared"]] or 0)
    a[2] += float(r[col["# Samples"]] or 0)
print(f"total warp-instructions {tot_inst:.4g}  shared wavefronts {tot_wf:.4g}  samples {tot_smp:.0f}")
print(f"{'opcode':28s} {'inst %':>8s} {'wavefront %':>12s} {'samples %':>10s}")
for op, a in sorted(by.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{op:28s} {100 * a[0] / tot_inst:8.2f} {100 * a[1] / max(tot_wf, 1):12.2f} {100 * a[2] / max(tot_smp, 1):10.2f}")
print("\nhottest instructions by samples:")
for r in sorted(data, key=lambda r: -float(r[col["# Samples"]] or 0))[:top]:
    print(f"{r[col['Address']][-6:]} {float(r[col['# Samples']] or 0) / max(tot_smp, 1) * 100:6.2f}%  inst {float(r[col['Instructions Executed']] or 0):.3g}  wf {float(r[col['L1 Wavefronts Shared']] or 0):.3g}  {r[col['Source']][:90]}")
