"""Per-kernel totals and shares of an ncu launch list:
    ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file launches.csv <command>
    python tools/ncu_launch_summary.py launches.csv "<header line>"
Per-launch times under ncu are serialised and cold-cache: compare SHARES with the CUDA-event numbers, not absolutes."""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 10 and r[0].isdigit()]
tot = collections.defaultdict(float)
cnt = collections.Counter()
for r in rows:
    if r[-3] != "gpu__time_duration.sum":
        continue
    unit, val = r[-2], float(r[-1].replace(",", ""))
    ms = val * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(unit, 1e-6)
    name = r[4]
    tot[name] += ms
    cnt[name] += 1
total = sum(tot.values())
if len(sys.argv) > 2:
    print(sys.argv[2])
print("per-launch times are serialised / cold-cache: compare shares, not absolutes, with the CUDA-event numbers")
print(f"total {total:.1f} ms over {sum(cnt.values())} launches\n")
for name, ms in sorted(tot.items(), key=lambda t: -t[1])[:25]:
    print(f"{ms:12.1f} ms {100 * ms / total:5.1f}%  n={cnt[name]:5d}  avg {ms / cnt[name]:10.3f} ms  {name[:110]}")
