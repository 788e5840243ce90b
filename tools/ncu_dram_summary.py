"""Summarise an `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv` log:
per launch of this library's kernels: ms, DRAM GB read / written, GB/s and fraction of the measured HBM peak.
usage: python tools/ncu_dram_summary.py log.csv [peak_GBps] [skip_first_n_launches]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
peak = float(sys.argv[2]) if len(sys.argv) > 2 else 6547.0
skip = int(sys.argv[3]) if len(sys.argv) > 3 else 0
hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
h = rows[hi]
ik, im, iv, iu, iid = (h.index(x) for x in ("Kernel Name", "Metric Name", "Metric Value", "Metric Unit", "ID"))
per = collections.OrderedDict()
for r in rows[hi + 1:]:
    if len(r) > iv:
        per.setdefault(r[iid], {"name": r[ik]})[r[im]] = (float(r[iv].replace(",", "")), r[iu])
m = {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1, "byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
print(f"{'kernel':74s} {'ms':>8s} {'rd GB':>8s} {'wr GB':>8s} {'GB/s':>8s} {'of peak':>8s}")
n = 0
for d in per.values():
    if "cxl::" not in d["name"]:
        continue
    n += 1
    if n <= skip:
        continue
    t, rd, wr = (d[k][0] * m[d[k][1]] for k in ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum"))
    bw = (rd + wr) / t / 1e9
    print(f"{d['name'][:74]:74s} {t * 1e3:8.3f} {rd / 1e9:8.3f} {wr / 1e9:8.3f} {bw:8.0f} {bw / peak:8.2f}")
