"""Print the hottest SASS instructions (by warp-stall samples) of an `ncu --page source --csv --print-source sass` dump."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = rows[1]
i_s = hdr.index("# Samples")
i_src = hdr.index("Source")
i_ex = hdr.index("Instructions Executed")
body = [r for r in rows[2:] if len(r) > i_s]


def f(x):
    try:
        return float(x)
    except ValueError:
        return 0.0


tot = sum(f(r[i_s]) for r in body)
print("total samples", tot, "instructions", len(body))
for idx, r in sorted(enumerate(body), key=lambda t: -f(t[1][i_s]))[:n]:
    print(f"{f(r[i_s]) / tot * 100:6.2f}%  #{idx:5d} exec={r[i_ex]:>10s}  {r[i_src].strip()[:110]}")
