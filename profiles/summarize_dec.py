#!/usr/bin/env python
"""gpurun_out/dec_launches.csv (ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum -k regex:k_dec
python scratch/dec_prof.py) -> profiles/r1_dec_launches.csv: the launches of ONE fixed-point decode (the last of three)."""
import collections
import csv
import re
import sys

src = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/dec_launches.csv"
rows = list(csv.reader(open(src)))
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
h = rows[hi]
kn, mn, mv, idc, gs, bs = (h.index(k) for k in ("Kernel Name", "Metric Name", "Metric Value", "ID", "Grid Size", "Block Size"))
d = collections.OrderedDict()
for r in rows[hi + 1:]:
    if len(r) > mv:
        d.setdefault(r[idc], {"k": r[kn], "g": r[gs], "b": r[bs]})[r[mn]] = float(r[mv].replace(",", ""))
ids = list(d)
n = len(ids) // 3
with open("profiles/r1_dec_launches.csv", "w") as f:
    f.write("# fixed-point decoder upsampling + synthesis (ccb_dec_int) on the kodim01-lmbda-0001 integers, 768x512:\n")
    f.write("# ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:k_dec "
            "python scratch/dec_prof.py\n# (third of three decodes; per-launch times are cold-cache and serialised)\n")
    f.write("order,kernel,grid,block,time_us,warp_inst\n")
    tot = 0.0
    for j, i in enumerate(ids[-n:]):
        x = d[i]
        m = re.search(r"(k_dec_[a-z0-9_]+(<[^>]*>)?)", x["k"])
        f.write(f'{j},{m.group(1) if m else x["k"]},"{x["g"]}","{x["b"]}",{x["gpu__time_duration.sum"] / 1e3:.2f},'
                f'{int(x["smsp__inst_executed.sum"])}\n')
        tot += x["gpu__time_duration.sum"] / 1e3
    f.write(f",TOTAL,,,{tot:.2f},\n")
print(open("profiles/r1_dec_launches.csv").read())
