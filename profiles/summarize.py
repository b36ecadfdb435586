#!/usr/bin/env python
"""Turn the raw ncu outputs of a round (gpurun_out/r1_launches.csv from the time-only pass, and
`ncu -i r1_full.ncu-rep --page raw --csv > /tmp/r1_full_raw.csv` from the --set full capture) into the committed
summaries: r1_launches_one_step.csv, r1_full_metrics.csv, r1_traffic.json."""
import collections
import csv
import json
import re
import sys

launch_csv = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/r1_launches.csv"
raw_csv = sys.argv[2] if len(sys.argv) > 2 else "/tmp/r1_full_raw.csv"
RND = sys.argv[3] if len(sys.argv) > 3 else "r1"  # prefix of the files written under profiles/


def short(name):
    name = re.sub(r"\(.*", "", name)
    return re.sub(r".*::", "", name)


rows = list(csv.reader(open(launch_csv)))
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
hdr = rows[hi]
kn, mv, gs, bs = (hdr.index(k) for k in ("Kernel Name", "Metric Value", "Grid Size", "Block Size"))
seq = [(short(r[kn]), r[gs], r[bs], float(r[mv].replace(",", "")) / 1e3) for r in rows[hi + 1:] if len(r) > mv]
idx = [i for i, s in enumerate(seq) if s[0].startswith("k_quant<0>")]
# which iteration of the capture: default the last complete one; bench.py's later loops time the step serially and the
# dominant kernel alone, so the round-2 summary passes 6 (a CUDA-graph replay of the two-stream step)
STEP = int(sys.argv[4]) if len(sys.argv) > 4 else len(idx) - 2
step = [s for s in seq[idx[STEP]:idx[STEP + 1]] if s[0].startswith("k_")]
tot = sum(s[3] for s in step)
with open(f"profiles/{RND}_launches_one_step.csv", "w") as f:
    f.write("# one iteration of bench.py (768x512 HOP) under: ncu --metrics gpu__time_duration.sum --clock-control none\n")
    f.write("# per-launch times are cold-cache and serialised (both streams): compare SHARES, not absolutes\n")
    f.write("order,kernel,grid,block,time_us,share_pct\n")
    for i, s in enumerate(step):
        f.write(f'{i},{s[0]},"{s[1]}","{s[2]}",{s[3]:.2f},{100 * s[3] / tot:.1f}\n')
    f.write(f",TOTAL,,,{tot:.2f},100.0\n")
agg = collections.OrderedDict()
for s in step:
    a = agg.setdefault(s[0], [0, 0.0])
    a[0] += 1
    a[1] += s[3]
print("one step:", len(step), "launches, sum", round(tot, 1), "us")
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"  {k:<28} x{n}  {t:7.2f} us  {100 * t / tot:5.1f}%")

rows = list(csv.reader(open(raw_csv)))
hdr, units = rows[0], rows[1]
want = collections.OrderedDict([
    ("gpu__time_duration.sum", "time_us"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
    ("launch__registers_per_thread", "regs"), ("launch__occupancy_limit_registers", "occ_lim_regs"),
    ("launch__occupancy_limit_shared_mem", "occ_lim_smem"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved_occ_pct"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_active_pct"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_throughput_pct"),
    ("smsp__inst_executed.sum", "warp_inst"), ("dram__bytes_read.sum", "dram_read_bytes"),
    ("dram__bytes_write.sum", "dram_write_bytes"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem_bank_conflicts"),
    ("smsp__pcsamp_warps_issue_stalled_short_scoreboard", "stall_short_scoreboard_samples"),
    ("smsp__pcsamp_warps_issue_stalled_long_scoreboard", "stall_long_scoreboard_samples"),
    ("smsp__pcsamp_warps_issue_stalled_mio_throttle", "stall_mio_throttle_samples"),
    ("smsp__pcsamp_warps_issue_stalled_math_pipe_throttle", "stall_math_pipe_samples"),
    ("smsp__pcsamp_warps_issue_stalled_barrier", "stall_barrier_samples"),
    ("smsp__pcsamp_warps_issue_stalled_selected", "issued_samples"),
    ("launch__waves_per_multiprocessor", "waves_per_sm")])
cols = {k: hdr.index(k) for k in want if k in hdr}
kn = hdr.index("Kernel Name")


def num(x):
    try:
        return float(x.replace(",", ""))
    except Exception:
        return None


out = []
for r in rows[2:]:
    d = {"kernel": short(r[kn])}
    for k, i in cols.items():
        v, u = num(r[i]), units[i]
        if k.startswith("dram__bytes"):
            v = None if v is None else v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
        if k == "gpu__time_duration.sum":
            v = None if v is None else v * {"ns": 1e-3, "us": 1, "ms": 1e3}.get(u, 1)
        d[want[k]] = v
    out.append(d)
with open(f"profiles/{RND}_full_metrics.csv", "w") as f:
    keys = ["kernel"] + [want[k] for k in cols]
    f.write("# ncu --set full --clock-control none --import-source on, one iteration of bench.py (768x512 HOP); "
            "units: us / bytes / counts / pc-sampling samples\n")
    f.write(",".join(keys) + "\n")
    for d in out:
        f.write(",".join("" if d.get(k) is None else (d[k] if isinstance(d[k], str) else f"{d[k]:.6g}") for k in keys) + "\n")
traffic = {}
for d in out:
    kk = re.sub(r"<.*", "", d["kernel"])
    t = traffic.setdefault(kk, {"dram_bytes": 0, "launches": 0})
    t["dram_bytes"] += (d["dram_read_bytes"] or 0) + (d["dram_write_bytes"] or 0)
    t["launches"] += 1
for k in traffic:
    traffic[k]["dram_bytes"] /= traffic[k]["launches"]
json.dump(traffic, open(f"profiles/{RND}_traffic.json", "w"), indent=1)
for d in out:
    print(f"{d['kernel'][:24]:<24} {d['time_us']:6.1f} us regs {d['regs']:.0f} occ {d['achieved_occ_pct']:.1f} "
          f"issue {d['issue_active_pct']:.1f} inst {d['warp_inst'] / 1e6:.2f}M dram {((d['dram_read_bytes'] or 0) / 1e6):.2f}MB")
