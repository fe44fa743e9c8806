"""Dev tool: copies one gpu_round.sh visit (gpurun_out/<tag>_*) into profiles/ as the round's judged summaries.
usage: python scripts/export_profiles.py <tag> [round-prefix]"""
import collections
import csv
import json
import shutil
import subprocess
import sys

tag = sys.argv[1]
rp = sys.argv[2] if len(sys.argv) > 2 else "r1"
g = f"gpurun_out/{tag}"
rep = f"{g}_panda_fkcc_mbm.ncu-rep"
subprocess.run(f"ncu -i {rep} --page details > profiles/{rp}_panda_fkcc_mbm_ncu_details.txt 2>&1", shell=True, check=False)
subprocess.run(f"bash scripts/ncu_summary.sh {rep} 60 > profiles/{rp}_panda_fkcc_mbm_lines.txt 2>&1", shell=True, check=False)
shutil.copy(f"{g}_bench_launches.csv", f"profiles/{rp}_bench_launches.csv")
shutil.copy(f"{g}_bench.json", f"profiles/{rp}_bench_line_n1.json")
shutil.copy(f"{g}_bench_ref.json", f"profiles/{rp}_bench_reference_line.json")
rows = list(csv.reader(open(f"{g}_bench_launches.csv")))
hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
h, data = rows[hi], rows[hi + 1:]
kn, mv, mu = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
tot, cnt = collections.Counter(), collections.Counter()
for r in data:
    if len(r) <= mv:
        continue
    v, u = float(r[mv].replace(",", "")), r[mu]
    tot[r[kn]] += v / 1e6 if u in ("ns", "nsecond") else v / 1e3 if u in ("us", "usecond") else v
    cnt[r[kn]] += 1
T = sum(tot.values())
out = ["ncu --metrics gpu__time_duration.sum --clock-control none -c 400: python bench.py --steps 2 --warmup 3 --configs 20000000",
       "(cold-cache, serialised per-launch times; share of the whole process: side metrics, work-count pass and the FP32-peak probe included)"]
for k, v in tot.most_common(12):
    out.append(f"{v:10.3f} ms  {100 * v / T:6.2f}%  x{cnt[k]:<3d} {k[:110]}")
open(f"profiles/{rp}_bench_launches_summary.txt", "w").write("\n".join(out) + "\n")
raw = subprocess.run(f"ncu -i {rep} --page raw --csv", shell=True, capture_output=True, text=True).stdout
r = list(csv.reader(raw.splitlines()))
hh, uu, vv = r[0], r[1], r[2]


def metric(k):
    i = hh.index(k)
    return float(vv[i].replace(",", "")) * {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1}.get(uu[i], 1)


rd, wr = metric("dram__bytes_read.sum"), metric("dram__bytes_write.sum")
fma = metric("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active")
n = 8_000_000
json.dump({"kernel": "fkcc_kernel<PandaTrace,0>", "configs_in_capture": n, "dram_bytes_read": rd, "dram_bytes_write": wr,
           "bytes_per_config": (rd + wr) / n, "algorithmic_bytes_per_config": 28.125, "fma_pipe_pct": fma,
           "issue_slots_busy_pct": metric("smsp__issue_active.avg.pct_of_peak_sustained_active"),
           "threads_per_instruction": metric("smsp__thread_inst_executed_per_inst_executed.ratio"),
           "source": "ncu --set full --clock-control none, one launch of bench.py --configs 8000000 "
                     f"(profiles/{rp}_panda_fkcc_mbm_ncu_details.txt)"},
          open(f"profiles/{rp}_panda_fkcc_mbm_traffic.json", "w"), indent=1)
print("\n".join(out[:6]))
print("traffic B/config", (rd + wr) / n)
