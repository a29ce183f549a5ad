#!/bin/bash
# The round's ncu evidence (run on the GPU box AFTER the same commands have exited 0 without ncu):
#   1. launch list of the bench command (gpu__time_duration.sum per launch, cold-cache, serialised)  -> gpurun_out/r2_launches.csv
#   2. --set full of the temporal k_lk launch, k_preprocess, k_ransac, k_subpix, k_dv_* of one step   -> gpurun_out/r2_step_full.ncu-rep
# and the text summaries read from them.  Copy what is to be kept into profiles/.
set -u
mkdir -p gpurun_out
python bench.py --steps 1 --warmup 1 --frames-per-step 4 --no-cpu-baseline > gpurun_out/r2_bench_small.json 2> gpurun_out/r2_bench_small.err || { echo "bench failed"; tail -3 gpurun_out/r2_bench_small.err; exit 1; }
ncu --metrics gpu__time_duration.sum --print-units base --clock-control none -c 1500 --csv --log-file gpurun_out/r2_launches.csv \
    python bench.py --steps 1 --warmup 1 --frames-per-step 4 --no-cpu-baseline > gpurun_out/r2_ncu_bench.log 2>&1
UVIO_KLT_GROUPS=1 python tools/profile_step.py > gpurun_out/r2_profile_step.log 2>&1 || { echo "profile_step failed"; exit 1; }
# every k_lk launch of the same workload: instructions and DRAM bytes summed over the run / features tracked in the run
UVIO_KLT_GROUPS=1 ncu --metrics smsp__inst_executed.sum,gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active \
    --print-units base --clock-control none -k regex:k_lk --csv --log-file gpurun_out/r2_lk_all.csv python tools/profile_step.py > gpurun_out/r2_ncu_lk_all.log 2>&1
python - <<'PY'
import csv, json, re, sys, os
sys.path.insert(0, os.getcwd())
import bench
log = open("gpurun_out/r2_profile_step.log").read()
pl = int(re.search(r"lk point-levels (\d+)", log).group(1))
cfg = bench.workload_config("euroc_mav")
feats = pl / len(bench._levels(cfg))
rows = [r for r in csv.reader(l for l in open("gpurun_out/r2_lk_all.csv") if l.startswith('"'))]
hdr = rows[0]
iN, iM, iV = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value")
iID = hdr.index("ID")
per = {}
for r in rows[1:]:
    per.setdefault(r[iID], {})[r[iM]] = float(r[iV].replace(",", ""))
tot = lambda m: sum(v.get(m, 0.0) for v in per.values())
big = max(per.values(), key=lambda v: v.get("gpu__time_duration.sum", 0))
out = {"capture": "ncu --metrics ... -k regex:k_lk over tools/profile_step.py (64 euroc_mav streams x 8 frames, UVIO_KLT_GROUPS=1): sums over all %d k_lk launches; pipe/issue figures of the longest launch" % len(per),
       "launches": len(per), "features_in_launch": feats, "smsp__inst_executed_sum": tot("smsp__inst_executed.sum"),
       "dram_bytes_read": tot("dram__bytes_read.sum"), "dram_bytes_write": tot("dram__bytes_write.sum"),
       "duration_us_sum": tot("gpu__time_duration.sum") / 1e3, "duration_us": big.get("gpu__time_duration.sum", 0) / 1e3,
       "fma_pipe_utilisation": big.get("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
       "alu_pipe_utilisation": big.get("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
       "issue_active": big.get("smsp__issue_active.avg.pct_of_peak_sustained_active")}
json.dump(out, open("gpurun_out/r2_lk_ncu.json", "w"), indent=1)
print("k_lk:", out["smsp__inst_executed_sum"] / feats, "warp instructions per feature;", out)
PY
# one steady-state step with every tracker in one group: each kernel of the chain once, all streams in the launch
UVIO_KLT_GROUPS=1 ncu --set full --clock-control none --import-source on -s 88 -c 16 -f -o gpurun_out/r2_step_full \
    python tools/profile_step.py > gpurun_out/r2_ncu_full.log 2>&1
ncu -i gpurun_out/r2_step_full.ncu-rep --page raw --csv --print-units base > gpurun_out/r2_step_full_raw.csv 2>/dev/null
python - <<'PY'
import csv, json
rows = list(csv.reader(open("gpurun_out/r2_step_full_raw.csv")))
hdr = rows[0]
want = ["Kernel Name", "gpu__time_duration.sum", "smsp__inst_executed.sum", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__inst_executed.avg.per_cycle_active", "l1tex__t_bytes.sum", "lts__t_bytes.sum"]
idx = {w: hdr.index(w) for w in want if w in hdr}
out = []
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    out.append({w: r[i] for w, i in idx.items()})
with open("gpurun_out/r2_step_full_summary.json", "w") as f:
    json.dump({"units_row": {w: rows[1][i] for w, i in idx.items()}, "kernels": out}, f, indent=1)
pre = [k for k in out if "k_preprocess" in k.get("Kernel Name", "")]
if pre:
    f = lambda x: float(str(x).replace(",", ""))
    units = {w: rows[1][i] for w, i in idx.items()}
    mul = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "nsecond": 1e-3, "second": 1e6}
    g = lambda w: f(pre[0][w]) * mul.get(units.get(w, ""), 1.0)          # bytes / microseconds whatever unit ncu chose to print
    json.dump({"capture": "ncu --set full of one k_preprocess launch of tools/profile_step.py (64 euroc_mav streams = 128 images)", "images_in_launch": 128,
               "dram_bytes_read": g("dram__bytes_read.sum"), "dram_bytes_write": g("dram__bytes_write.sum"),
               "duration_us": g("gpu__time_duration.sum"), "units": {w: rows[1][i] for w, i in idx.items() if w.startswith("dram__bytes") or w.startswith("gpu__time")}},
              open("gpurun_out/r2_preprocess_traffic.json", "w"), indent=1)
for k in out:
    print(k.get("Kernel Name", "?")[:40], k.get("gpu__time_duration.sum"), k.get("smsp__inst_executed.sum"), k.get("smsp__issue_active.avg.pct_of_peak_sustained_active"))
PY
