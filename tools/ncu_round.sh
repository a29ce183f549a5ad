#!/bin/bash
# The round's ncu evidence (run on the GPU box AFTER the same commands have exited 0 without ncu):
#   1. launch list of the bench command (gpu__time_duration.sum per launch, cold-cache, serialised)  -> gpurun_out/r2_launches.csv
#   2. --set full of the temporal k_lk launch, k_preprocess, k_ransac, k_subpix, k_dv_* of one step   -> gpurun_out/r2_step_full.ncu-rep
# and the text summaries read from them.  Copy what is to be kept into profiles/.
set -u
mkdir -p gpurun_out
python bench.py --steps 1 --warmup 1 --frames-per-step 4 --no-cpu-baseline > gpurun_out/r2_bench_small.json 2> gpurun_out/r2_bench_small.err || { echo "bench failed"; tail -3 gpurun_out/r2_bench_small.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2_launches.csv \
    python bench.py --steps 1 --warmup 1 --frames-per-step 4 --no-cpu-baseline > gpurun_out/r2_ncu_bench.log 2>&1
UVIO_KLT_GROUPS=1 python tools/profile_step.py > gpurun_out/r2_profile_step.log 2>&1 || { echo "profile_step failed"; exit 1; }
# one steady-state step with every tracker in one group: each kernel of the chain once, all streams in the launch
UVIO_KLT_GROUPS=1 ncu --set full --clock-control none --import-source on -s 66 -c 12 -f -o gpurun_out/r2_step_full \
    python tools/profile_step.py > gpurun_out/r2_ncu_full.log 2>&1
ncu -i gpurun_out/r2_step_full.ncu-rep --page raw --csv > gpurun_out/r2_step_full_raw.csv 2>/dev/null
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
for k in out:
    print(k.get("Kernel Name", "?")[:40], k.get("gpu__time_duration.sum"), k.get("smsp__inst_executed.sum"), k.get("smsp__issue_active.avg.pct_of_peak_sustained_active"))
PY
