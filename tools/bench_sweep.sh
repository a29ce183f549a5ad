#!/bin/bash
# usage: tools/bench_sweep.sh VAR v1 v2 ...   -- bench.py (no CPU arm) once per value of an environment variable; prints value / e2e
VAR=$1; shift
for v in "$@"; do
  env $VAR=$v python bench.py --no-cpu-baseline $BENCH_ARGS > gpurun_out/sweep_${VAR}_$v.json 2> gpurun_out/sweep_${VAR}_$v.err || { echo "$VAR=$v FAILED"; tail -3 gpurun_out/sweep_${VAR}_$v.err; continue; }
  python - "$VAR" "$v" <<'PY'
import json, sys
d = json.load(open(f"gpurun_out/sweep_{sys.argv[1]}_{sys.argv[2]}.json"))
r = d["roofline"]
sh = r["kernel_time_share_of_gpu_time"]
busy = r["gpu_busy_ms_per_feed"]
print(f"{sys.argv[1]}={sys.argv[2]}: resident {d['value']:.0f} e2e {d['e2e']['value']:.0f} ms/feed {d['config']['ms_per_feed']:.3f} e2e ms/feed {d['e2e']['ms_per_feed']:.3f} "
      f"launches {d['gpu_launches']} lk avg {r['avg_launch_ms']:.3f} frac {r.get('frac')} shares {sh} host {d['host']['resident_us_per_group_pass']} "
      f"calls {d['host']['resident_call_us_per_feed']} e2e calls {d['host']['e2e_call_us_per_feed']} clocks {d['clocks']}")
PY
done
