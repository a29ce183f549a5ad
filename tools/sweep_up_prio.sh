for prio in low high; do for grp in 0 32; do
  UVIO_KLT_UP_PRIO=$prio UVIO_KLT_SUBMIT_GROUP=$grp python bench.py --no-cpu-baseline --steps 10 --warmup 3 > gpurun_out/sw_${prio}_$grp.json 2> gpurun_out/sw_${prio}_$grp.err || { echo FAIL; tail -3 gpurun_out/sw_${prio}_$grp.err; continue; }
  python - $prio $grp <<'PY'
import json, sys
d=json.loads([l for l in open(f"gpurun_out/sw_{sys.argv[1]}_{sys.argv[2]}.json") if l.startswith("{")][-1])
print(f"prio={sys.argv[1]} group={sys.argv[2]}: resident {d['value']:.0f} e2e {d['e2e']['value']:.0f} e2e ms/feed {d['e2e']['ms_per_feed']:.3f}")
PY
done; done
