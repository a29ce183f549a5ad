set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/t_gpu_r2a.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/t_gpu_r2a.log
python bench.py --steps 20 --warmup 3 > gpurun_out/bench_r2a.json 2> gpurun_out/bench_r2a.err; echo "bench rc=$?"
taskset -c 0-3 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r2a_4c.json 2> gpurun_out/bench_r2a_4c.err; echo "bench4c rc=$?"
taskset -c 0-1 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r2a_2c.json 2> gpurun_out/bench_r2a_2c.err; echo "bench2c rc=$?"
bash tools/ncu_round.sh > gpurun_out/ncu_round.log 2>&1; echo "ncu rc=$?"; tail -20 gpurun_out/ncu_round.log
python - <<'PY'
import json
for f in ("bench_r2a","bench_r2a_4c","bench_r2a_2c"):
    try:
        d=json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith("{")][-1])
        print(f, round(d["value"]), round(d["e2e"]["value"]), d["e2e"].get("upload_only"), d["e2e"].get("host_numa"), d["roofline"].get("frac"))
    except Exception as e: print(f, "ERR", e)
PY
