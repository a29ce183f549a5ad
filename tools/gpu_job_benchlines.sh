# bench lines of every BASELINE configuration + the CPU arm (no ncu, no tests): gpurun_out/bench_final*.json
set -u
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 3 > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo "bench rc=$?"
for c in tum_vi rs_d455 hd_1000 uzhfpv_8; do
  python bench.py --config $c --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/bench_final_$c.json 2> gpurun_out/bench_final_$c.err; echo "bench $c rc=$?"
done
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_final_reference.json 2> gpurun_out/bench_final_reference.err; echo "reference rc=$?"
python tools/desc_bench.py > gpurun_out/desc_bench.json 2> gpurun_out/desc_bench.err; echo "desc rc=$?"
python - <<'PY'
import json
for f in ("bench_final","bench_final_tum_vi","bench_final_rs_d455","bench_final_hd_1000","bench_final_uzhfpv_8","bench_final_reference"):
    try:
        d=json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith("{")][-1])
        r=d.get("roofline") or {}; ip=d.get("roofline_image_passes") or {}
        print(f, round(d["value"]), round(d["e2e"]["value"]), "frac", r.get("frac"), "img", ip.get("frac"), d["config"].get("streams_per_gpu"), d["config"].get("ms_per_feed"), d["e2e"].get("ms_per_feed"), (d["config"].get("single_stream_frame_latency_ms") or {}).get("p50"), d.get("clocks"))
    except Exception as e: print(f, "ERR", e)
print(open("gpurun_out/desc_bench.json").read()[:600])
PY
