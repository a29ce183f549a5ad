set -x
python bench.py > gpurun_out/r1_bench.json 2>gpurun_out/b.err; echo "bench exit $?"
python bench.py --impl reference > gpurun_out/r1_bench_reference.json 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r1_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_l.log 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
