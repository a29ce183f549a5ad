set -x
python -m pytest tests -m gpu -x -q > gpurun_out/t.log 2>&1; echo "tests exit $?" >> gpurun_out/t.log
UVIO_KLT_AHEAD=0 UVIO_KLT_LANES=1 UVIO_KLT_DEPTH=1 ncu --metrics smsp__inst_executed.sum,gpu__time_duration.sum,launch__registers_per_thread --clock-control none -k regex:k_lk -s 8 -c 4 --csv --log-file gpurun_out/lk_jw24.csv python bench.py --streams 32 --steps 3 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
python bench.py --streams 64 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_s64.log 2>&1
python bench.py --streams 32 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_s32.log 2>&1
