set -x
python -m pytest tests -m gpu -x -q > gpurun_out/t.log 2>&1; echo "tests exit $?" >> gpurun_out/t.log
python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_s64.log 2>&1
