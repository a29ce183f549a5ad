set -x
python -m pytest tests -m gpu -x -q > gpurun_out/t.log 2>&1; echo "tests exit $?" >> gpurun_out/t.log
python bench.py --streams 64 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_e1_s64.log 2>&1
UVIO_KLT_EARLY=0 python bench.py --streams 64 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_e0_s64.log 2>&1
python bench.py --streams 32 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_e1_s32.log 2>&1
UVIO_KLT_EARLY=0 python bench.py --streams 32 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_e0_s32.log 2>&1
