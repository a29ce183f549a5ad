set -x
python -m pytest tests -m gpu -x -q > gpurun_out/t.log 2>&1; echo "tests exit $?" >> gpurun_out/t.log
python bench.py --streams 32 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_ahead1_s32.log 2>&1
UVIO_KLT_AHEAD=0 python bench.py --streams 32 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_ahead0_s32.log 2>&1
python bench.py --streams 64 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_ahead1_s64.log 2>&1
UVIO_KLT_AHEAD=0 python bench.py --streams 64 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_ahead0_s64.log 2>&1
python bench.py --streams 96 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_ahead1_s96.log 2>&1
