set -x
python -m pytest tests -m gpu -x -q > gpurun_out/t.log 2>&1; echo "tests exit $?" >> gpurun_out/t.log
python bench.py --streams 64 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_s64.log 2>&1
UVIO_KLT_UPLOAD_PIECE=4194304 python bench.py --streams 64 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_s64p4m.log 2>&1
UVIO_KLT_UPLOAD_PIECE=268435456 python bench.py --streams 64 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_s64pall.log 2>&1
python bench.py --streams 32 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_s32.log 2>&1
UVIO_KLT_JOBS=copy python bench.py --streams 64 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_s64copy.log 2>&1
python bench.py --streams 96 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_s96.log 2>&1
