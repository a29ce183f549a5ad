set -x
python -m pytest tests -m gpu -x -q > gpurun_out/t.log 2>&1; echo "tests exit $?" >> gpurun_out/t.log
python bench.py --streams 32 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_base.log 2>&1
python bench.py --streams 64 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_s64.log 2>&1
UVIO_KLT_UPLOAD_PIECE=524288 python bench.py --streams 32 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_p512k.log 2>&1
UVIO_KLT_UPLOAD_PIECE=262144 python bench.py --streams 32 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_p256k.log 2>&1
UVIO_KLT_SUBMIT_GROUP=16 python bench.py --streams 32 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_g16.log 2>&1
