set -x
UVIO_KLT_SUBMIT_GROUP=64 UVIO_KLT_LIB=$PWD/tools/libuvio_klt_prof.so python bench.py --streams 32 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_prof.log 2>&1
UVIO_KLT_SUBMIT_GROUP=64 UVIO_KLT_LANES=8 UVIO_KLT_DEPTH=1 python bench.py --streams 32 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_l8.log 2>&1
UVIO_KLT_SUBMIT_GROUP=64 UVIO_KLT_LANES=12 UVIO_KLT_DEPTH=1 python bench.py --streams 32 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_l12.log 2>&1
UVIO_KLT_SUBMIT_GROUP=64 UVIO_KLT_LANES=16 UVIO_KLT_DEPTH=1 python bench.py --streams 32 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_l16.log 2>&1
UVIO_KLT_SUBMIT_GROUP=64 python bench.py --streams 64 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_s64.log 2>&1
UVIO_KLT_SUBMIT_GROUP=128 UVIO_KLT_LANES=8 UVIO_KLT_DEPTH=1 python bench.py --streams 64 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/b_s64l8.log 2>&1
nproc; lscpu | head -20
