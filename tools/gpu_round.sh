UVIO_KLT_AHEAD=0 UVIO_KLT_LANES=1 UVIO_KLT_DEPTH=1 ncu --set full --clock-control none --import-source on -k regex:'k_(lk|preprocess|ransac|subpix|fast)' -s 36 -c 6 -o gpurun_out/step_full python bench.py --streams 32 --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_f.log 2>&1
ls -la gpurun_out
