i=0
for v in "" tools/liblk28.so tools/liblk_old.so; do
  i=$((i+1))
  UVIO_KLT_LIB=${v:+$PWD/$v} UVIO_KLT_LANES=1 UVIO_KLT_DEPTH=1 ncu --metrics smsp__inst_executed.sum,gpu__time_duration.sum,launch__registers_per_thread --clock-control none -k regex:k_lk -s 8 -c 4 --csv --log-file gpurun_out/lk_v$i.csv python bench.py --streams 32 --steps 3 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
done
