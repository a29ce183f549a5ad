for v in base spread fast both first4 first16; do echo $v; UVIO_KLT_LIB=$PWD/tools/librs_$v.so python tools/ransac_bench.py 2>&1 | tail -1; done
