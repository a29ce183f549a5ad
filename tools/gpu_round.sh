set -x
python tools/h2d_probe.py > gpurun_out/h2d2.json 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/t.log 2>&1; echo "tests exit $?" >> gpurun_out/t.log
