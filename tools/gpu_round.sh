for i in 1 2; do python -m pytest tests -m gpu -x -q 2>&1 | tail -1; done
SKIP_SINGLE=1 REPS=100 python tools/determinism_probe.py 2>&1 | tail -2
UVIO_KLT_BACKGROUND=1 python -m pytest tests/test_ahead.py tests/test_submit.py -m gpu -x -q 2>&1 | tail -1
