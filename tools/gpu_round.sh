export SKIP_SINGLE=1 REPS=120
python tools/determinism_probe.py 2>&1 | tail -4
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
