#!/bin/bash
# compute-sanitizer over the stage parity tests, the feature store / triangulation tests and one batched sequence (19 trackers
# through the device-side pipeline).  Run on the GPU box:  tools/sanitize.sh [tool ...]   (default: memcheck racecheck initcheck synccheck)
# Summaries land in gpurun_out/sanitizer_<tool>.txt; copy them to profiles/ when they are the ones to keep.
set -u
TOOLS=${@:-memcheck racecheck initcheck synccheck}
mkdir -p gpurun_out
cat > /tmp/sanitize_seq.py <<'PY'
import sys, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import numpy as np
from uvio_b200 import configs
from uvio_b200.track import CameraData, feed_batch, make_tracker
from util import frames_for, masks_for
cfg = configs.EUROC_MAV.with_(num_features=400, min_px_dist=6)
n_trk, n = 7, 4
masks = masks_for(cfg)
seqs = [frames_for(cfg, n, stream_id=60 + s % 3) for s in range(n_trk)]
trks = [make_tracker(cfg) for _ in range(n_trk)]
tot = 0
for k in range(n):
    feed_batch(trks, [CameraData(0.05 * k, [0, 1], list(seqs[s][k]), masks) for s in range(n_trk)])
    tot += sum(len(t.new_observations()) for t in trks)
print("observations", tot)
assert tot > 500
PY
for t in $TOOLS; do
  out=gpurun_out/sanitizer_$t.txt
  echo "== compute-sanitizer --tool $t ==" > $out
  targets=("python /tmp/sanitize_seq.py")
  # the stage / feature store / triangulation parity tests run under memcheck only (racecheck of a whole pytest run takes too long)
  if [ "$t" = memcheck ]; then targets+=("python -m pytest tests/test_gpu_stages.py tests/test_featdb.py tests/test_tri.py tests/test_zz_no_alloc.py -m gpu -x -q"); fi
  for target in "${targets[@]}"; do
    echo "-- $target" >> $out
    timeout 600 compute-sanitizer --tool $t --print-limit 20 $target > gpurun_out/sanitizer_raw.txt 2>&1
    echo "exit code $?" >> $out
    grep -E "ERROR SUMMARY|Error|error|RACECHECK SUMMARY|hazard|passed|failed|observations|Uninitialized|Invalid|at .*\.(cu|h):" gpurun_out/sanitizer_raw.txt | head -60 >> $out
    tail -5 gpurun_out/sanitizer_raw.txt >> $out
  done
  tail -6 $out
done
