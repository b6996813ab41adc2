#!/bin/bash
# Batch-1 latency of a KITTI stereo pair through the C++ drop-in, three ways (tools/latency_bench.cpp): two threads + orb_stereo_match,
# one thread + orb_stereo_match, ORBextractor::ExtractStereo; then one TUM frame (mono).  Usage: tools/latency_run.sh [out.json]
set -e
cd "$(dirname "$0")/.."
python - <<'PY'
import importlib, numpy as np
synth = importlib.import_module("orb_slam2-chinese-annotation_b200.synth")
L, R = zip(*[synth.make_stereo_pair(1241, 376, i, fx=718.856) for i in range(8)])
np.concatenate([np.stack(L).ravel(), np.stack(R).ravel()]).tofile("/tmp/orb_lat_kitti.bin")
np.stack([synth.make_image(640, 480, i) for i in range(8)]).tofile("/tmp/orb_lat_tum.bin")
PY
out=${1:-/dev/stdout}
for t in 1 0 2; do
  tools/build/latency_bench stereo /tmp/orb_lat_kitti.bin 1241 376 8 2000 386.1448 718.856 300 30 0 $t
done > "$out"
tools/build/latency_bench mono /tmp/orb_lat_tum.bin 640 480 8 1000 0 0 300 30 0 1 >> "$out"
rm -f /tmp/orb_lat_kitti.bin /tmp/orb_lat_tum.bin
