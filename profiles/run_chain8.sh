#!/bin/bash
# BASELINE configs[4], device-resident mode, on every GPU of the box at once (8 streams per GPU),
# once per host-wait policy (NNSB_WAIT, nnsb_runtime.cu: host_wait).  One tools/stream_bench process
# per GPU behind a file barrier; prints each rank's line.  usage: bash profiles/run_chain8.sh [GPUS]
N=${1:-8}
mkdir -p gpurun_out
for policy in spin yield sleep; do
  for width in 1920 1918; do
    dir=$(mktemp -d /tmp/nnsb_barrier_XXXXXX)
    for r in $(seq 0 $((N - 1))); do
      NNSB_WAIT=$policy tools/stream_bench $r 8 300 resident $width - barrier:$dir:$r:$N > gpurun_out/chain8_${policy}_${width}_$r.json 2> gpurun_out/chain8_${policy}_${width}_$r.err &
    done
    wait
    python - "$policy" "$width" "$N" <<'PY'
import json, sys
policy, width, n = sys.argv[1], sys.argv[2], int(sys.argv[3])
rs = [json.loads(open("gpurun_out/chain8_%s_%s_%d.json" % (policy, width, r)).read().strip().splitlines()[-1]) for r in range(n)]
slow = max(r["seconds"] for r in rs)
print("%-5s %s: %d GPUs, %.0f frames/s (slowest rank %.4f s, fastest %.4f s)" % (
    policy, width, n, sum(r["streams"] * r["frames_per_stream"] for r in rs) / slow, slow, min(r["seconds"] for r in rs)))
PY
    rm -rf $dir
  done
done
