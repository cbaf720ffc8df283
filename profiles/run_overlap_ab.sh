#!/bin/bash
# profiles/run_overlap_ab.sh <N>: A/B of the side-stream optimiser step (default) against --no-overlap (in line), c4 weak and
# c2 strong at N ranks, A-B-A order; one torchrun (or plain python at N = 1) per line under its own timeout -> gpurun_out/r02_overlap_*.json
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
port=29700
run() {  # tag workload scaling N extra
  port=$((port + 1))
  out=gpurun_out/r02_overlap_$2_$3_$4_$1.json
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $4 --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus $4 --workload $2 --scaling $3 --steps 30 --warmup 5 --no-cpu-baseline $5 > $out 2> ${out%.json}.err
  python - "$out" "$@" <<'PY' || { grep -v "^W1105\|^  File\|^    \|frame #" ${out%.json}.err | tail -15; }
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(*sys.argv[2:6], "ms/step %.3f" % d["ms_per_step"], "e2e ms %.3f" % d["e2e"]["ms_per_step"], "collectives", d.get("collectives_ms_per_step"))
PY
}
n=${1:-2}
run overlap_a c4 weak $n ""
run sync_a c4 weak $n --no-overlap
run overlap_b c4 weak $n ""
run overlap_a c2 strong $n ""
run sync_a c2 strong $n --no-overlap
run overlap_b c2 strong $n ""
