#!/bin/bash
# profiles/run_scaling.sh <N...>: multi-GPU bench lines of round 2 (c4 weak, c2 strong, c5) for the given rank counts,
# one torchrun per line, each under its own timeout; results in gpurun_out/r02_scale_<workload>_<scaling>_<N>.json
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
port=29600
run() {  # workload scaling N
  port=$((port + 1))
  out=gpurun_out/r02_scale_$1_$2_$3.json
  timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $3 --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus $3 --workload $1 --scaling $2 --steps ${STEPS:-20} --warmup 5 > $out 2> ${out%.json}.err
  python - "$out" "$1" "$2" "$3" <<'PY' || { grep -v "^W1105\|^  File\|^    \|frame #" ${out%.json}.err | tail -15; }
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[2], sys.argv[3], "N=" + sys.argv[4], "value %.4g %s" % (d["value"], d["unit"]), "ms/step %.3f" % d["ms_per_step"],
      "e2e %.4g" % d["e2e"]["value"], "collectives", d.get("collectives_ms_per_step"), "kernels", d["roofline"].get("kernels_ms_per_step"),
      "other", d.get("other_renders_ms"), "mem", d.get("peak_mem_gib"))
PY
}
# WORKLOADS="c4:weak c3:strong" selects a subset (default: all four)
for n in "$@"; do
  for ws in ${WORKLOADS:-c4:weak c2:strong c3:strong c5:weak}; do
    w=${ws%%:*}; s=${ws##*:}
    if [ $w = c5 ]; then STEPS=3 run $w $s $n; else run $w $s $n; fi
  done
done
