#!/bin/bash
# profiles/run_c3_strong_ab.sh <N>: c3 strong at N ranks: default vs --no-overlap vs --no-jvp (where does the step time go?)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
port=29800
for v in default no-overlap no-jvp default; do
  port=$((port + 1))
  out=gpurun_out/r02_c3s_$1_$v.json
  flag=""; [ $v != default ] && flag="--$v"
  timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus $1 --workload c3 --scaling strong --steps 30 --warmup 5 --no-cpu-baseline $flag > $out 2> ${out%.json}.err
  python - $out $v <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
k = d["roofline"]["kernels_ms_per_step"]
print(sys.argv[2], "ms/step %.3f" % d["ms_per_step"], "e2e %.3f" % d["e2e"]["ms_per_step"], "kernel sum %.3f" % sum(k.values()), {x: k[x] for x in k if "sigma" in x or "jvp" in x})
PY
done
