#!/bin/bash
# GPU-box script: c3 bench per jvp gather-warp variant (profiles/build_variant.sh), then the default library again
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
cp loner_b200/lib/libloner_b200.so /tmp/lib_default.so
show() { python - "$1" <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
k = d["roofline"]["kernels_ms_per_step"]
print(sys.argv[1], "ms/step %.3f" % d["ms_per_step"], {x: k[x] for x in k if "sigma" in x})
PY
}
for v in ${VARIANTS:-default g16 g24 default}; do
  [ $v = default ] && cp /tmp/lib_default.so loner_b200/lib/libloner_b200.so || cp profiles/variants/lib_$v.so loner_b200/lib/libloner_b200.so
  out=gpurun_out/r02_c3_jvp_$v.json
  timeout 150 python bench.py --workload c3 --steps 20 --warmup 5 --no-cpu-baseline > $out 2> ${out%.json}.err && show $out
done
cp /tmp/lib_default.so loner_b200/lib/libloner_b200.so
