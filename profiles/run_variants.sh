#!/bin/bash
# GPU-box script: profiles/run_variants.sh <workload> <variant...>: bench step per library variant (profiles/build_variant.sh;
# "default" = the shipped library), A-B-...-A order; prints ms per step and the sigma kernels' ms per step
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
cp loner_b200/lib/libloner_b200.so /tmp/lib_default.so
w=$1; shift
for v in "$@"; do
  [ $v = default ] && cp /tmp/lib_default.so loner_b200/lib/libloner_b200.so || cp profiles/variants/lib_$v.so loner_b200/lib/libloner_b200.so
  out=gpurun_out/r02_${w}_variant_$v.json
  timeout 150 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline > $out 2> ${out%.json}.err && python - $out <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
k = d["roofline"]["kernels_ms_per_step"]
print(sys.argv[1], "ms/step %.3f" % d["ms_per_step"], {x: k[x] for x in k if "sigma" in x or "adam" in x})
PY
done
cp /tmp/lib_default.so loner_b200/lib/libloner_b200.so
