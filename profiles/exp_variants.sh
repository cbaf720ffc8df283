#!/bin/bash
# bench step + kernel micro-benchmarks for every library variant under loner_b200/lib/variants (build-time parameter sweeps)
cp loner_b200/lib/libloner_b200.so /tmp/lib_default.so
for v in loner_b200/lib/variants/*.so; do
  cp $v loner_b200/lib/libloner_b200.so
  python bench.py --no-cpu-baseline --steps 20 --warmup 5 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);k=d['roofline']['kernels_ms_per_step'];print('$(basename $v)', 'step', round(d['ms_per_step'],3), 'sigma_bwd', k['lb_sigma_bwd'], 'sigma_fwd', k['lb_sigma_fwd'])"
  python profiles/bench_levels_z.py 2>&1 | sed -n 2,3p
done
cp /tmp/lib_default.so loner_b200/lib/libloner_b200.so
