#!/bin/bash
# development: bench value against the scatter's row length (D3GA_DIRECT_CHUNK; unset = the library's own choice)
for rep in 1 2; do
  for c in "" 1152 1728 2304 3456 4608 6912; do
    D3GA_DIRECT_CHUNK=$c timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ab.json 2> gpurun_out/ab.err
    python -c "import json; d=json.loads(open('gpurun_out/ab.json').read().strip().splitlines()[-1]); k=d['roofline']['kernels']; print('chunk=${c:-auto}', round(d['value'],1), round(k['direct_scatter_kernel']['avg_ms']*1e3,1), round(k['direct_count_kernel']['avg_ms']*1e3,1))" || tail -3 gpurun_out/ab.err
  done
done
