#!/bin/bash
# development: bench value of library variants on the same box, alternating;  usage: ab_bench.sh lib1.so lib2.so ...
for rep in 1 2; do
  for L in "$@"; do
    D3GA_LIB=$PWD/$L timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ab.json 2> gpurun_out/ab.err
    python -c "import json; d=json.loads(open('gpurun_out/ab.json').read().strip().splitlines()[-1]); k=d['roofline']['kernels']; print('$L', round(d['value'],1), {n: round(v['avg_ms']*1e3,1) for n,v in k.items() if 'blend' in n})" || tail -3 gpurun_out/ab.err
  done
done
