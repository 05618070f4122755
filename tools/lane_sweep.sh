#!/bin/bash
# development: views/s for several lane counts and for the round-1 library on the same box
for L in 3 4 5 6; do
  timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --lanes $L > gpurun_out/sweep_$L.json 2> gpurun_out/sweep_$L.err
  python -c "import json; d=json.loads(open('gpurun_out/sweep_$L.json').read().strip().splitlines()[-1]); print('lanes', $L, d['value'])" || tail -3 gpurun_out/sweep_$L.err
done
D3GA_LIB=$PWD/variants/libd3ga_r1.so timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/sweep_r1.json 2> gpurun_out/sweep_r1.err
python -c "import json; d=json.loads(open('gpurun_out/sweep_r1.json').read().strip().splitlines()[-1]); print('r1lib', d['value'])" || tail -3 gpurun_out/sweep_r1.err
