#!/bin/bash
# development: views/s for several lane counts (views in flight) on the same box, twice
for rep in 1 2; do
for L in 2 3 4 5 6 8; do
  timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --lanes $L > gpurun_out/sweep_$L.json 2> gpurun_out/sweep_$L.err
  python -c "import json; d=json.loads(open('gpurun_out/sweep_$L.json').read().strip().splitlines()[-1]); print('lanes', $L, round(d['value'], 1))" || tail -3 gpurun_out/sweep_$L.err
done
done
