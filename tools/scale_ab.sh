#!/bin/bash
# development: N-GPU bench with different all-reduce chunkings on the same box;  usage: scale_ab.sh N chunks...
N=$1; shift
for c in "$@"; do
  D3GA_REDUCE_CHUNKS=$c timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2955$c bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/scale_$c.json 2> gpurun_out/scale_$c.err
  python -c "import json; d=json.loads(open('gpurun_out/scale_$c.json').read().strip().splitlines()[-1]); print('chunks', $c, d['n_gpus'], d['value'], d['ms_per_step'])" || tail -3 gpurun_out/scale_$c.err
done
timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/scale_n1.json 2>/dev/null
python -c "import json; d=json.loads(open('gpurun_out/scale_n1.json').read().strip().splitlines()[-1]); print('n1', d['value'], d['ms_per_step'])"
