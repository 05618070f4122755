#!/bin/bash
# round-end scaling evidence on ONE 8-GPU box: the default bench line at N = 8 (with the host feed) and at N = 1
N=${1:-8}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus $N --no-cpu-baseline > gpurun_out/r2_bench_metric_n$N.json 2> gpurun_out/r2_bench_metric_n$N.err
python -c "import json; d=json.loads(open('gpurun_out/r2_bench_metric_n$N.json').read().strip().splitlines()[-1]); print('n$N', d['value'], d['e2e']['value'], d['ms_per_step'], d['clocks'])" || tail -5 gpurun_out/r2_bench_metric_n$N.err
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/r2_bench_metric_n1_samebox.json 2> gpurun_out/r2_bench_metric_n1_samebox.err
python -c "import json; d=json.loads(open('gpurun_out/r2_bench_metric_n1_samebox.json').read().strip().splitlines()[-1]); print('n1', d['value'], d['e2e']['value'], d['ms_per_step'])"
