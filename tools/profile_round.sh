#!/bin/bash
# Round-2 evidence run (one B200): tests + parity log, the bench line and the reference arm, the ncu launch list of the bench
# command and `--set full` captures of the hot kernels.  Everything lands in gpurun_out/; profiles/ gets the summaries.
set -u
export D3GA_PARITY_LOG=$PWD/gpurun_out/r2_parity.jsonl
rm -f $D3GA_PARITY_LOG
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r2_tests.log 2>&1; tail -3 gpurun_out/r2_tests.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_reference.json 2> gpurun_out/r2_bench_reference.err
timeout 600 python bench.py > gpurun_out/r2_bench_metric.json 2> gpurun_out/r2_bench_metric.err; tail -2 gpurun_out/r2_bench_metric.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_launches.csv \
    python bench.py --steps 1 --warmup 1 --views 2 --no-e2e --no-cpu-baseline > gpurun_out/r2_launches.log 2>&1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"blend_" --launch-skip 4 -c 2 -o gpurun_out/r2_blend \
    python tools/one_view.py metric 3 20 > gpurun_out/r2_ncu_blend.log 2>&1
timeout 900 ncu --set full --clock-control none -k regex:"preprocess|depth_|direct_|sort_pass" --launch-skip 24 -c 12 -o gpurun_out/r2_small \
    python tools/one_view.py metric 3 20 > gpurun_out/r2_ncu_small.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:"deform_" --launch-skip 4 -c 2 -o gpurun_out/r2_deform \
    python tools/deform_probe.py metric sorted full > gpurun_out/r2_ncu_deform.log 2>&1
ls -la gpurun_out/ | tail -20
