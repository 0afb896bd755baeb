#!/bin/bash
# GPU tests + bench lines of cfg2 (both modes) + ncu launch list + one full capture, written to gpurun_out/:  bash tools/record_quick.sh r01_i
tag=${1:-rec}
o=gpurun_out
mkdir -p $o
( time timeout 600 python -m pytest tests -m gpu -x -q ) > $o/${tag}_pytest_gpu.log 2>&1
tail -3 $o/${tag}_pytest_gpu.log
python bench.py > $o/${tag}_bench_cfg2_analytic.json 2> $o/${tag}_bench_cfg2_analytic.err
python bench.py --mode numerical --no-cpu-baseline > $o/${tag}_bench_cfg2_numerical.json 2>/dev/null
for c in cfg3 cfg4; do
  python bench.py --config $c --mode analytic --steps 5 --warmup 3 --no-cpu-baseline > $o/${tag}_bench_${c}_analytic.json 2>/dev/null
done
python tools/prof_one.py north_star analytic 2 > $o/${tag}_north_star_analytic.txt 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/${tag}_launches_ana.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $o/${tag}_ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:transient_kernel -c 2 -o $o/${tag}_prof_ana -f python tools/prof_one.py cfg2 analytic 1 > $o/${tag}_ncu_full.log 2>&1
cut -c1-400 $o/${tag}_bench_cfg2_analytic.json
ls -la $o | tail -12
