#!/bin/bash
# GPU tests, bench lines of every single-GPU configuration, ncu launch list + full capture, reference arm -> gpurun_out/ (essentials first):
#   bash tools/record_full.sh r01_k
tag=${1:-rec}
o=gpurun_out
mkdir -p $o
( time timeout 600 python -m pytest tests -m gpu -x -q ) > $o/${tag}_pytest_gpu.log 2>&1
tail -3 $o/${tag}_pytest_gpu.log
python bench.py > $o/${tag}_bench_cfg2_analytic.json 2> $o/${tag}_bench_cfg2_analytic.err
python bench.py --mode numerical --no-cpu-baseline > $o/${tag}_bench_cfg2_numerical.json 2>/dev/null
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/${tag}_launches_ana.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $o/${tag}_ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:transient_kernel -c 2 -o $o/${tag}_prof_ana -f python tools/prof_one.py cfg2 analytic 1 > $o/${tag}_ncu_full.log 2>&1
for c in cfg3 cfg4 cfg1; do
  for m in analytic numerical; do
    python bench.py --config $c --mode $m --steps 5 --warmup 3 --no-cpu-baseline > $o/${tag}_bench_${c}_${m}.json 2>/dev/null
  done
done
python tools/prof_one.py north_star numerical 2 > $o/${tag}_north_star_numerical.txt 2>&1
python tools/prof_one.py north_star analytic 2 > $o/${tag}_north_star_analytic.txt 2>&1
python bench.py --impl reference --steps 2 --warmup 1 > $o/${tag}_bench_reference_arm.json 2>/dev/null
cut -c1-300 $o/${tag}_bench_cfg2_analytic.json
ls $o | grep ${tag} | wc -l
