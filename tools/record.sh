#!/bin/bash
# Bench lines + ncu artefacts of one tag, written to gpurun_out/ (run under gpurun from the repo root):  bash tools/record.sh r01_h
tag=${1:-rec}
o=gpurun_out
mkdir -p $o
python bench.py > $o/${tag}_bench_cfg2_analytic.json 2> $o/${tag}_bench_cfg2_analytic.err
python bench.py --mode numerical --no-cpu-baseline > $o/${tag}_bench_cfg2_numerical.json 2>/dev/null
for c in cfg1 cfg3 cfg4; do
  for m in analytic numerical; do
    python bench.py --config $c --mode $m --steps 5 --warmup 3 --no-cpu-baseline > $o/${tag}_bench_${c}_${m}.json 2>/dev/null
  done
done
python bench.py --impl reference --steps 2 --warmup 1 > $o/${tag}_bench_reference_arm.json 2>/dev/null
python tools/prof_one.py north_star numerical 2 > $o/${tag}_north_star_numerical.txt 2>&1
python tools/prof_one.py north_star analytic 2 > $o/${tag}_north_star_analytic.txt 2>&1
# ncu: launch list of the bench command, then one full capture of the two hot kernels (never a bench value)
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/${tag}_launches_ana.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $o/${tag}_ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:transient_kernel -c 2 -o $o/${tag}_prof_ana -f python tools/prof_one.py cfg2 analytic 1 > $o/${tag}_ncu_full.log 2>&1
ls -la $o | tail -30
