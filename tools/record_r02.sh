#!/bin/bash
# Round-2 record: GPU tests, the driver's bench line, per-config lines, ncu launch list + full captures -> gpurun_out/ :  bash tools/record_r02.sh r02_k
tag=${1:-r02_x}
o=gpurun_out
mkdir -p $o
( time timeout 1500 python -m pytest tests -m gpu -q ) > $o/${tag}_pytest_gpu.log 2>&1
tail -3 $o/${tag}_pytest_gpu.log
python bench.py --occlusion > $o/${tag}_bench_cfg4_analytic.json 2> $o/${tag}_bench_cfg4_analytic.err
python bench.py --steps 2 --warmup 3 --no-extra --no-cpu-baseline > $o/${tag}_plain.json 2>/dev/null && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"transient|shell_|scan_c|preprocess|mse_loss|reduce_tile|analytic_filter" -c 120 --csv \
    --log-file $o/${tag}_launches_cfg4_analytic.csv python bench.py --steps 2 --warmup 3 --no-extra --no-cpu-baseline > $o/${tag}_ncu_launch.log 2>&1
python tools/prof_one.py cfg4 analytic 1 rows=16 > $o/${tag}_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"transient_(narrow_)?kernel" -c 8 -o $o/${tag}_prof_cfg4_ana -f python tools/prof_one.py cfg4 analytic 1 rows=16 > $o/${tag}_ncu_full.log 2>&1
for c in cfg4 cfg3 cfg2 cfg1; do
  for m in analytic numerical; do
    [ "$c$m" = "cfg4analytic" ] && continue
    python bench.py --config $c --mode $m --steps 5 --warmup 3 --no-cpu-baseline --no-extra > $o/${tag}_bench_${c}_${m}.json 2>/dev/null
  done
done
python bench.py --impl reference --steps 2 --warmup 1 > $o/${tag}_bench_reference_arm.json 2>/dev/null
cut -c1-400 $o/${tag}_bench_cfg4_analytic.json
ls $o | grep ${tag} | wc -l
