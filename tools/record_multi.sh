#!/bin/bash
# Multi-GPU bench lines (run under gpurun --gpus N from the repo root):  bash tools/record_multi.sh r01_i 2
tag=${1:-rec}; n=${2:-2}
o=gpurun_out
mkdir -p $o
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n "$@"; }
run --steps 10 --warmup 3 --no-cpu-baseline > $o/${tag}_bench_cfg2_analytic_${n}gpu.json 2> $o/${tag}_multi_${n}gpu.err
run --steps 10 --warmup 3 --mode numerical --no-cpu-baseline > $o/${tag}_bench_cfg2_numerical_${n}gpu.json 2>> $o/${tag}_multi_${n}gpu.err
run --config cfg4 --scaling strong --steps 4 --warmup 3 --no-cpu-baseline > $o/${tag}_bench_cfg4_analytic_${n}gpu_strong.json 2>> $o/${tag}_multi_${n}gpu.err
if [ "$n" -ge 8 ]; then
  run --config cfg5 --mode numerical --scaling strong --steps 3 --warmup 3 --no-cpu-baseline > $o/${tag}_bench_cfg5_numerical_${n}gpu_strong.json 2>> $o/${tag}_multi_${n}gpu.err
fi
tail -5 $o/${tag}_multi_${n}gpu.err
for f in $o/${tag}_bench_*_${n}gpu*.json; do echo $f; cut -c1-330 $f; done
