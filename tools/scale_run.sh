#!/bin/bash
# 1 -> 8 GPU runs of the headline bench (atmosphere sharding, weak) and the band-sharded variants.
# usage (on the GPU box): bash tools/scale_run.sh <tag> [maxN]
tag=${1:-scale}; maxn=${2:-8}
run() { # n, extra args..., output suffix last
  local n=$1; shift; local out=${@: -1}; set -- "${@:1:$(($#-1))}"
  if [ "$n" = 1 ]; then python bench.py --gpus 1 "$@" > gpurun_out/${tag}_${out}_n1.json 2> gpurun_out/${tag}_${out}_n1.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+RANDOM%400)) \
       bench.py --gpus $n "$@" > gpurun_out/${tag}_${out}_n$n.json 2> gpurun_out/${tag}_${out}_n$n.err; fi
  grep '^{' gpurun_out/${tag}_${out}_n$n.json | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$out n=$n', round(d['value'],2), d['unit'], 'ms/step', round(d['ms_per_step'],3), d.get('bit_identical_to_single_gpu',''))" || tail -3 gpurun_out/${tag}_${out}_n$n.err
}
for n in 1 2 4 8; do [ $n -le $maxn ] && run $n --steps 20 --warmup 3 --no-cpu-baseline atm; done
for n in 2 4 8; do [ $n -le $maxn ] && run $n --steps 20 --warmup 3 --sharding bands --verify bands; done
for n in 8; do [ $n -le $maxn ] && [ -z "$SKIP_C5" ] && run $n --workload C5 --steps 5 --warmup 3 c5; done
for n in 8; do [ $n -le $maxn ] && run $n --workload C3 --scale 1.0 --steps 2 --warmup 3 c3; done
