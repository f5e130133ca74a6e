#!/bin/bash
# Strong-scaling sweep on one box: N = 1,2,4,8 for the fused path, the NCCL-halo baseline and
# the convergence variant (BASELINE configs 3 and 4).  Writes one JSON line per run.
set -u
OUT=gpurun_out/scaling_$(date +%H%M%S).jsonl
NG=${1:-8}
STEPS=${STEPS:-1000}
run() { # n impl extra...
  local n=$1; shift; local impl=$1; shift
  if [ "$n" = 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps $STEPS --warmup 24 --impl $impl "$@" 2>gpurun_out/err_${impl}_$n.log | tail -1 >> $OUT
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 \
      --master-port $((29600 + n)) bench.py --gpus $n --steps $STEPS --warmup 24 --impl $impl "$@" 2>gpurun_out/err_${impl}_$n.log | tail -1 >> $OUT
  fi
}
for n in 1 2 4 8; do
  [ $n -le $NG ] || continue
  run $n ours
  run $n ours --conv --no-e2e
done
for n in 2 4 8; do
  [ $n -le $NG ] || continue
  STEPS=200 run $n nccl-halo
done
run 1 reference
cat $OUT | cut -c1-260
