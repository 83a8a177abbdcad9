#!/bin/bash
# Driver-reproducible bench lines for profiles/: every BASELINE.json config at N ranks.
#   bash scripts/sweep.sh <N> [out_dir] [tag]
N=${1:-1}; OUT=${2:-gpurun_out}; TAG=${3:-r02}
run() { name=$1; shift
  if [ "$N" = "1" ]; then timeout 600 python bench.py "$@" > $OUT/${TAG}_bench_${name}_n1.json 2> $OUT/${TAG}_bench_${name}_n1.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N "$@" > $OUT/${TAG}_bench_${name}_n$N.json 2> $OUT/${TAG}_bench_${name}_n$N.err; fi
  echo "$name n=$N rc=$?"; }
FAST="--steps 50 --warmup 10 --no-e2e --no-cpu-baseline"
if [ "$N" = "1" ]; then
  run c3 --steps 100 --warmup 20
  run c1 --config C1 $FAST
  run c2 --config C2 $FAST
  run c4 --config C4 $FAST
  for a in 0 0.4 0.8 1.0 1.2; do run c5_zipf$a --config C5 --dist zipf --alpha $a $FAST; done
  run c3_zipf1.05 --dist zipf --alpha 1.05 $FAST
else
  run c3 --steps 50 --warmup 10 --no-cpu-baseline
  run c3_plan --placement plan $FAST
  run c3_noreplicas --replicate off $FAST
  run c4 --config C4 $FAST
  # BASELINE.json config 5: the Zipf sweep at every N (FULL=1 for all five alphas)
  if [ "$FULL" = "1" ]; then ALPHAS="0 0.4 0.8 1.0 1.2"; else ALPHAS="0 1.2"; fi
  for a in $ALPHAS; do run c5_zipf$a --config C5 --dist zipf --alpha $a $FAST; done
fi
