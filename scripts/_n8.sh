N=$1
timeout 900 python -m pytest tests/test_gpu_model_sharded.py tests/test_gpu_sharded.py -q > gpurun_out/r2_n${N}_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_n${N}_tests.log; tail -4 gpurun_out/r2_n${N}_tests.log
run() { n=$1; name=$2; shift; shift; timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 50 --warmup 10 --no-cpu-baseline "$@" > gpurun_out/r2_n${n}_$name.json 2> gpurun_out/r2_n${n}_$name.err; echo "n=$n $name rc=$?"; }
for n in 2 4 8; do
  if [ $n -le $N ]; then
    run $n auto
    run $n off --replicate off --no-e2e
    FFEMB_TRACE=1 run $n auto_trace --graph off --no-e2e
    FFEMB_TRACE=1 run $n off_trace --graph off --replicate off --no-e2e
  fi
done
