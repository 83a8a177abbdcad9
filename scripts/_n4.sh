timeout 300 python -m pytest tests/test_gpu_model_sharded.py -q -k "world1 or (matches_oracle and push)" 2>&1 | tail -3
run() { n=$1; name=$2; shift; shift; timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 50 --warmup 10 --no-cpu-baseline --no-e2e "$@" > gpurun_out/r2_n${n}_$name.json 2> gpurun_out/r2_n${n}_$name.err; echo "n=$n $name rc=$?"; }
run 4 auto_b
FFEMB_BWD=pull run 4 pull_b
run 2 auto_b
FFEMB_BWD=push run 2 push_b
