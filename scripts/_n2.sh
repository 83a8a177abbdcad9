timeout 500 python -m pytest tests/test_gpu_model_sharded.py -x -q 2>&1 | tail -5
run() { name=$1; shift; timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 50 --warmup 10 --no-cpu-baseline --no-e2e "$@" > gpurun_out/r2_n2_$name.json 2> gpurun_out/r2_n2_$name.err; echo "$name rc=$?"; }
run auto2
FFEMB_BWD=push run push2
FFEMB_TRACE=1 run auto2_trace --graph off
FFEMB_BWD=push FFEMB_TRACE=1 run push2_trace --graph off
grep -A22 "^TRACE rank 0" gpurun_out/r2_n2_auto2_trace.err | head -30
grep -A22 "^TRACE rank 0" gpurun_out/r2_n2_push2_trace.err | head -30
