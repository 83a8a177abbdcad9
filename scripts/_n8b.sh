bash scripts/sweep.sh 8
bash scripts/sweep.sh 4
bash scripts/sweep.sh 2
N=8 FFEMB_TRACE=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 8 --steps 50 --warmup 10 --no-e2e --no-cpu-baseline --graph off > gpurun_out/r02_trace_c3_n8.json 2> gpurun_out/r02_trace_c3_n8.err
FFEMB_TRACE=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29535 bench.py --gpus 8 --steps 50 --warmup 10 --no-e2e --no-cpu-baseline --graph off --placement plan > gpurun_out/r02_trace_c3_plan_n8.json 2> gpurun_out/r02_trace_c3_plan_n8.err
