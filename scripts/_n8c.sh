run() { name=$1; shift; timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 50 --warmup 10 --no-cpu-baseline --no-e2e "$@" > gpurun_out/r2_n8_$name.json 2> gpurun_out/r2_n8_$name.err; echo "$name rc=$?"; }
run cs8_bwd
FFEMB_SHARDED_SORT=fwd run cs8_fwd
FFEMB_TRACE=1 run cs8_bwd_trace --graph off
