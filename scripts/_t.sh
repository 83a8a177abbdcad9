python -m pytest tests/test_gpu_sort.py tests/test_gpu_parity.py tests/test_gpu_model_sharded.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -2
python scripts/check_own_sort.py 2>&1 | tail -2
for c in C3 C2 C1 C5; do python bench.py --config $c --steps 100 --warmup 20 --no-e2e --no-cpu-baseline > gpurun_out/r2_${c}_cs8.json 2>/dev/null; done
