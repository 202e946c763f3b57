mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sampling.py tests/test_gpu_model.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -n 4
LLB200_BENCH_REFGPU=0 LLB200_BENCH_CPU=0 timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/c45_bench.json 2> gpurun_out/c45_bench.err; cut -c1-200 gpurun_out/c45_bench.json; grep -o '"decode_step": {[^}]*}' gpurun_out/c45_bench.json
