mkdir -p gpurun_out
bash run_gpu_tests.sh 600 1700 > /dev/null 2>&1; tail -n 6 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/c34_bench.json 2> gpurun_out/c34_bench.err; tail -c 2500 gpurun_out/c34_bench.json
