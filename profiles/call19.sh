mkdir -p gpurun_out
bash run_gpu_tests.sh 600 1500 > gpurun_out/c19_pytest_tail.txt 2>&1; tail -8 gpurun_out/c19_pytest_tail.txt
timeout 600 python profiles/batch_probe.py 9,32,64 > gpurun_out/c19_batch.txt 2>&1; tail -3 gpurun_out/c19_batch.txt
LLB200_BENCH_CPU=0 timeout -s KILL 600 python bench.py --steps 5 --warmup 3 > gpurun_out/c19_bench.json 2> gpurun_out/c19_bench.err; tail -c 1500 gpurun_out/c19_bench.json; tail -3 gpurun_out/c19_bench.err
