mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider 2>&1 | tail -n 5
bash run_gpu_tests.sh 600 1700 > /dev/null 2>&1; tail -n 8 gpurun_out/pytest_gpu.log
PROBE_SEQS=qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -n 1
LLB200_GEMV_PARITY=0 PROBE_SEQS=qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -n 1
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/c22_bench.json 2> gpurun_out/c22_bench.err; tail -c 3000 gpurun_out/c22_bench.json
LLB200_GEMV_PARITY=0 LLB200_BENCH_CPU=0 LLB200_BENCH_REFGPU=0 timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/c22_bench_noparity.json 2>> gpurun_out/c22_bench.err; tail -c 1500 gpurun_out/c22_bench_noparity.json
L=$PWD/lit-llama-ja_b200/lit_llama
LLB200_LIB=$L/libllb200_loopclk.so timeout 300 python profiles/attn_trace.py > gpurun_out/c22_attn_parity.txt 2>&1; head -n 8 gpurun_out/c22_attn_parity.txt
