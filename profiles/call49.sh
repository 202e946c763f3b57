mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tp.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -n 2
TP_CHECK_PDL=1 TAIL=6 ./run_tp.sh 2 240 c49_tp_check_ll profiles/tp_check.py ll
LLB200_BENCH_65B=0 TAIL=3 ./run_tp.sh 2 500 c49_bench_n2 bench.py --gpus 2 --steps 3 --warmup 3
