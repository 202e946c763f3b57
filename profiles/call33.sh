mkdir -p gpurun_out
TP_CHECK_PDL=1 TAIL=8 ./run_tp.sh 2 240 c33_tp_check_ll profiles/tp_check.py ll
LLB200_BENCH_65B=0 TAIL=3 ./run_tp.sh 2 500 c33_bench_n2 bench.py --gpus 2 --steps 3 --warmup 3
