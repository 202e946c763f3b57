mkdir -p gpurun_out
TAIL=3 ./run_tp.sh 8 700 c43_bench_n8 bench.py --gpus 8 --steps 2 --warmup 3
