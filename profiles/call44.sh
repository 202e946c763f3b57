mkdir -p gpurun_out
TAIL=3 ./run_tp.sh 4 700 c44_bench_n4 bench.py --gpus 4 --steps 2 --warmup 3
