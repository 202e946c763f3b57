mkdir -p gpurun_out
nvidia-smi -L | head -4
timeout -s KILL 900 python -m pytest tests/test_gpu_tp.py -m gpu -x -q --timeout=600 -p no:cacheprovider > gpurun_out/c13_tp_tests.log 2>&1
echo "rc=$?" >> gpurun_out/c13_tp_tests.log; tail -15 gpurun_out/c13_tp_tests.log
TAIL=5 ./run_tp.sh 2 900 c13_bench_tp2 bench.py --gpus 2 --steps 3 --warmup 3
