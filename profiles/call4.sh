mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests/test_gpu_config2.py -m gpu -x -q -s --timeout=600 -p no:cacheprovider > gpurun_out/c4_config2.log 2>&1
echo "rc=$?" >> gpurun_out/c4_config2.log
grep -E "config2 parity|passed|failed|Error|rc=" gpurun_out/c4_config2.log | tail -20
timeout -s KILL 900 python profiles/parity_7b.py --out gpurun_out/c4_parity_7b_g128.json > gpurun_out/c4_parity_7b.log 2>&1
tail -25 gpurun_out/c4_parity_7b.log
timeout -s KILL 600 python bench.py --steps 5 --warmup 3 > gpurun_out/c4_bench.json 2> gpurun_out/c4_bench.err
tail -c 3000 gpurun_out/c4_bench.json; tail -5 gpurun_out/c4_bench.err
