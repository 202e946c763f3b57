mkdir -p gpurun_out
timeout -s KILL 1200 python -m pytest tests/test_gpu_config2.py -m gpu -x -q -s --timeout=600 -p no:cacheprovider > gpurun_out/c3_config2.log 2>&1
echo "rc=$?" >> gpurun_out/c3_config2.log
grep -E "config2 parity|passed|failed|Error|error|rc=" gpurun_out/c3_config2.log | tail -40
