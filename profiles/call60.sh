mkdir -p gpurun_out
timeout -s KILL 150 python -m pytest tests/test_gpu_config2.py tests/test_checkpoint_stream.py tests/test_gpu_int8.py -m gpu -v --timeout=80 --timeout-method=thread -p no:cacheprovider > gpurun_out/c60_tests.log 2>&1
echo "rc=$?" >> gpurun_out/c60_tests.log
grep -c PASSED gpurun_out/c60_tests.log; grep -E "FAILED|ERROR|rc=|passed|failed" gpurun_out/c60_tests.log | tail -n 12
