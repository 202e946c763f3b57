mkdir -p gpurun_out
timeout 115 python -m pytest tests/test_adapter.py tests/test_gpu_model.py -m gpu -q -p no:cacheprovider 2>&1 | tail -n 40 | tee gpurun_out/c58_tests.log
