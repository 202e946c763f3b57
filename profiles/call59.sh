mkdir -p gpurun_out
timeout 70 python -m pytest tests/test_adapter.py tests/test_callers.py tests/test_gpu_sampling.py tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider 2>&1 | tail -n 30 | tee gpurun_out/c59_tests.log
timeout 40 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 3 | tee gpurun_out/c59_smoke.log
