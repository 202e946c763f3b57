mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_config2.py tests/test_callers.py -m gpu -q -x -p no:cacheprovider -k "q8 or gptq_int8 or perplexity or specialised" 2>&1 | tail -n 12
timeout 600 python profiles/q8_probe.py 2>&1 | tail -n 2 | tee gpurun_out/c39_q8_probe.txt
