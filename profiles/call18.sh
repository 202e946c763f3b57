mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests/test_gpu_quant.py -m gpu -x -q --timeout=300 -p no:cacheprovider -k "fused_dequant or large_m" > gpurun_out/c18_fused.log 2>&1; tail -3 gpurun_out/c18_fused.log
timeout 600 python profiles/batch_probe.py 9,16,32,64,128,256 > gpurun_out/c18_batch_fused.txt 2>&1; tail -7 gpurun_out/c18_batch_fused.txt
