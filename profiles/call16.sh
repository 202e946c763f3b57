mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests/test_gpu_quant.py -m gpu -x -q --timeout=300 -p no:cacheprovider -k "fused_dequant or large_m" > gpurun_out/c16_fused.log 2>&1; tail -2 gpurun_out/c16_fused.log
timeout 600 python profiles/batch_probe.py 1,8,9,16,32,64,128 > gpurun_out/c16_batch_fused.txt 2>&1; cat gpurun_out/c16_batch_fused.txt | tail -8
LLB200_FUSED_GEMM=0 timeout 600 python profiles/batch_probe.py 9,16,32,64,128 > gpurun_out/c16_batch_old.txt 2>&1; tail -6 gpurun_out/c16_batch_old.txt
LLB200_FUSED_GEMM_MAX_M=1024 timeout 600 python profiles/batch_probe.py 256,512,1024 > gpurun_out/c16_batch_loop.txt 2>&1; tail -4 gpurun_out/c16_batch_loop.txt
timeout 600 python profiles/batch_probe.py 256,512,1024 > gpurun_out/c16_batch_twopass.txt 2>&1; tail -4 gpurun_out/c16_batch_twopass.txt
