mkdir -p gpurun_out
timeout 600 python profiles/batch_probe.py 9,16,32,64,128 > gpurun_out/c17_batch_fused.txt 2>&1; tail -6 gpurun_out/c17_batch_fused.txt
LLB200_FUSED_GEMM=0 timeout 600 python profiles/batch_probe.py 32,128 > gpurun_out/c17_batch_old.txt 2>&1; tail -3 gpurun_out/c17_batch_old.txt
LLB200_FUSED_GEMM_MAX_M=1024 timeout 600 python profiles/batch_probe.py 256,512,1024 > gpurun_out/c17_batch_loop.txt 2>&1; tail -4 gpurun_out/c17_batch_loop.txt
timeout 600 python profiles/batch_probe.py 256,512,1024 > gpurun_out/c17_batch_twopass.txt 2>&1; tail -4 gpurun_out/c17_batch_twopass.txt
