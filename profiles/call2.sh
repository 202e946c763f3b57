mkdir -p gpurun_out
L=$PWD/lit-llama-ja_b200/lit_llama
LLB200_LIB=$L/libllb200_loopclk.so timeout 300 python profiles/gemv_loopclk.py qpfd > gpurun_out/c2_loopclk.txt 2>&1
LLB200_LIB=$L/libllb200_loopclk.so timeout 300 python profiles/gemv_trace.py qpfd > gpurun_out/c2_trace.txt 2>&1
tail -12 gpurun_out/c2_loopclk.txt; tail -10 gpurun_out/c2_trace.txt
