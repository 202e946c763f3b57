mkdir -p gpurun_out
L=$PWD/lit-llama-ja_b200/lit_llama
export LLB200_LIB=$L/libllb200_loopclk.so
timeout 300 python profiles/attn_trace.py > gpurun_out/c10_attn_plain.txt 2>&1
LLB200_PDL_ATTN=1 timeout 300 python profiles/attn_trace.py > gpurun_out/c10_attn_pdl.txt 2>&1
LLB200_ATTN_TRIGGER=1 timeout 300 python profiles/attn_trace.py > gpurun_out/c10_attn_trig1.txt 2>&1
LLB200_ATTN_TRIGGER=2 timeout 300 python profiles/attn_trace.py > gpurun_out/c10_attn_trig2.txt 2>&1
LLB200_PDL_ATTN=1 LLB200_ATTN_CLUSTER=0 timeout 300 python profiles/attn_trace.py > gpurun_out/c10_attn_pdl_nocl.txt 2>&1
tail -12 gpurun_out/c10_attn_*.txt
