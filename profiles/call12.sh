mkdir -p gpurun_out
L=$PWD/lit-llama-ja_b200/lit_llama
export LLB200_LIB=$L/libllb200_loopclk.so
timeout 300 python profiles/attn_trace.py > gpurun_out/c12_attn_plain.txt 2>&1
LLB200_PDL_ATTN=1 timeout 300 python profiles/attn_trace.py > gpurun_out/c12_attn_pdl.txt 2>&1
LLB200_ATTN_TRIGGER=2 timeout 300 python profiles/attn_trace.py > gpurun_out/c12_attn_trig2.txt 2>&1
grep -h "mode\|CTAs per SM" gpurun_out/c12_attn_*.txt
