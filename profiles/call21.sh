mkdir -p gpurun_out
PROBE_SEQS=qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -1
LLB200_GEMV_PARITY=1 PROBE_APDL=1 PROBE_SEQS=qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -1
LLB200_GEMV_PARITY=1 LLB200_ATTN_TRIGGER=2 PROBE_APDL=1 PROBE_SEQS=qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -1
L=$PWD/lit-llama-ja_b200/lit_llama
LLB200_LIB=$L/libllb200_loopclk.so LLB200_GEMV_PARITY=1 LLB200_PDL_ATTN=1 timeout 300 python profiles/attn_trace.py > gpurun_out/c21_attn_parity_pdl.txt 2>&1; head -8 gpurun_out/c21_attn_parity_pdl.txt
