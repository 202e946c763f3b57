mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests/test_gpu_config2.py -m gpu -x -q --timeout=600 -p no:cacheprovider -k "writes_nothing or fused_decode_chain_at" > gpurun_out/c20_tests.log 2>&1; tail -3 gpurun_out/c20_tests.log
PROBE_SEQS=qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -1
LLB200_GEMV_PARITY=1 PROBE_APDL=1 PROBE_SEQS=qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -1
LLB200_GEMV_PARITY=1 PROBE_APDL=0 LLB200_PDL_ATTN=0 PROBE_SEQS=qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -1
L=$PWD/lit-llama-ja_b200/lit_llama
LLB200_LIB=$L/libllb200_loopclk.so LLB200_GEMV_PARITY=1 LLB200_PDL_ATTN=1 timeout 300 python profiles/attn_trace.py > gpurun_out/c20_attn_parity_pdl.txt 2>&1; head -8 gpurun_out/c20_attn_parity_pdl.txt
