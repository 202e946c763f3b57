mkdir -p gpurun_out
PROBE_SEQS=qpfd,qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -n 1
LLB200_GEMV_WARM=0 PROBE_SEQS=qpfd,qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -n 1
LLB200_LIB=$PWD/lit-llama-ja_b200/lit_llama/libllb200_trace.so timeout 400 python profiles/gemv_trace.py qpfd 2>&1 | tail -n 5 | tee gpurun_out/c28_gemv_trace_warm.txt
timeout 600 python -m pytest tests/test_gpu_config2.py -m gpu -q -x -p no:cacheprovider -k "fused_decode_chain or specialised or writes_nothing" 2>&1 | tail -n 3
