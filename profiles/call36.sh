mkdir -p gpurun_out
L=$PWD/lit-llama-ja_b200/lit_llama
LLB200_LIB=$L/libllb200_prev.so PROBE_SEQS=q,p,f,d,qpfd,qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -n 1
PROBE_SEQS=q,p,f,d,qpfd,qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -n 1
LLB200_GEMV_CLUSTER=0 PROBE_SEQS=q,p,f,d,qpfd,qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -n 1
LLB200_LIB=$L/libllb200_prev.so PROBE_SEQS=q,p,f,d,qpfd,qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -n 1
