mkdir -p gpurun_out
L=$PWD/lit-llama-ja_b200/lit_llama
LLB200_LIB=$L/libllb200_prev.so PROBE_SEQS=q,p,f,d,qpfd,qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -n 1
PROBE_SEQS=q,p,f,d,qpfd,qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -n 1
LLB200_GEMV_CLUSTER=0 PROBE_SEQS=q,p,f,d,qpfd,qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -n 1
timeout 900 python -m pytest tests/test_gpu_config2.py tests/test_gpu_quant.py -m gpu -q -x -p no:cacheprovider -k "not attention_hs128 and not staged_kv" 2>&1 | tail -n 3
