mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_config2.py -m gpu -q -x -p no:cacheprovider -k "digit_planes or fused_decode_chain_at_7b or gptq_int8" 2>&1 | tail -n 3
PROBE_SEQS=qAp,qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -n 1
LLB200_ATTN_DIGITS=0 PROBE_SEQS=qAp,qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -n 1
