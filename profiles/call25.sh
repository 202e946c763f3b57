mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_config2.py -m gpu -q -x -p no:cacheprovider -k "attention or fused_decode_chain" 2>&1 | tail -n 8
PROBE_SEQS=A,qAp,qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -n 1
LLB200_ATTN_STAGE=0 PROBE_SEQS=A,qAp,qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -n 1
LLB200_ATTN_TRIGGER=2 PROBE_SEQS=A,qAp,qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -n 1
