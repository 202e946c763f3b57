mkdir -p gpurun_out
PROBE_SEQS=q,f,qpfd,qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -n 1
timeout 900 python -m pytest tests/test_gpu_config2.py tests/test_gpu_quant.py -m gpu -q -x -p no:cacheprovider -k "fused_decode_chain or specialised or writes_nothing or q8 or gptq_int8 or programmatic or gemv" 2>&1 | tail -n 3
PROBE_SEQS=q,f,qpfd,qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -n 1
