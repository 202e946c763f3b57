mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_config2.py -m gpu -q -x -p no:cacheprovider -k "l2_prefetch or fused_decode_chain_at_7b" 2>&1 | tail -n 6
timeout 900 python profiles/l2pf_probe.py > gpurun_out/c23_l2pf.txt 2>&1; cat gpurun_out/c23_l2pf.txt | tail -n 12
