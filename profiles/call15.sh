mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests/test_gpu_quant.py -m gpu -x -q --timeout=300 -p no:cacheprovider -k "fused_dequant or large_m or gemv_vs_oracle or repack" > gpurun_out/c15_fused.log 2>&1
echo "rc=$?" >> gpurun_out/c15_fused.log; tail -25 gpurun_out/c15_fused.log
