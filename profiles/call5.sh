mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests/test_gpu_quant.py tests/test_gpu_model.py -m gpu -x -q --timeout=300 -p no:cacheprovider > gpurun_out/c5_quant.log 2>&1
echo "rc=$?" >> gpurun_out/c5_quant.log; tail -15 gpurun_out/c5_quant.log
timeout -s KILL 900 python -m pytest tests/test_gpu_config2.py -m gpu -x -q -s --timeout=600 -p no:cacheprovider > gpurun_out/c5_config2.log 2>&1
echo "rc=$?" >> gpurun_out/c5_config2.log
grep -E "config2 parity|passed|failed|Error|rc=" gpurun_out/c5_config2.log | tail -20
PROBE_SEQS=q,A,p,f,d,qpfd,qApfd timeout 300 python profiles/step_probe.py > gpurun_out/c5_probe.txt 2>&1; tail -3 gpurun_out/c5_probe.txt
LLB200_GEMV_KIND=0 PROBE_SEQS=q,p,f,d,qpfd timeout 300 python profiles/step_probe.py > gpurun_out/c5_probe_bf16.txt 2>&1; tail -2 gpurun_out/c5_probe_bf16.txt
