mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests/test_gpu_quant.py tests/test_gpu_model.py tests/test_gpu_config2.py -m gpu -x -q --timeout=600 -p no:cacheprovider > gpurun_out/c7_tests.log 2>&1
echo "rc=$?" >> gpurun_out/c7_tests.log; tail -6 gpurun_out/c7_tests.log
PROBE_SEQS=q,A,p,f,d,qpfd,qApfd timeout 300 python profiles/step_probe.py > gpurun_out/c7_probe.txt 2>&1; tail -2 gpurun_out/c7_probe.txt
L=$PWD/lit-llama-ja_b200/lit_llama
LLB200_LIB=$L/libllb200_loopclk.so timeout 300 python profiles/gemv_trace.py qpfd > gpurun_out/c7_trace.txt 2>&1; tail -5 gpurun_out/c7_trace.txt
