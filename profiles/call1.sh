mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/c1_smi.txt
( cd profiles/micro && timeout 120 ./loop_mix ) > gpurun_out/c1_loop_mix.txt 2>&1
( cd profiles/micro && timeout 60 ./tmem_a_mma 16 0; timeout 60 ./tmem_a_mma 8 0; timeout 60 ./tmem_a_mma 16 128 ) > gpurun_out/c1_tmem.txt 2>&1
bash run_gpu_tests.sh 120 900 > gpurun_out/c1_pytest_tail.txt 2>&1
PROBE_SEQS=q,A,p,f,d,qpfd,qApfd timeout 300 python profiles/step_probe.py > gpurun_out/c1_probe_base.txt 2>&1
LLB200_GEMV_STREAMK=1 PROBE_SEQS=q,p,f,d,qpfd timeout 300 python profiles/step_probe.py > gpurun_out/c1_probe_sk.txt 2>&1
LLB200_GEMV_STREAMK=1 timeout 300 python -m pytest tests/test_gpu_quant.py -m gpu -q -k "streamk or stream_k or STREAMK or split" -p no:cacheprovider > gpurun_out/c1_sk_tests.txt 2>&1
tail -3 gpurun_out/c1_*.txt
