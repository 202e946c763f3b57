mkdir -p gpurun_out
TAIL=2 ./run_tp.sh 2 240 c30_tp2_7b_pdl profiles/tp_probe.py 7B
LLB200_TP_PDL=0 TAIL=2 ./run_tp.sh 2 240 c30_tp2_7b_plain profiles/tp_probe.py 7B
LLB200_GEMV_KIND=0 TAIL=2 ./run_tp.sh 2 240 c30_tp2_7b_bf16kernels profiles/tp_probe.py 7B
TAIL=2 ./run_tp.sh 2 400 c30_tp2_65b_pdl profiles/tp_probe.py 65B
TP_CHECK_PDL=1 TAIL=12 ./run_tp.sh 2 240 c30_tp_check_pdl profiles/tp_check.py
timeout 600 python -m pytest tests/test_gpu_tp.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -n 2
