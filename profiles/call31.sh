mkdir -p gpurun_out
TP_CHECK_PDL=1 TAIL=16 ./run_tp.sh 2 240 c31_tp_check_ll profiles/tp_check.py ll
TAIL=2 ./run_tp.sh 2 240 c31_tp2_7b_ll profiles/tp_probe.py 7B
LLB200_TP_EXCHANGE=peer TAIL=2 ./run_tp.sh 2 240 c31_tp2_7b_peer profiles/tp_probe.py 7B
TAIL=2 ./run_tp.sh 2 400 c31_tp2_65b_ll profiles/tp_probe.py 65B
timeout 600 python -m pytest tests/test_gpu_tp.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -n 2
