mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tp.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -n 4
TAIL=2 ./run_tp.sh 2 240 c29_tp2_7b_pdl profiles/tp_probe.py 7B
LLB200_TP_PDL=0 TAIL=2 ./run_tp.sh 2 240 c29_tp2_7b_plain profiles/tp_probe.py 7B
TAIL=2 ./run_tp.sh 2 400 c29_tp2_65b_pdl profiles/tp_probe.py 65B
LLB200_TP_PDL=0 TAIL=2 ./run_tp.sh 2 400 c29_tp2_65b_plain profiles/tp_probe.py 65B
TAIL=12 ./run_tp.sh 2 240 c29_tp_check profiles/tp_check.py
