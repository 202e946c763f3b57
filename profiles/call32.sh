mkdir -p gpurun_out
TP_CHECK_PDL=1 TAIL=30 ./run_tp.sh 8 300 c32_tp8_check_ll profiles/tp_check.py ll
TAIL=2 ./run_tp.sh 8 300 c32_tp8_65b_ll profiles/tp_probe.py 65B
TAIL=2 ./run_tp.sh 8 200 c32_tp8_7b_ll profiles/tp_probe.py 7B
