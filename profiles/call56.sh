mkdir -p gpurun_out
bash run_gpu_tests.sh 600 2400 > /dev/null 2>&1; tail -n 4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/c56_bench.json 2> gpurun_out/c56_bench.err; cut -c1-200 gpurun_out/c56_bench.json; grep -o '"roofline": {[^}]*}' gpurun_out/c56_bench.json; grep -o '"decode_step": {[^}]*}' gpurun_out/c56_bench.json; grep -o '"e2e": {[^}]*}' gpurun_out/c56_bench.json
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 1
