mkdir -p gpurun_out
bash run_gpu_tests.sh 600 2400 > /dev/null 2>&1; tail -n 4 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 1
