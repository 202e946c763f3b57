mkdir -p gpurun_out
S=/usr/local/cuda/bin/compute-sanitizer
timeout -s KILL 900 $S --tool memcheck --print-limit 20 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c11_memcheck_smoke.txt 2>&1; tail -4 gpurun_out/c11_memcheck_smoke.txt
timeout -s KILL 900 $S --tool racecheck --print-limit 20 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c11_racecheck_smoke.txt 2>&1; tail -4 gpurun_out/c11_racecheck_smoke.txt
timeout -s KILL 1200 $S --tool memcheck --print-limit 20 python -m pytest tests/test_gpu_config2.py -m gpu -x -q -k "specialised or attention_hs128" -p no:cacheprovider > gpurun_out/c11_memcheck_config2.txt 2>&1; tail -4 gpurun_out/c11_memcheck_config2.txt
timeout -s KILL 1200 $S --tool racecheck --print-limit 20 python -m pytest tests/test_gpu_config2.py -m gpu -x -q -k "specialised and 1-" -p no:cacheprovider > gpurun_out/c11_racecheck_config2.txt 2>&1; tail -4 gpurun_out/c11_racecheck_config2.txt
