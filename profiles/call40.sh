mkdir -p gpurun_out
timeout 900 python profiles/parity_7b.py --steps 256 --out gpurun_out/r02_parity_config2_g128.json 2>&1 | tail -n 6
timeout 1200 python profiles/parity_7b.py --per-row --steps 64 --no-fp32 --out gpurun_out/r02_parity_config2_perrow.json 2>&1 | tail -n 6
