mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_checkpoint_stream.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -n 3
export LLB200_BENCH_REFGPU=0 LLB200_BENCH_CPU=0
python bench.py --steps 1 --warmup 3 > gpurun_out/c38_plain.log 2>&1 &&
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -s 1200 -c 500 --csv --log-file gpurun_out/r02_bench_decode_launches.csv python bench.py --steps 1 --warmup 3 > gpurun_out/c38_ncu1.log 2>&1
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:"gemv_w4a8|decode_attn_pf" -s 1200 -c 6 -o gpurun_out/r02_decode_full python bench.py --steps 1 --warmup 3 > gpurun_out/c38_ncu2.log 2>&1
tail -2 gpurun_out/c38_plain.log | cut -c1-300; tail -2 gpurun_out/c38_ncu1.log | cut -c1-200; tail -2 gpurun_out/c38_ncu2.log | cut -c1-200; ls -la gpurun_out/r02_*
