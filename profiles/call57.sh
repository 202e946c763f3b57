mkdir -p gpurun_out
export LLB200_BENCH_REFGPU=0 LLB200_BENCH_CPU=0
python bench.py --steps 1 --warmup 3 > gpurun_out/c57_plain.log 2>&1 &&
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -s 1200 -c 500 --csv --log-file gpurun_out/r02_bench_decode_launches_final.csv python bench.py --steps 1 --warmup 3 > gpurun_out/c57_ncu1.log 2>&1
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:"gemv_w4a8|decode_attn_pf" -s 1200 -c 6 -o gpurun_out/r02_decode_full_final python bench.py --steps 1 --warmup 3 > gpurun_out/c57_ncu2.log 2>&1
tail -2 gpurun_out/c57_plain.log | cut -c1-300; tail -2 gpurun_out/c57_ncu1.log | cut -c1-200; tail -2 gpurun_out/c57_ncu2.log | cut -c1-200; ls -la gpurun_out/r02_*final*
