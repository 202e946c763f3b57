mkdir -p gpurun_out
export LLB200_BENCH_REFGPU=0 LLB200_BENCH_CPU=0
python bench.py --steps 1 --warmup 3 > gpurun_out/c14_plain.log 2>&1 &&
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_bench_launches.csv python bench.py --steps 1 --warmup 3 > gpurun_out/c14_ncu1.log 2>&1
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:gemv_w4a8 -s 300 -c 4 -o gpurun_out/r02_gemv_i8_full python bench.py --steps 1 --warmup 3 > gpurun_out/c14_ncu2.log 2>&1
tail -3 gpurun_out/c14_plain.log gpurun_out/c14_ncu1.log gpurun_out/c14_ncu2.log; ls -la gpurun_out/r02_*
