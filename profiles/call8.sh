mkdir -p gpurun_out
( cd profiles/micro && RATES_ONLY=1 timeout 120 ./loop_mix ) 2>&1 | grep -E "F2I|I2F|HMNMX|HMUL2" > gpurun_out/c8_rates.txt; cat gpurun_out/c8_rates.txt
L=$PWD/lit-llama-ja_b200/lit_llama
for sp in qq ff dd pp; do LLB200_LIB=$L/libllb200_loopclk.so timeout 300 python profiles/gemv_trace.py $sp 2>&1 | tail -4 > gpurun_out/c8_trace_$sp.txt; cat gpurun_out/c8_trace_$sp.txt; done
