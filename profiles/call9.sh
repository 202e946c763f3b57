mkdir -p gpurun_out
timeout -s KILL 600 python bench.py --steps 5 --warmup 3 > gpurun_out/c9_bench.json 2> gpurun_out/c9_bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/c9_bench.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step')}, d['e2e']['value'], d['roofline']['frac'], d['roofline']['avg_launch_us'], d['decode_step']['ms_per_token'], d['decode_step']['frac_of_peak'], d.get('reference_gpu'), d['cpu_baseline']['value'])
PY
PROBE_APDL=1 PROBE_SEQS=qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -1
LLB200_ATTN_TRIGGER=1 PROBE_SEQS=qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -1
LLB200_ATTN_TRIGGER=2 PROBE_SEQS=qApfd timeout 300 python profiles/step_probe.py 2>&1 | tail -1
