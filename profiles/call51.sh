mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_quant.py tests/test_gpu_model.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -n 2
timeout 600 python -m pytest tests/test_gpu_config2.py -m gpu -q -x -p no:cacheprovider -k "fused_decode_chain_at_7b" 2>&1 | tail -n 2
python - <<'PY'
import sys, torch
sys.path[:0] = ['.', 'lit-llama-ja_b200']
import bench
from lit_llama import _native as N
dev = torch.device('cuda:0')
for (n, k) in ((12288, 4096), (11008, 4096), (4096, 11008)):
    from lit_llama.quantization import ColBlockQuantizedLinear
    torch.set_default_dtype(torch.bfloat16)
    lin = ColBlockQuantizedLinear(k, n, False, bits=4, tile_cols=128).to(dev)
    torch.set_default_dtype(torch.float32)
    lin.quant_weight.random_(0, 255); lin.scales.fill_(0.01); lin.zeros.fill_(8.0)
    w = lin.get_weight(torch.bfloat16)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ts = []
    for _ in range(5):
        flush.zero_(); torch.cuda.synchronize(); e0.record(); w = lin.get_weight(torch.bfloat16); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    t = min(ts)
    print(f"unpack {n}x{k}: {t:.1f} us, {(n * k * 2.5 + n * k / 32) / t / 1e6:.2f} TB/s (read packed + write bf16)")
PY
LLB200_BENCH_REFGPU=0 LLB200_BENCH_CPU=0 timeout 600 python bench.py --steps 5 --warmup 3 2>/dev/null | cut -c1-200
