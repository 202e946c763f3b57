"""Batched int4 decode (7B, group 128, bf16): ms per step and tokens/s for batch 1..128 at position 640 of a 768-slot cache.
Batch <= 8 runs the fused CUDA-graph engine (int4 GEMV); 9..128 the per-module path whose Linears are the fused tcgen05
de-quantise + GEMM (ll_int4_gemm_tc); LLB200_FUSED_GEMM=0 forces the old route (SIMT <= 16 rows, unpack + GEMM above)."""
import json
import os
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "lit-llama-ja_b200")]
import bench  # noqa: E402
from lit_llama.quantization import ColBlockQuantizedLinear  # noqa: E402

dev = torch.device("cuda:0")
model = bench.build_model(dev)
if os.environ.get("LLB200_FUSED_GEMM") == "0":
    ColBlockQuantizedLinear.FUSED_GEMM_MAX_M = 0
    ColBlockQuantizedLinear.FUSED_GEMM_LOOP_MAX_M = 0
S, POS = 768, 640
wbytes, _ = bench.algorithmic_bytes(model, 0)
out = []
for B in [int(b) for b in (sys.argv[1].split(",") if len(sys.argv) > 1 else "1,8,9,16,32,64,128".split(","))]:
    model.reset_cache()
    model._engine = None
    idx = torch.randint(0, 32000, (B, 1), device=dev)
    ms = float("nan")
    with torch.no_grad():
      if B <= 128:      # (larger batches: the Linears alone - the KV cache of a 768-slot batch-1024 decode does not fit)
        model(torch.randint(0, 32000, (B, 4), device=dev), S, torch.arange(4, device=dev))   # caches
        pos = torch.tensor([POS], device=dev)
        for _ in range(3):
            model(idx, S, pos)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        reps = 10
        for _ in range(reps):
            model(idx, S, pos)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # the quantized Linears alone (one token's worth of layers at this batch)
    x = {k: torch.randn((B, k), device=dev).bfloat16() for k in (4096, 11008)}
    lins = [m for m in model.modules() if isinstance(m, ColBlockQuantizedLinear)]
    with torch.no_grad():
        def run_lins():
            for m in lins:
                m(x[m.in_features])
        st = torch.cuda.Stream()
        st.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(st):
            run_lins()
            run_lins()
        torch.cuda.current_stream().wait_stream(st)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()          # one CUDA graph of the 161 Linears: device time without host launch gaps
        with torch.cuda.graph(g):
            run_lins()
        g.replay()
        torch.cuda.synchronize(); e0.record()
        for _ in range(5):
            g.replay()
        e1.record(); torch.cuda.synchronize()
    lin_ms = e0.elapsed_time(e1) / 5
    row = {"batch": B, "ms_per_step": ms, "tokens_per_s": 1e3 * B / ms, "engine": model._engine is not None,
           "linears_ms": lin_ms, "linears_gbs_on_packed_bytes": wbytes / (lin_ms * 1e-3) / 1e9}
    out.append(row)
    print(json.dumps(row), flush=True)
