"""gptq.int8 (8-bit GPTQ weights, group 128): device time of one fused 7B decode step (CUDA graph, position 640) and of
the GEMV chain alone, against the HBM rate on 6.8 GB of weights; the reference-semantics SIMT route for comparison."""
import json
import os
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "lit-llama-ja_b200")]
import bench  # noqa: E402
from lit_llama import _native as NV  # noqa: E402

dev = torch.device("cuda:0")
model = bench.build_model(dev, mode="gptq.int8")
S, POS = 768, 640
with torch.no_grad():
    model(torch.zeros((1, 4), dtype=torch.int64, device=dev), S, torch.arange(4, device=dev))
eng = model._fused_engine(1, S)
assert eng is not None, "no fused engine for gptq.int8"
eng.pos.copy_(torch.tensor([POS, POS % S, min(POS + 1, S)], dtype=torch.int32))


def timed_graph(fn, reps=20):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    for _ in range(3):
        g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(reps):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    return 1e3 * e0.elapsed_time(e1) / reps


def gemvs():
    for layer in eng.layers:
        for i in (0, 3, 4, 5):
            NV.gemv(layer[i])
    NV.gemv(eng.a_head)


wbytes = sum(m.stream_for(1)[0].numel() for m in model.modules() if hasattr(m, "stream_for"))
peak = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"] if (ROOT / "MEASURED_PEAKS.json").exists() else 6650.0
t_step, t_gemv = timed_graph(eng._run_step), timed_graph(gemvs)
print(json.dumps({"mode": "gptq.int8 g128 7B batch 1", "step_us": t_step, "tokens_per_s": 1e6 / t_step, "gemv_chain_us": t_gemv,
                  "gemv_launches": 129, "avg_launch_us": t_gemv / 129, "weight_stream_bytes": wbytes,
                  "gemv_gbs": wbytes / t_gemv / 1e3, "gemv_frac_of_hbm_peak": wbytes / t_gemv / 1e3 / peak}))
