"""Device time of the 512-token prefill of the 7B int4 model (module path: unpack + tcgen05 GEMM per Linear, library SDPA)."""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "lit-llama-ja_b200")]
import bench  # noqa: E402

dev = torch.device("cuda:0")
model = bench.build_model(dev)
idx = torch.randint(0, 32000, (1, 512), device=dev)
pos = torch.arange(512, device=dev)
ts = []
with torch.no_grad():
    for i in range(6):
        model.reset_cache()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        model(idx, 768, pos)
        e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
print(f"prefill 512 tokens: {min(ts[2:]):.2f} ms (runs: {' '.join(f'{t:.2f}' for t in ts)})")

# the same forward captured in a CUDA graph: device time without the host's launch gaps
s = torch.cuda.Stream()
s.wait_stream(torch.cuda.current_stream())
with torch.no_grad(), torch.cuda.stream(s):
    model(idx, 768, pos)
torch.cuda.current_stream().wait_stream(s)
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.no_grad(), torch.cuda.graph(g):
    out = model(idx, 768, pos)
ts = []
for i in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
print(f"prefill 512 tokens as one graph replay: {min(ts):.2f} ms")
