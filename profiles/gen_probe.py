"""Where does generate() time go: prefill vs per-token loop (7B int4, 512 + N new tokens)."""
import sys
import time
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "lit-llama-ja_b200")]
import bench  # noqa: E402
import generate as G  # noqa: E402

dev = torch.device("cuda:0")
model = bench.build_model(dev)
prompt = torch.randint(0, 32000, (512,), dtype=torch.int32, device=dev)


def run(n):
    model.reset_cache()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    G.generate(model, prompt, n, max_seq_length=768, temperature=0.8, top_k=200)
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) * 1e3


for _ in range(2):
    run(8)
t1, t2, t3 = min(run(1) for _ in range(3)), min(run(65) for _ in range(3)), min(run(257) for _ in range(2))
print(f"prefill+1 sample: {t1:.1f} ms | 65 tokens: {t2:.1f} ms | 257 tokens: {t3:.1f} ms | per decode token: {(t3 - t2) / 192:.3f} ms")
# prefill breakdown with the torch profiler-free approach: time model() alone
model.reset_cache()
x = prompt.view(1, -1)
pos = torch.arange(512, device=dev)
for _ in range(2):
    model.reset_cache(); model(x, 768, pos)
torch.cuda.synchronize(); t0 = time.perf_counter()
model.reset_cache(); model(x, 768, pos)
torch.cuda.synchronize()
print(f"prefill forward alone: {(time.perf_counter() - t0) * 1e3:.1f} ms")
