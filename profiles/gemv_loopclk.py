"""Where the consumer warps of the int4 GEMV spend the main loop: cycles waiting for a chunk (full barrier) vs total,
per kernel of the q,p,f,d chain (7B, B=1), from clock64 sums compiled in with -DLLB_LOOPCLK.

    python lit-llama-ja_b200/build.py --variant loopclk LLB_LOOPCLK=1 LLB_TRACE=1
    LLB200_LIB=$PWD/lit-llama-ja_b200/lit_llama/libllb200_loopclk.so python profiles/gemv_loopclk.py qpfd
"""
import ctypes as C
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "lit-llama-ja_b200")]
import bench  # noqa: E402
from lit_llama import _native as NV  # noqa: E402

dev = torch.device("cuda:0")
model = bench.build_model(dev)
S = 768
with torch.no_grad():
    model(torch.zeros((1, 4), dtype=torch.int64, device=dev), S, torch.arange(4, device=dev))
eng = model._fused_engine(1, S)
eng.pos.copy_(torch.tensor([640, 640, 641], dtype=torch.int32))
spec = sys.argv[1] if len(sys.argv) > 1 else "qpfd"
NL = len(spec) * 8
buf = torch.zeros((NL, 256, 16), dtype=torch.int64, device=dev)
NV.lib.ll_debug_gemv_trace.argtypes = [C.c_void_p, C.c_int]


def seq():
    for layer in eng.layers[:8]:
        for c in spec:
            NV.gemv(layer[{"q": 0, "p": 3, "f": 4, "d": 5}[c]])


s = torch.cuda.Stream()
with torch.cuda.stream(s):
    seq()
torch.cuda.synchronize()
NV.lib.ll_debug_gemv_trace(C.c_void_p(buf.data_ptr()), NL)
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    seq()
NV.lib.ll_debug_gemv_trace(None, 0)
REPLAYS = 3
for _ in range(REPLAYS):
    g.replay()
torch.cuda.synchronize()
t = buf.cpu().numpy().astype(np.float64)[:, :148, :]
W = 8
print("kernel | chunks/CTA (max) | loop clk per chunk per warp | of which waiting for the chunk | words/lane/chunk=16 -> clk per word per SMSP (2 warps)")
for i in range(len(spec) * 2, len(spec) * 4):
    k = spec[i % len(spec)]
    nch = t[i, :, 13]
    busy = nch > 0
    loop = t[i, busy, 12] / (REPLAYS * W)
    wait = t[i, busy, 11] / (REPLAYS * W)
    per_chunk = loop.sum() / nch[busy].sum()
    wait_chunk = wait.sum() / nch[busy].sum()
    print(f"{i:3d} {k} | {nch.max():4.0f} | {per_chunk:7.1f} | {wait_chunk:7.1f} ({100 * wait_chunk / per_chunk:4.1f} %) | {2 * per_chunk / 16:6.1f}")
