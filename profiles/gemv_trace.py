"""Stage time stamps (globaltimer) of the GEMV kernels inside a captured q,p,f,d x32 chain (7B, B=1).

Needs the tuning build with the stamps compiled in:
    python lit-llama-ja_b200/build.py --variant trace LLB_TRACE=1
    LLB200_LIB=$PWD/lit-llama-ja_b200/lit_llama/libllb200_trace.so python profiles/gemv_trace.py qpfd
"""
import ctypes as C
import os
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
for _ in range(3):
    g.replay()
torch.cuda.synchronize()
t = buf.cpu().numpy().astype(np.float64)[:, :148, :]  # [launch][cta][stage] ns
names = ["start", "pre-wait", "wait-ret", "x-staged", "loop-end", "reduced", "end", "rnorm", "nw-landed", "x-landed", "x-math"]
order = [0, 1, 2, 7, 8, 9, 10, 3, 4, 5, 6]
base = t[:, :, 0].min()
print("launch kind | kernel span (min start -> max end) | per-stage mean over CTAs relative to this kernel's min start (us)")
for i in range(len(spec) * 2, len(spec) * 4):
    k = spec[i % len(spec)]
    st0 = t[i, :, 0].min()
    rel = (t[i] - st0) / 1e3
    prev_end = (t[i - 1, :, 6].max() - st0) / 1e3
    row = " ".join(f"{names[j]}={rel[:, j].mean():6.2f}" for j in order if t[i, 0, j] > 0)
    print(f"{i:3d} {k} | start@{(st0 - base) / 1e3:8.2f} prev_end@{prev_end:6.2f} | {row}")
