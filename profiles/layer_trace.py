"""Timeline of one decoder block inside the real fused decode step (7B, batch 1, position 640): %globaltimer stamps per CTA
of the four GEMVs and the attention kernel (min / mean / max over CTAs, us after the block's first CTA started).

    python lit-llama-ja_b200/build.py --variant trace LLB_TRACE=1
    LLB200_LIB=$PWD/lit-llama-ja_b200/lit_llama/libllb200_trace.so python profiles/layer_trace.py
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
NLAY = 8
gbuf = torch.zeros((NLAY * 4, 256, 16), dtype=torch.int64, device=dev)
abuf = torch.zeros((NLAY, 512, 4), dtype=torch.int64, device=dev)
NV.lib.ll_debug_gemv_trace.argtypes = [C.c_void_p, C.c_int]
NV.lib.ll_debug_attn_trace.argtypes = [C.c_void_p, C.c_int]

s = torch.cuda.Stream()
with torch.cuda.stream(s):
    eng._run_step()
torch.cuda.synchronize()
NV.lib.ll_debug_gemv_trace(C.c_void_p(gbuf.data_ptr()), NLAY * 4)
NV.lib.ll_debug_attn_trace(C.c_void_p(abuf.data_ptr()), NLAY)
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    eng._run_step()
NV.lib.ll_debug_gemv_trace(None, 0)
NV.lib.ll_debug_attn_trace(None, 0)
for _ in range(3):
    g.replay()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); e0.record()
for _ in range(10):
    g.replay()
e1.record(); torch.cuda.synchronize()
print(f"full step {1e3 * e0.elapsed_time(e1) / 10:.1f} us (stamps compiled in)")
gt = gbuf.cpu().numpy().astype(np.float64)[:, :148, :]
at = abuf.cpu().numpy().astype(np.float64)[:, :256, :]


def rng(a):
    a = a[a > 0]
    return f"{a.min():6.2f}/{a.mean():6.2f}/{a.max():6.2f}" if a.size else "   -"


for L in (4, 5):
    ks = {n: gt[4 * L + i] for i, n in enumerate("qpfd")}
    a = at[L]
    t0 = ks["q"][:, 0][ks["q"][:, 0] > 0].min()
    f = lambda x: np.where(x > 0, (x - t0) / 1e3, 0)
    print(f"block {L} (us after its first c_attn CTA started; min/mean/max over CTAs)")
    for n in "qApfd":
        if n == "A":
            print(f"  A  start {rng(f(a[:, 0]))}  after-wait {rng(f(a[:, 1]))}  rows-done {rng(f(a[:, 2]))}  combined {rng(f(a[:, 3]))}")
        else:
            k = ks[n]
            print(f"  {n}  start {rng(f(k[:, 0]))}  wait-ret {rng(f(k[:, 2]))}  x-staged {rng(f(k[:, 3]))}  loop-end {rng(f(k[:, 4]))}  end {rng(f(k[:, 6]))}")
    nq = gt[4 * L + 4]
    print(f"  next q  start {rng(f(nq[:, 0]))}  wait-ret {rng(f(nq[:, 2]))}   -> block time {(nq[:, 2][nq[:, 2] > 0].min() - ks['q'][:, 2][ks['q'][:, 2] > 0].min()) / 1e3:.2f} us (wait-ret to wait-ret)")
