"""Where the attention kernel and its neighbouring GEMVs are on the device, per CTA (%globaltimer), in the chain
q (c_attn GEMV) -> A (decode attention) -> p (attn.c_proj GEMV) of the 7B model - to explain the cost of overlapping the
attention kernel with its neighbours.  Needs the tuning build with the stamps:

    python lit-llama-ja_b200/build.py --variant loopclk LLB_LOOPCLK=1 LLB_TRACE=1
    LLB200_LIB=.../libllb200_loopclk.so [LLB200_PDL_ATTN=1 | LLB200_ATTN_TRIGGER=1|2] python profiles/attn_trace.py
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
cfg = model.config
apdl = os.environ.get("LLB200_PDL_ATTN", "0") == "1"
flags = NV.GEMV_PDL if apdl else 0
NLAY = 8
gbuf = torch.zeros((NLAY * 2, 256, 16), dtype=torch.int64, device=dev)
abuf = torch.zeros((NLAY, 512, 4), dtype=torch.int64, device=dev)
NV.lib.ll_debug_gemv_trace.argtypes = [C.c_void_p, C.c_int]
NV.lib.ll_debug_attn_trace.argtypes = [C.c_void_p, C.c_int]


def seq():
    for layer in eng.layers[:NLAY]:
        NV.gemv(layer[0])
        NV.call("ll_decode_attn", NV.ptr(eng.q), NV.ptr(layer[1]), NV.ptr(layer[2]), NV.ptr(eng.pos), NV.ptr(eng.att), NV.LL_BF16,
                1, cfg.n_head, cfg.head_size, S, eng.splits, NV.ptr(eng.attn_scratch), NV.ptr(eng.attn_counters), flags, NV.stream())
        NV.gemv(layer[3])


s = torch.cuda.Stream()
with torch.cuda.stream(s):
    seq()
torch.cuda.synchronize()
NV.lib.ll_debug_gemv_trace(C.c_void_p(gbuf.data_ptr()), NLAY * 2)
NV.lib.ll_debug_attn_trace(C.c_void_p(abuf.data_ptr()), NLAY)
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    seq()
NV.lib.ll_debug_gemv_trace(None, 0)
NV.lib.ll_debug_attn_trace(None, 0)
for _ in range(3):
    g.replay()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); e0.record()
for _ in range(10):
    g.replay()
e1.record(); torch.cuda.synchronize()
print(f"mode: attention PDL={apdl} trigger={os.environ.get('LLB200_ATTN_TRIGGER', '0')}: {1e3 * e0.elapsed_time(e1) / 10 / NLAY:.2f} us per layer (q, A, p)")
gt = gbuf.cpu().numpy().astype(np.float64)[:, :148, :]
at = abuf.cpu().numpy().astype(np.float64)[:, :256, :]


def rng(a):
    a = a[a > 0]
    return f"{a.min():6.2f}/{a.mean():6.2f}/{a.max():6.2f}" if a.size else "   -"


for L in (4, 5):
    q, p, a = gt[2 * L], gt[2 * L + 1], at[L]
    t0 = q[:, 0].min()
    f = lambda x: (x - t0) / 1e3
    print(f"layer {L} (us after the first q CTA started; min/mean/max over CTAs)")
    print(f"  q  start {rng(f(q[:, 0]))}  wait-ret {rng(f(q[:, 2]))}  x-staged {rng(f(q[:, 3]))}  loop-end {rng(f(q[:, 4]))}  end {rng(f(q[:, 6]))}")
    print(f"  A  start {rng(f(a[:, 0]))}  after-wait {rng(f(a[:, 1]))}  kv-done {rng(f(a[:, 2]))}  combined {rng(f(a[:, 3]))}")
    print(f"  p  start {rng(f(p[:, 0]))}  wait-ret {rng(f(p[:, 2]))}  x-staged {rng(f(p[:, 3]))}  loop-end {rng(f(p[:, 4]))}  end {rng(f(p[:, 6]))}")
    for name, k in (("q", q), ("p", p)):
        sm = k[:, 14].astype(int)
        sm = sm[sm > 0] - 1
        cnt = np.bincount(sm, minlength=148)
        print(f"  {name}: CTAs per SM: {int((cnt == 0).sum())} SMs with none, {int((cnt == 1).sum())} with one, {int((cnt >= 2).sum())} with two or more")
    nq = gt[2 * L + 2] if 2 * L + 2 < gt.shape[0] else None
    if nq is not None:
        print(f"  next q  start {rng(f(nq[:, 0]))}  wait-ret {rng(f(nq[:, 2]))}")
