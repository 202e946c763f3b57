"""Standalone int4-GEMV microbenchmark / ncu target: the five 7B layer shapes at M=1, random packed
weights (group 128).  Prints per-shape device time and achieved GB/s (CUDA events, L2 flushed by
cycling through enough distinct weight copies)."""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "lit-llama-ja_b200"))
from lit_llama import _native as NV  # noqa: E402

dev = torch.device("cuda:0")
SHAPES = {"c_attn": (12288, 4096, 1), "attn.c_proj": (4096, 4096, 1), "fc1|fc2": (11008, 4096, 2),
          "mlp.c_proj": (4096, 11008, 1), "lm_head": (32000, 4096, 1)}
GROUP = 128
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
only = sys.argv[2] if len(sys.argv) > 2 else None
M = int(sys.argv[3]) if len(sys.argv) > 3 else 1

for name, (n, k, nmat) in SHAPES.items():
    if only and only != name:
        continue
    bytes_one = nmat * (n * k // 2 + 4 * n * (k // GROUP))
    copies = max(2, int(400e6 // bytes_one) + 1)          # > 3x L2 worth of distinct weights
    qws = [torch.randint(0, 256, (nmat, k // 2, n), dtype=torch.uint8, device=dev) for _ in range(copies)]
    s = (torch.rand((nmat, k // GROUP, n), device=dev) * 0.01 + 0.001).bfloat16().view(torch.int16).int() & 0xffff
    z = (torch.randint(0, 16, (nmat, k // GROUP, n), device=dev) + 128).bfloat16().view(torch.int16).int() & 0xffff
    sz = (s | (z << 16)).int().contiguous()
    x = torch.randn((M, k), device=dev).bfloat16()
    y = torch.empty((M, n), device=dev, dtype=torch.bfloat16)
    ks, pb, cb = NV.gemv_plan(n, k, GROUP, nmat, M)
    ws = NV.Workspace.get(dev).ensure(pb, cb)
    args = []
    for qw in qws:
        a = NV.GemvArgs()
        a.qw, a.sz = qw[0].data_ptr(), sz[0].data_ptr()
        if nmat == 2:
            a.qw2, a.sz2, a.epilogue = qw[1].data_ptr(), sz[1].data_ptr(), NV.EPI_SWIGLU
        a.N, a.K, a.group, a.M, a.x, a.y = n, k, GROUP, M, x.data_ptr(), y.data_ptr()
        a.partials, a.counters, a.ksplit = ws.partials.data_ptr(), ws.counters.data_ptr(), ks
        args.append(a)
    for a in args[:2]:
        NV.gemv(a)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        NV.gemv(args[i % copies])
    e1.record()
    torch.cuda.synchronize()
    us = 1e3 * e0.elapsed_time(e1) / reps
    print(f"{name:12s} N={n:6d} K={k:6d} nmat={nmat} M={M} ksplit={ks:2d}  {us:8.2f} us  {bytes_one / us / 1e3:8.1f} GB/s", flush=True)
