"""Standalone int4-GEMV microbenchmark / ncu target: the five 7B layer shapes at M=1, random packed
weights (group 128).  Prints per-shape device time and achieved GB/s (CUDA events, L2 defeated by
cycling through >400 MB of distinct weight streams).  argv: reps [shape-name] [M] [grid]"""
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
import numpy as np  # noqa: E402
TAIL = torch.from_numpy(np.array([0xC1C03C24] * 16, dtype=np.uint32).view(np.uint8).copy()).to(dev)  # bf16 (0.01, -24.0)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
only = sys.argv[2] if len(sys.argv) > 2 and sys.argv[2] != "all" else None
M = int(sys.argv[3]) if len(sys.argv) > 3 else 1
grid = int(sys.argv[4]) if len(sys.argv) > 4 else 0

for name, (n, k, nmat) in SHAPES.items():
    if only and only != name:
        continue
    sb = NV.stream_bytes(n, k, GROUP)
    alg = nmat * (n * k // 2 + 4 * n * (k // GROUP)) + 2 * (n + k)
    copies = max(2, int(400e6 // (nmat * sb)) + 1)
    # random stream bytes: nibbles random; tails = (scale 0.01, -(16+8)) so results stay finite
    ws_ = []
    for _ in range(copies):
        w = torch.randint(0, 256, (nmat, sb), dtype=torch.uint8, device=dev)
        blk = w.view(nmat, -1, 1088)
        blk[:, :, 1024:] = TAIL
        ws_.append(w)
    x = torch.randn((M, k), device=dev).bfloat16()
    y = torch.empty((M, n), device=dev, dtype=torch.bfloat16)
    args = []
    for w in ws_:
        a = NV.GemvArgs()
        a.w = w[0].data_ptr()
        if nmat == 2:
            a.w2, a.epilogue = w[1].data_ptr(), NV.EPI_SWIGLU
        a.N, a.K, a.group, a.M, a.x, a.y = n, k, GROUP, M, x.data_ptr(), y.data_ptr()
        a.grid = grid
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
    print(f"{name:12s} N={n:6d} K={k:6d} nmat={nmat} M={M} grid={grid or 'sm'}  {us:8.2f} us  "
          f"{alg / us / 1e3:8.1f} GB/s algorithmic ({nmat * sb / us / 1e3:8.1f} GB/s streamed)", flush=True)
