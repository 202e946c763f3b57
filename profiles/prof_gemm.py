"""tcgen05 GEMM / prefill-linear microbenchmark: TFLOP/s of ll_gemm_bf16_tc and of the whole int4 large-M linear
(unpack + GEMM) on the 7B / 13B layer shapes.  argv: M [reps]"""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "lit-llama-ja_b200"))
from lit_llama import _native as NV  # noqa: E402

dev = torch.device("cuda:0")
M = int(sys.argv[1]) if len(sys.argv) > 1 else 512
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
E, H = (4096, 11008) if len(sys.argv) < 4 else (5120, 13824)
shapes = {"c_attn": (3 * E, E), "c_proj": (E, E), "fc": (H, E), "down": (E, H), "lm_head": (32000, E)}


def timed(fn):
    for _ in range(2):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for name, (n, k) in shapes.items():
    x = torch.randn((M, k), device=dev).bfloat16()
    w = (torch.randn((n, k), device=dev) * 0.02).bfloat16()
    y = torch.empty((M, n), device=dev, dtype=torch.bfloat16)
    qw = torch.randint(0, 256, (k // 2, n), dtype=torch.uint8, device=dev)
    sc = (torch.rand((n, k // 128), device=dev) * 0.01).bfloat16()
    zr = torch.randint(0, 16, (n, k // 128), device=dev).bfloat16()
    t_g = timed(lambda: NV.call("ll_gemm_bf16_tc", NV.ptr(x), NV.ptr(w), None, NV.ptr(y), M, n, k, NV.stream()))
    t_u = timed(lambda: NV.call("ll_get_weight", NV.ptr(qw), NV.ptr(sc), NV.ptr(zr), NV.LL_BF16, NV.ptr(w), NV.LL_BF16, n, k, 4, 128, NV.stream()))
    t_c = timed(lambda: torch.matmul(x, w.t()))
    fl = 2.0 * M * n * k
    print(f"{name:8s} M={M} N={n} K={k}: tcgen05 GEMM {t_g * 1e3:8.1f} us {fl / t_g / 1e9:7.1f} TFLOP/s | cuBLAS {t_c * 1e3:8.1f} us "
          f"{fl / t_c / 1e9:7.1f} | unpack {t_u * 1e3:7.1f} us {(n * k * 2.5) / t_u / 1e6:7.1f} GB/s", flush=True)
