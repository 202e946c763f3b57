"""GPTQ solver timing on B200: one 7B-sized Linear, 128 calibration samples of 2048 tokens folded into the Hessian
in chunks; reports the Hessian GEMM, the cuSOLVER factorisation and the column walk (ll_gptq_block + cross-block GEMM)."""
import sys
import time
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "lit-llama-ja_b200")]
from lit_llama import quantization as Qz  # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(0)


def sync_time(fn):
    torch.cuda.synchronize(); t0 = time.perf_counter(); r = fn(); torch.cuda.synchronize()
    return r, (time.perf_counter() - t0) * 1e3


for N_, K, groupsize in ((4096, 4096, -1), (4096, 4096, 128), (11008, 4096, 128), (4096, 11008, 128)):
    lin = torch.nn.Linear(K, N_, bias=False, device=dev)
    with torch.no_grad():
        lin.weight.mul_(0.3)
    q = Qz.GPTQQuantizer(lin, bits=4, groupsize=groupsize, actorder=(groupsize == -1))
    x = torch.randn(8, 2048, K, device=dev)
    _, t_h = sync_time(lambda: [q.collect_input_stats(None, (x[i:i + 1],), None) for i in range(8)])
    # time the pieces of quantize() separately on copies
    H = q.H.clone()
    def factor():
        Hd = H.clone()
        Hd.diagonal().add_(0.01 * Hd.diagonal().mean())
        return torch.linalg.cholesky(torch.cholesky_inverse(torch.linalg.cholesky(Hd)), upper=True).contiguous()
    factor()
    Hinv, t_f = sync_time(factor)
    W = lin.weight.detach().float().clone()
    G = 1 if groupsize == -1 else K // groupsize
    s, z = Qz.find_params_weight(W, 15)
    sc = torch.zeros((N_, G), device=dev); zr = torch.zeros((N_, G), device=dev); sc[:] = s; zr[:] = z
    Qz.gptq_solve(W.clone(), Hinv, sc.clone(), zr.clone(), bits=4, groupsize=groupsize)
    _, t_w = sync_time(lambda: Qz.gptq_solve(W.clone(), Hinv, sc, zr, bits=4, groupsize=groupsize))
    (qm, err), t_all = sync_time(q.quantize)
    print(f"[{N_},{K}] group {groupsize}: Hessian of 8x2048 tokens {t_h:.1f} ms | factor {t_f:.1f} ms | column walk {t_w:.1f} ms "
          f"({K // 128} blocks) | quantize() total {t_all:.1f} ms | error {err:.3g}", flush=True)
