"""The reference's CPU path for a quantized Linear, restated with torch CPU ops so that it runs
multi-threaded exactly like the reference does on a host (TEST/BENCH INFRASTRUCTURE ONLY).

ref: ColBlockQuantizedLinear.get_weight + forward fallback, lit_llama/quantization.py:390-421 - the
same op sequence: per-nibble strided writes, then per group ``-= zeros`` and ``*= scales`` in place,
then ``F.linear``.  Used by bench.py for ``cpu_baseline`` / ``--impl reference``; checked against
oracle/quant.py in tests/test_oracle_golden.py."""
import torch


def get_weight(qw: torch.Tensor, scales: torch.Tensor, zeros: torch.Tensor, bits: int, tile: int, dtype=torch.float32):
    n, kp = qw.shape
    epb = 8 // bits
    w = torch.empty((n, kp * epb), dtype=dtype)
    mask = (1 << bits) - 1
    for nr in range(epb):
        w[:, nr::epb] = ((qw >> (nr * bits)) & mask).float()
    for j in range(scales.shape[1]):
        w[:, j * tile:(j + 1) * tile] -= zeros[:, j:j + 1]
        w[:, j * tile:(j + 1) * tile] *= scales[:, j:j + 1]
    return w


def linear_forward(x, qw, scales, zeros, bits: int, tile: int, bias=None):
    return torch.nn.functional.linear(x, get_weight(qw, scales, zeros, bits, tile, x.dtype), bias)


def decode_token_seconds(n_embd: int, n_hidden: int, vocab: int, n_layer: int, group: int, repeats: int = 1, seed: int = 0):
    """Wall time of the quantized Linears of ONE decoded token (M=1, fp32 as the reference runs on
    a CPU: generate.py:121 picks 32-true without CUDA), measured on one block + lm_head and scaled
    to ``n_layer`` blocks.  Returns (seconds_per_token, sample description)."""
    import time

    g = torch.Generator().manual_seed(seed)
    shapes = {"c_attn": (3 * n_embd, n_embd), "attn.c_proj": (n_embd, n_embd), "c_fc1": (n_hidden, n_embd),
              "c_fc2": (n_hidden, n_embd), "mlp.c_proj": (n_embd, n_hidden), "lm_head": (vocab, n_embd)}
    layers = {}
    for name, (n, k) in shapes.items():
        qw = torch.randint(0, 256, (k // 2, n), dtype=torch.uint8, generator=g).t()   # the (1, N) strides
        s = torch.rand((n, k // group), generator=g) * 0.01 + 0.001
        z = torch.randint(0, 16, (n, k // group), generator=g).float()
        layers[name] = (qw, s, z, torch.randn((1, k), generator=g))
    best_block = best_head = float("inf")
    for _ in range(repeats):
        t0 = time.perf_counter()
        for name in ("c_attn", "attn.c_proj", "c_fc1", "c_fc2", "mlp.c_proj"):
            qw, s, z, x = layers[name]
            linear_forward(x, qw, s, z, 4, group)
        t1 = time.perf_counter()
        qw, s, z, x = layers["lm_head"]
        linear_forward(x, qw, s, z, 4, group)
        t2 = time.perf_counter()
        best_block, best_head = min(best_block, t1 - t0), min(best_head, t2 - t1)
    sample = (f"quantized Linears of one decoded token: 1 transformer block (5 layers) + lm_head timed "
              f"{repeats}x (best), block time x{n_layer}; attention/norm/sampling excluded (favours the CPU)")
    return best_block * n_layer + best_head, sample
