"""Shared builders for the tiny LLaMA (BASELINE.json configs[0]) used by CPU and GPU tests.
``tiny_weights`` replays tests/golden/make_golden.py's generator (numpy legacy RNG)."""
import numpy as np

from oracle import model as OM

TINY = dict(block_size=64, vocab_size=100, n_layer=2, n_head=4, n_embd=128)


def tiny_weights(cfg=TINY):
    r = np.random.RandomState(1234)
    E, L = cfg["n_embd"], cfg["n_layer"]
    V = cfg["vocab_size"] + (-cfg["vocab_size"]) % 64
    H = int(2 * 4 * E / 3)
    H = H + (-H) % 256
    sd = {"lm_head.weight": (V, E), "transformer.wte.weight": (V, E)}
    for i in range(L):
        p = f"transformer.h.{i}."
        sd[p + "attn.c_attn.weight"] = (3 * E, E)
        sd[p + "attn.c_proj.weight"] = (E, E)
        sd[p + "mlp.c_fc1.weight"] = (H, E)
        sd[p + "mlp.c_fc2.weight"] = (H, E)
        sd[p + "mlp.c_proj.weight"] = (E, H)
    out = {k: (r.randn(*s) * (0.5 if "wte" in k else 0.08)).astype(np.float32) for k, s in sd.items()}
    for i in range(L):
        for n in ("rms_1", "rms_2"):
            out[f"transformer.h.{i}.{n}.scale"] = (1 + 0.1 * r.randn(E)).astype(np.float32)
    out["transformer.ln_f.scale"] = (1 + 0.1 * r.randn(E)).astype(np.float32)
    return out


def linear_names(cfg=TINY):
    names = ["lm_head"]
    for i in range(cfg["n_layer"]):
        p = f"transformer.h.{i}."
        names += [p + "attn.c_attn", p + "attn.c_proj", p + "mlp.c_fc1", p + "mlp.c_fc2", p + "mlp.c_proj"]
    return names


def oracle_fp_model(dtype="fp32"):
    W = tiny_weights()
    lin = {n: OM.FLinear(W[n + ".weight"]) for n in linear_names()}
    norms = {k: v for k, v in W.items() if k.endswith(".scale")}
    return OM.LLaMA(OM.Config(**TINY), lin, W["transformer.wte.weight"], norms, dtype)


def oracle_q_model(g, prefix, tile_cols, dtype):
    """Oracle model over quantized state stored in the golden npz under ``prefix``."""
    W = tiny_weights()
    lin = {}
    for n in linear_names():
        if prefix.startswith("gptq"):
            qw, s, z = g[f"{prefix}{n}.quant_weight"], g[f"{prefix}{n}.scales"], g[f"{prefix}{n}.zeros"]
        else:
            qw, s, z = g[f"{prefix}{n}.packed"], g[f"{prefix}{n}.scales"], g[f"{prefix}{n}.zeros"]
        lin[n] = OM.QLinear(qw, s, z, 4, tile_cols)
    norms = {k: v for k, v in W.items() if k.endswith(".scale")}
    return OM.LLaMA(OM.Config(**TINY), lin, W["transformer.wte.weight"], norms, dtype)
