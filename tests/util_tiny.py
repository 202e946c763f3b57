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


# --------------------------------------------------------------------------- LLaMA-Adapter
ADAPTER_TINY = dict(block_size=64, vocab_size=128, n_layer=3, n_head=4, n_embd=128, adapter_prompt_length=10,
                    adapter_start_layer=1)


def adapter_tiny_weights(cfg=ADAPTER_TINY, v2=False):
    """State dict (reference key names) of the tiny adapter model: the plain model's tensors, the adaption prompts and
    NON-zero gating factors for the adapted layers, and - for v2 - a bias and a scale per Linear.
    tests/golden/make_golden_adapter.py loads exactly this into the reference."""
    base = {k: cfg[k] for k in ("block_size", "vocab_size", "n_layer", "n_head", "n_embd")}
    W = tiny_weights(base)
    r = np.random.RandomState(4321)
    for i in range(cfg["adapter_start_layer"], cfg["n_layer"]):
        p = f"transformer.h.{i}.attn."
        W[p + "adapter_wte.weight"] = (r.randn(cfg["adapter_prompt_length"], cfg["n_embd"]) * 0.5).astype(np.float32)
        W[p + "gating_factor"] = (r.randn(1, cfg["n_head"], 1, 1) * 0.7).astype(np.float32)
    if v2:
        for n in linear_names(base):
            out = W[n + ".weight"].shape[0]
            W[n + ".adapter_bias"] = (r.randn(out) * 0.05).astype(np.float32)
            W[n + ".adapter_scale"] = (1 + 0.1 * r.randn(out)).astype(np.float32)
    return W


def oracle_adapter_model(dtype="fp32", v2=False, linears=None, cfg=ADAPTER_TINY):
    """Oracle adapter model over ``adapter_tiny_weights``; ``linears`` (name -> oracle linear) replaces the fp Linears,
    e.g. by quantized ones."""
    from oracle import adapter as OA

    W = adapter_tiny_weights(cfg, v2)
    base = {k: cfg[k] for k in ("block_size", "vocab_size", "n_layer", "n_head", "n_embd")}
    lin = dict(linears) if linears is not None else {n: OM.FLinear(W[n + ".weight"]) for n in linear_names(base)}
    if v2:
        lin = {n: OA.V2Linear(l, W[n + ".adapter_bias"], W[n + ".adapter_scale"]) for n, l in lin.items()}
    norms = {k: v for k, v in W.items() if k.endswith(".scale")}
    start = cfg["adapter_start_layer"]
    return OA.AdapterLLaMA(OM.Config(**base), lin, W["transformer.wte.weight"], norms, dtype,
                           adapter_wte={i: W[f"transformer.h.{i}.attn.adapter_wte.weight"] for i in range(start, cfg["n_layer"])},
                           gating_factor={i: W[f"transformer.h.{i}.attn.gating_factor"].reshape(-1) for i in range(start, cfg["n_layer"])},
                           adapter_start_layer=start)
