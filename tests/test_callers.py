"""The other callers of the quantized-inference boundary: generate/full.py and evaluate/full.py."""
import importlib.util
import math
from pathlib import Path

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
PKG = ROOT / "lit-llama-ja_b200"


def _load(rel, name):
    spec = importlib.util.spec_from_file_location(name, PKG / rel)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_generate_full_surface_and_prompt_template():
    """Flags of the reference's generate/full.py:21-33 and its Alpaca template (scripts/prepare_alpaca.py:111-125)."""
    import inspect

    gf = _load("generate/full.py", "llb_generate_full")
    params = inspect.signature(gf.main).parameters
    for flag in ("prompt", "num_samples", "max_new_tokens", "top_k", "temperature", "checkpoint_path", "tokenizer_path",
                 "model_size", "quantize"):
        assert flag in params, flag
    assert params["top_k"].default == 200 and params["temperature"].default == 0.8 and params["model_size"].default == "7B"
    p = gf.generate_prompt({"instruction": "Say hi", "input": ""})
    assert p.startswith("Below is an instruction that describes a task. ") and p.endswith("### Instruction:\nSay hi\n\n### Response:")
    p2 = gf.generate_prompt({"instruction": "Add", "input": "1 2"})
    assert "paired with an input" in p2 and "### Input:\n1 2\n\n### Response:" in p2
    ev = _load("evaluate/full.py", "llb_evaluate_full")
    for flag in ("checkpoint_path", "tokenizer_path", "model_size", "dtype", "quantize"):
        assert flag in inspect.signature(ev.main).parameters, flag


@pytest.mark.gpu
def test_perplexity_windows_match_the_oracle():
    """evaluate/full.py's loop on the tiny int4 model (windows of 24 tokens, no KV cache: the large-M Linear path) against the
    same windows through the numpy oracle: ref evaluate/full.py:104-125."""
    from oracle import model as OM  # noqa: F401
    from tests.test_gpu_model import build_q_model
    from tests.util_tiny import oracle_q_model

    ev = _load("evaluate/full.py", "llb_evaluate_full")
    g = np.load(ROOT / "tests" / "golden" / "model_tiny.npz")
    m = build_q_model(g, "rtn32bf16.", torch.bfloat16)
    o = oracle_q_model(g, "rtn32bf16.", tile_cols=32, dtype="bf16")
    r = np.random.RandomState(0)
    ids = r.randint(0, 100, size=(1, 61))
    ppl, toks = ev.perplexity(m, torch.from_numpy(ids).to("cuda"), block_size=24)
    nll, n = 0.0, 0
    for i in range(0, 61, 24):
        w = ids[:, i:i + 24]
        if w.shape[1] < 2:
            break
        logits = o.forward(w)[0].astype(np.float64)
        lse = np.log(np.exp(logits[:-1] - logits[:-1].max(-1, keepdims=True)).sum(-1)) + logits[:-1].max(-1)
        nll += float((lse - logits[np.arange(w.shape[1] - 1), w[0, 1:]]).sum())
        n += w.shape[1] - 1
    assert toks == n == 58
    want = math.exp(nll / n)
    assert abs(ppl - want) <= 0.02 * want, (ppl, want)
