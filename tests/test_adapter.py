"""LLaMA-Adapter callers of the quantized-inference boundary: lit_llama/adapter.py, adapter_v2.py and
generate/adapter*.py.  The oracle (oracle/adapter.py) is pinned against outputs of the unmodified reference
(tests/golden/adapter_tiny.npz, written by tests/golden/make_golden_adapter.py); the CUDA path is compared with the
goldens (fp32) and with the oracle (int4 weights, bf16 activations)."""
import importlib.util
import inspect
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import model as OM
from oracle import quant as OQ
from tests.util_tiny import ADAPTER_TINY, adapter_tiny_weights, linear_names, oracle_adapter_model

ROOT = Path(__file__).resolve().parent.parent
PKG = ROOT / "lit-llama-ja_b200"
DEV = "cuda"
BASE = {k: ADAPTER_TINY[k] for k in ("block_size", "vocab_size", "n_layer", "n_head", "n_embd")}


@pytest.fixture(scope="module")
def g():
    return np.load(ROOT / "tests" / "golden" / "adapter_tiny.npz")


def _steps(forward, idx, B, S=16):
    """Prefill 8 tokens, then 4 single-token steps: last-position logits [B, 5, V] (the layout of the goldens)."""
    got = [forward(idx[:B, :8], S, np.arange(8))[:, -1]]
    for t in range(8, 12):
        got.append(forward(idx[:B, t:t + 1], S, np.array([t]))[:, -1])
    return np.stack(got, axis=1)


# ------------------------------------------------------------------------------------------ CPU: oracle and host logic
@pytest.mark.parametrize("v2", [False, True])
def test_oracle_matches_the_reference_goldens(g, v2):
    tag = "v2." if v2 else "v1."
    m = oracle_adapter_model(v2=v2)
    np.testing.assert_allclose(m.forward(g["idx"]), g[tag + "nocache"], rtol=0, atol=2e-5)
    for B in (1, 2):
        m.reset_cache()
        np.testing.assert_allclose(_steps(m.forward, g["idx"], B), g[f"{tag}cached_b{B}"], rtol=0, atol=2e-5)
    # the fixture exercises what it is meant to: the prompts and the v2 parameters move the logits
    plain = OM.LLaMA(OM.Config(**BASE), {n: OM.FLinear(adapter_tiny_weights()[n + ".weight"]) for n in linear_names(BASE)},
                     adapter_tiny_weights()["transformer.wte.weight"],
                     {k: v for k, v in adapter_tiny_weights().items() if k.endswith(".scale")})
    assert np.abs(plain.forward(g["idx"]) - g["v1.nocache"]).max() > 0.1
    assert np.abs(g["v1.nocache"] - g["v2.nocache"]).max() > 0.1


def test_module_tree_and_state_helpers(g):
    """State-dict keys of the reference (adapter.py:58-75, 226-238; adapter_v2.py:35-45), the zero-initialised gate, the
    trainable-parameter filters and the single-gate checkpoint rule (adapter.py:171-186)."""
    from lit_llama import adapter, adapter_v2

    m = adapter.LLaMA(adapter.LLaMAConfig(**ADAPTER_TINY))
    want = set(adapter_tiny_weights())
    assert set(m.state_dict()) == want
    assert "transformer.h.0.attn.gating_factor" not in want and "transformer.h.1.attn.gating_factor" in want
    assert float(m.transformer.h[1].attn.gating_factor.detach().abs().max()) == 0.0
    assert m.lm_head.weight.shape == (ADAPTER_TINY["vocab_size"], ADAPTER_TINY["n_embd"])
    assert adapter.LLaMAConfig().adapter_prompt_length == 10 and adapter.LLaMAConfig().adapter_start_layer == 2
    assert adapter.LLaMAConfig.from_name("7B").n_embd == 4096

    adapter.mark_only_adapter_as_trainable(m)
    trainable = {n for n, p in m.named_parameters() if p.requires_grad}
    assert trainable == {k for k in want if "adapter_wte" in k or "gating_factor" in k}
    assert set(adapter.adapter_state_from_state_dict(m.state_dict())) == trainable

    m.load_state_dict({"transformer.h.1.attn.gating_factor": torch.tensor([0.25])}, strict=False)
    np.testing.assert_array_equal(m.transformer.h[1].attn.gating_factor.detach().numpy(), g["old_gate"])

    adapter_v2.add_adapter_v2_parameters_to_linear_layers(m)
    want2 = set(adapter_tiny_weights(v2=True))
    assert set(m.state_dict()) == want2
    assert float(m.lm_head.adapter_bias.detach().abs().max()) == 0.0 and float(m.lm_head.adapter_scale.detach().min()) == 1.0
    adapter_v2.mark_only_adapter_v2_as_trainable(m)
    trainable = {n for n, p in m.named_parameters() if p.requires_grad}
    subs = adapter_v2.get_adapter_substrings()
    assert subs == ["adapter_wte", "gating_factor", "adapter_scale", "adapter_bias", "rms_1", "rms_2", "ln_f"]
    assert trainable == {k for k in want2 if any(s in k for s in subs)}
    assert set(adapter_v2.adapter_v2_state_from_state_dict(m.state_dict())) == trainable


def test_v2_parameters_on_quantized_linears():
    """The v2 patch also takes this package's quantized Linears (in the reference it cannot: adapter_v2.py:45 tests
    ``isinstance(module, nn.Linear)`` against the partial that quantization("gptq.int4") installs)."""
    from lit_llama import adapter, adapter_v2
    from lit_llama.quantization import ColBlockQuantizedLinear
    from lit_llama.utils import quantization

    with quantization("gptq.int4"):
        m = adapter.LLaMA(adapter.LLaMAConfig(**ADAPTER_TINY))
        adapter_v2.add_adapter_v2_parameters_to_linear_layers(m)      # inside the context, as generate/adapter_v2.py:71-73
    lin = m.transformer.h[0].attn.c_attn
    assert isinstance(lin, ColBlockQuantizedLinear)
    assert lin.adapter_bias.shape == (3 * ADAPTER_TINY["n_embd"],) and lin.adapter_scale.dtype == lin.scales.dtype
    assert "lm_head.adapter_scale" in m.state_dict()


def _load(rel, name):
    spec = importlib.util.spec_from_file_location(name, PKG / rel)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.mark.parametrize("script,default_path", [("generate/adapter.py", "out/adapter/alpaca/lit-llama-adapter-finetuned.pth"),
                                                 ("generate/adapter_v2.py", "out/adapter_v2/alpaca/lit-llama-adapter-finetuned.pth")])
def test_generate_adapter_surface(script, default_path):
    """Flags and defaults of the reference's generate/adapter.py:20-30 and generate/adapter_v2.py:22-32."""
    mod = _load(script, "llb_" + script.replace("/", "_").replace(".py", ""))
    params = inspect.signature(mod.main).parameters
    assert list(params) == ["prompt", "input", "adapter_path", "pretrained_path", "tokenizer_path", "quantize",
                            "max_new_tokens", "top_k", "temperature"]
    assert params["prompt"].default == "What food do lamas eat?" and params["adapter_path"].default == Path(default_path)
    assert params["max_new_tokens"].default == 100 and params["top_k"].default == 200 and params["temperature"].default == 0.8
    assert params["pretrained_path"].default == Path("checkpoints/lit-llama/7B/lit-llama.pth")


@pytest.mark.parametrize("script", ["evaluate/adapter.py", "evaluate/adapter_v2.py"])
def test_evaluate_adapter_surface(script):
    """Flags of the reference's evaluate/adapter.py:50-62 (dataset names become local text files: no network)."""
    mod = _load(script, "llb_" + script.replace("/", "_").replace(".py", ""))
    params = inspect.signature(mod.main).parameters
    assert list(params) == ["data_paths", "accelerator", "adapter_path", "checkpoint_path", "tokenizer_path", "dtype", "quantize"]
    assert params["dtype"].default == "float32" and params["quantize"].default is None
    assert "adapter_v2" in str(params["adapter_path"].default) if script.endswith("v2.py") else "out/adapter/" in str(params["adapter_path"].default)


@pytest.mark.parametrize("v2", [False, True])
def test_load_adapter_model_from_two_checkpoints(tmp_path, monkeypatch, v2):
    """generate/adapter.py:57-68 (v2: generate/adapter_v2.py:64-76): the model size comes from the pretrained checkpoint's
    embedding width, the pretrained and the adapter checkpoint are loaded one after the other through lazy_load, both
    non-strictly; afterwards every tensor is the checkpoint's (host logic only: built on the CPU, nothing is computed)."""
    import lit_llama.model as LM
    import lit_llama.utils as LU

    ga = _load("generate/adapter.py", "llb_generate_adapter_t")
    W = adapter_tiny_weights(v2=v2)
    subs = ("adapter_wte", "gating_factor", "adapter_scale", "adapter_bias")
    torch.save({k: torch.from_numpy(v) for k, v in W.items() if not any(s in k for s in subs)}, tmp_path / "base.pth")
    torch.save({k: torch.from_numpy(v) for k, v in W.items() if any(s in k for s in subs)}, tmp_path / "adapter.pth")
    monkeypatch.setitem(LM.llama_configs, "adapter-tiny", dict(ADAPTER_TINY))
    monkeypatch.setitem(LU.llama_model_sizes, ADAPTER_TINY["n_embd"], "adapter-tiny")
    m = ga.load_adapter_model(tmp_path / "base.pth", tmp_path / "adapter.pth", None, "cpu", torch.float32, v2)
    assert not m.training and type(m).__module__ == "lit_llama.adapter"
    sd = m.state_dict()
    assert set(sd) == set(W)
    for k, v in W.items():
        np.testing.assert_array_equal(sd[k].numpy(), v, err_msg=k)


# ------------------------------------------------------------------------------------------ GPU: kernels
@pytest.mark.gpu
@pytest.mark.parametrize("dtype,hs,batched", [(torch.bfloat16, 128, 0), (torch.float32, 32, 1), (torch.bfloat16, 64, 1)])
def test_adapter_attn_add_kernel_vs_oracle(dtype, hs, batched):
    from lit_llama import _native as N

    r = np.random.RandomState(3)
    B, nh, T, aT = 2, 8, 3, 10
    od = "bf16" if dtype == torch.bfloat16 else "fp32"
    q = OQ.rd(r.randn(B, nh, T, hs).astype(np.float32), od)
    ak = OQ.rd(r.randn(B if batched else 1, nh, aT, hs).astype(np.float32), od)
    av = OQ.rd(r.randn(B if batched else 1, nh, aT, hs).astype(np.float32), od)
    gate = OQ.rd((r.randn(nh) * 0.7).astype(np.float32), od)
    y = OQ.rd(r.randn(B, T, nh * hs).astype(np.float32), od)
    rep = 1 if batched else B
    ay = OM.sdpa(q, np.repeat(ak, rep, 0), np.repeat(av, rep, 0), np.ones((T, aT), bool), od)       # [B, nh, T, hs]
    want = OQ.rd(y + OQ.rd(gate.reshape(1, -1, 1, 1) * ay, od).transpose(0, 2, 1, 3).reshape(B, T, nh * hs), od)
    t = lambda a: torch.from_numpy(a).to(DEV, dtype).contiguous()
    tq, tk, tv, tg, ty = t(q), t(ak), t(av), t(gate), t(y)
    N.call("ll_adapter_attn_add", N.ptr(tq), N.ptr(tk), N.ptr(tv), N.ptr(tg), N.ptr(ty), N.dtype_code(dtype), B, nh, T, hs, aT,
           batched, N.stream())
    got = ty.float().cpu().numpy()
    # fp32 softmax on both sides; a result within one rounding step of the activation dtype where exp / the sum order differ
    tol = 2.0**-7 if dtype == torch.bfloat16 else 1e-5
    assert np.abs(got - want).max() <= tol * max(1.0, np.abs(want).max()), np.abs(got - want).max()
    if dtype == torch.bfloat16:
        assert (got == want).mean() > 0.95


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_bias_scale_kernel_is_bit_exact(dtype):
    from lit_llama import _native as N

    r = np.random.RandomState(4)
    od = "bf16" if dtype == torch.bfloat16 else "fp32"
    rows, n = 37, 384
    y = OQ.rd(r.randn(rows, n).astype(np.float32), od)
    b = OQ.rd((r.randn(n) * 0.05).astype(np.float32), od)
    s = OQ.rd((1 + 0.1 * r.randn(n)).astype(np.float32), od)
    want = OQ.rd(s * OQ.rd(y + b, od), od)
    t = lambda a: torch.from_numpy(a).to(DEV, dtype).contiguous()
    ty, tb, ts, out = t(y), t(b), t(s), torch.empty(rows, n, device=DEV, dtype=dtype)     # named: the pointers must stay valid
    N.call("ll_bias_scale", N.ptr(ty), N.ptr(tb), N.ptr(ts), N.ptr(out), N.dtype_code(dtype), rows, n, N.stream())
    np.testing.assert_array_equal(out.float().cpu().numpy(), want)
    N.call("ll_bias_scale", N.ptr(ty), N.ptr(tb), N.ptr(ts), N.ptr(ty), N.dtype_code(dtype), rows, n, N.stream())   # in place
    np.testing.assert_array_equal(ty.float().cpu().numpy(), want)


# ------------------------------------------------------------------------------------------ GPU: model level
def _product_model(v2, dtype=torch.float32, quantized=False):
    from lit_llama import adapter, adapter_v2
    from lit_llama.utils import quantization

    W = adapter_tiny_weights(v2=v2)
    sd = {k: torch.from_numpy(v) for k, v in W.items()}
    olin = None
    torch.set_default_dtype(dtype)
    try:
        if quantized:
            olin = {}
            for n in linear_names(BASE):
                Qw, s, z = OQ.rtn_quantize(sd.pop(n + ".weight").numpy(), 4, 32)
                s, z = OQ.bf16_round(s), OQ.bf16_round(z)
                packed = OQ.pack_weight(Qw, s, z, 4, 32, "bf16")
                olin[n] = OM.QLinear(packed, s, z, 4, 32)
                sd[n + ".quant_weight"] = torch.from_numpy(packed).t().contiguous().t()
                sd[n + ".scales"], sd[n + ".zeros"] = torch.from_numpy(s), torch.from_numpy(z)
            with quantization("gptq.int4"):
                m = adapter.LLaMA(adapter.LLaMAConfig(**ADAPTER_TINY))
        else:
            m = adapter.LLaMA(adapter.LLaMAConfig(**ADAPTER_TINY))
    finally:
        torch.set_default_dtype(torch.float32)
    if v2:
        adapter_v2.add_adapter_v2_parameters_to_linear_layers(m)
    m.load_state_dict(sd)
    return m.to(DEV).to(dtype).eval() if not quantized else m.to(DEV).eval(), olin


def _gpu_forward(m):
    def f(idx, S=None, pos=None):
        with torch.no_grad():
            ids = torch.from_numpy(np.ascontiguousarray(idx)).to(DEV)
            if pos is None:
                return m(ids).float().cpu().numpy()
            return m(ids, S, torch.from_numpy(pos).to(DEV)).float().cpu().numpy()
    return f


@pytest.mark.gpu
@pytest.mark.parametrize("v2", [False, True])
def test_fp32_model_vs_reference_goldens(g, v2):
    """The CUDA path on the reference's own outputs: no-cache forward, and prefill + single-token steps with a KV cache at
    batch 1 and 2 - launched one by one and as graph replays (the adapter's decode path)."""
    tag = "v2." if v2 else "v1."
    m, _ = _product_model(v2)
    f = _gpu_forward(m)
    np.testing.assert_allclose(f(g["idx"]), g[tag + "nocache"], rtol=2e-3, atol=3e-4)
    for graph in (False, True):
        m.use_step_graph = graph
        for B in (1, 2):
            m.reset_cache()
            np.testing.assert_allclose(_steps(f, g["idx"], B), g[f"{tag}cached_b{B}"], rtol=2e-3, atol=3e-4)
            assert bool(m._step_graphs) == graph, "the single-token steps did not take the expected path"
            assert m.adapter_kv_caches[0] is None and m.adapter_kv_caches[1][0].shape == (1, 4, 10, 32)
    m.reset_cache()
    assert not m.adapter_kv_caches and not m.kv_caches


@pytest.mark.gpu
@pytest.mark.parametrize("v2", [False, True])
def test_int4_bf16_adapter_model_vs_oracle(v2):
    """int4 group-32 weights (bytes packed by the oracle), bf16 activations: the quantized Linears under the adapter and
    under the v2 bias/scale, against the oracle with the same bytes and CUDA rounding rules."""
    g = np.load(ROOT / "tests" / "golden" / "adapter_tiny.npz")
    m, olin = _product_model(v2, torch.bfloat16, quantized=True)
    o = oracle_adapter_model("bf16", v2=v2, linears=olin)
    f = _gpu_forward(m)
    want = o.forward(g["idx"])
    got = f(g["idx"])
    assert np.abs(got - want).max() <= 0.03 * np.abs(want).max(), np.abs(got - want).max()
    for B in (1, 2):
        m.reset_cache(); o.reset_cache()
        want = _steps(o.forward, g["idx"], B)
        got = _steps(f, g["idx"], B)
        assert np.abs(got - want).max() <= 0.03 * np.abs(want).max(), (B, np.abs(got - want).max())
