"""Size-independent properties of the oracle (CPU): what must hold for ANY shape, next to the fixed golden vectors of
tests/test_oracle_golden.py - the pack / unpack round trip (ref quantization.py:374-409), code range, linearity of the
de-quantised matmul, the empty batch, ragged last groups; and the zero-init rules of the adapter family (ref
adapter.py:72-74, adapter_v2.py:36-37): an untrained adapter computes exactly what the plain model computes."""
import numpy as np
import pytest
from hypothesis import given, settings
from hypothesis import strategies as st

from oracle import adapter as OA
from oracle import model as OM
from oracle import quant as OQ
from tests.util_tiny import ADAPTER_TINY, adapter_tiny_weights, linear_names

BASE = {k: ADAPTER_TINY[k] for k in ("block_size", "vocab_size", "n_layer", "n_head", "n_embd")}


@st.composite
def layer(draw):
    bits = draw(st.sampled_from([4, 8]))
    epb = 8 // bits
    N = draw(st.integers(1, 24))
    K = epb * draw(st.integers(1, 48))
    tile = draw(st.sampled_from([-1, 16, 32, 128]))     # 16 / 32 leave a ragged last group for most K
    seed = draw(st.integers(0, 2**31 - 1))
    return bits, N, K, tile, seed


def _quantized(bits, N, K, tile, seed):
    r = np.random.RandomState(seed)
    W = (r.randn(N, K) * 0.05).astype(np.float32)
    if N > 1:
        W[0] = 0.0                                       # the all-zero row rule of find_params_weight (quantization.py:476-479)
    Q, s, z = OQ.rtn_quantize(W, bits, tile)
    tc = K if tile == -1 else tile
    return r, Q, s, z, OQ.pack_weight(Q, s, z, bits, tc, "fp32"), tc


@settings(max_examples=60, deadline=None)
@given(layer())
def test_pack_unpack_round_trip_and_code_range(case):
    bits, N, K, tile, seed = case
    _, Q, s, z, packed, tc = _quantized(bits, N, K, tile, seed)
    assert packed.dtype == np.uint8 and packed.shape == (N, K * bits // 8)
    codes = OQ.unpack_codes(packed, bits)
    assert codes.shape == (N, K) and codes.min() >= 0 and codes.max() <= 2**bits - 1
    w = OQ.get_weight(packed, s, z, bits, tc, "fp32")
    # pack(get_weight(p)): pack_weight TRUNCATES (quantization.py:383), so a code whose fp32 `(c - z) * s / s + z` lands
    # at c - eps comes back as c - 1; never upward, never by more than one
    again = OQ.unpack_codes(OQ.pack_weight(w, s, z, bits, tc, "fp32"), bits).astype(np.int32)
    d = codes.astype(np.int32) - again
    assert d.min() >= 0 and d.max() <= 1
    # on-grid weights come back to within one step of the grid (exactly, but for the truncation quirk of pack_weight)
    G = s.shape[1]
    step = np.repeat(s, tc, axis=1)[:, :K] if G > 1 else s
    assert (np.abs(w - Q) <= step * 1.0001 + 1e-12).all()


@settings(max_examples=40, deadline=None)
@given(layer(), st.integers(0, 5))
def test_forward_is_linear_in_the_activations_and_handles_the_empty_batch(case, M):
    bits, N, K, tile, seed = case
    r, _, s, z, packed, tc = _quantized(bits, N, K, tile, seed)
    x, y = r.randn(M, K).astype(np.float32), r.randn(M, K).astype(np.float32)
    f = lambda v: OQ.linear_forward(v, packed, s, z, None, bits, tc, "fp32")
    out = f(x)
    assert out.shape == (M, N) and out.dtype == np.float32
    if M == 0:
        return
    mag = (np.abs(x) + np.abs(y)) @ np.abs(OQ.get_weight(packed, s, z, bits, tc, "fp32")).T
    assert (np.abs(f(2.0 * x - 0.5 * y) - (2.0 * out - 0.5 * f(y))) <= 1e-5 * mag + 1e-6).all()
    exact = OQ.linear_forward_exact(x, packed, s, z, bits, tc)
    assert (np.abs(out - exact) <= 1e-5 * mag + 1e-6).all()
    b = r.randn(N).astype(np.float32)
    assert np.allclose(OQ.linear_forward(x, packed, s, z, b, bits, tc, "fp32"), out + b, rtol=0, atol=1e-6)


def _plain(W, dtype="fp32"):
    return OM.LLaMA(OM.Config(**BASE), {n: OM.FLinear(W[n + ".weight"]) for n in linear_names(BASE)},
                    W["transformer.wte.weight"], {k: v for k, v in W.items() if k.endswith(".scale")}, dtype)


def _adapter(W, v2, dtype="fp32"):
    lin = {n: OM.FLinear(W[n + ".weight"]) for n in linear_names(BASE)}
    if v2:
        lin = {n: OA.V2Linear(l, W[n + ".adapter_bias"], W[n + ".adapter_scale"]) for n, l in lin.items()}
    start, L = ADAPTER_TINY["adapter_start_layer"], ADAPTER_TINY["n_layer"]
    return OA.AdapterLLaMA(OM.Config(**BASE), lin, W["transformer.wte.weight"], {k: v for k, v in W.items() if k.endswith(".scale")},
                           dtype, adapter_wte={i: W[f"transformer.h.{i}.attn.adapter_wte.weight"] for i in range(start, L)},
                           gating_factor={i: W[f"transformer.h.{i}.attn.gating_factor"].reshape(-1) for i in range(start, L)},
                           adapter_start_layer=start)


@pytest.mark.parametrize("dtype", ["fp32", "bf16"])
@pytest.mark.parametrize("v2", [False, True])
def test_untrained_adapter_is_the_plain_model(v2, dtype):
    """Zero gates (adapter.py:72-74) and zero bias / unit scale (adapter_v2.py:36-37) change nothing - bit for bit, with and
    without the KV cache; a trained gate on ONE head changes that head's contribution only through the sum before c_proj."""
    W = adapter_tiny_weights(v2=v2)
    for k in W:
        if k.endswith("gating_factor") or k.endswith("adapter_bias"):
            W[k] = np.zeros_like(W[k])
        if k.endswith("adapter_scale"):
            W[k] = np.ones_like(W[k])
    idx = np.random.RandomState(5).randint(0, BASE["vocab_size"], size=(2, 9))
    p, a = _plain(W, dtype), _adapter(W, v2, dtype)
    assert np.array_equal(p.forward(idx), a.forward(idx))
    want = [p.forward(idx[:, :6], 12, np.arange(6))[:, -1]] + [p.forward(idx[:, t:t + 1], 12, np.array([t]))[:, -1] for t in (6, 7, 8)]
    got = [a.forward(idx[:, :6], 12, np.arange(6))[:, -1]] + [a.forward(idx[:, t:t + 1], 12, np.array([t]))[:, -1] for t in (6, 7, 8)]
    assert np.array_equal(np.stack(want), np.stack(got))
    W2 = dict(W)
    W2["transformer.h.2.attn.gating_factor"] = np.array([0, 0.5, 0, 0], np.float32).reshape(1, 4, 1, 1)
    assert not np.array_equal(_adapter(W2, v2, dtype).forward(idx), p.forward(idx))


def test_adapter_prefix_cache_is_batch_independent_and_reset():
    """The prompt's keys / values do not depend on the batch (adapter.py:151-156 repeats one tensor) and are dropped by
    reset_cache (adapter.py:245-247): a new gate takes effect after a reset, as in the reference."""
    W = adapter_tiny_weights()
    a = _adapter(W, False)
    idx = np.random.RandomState(6).randint(0, BASE["vocab_size"], size=(3, 5))
    full = a.forward(idx, 8, np.arange(5))
    assert set(a.adapter_kv) == {1, 2} and a.adapter_kv[1][0].shape == (1, 4, 10, 32)
    a.reset_cache()
    assert not a.adapter_kv and not a.kv_caches
    one = a.forward(idx[1:2], 8, np.arange(5))
    np.testing.assert_allclose(one, full[1:2], rtol=0, atol=1e-5)
