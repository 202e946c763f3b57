"""GPU parity: CUDA kernels (through the C-ABI) vs the numpy oracle and the reference's goldens."""
import numpy as np
import pytest
import torch

from oracle import quant as OQ

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _cases(g):
    for c in g["cases"]:
        name, N, K, bits, tile = str(c).split(",")
        yield name, int(N), int(K), int(bits), int(tile)


def _tdt(dt):
    return torch.float32 if dt == "fp32" else torch.bfloat16


def make_lin(N, K, bits, tile, scales, zeros, dt, packed=None):
    from lit_llama.quantization import ColBlockQuantizedLinear

    torch.set_default_dtype(_tdt(dt))
    try:
        lin = ColBlockQuantizedLinear(K, N, False, bits=bits, tile_cols=tile)
    finally:
        torch.set_default_dtype(torch.float32)
    lin = lin.to(DEV)
    assert lin.quant_weight.stride() == (1, N)
    lin.scales.copy_(torch.from_numpy(np.ascontiguousarray(scales)))
    lin.zeros.copy_(torch.from_numpy(np.ascontiguousarray(zeros)))
    if packed is not None:
        lin.quant_weight.copy_(torch.from_numpy(packed))
    return lin


def ulp_close(y, ref, frac=0.02):
    """bf16 results whose only difference is fp32 summation order: <= 1 ulp, on few elements."""
    y, ref = np.asarray(y, np.float32), np.asarray(ref, np.float32)
    ulp = np.maximum(np.abs(ref), 1e-30) * 2.0**-7
    assert (np.abs(y - ref) <= ulp).all(), float(np.abs(y - ref).max())
    assert (y != ref).mean() <= frac, float((y != ref).mean())


@pytest.mark.parametrize("dt", ["fp32", "bf16"])
def test_pack_unpack_bit_exact_vs_reference_golden(golden, dt):
    g = golden["quant"]
    for name, N, K, bits, tile in _cases(g):
        s, z = g[f"{name}.{dt}.scales"], g[f"{name}.{dt}.zeros"]
        lin = make_lin(N, K, bits, tile, s, z, dt)
        lin.pack_weight(torch.from_numpy(g[f"{name}.Q"]).to(_tdt(dt)))
        assert np.array_equal(lin.quant_weight.cpu().numpy(), g[f"{name}.{dt}.packed"]), (name, dt)
        w = lin.get_weight(dtype=_tdt(dt)).float().cpu().numpy()
        assert np.array_equal(w, g[f"{name}.{dt}.get_weight"]), (name, dt)
        y = lin(torch.from_numpy(g[f"{name}.x"]).to(DEV, _tdt(dt))).float().cpu().numpy()
        ref = g[f"{name}.{dt}.forward"]
        if dt == "fp32":
            np.testing.assert_allclose(y, ref, rtol=1e-5, atol=1e-6)
        else:
            ulp_close(y, ref)


@pytest.mark.parametrize("bits,tile", [(4, 128), (4, -1), (8, 64), (4, 32)])
def test_pack_unpack_random_large_vs_oracle(bits, tile):
    r = np.random.RandomState(bits * 100 + (tile if tile > 0 else 7))
    N, K = 384, 1024
    W = (r.randn(N, K) * 0.02).astype(np.float32)
    Q, s, z = OQ.rtn_quantize(W, bits, tile)
    lin = make_lin(N, K, bits, tile, s, z, "fp32")
    lin.pack_weight(torch.from_numpy(Q))
    ref = OQ.pack_weight(Q, s, z, bits, tile)
    assert np.array_equal(lin.quant_weight.cpu().numpy(), ref)
    assert np.array_equal(lin.get_weight().cpu().numpy(), OQ.get_weight(ref, s, z, bits, tile))
    # round trip pack(get_weight()): same bytes as the oracle's round trip (the reference's truncation
    # makes it non-idempotent for some 8-bit codes - reproduced, not "fixed")
    lin2 = make_lin(N, K, bits, tile, s, z, "fp32")
    lin2.pack_weight(lin.get_weight())
    rt = OQ.pack_weight(OQ.get_weight(ref, s, z, bits, tile), s, z, bits, tile)
    assert np.array_equal(lin2.quant_weight.cpu().numpy(), rt)
    if bits == 4:
        assert np.array_equal(rt, ref)


def _bf16_layer(N, K, tile, seed):
    r = np.random.RandomState(seed)
    W = (r.randn(N, K) * 0.02).astype(np.float32)
    Q, s, z = OQ.rtn_quantize(W, 4, tile)
    s, z = OQ.bf16_round(s), OQ.bf16_round(z)
    packed = OQ.pack_weight(Q, s, z, 4, tile, "bf16")
    return packed, s, z


@pytest.mark.parametrize("N,K,tile,M", [
    (384, 512, 128, 1), (4096, 4096, 128, 1), (1024, 11008, 128, 1), (128, 128, 32, 1), (512, 128, 32, 3),
    (256, 512, -1, 1), (2048, 4096, -1, 8), (1008, 256, 64, 2), (32000, 1024, 128, 1), (384, 2048, 128, 5),
])
def test_int4_gemv_vs_oracle(N, K, tile, M):
    """TMA/mma GEMV == get_weight(bf16) + fp32-accumulated matmul, up to summation order."""
    packed, s, z = _bf16_layer(N, K, tile, N + K + M)
    lin = make_lin(N, K, 4, tile, s, z, "bf16", packed)
    assert lin.prepare() is not None
    x = OQ.bf16_round(np.random.RandomState(1).randn(M, K).astype(np.float32))
    y = lin(torch.from_numpy(x).to(DEV, torch.bfloat16)).float().cpu().numpy()
    ref = OQ.linear_forward(x, packed, s, z, None, 4, tile, "bf16")
    ulp_close(y, ref)
    exact = OQ.linear_forward_exact(x, packed, s, z, 4, tile)
    assert np.abs(y - exact).max() <= 0.02 * np.abs(exact).max()
    # deterministic: same bits on a second run (fixed-order split-K combine)
    y2 = lin(torch.from_numpy(x).to(DEV, torch.bfloat16)).float().cpu().numpy()
    assert np.array_equal(y, y2)


def test_int4_gemv_forced_ksplit_matches():
    from lit_llama import _native as NV

    N_, K, tile = 512, 2048, 128
    packed, s, z = _bf16_layer(N_, K, tile, 5)
    lin = make_lin(N_, K, 4, tile, s, z, "bf16", packed)
    lin.prepare()
    x = torch.from_numpy(OQ.bf16_round(np.random.RandomState(2).randn(2, K).astype(np.float32))).to(DEV, torch.bfloat16)
    outs = []
    for ks in (1, 2, 3, 8, 16):
        y = torch.empty((2, N_), device=DEV, dtype=torch.bfloat16)
        ws = NV.Workspace.get(x.device).ensure(ks * 2 * N_ * 4 + 64, 4096)
        a = NV.GemvArgs()
        a.qw, a.sz, a.N, a.K, a.group, a.M = lin.quant_weight.data_ptr(), lin._sz.data_ptr(), N_, K, tile, 2
        a.x, a.y, a.epilogue, a.ksplit = x.data_ptr(), y.data_ptr(), NV.EPI_STORE, ks
        a.partials, a.counters = ws.partials.data_ptr(), ws.counters.data_ptr()
        NV.gemv(a)
        outs.append(y.float().cpu().numpy())
        assert int(ws.counters.abs().sum()) == 0  # counters are left zero
    ref = OQ.linear_forward(x.float().cpu().numpy(), packed, s, z, None, 4, tile, "bf16")
    for o in outs:
        ulp_close(o, ref)


def test_simt_linear_fp32_8bit_and_bias():
    r = np.random.RandomState(3)
    N_, K, bits, tile = 96, 160, 8, 32
    W = (r.randn(N_, K) * 0.05).astype(np.float32)
    Q, s, z = OQ.rtn_quantize(W, bits, tile)
    packed = OQ.pack_weight(Q, s, z, bits, tile)
    from lit_llama.quantization import ColBlockQuantizedLinear

    lin = ColBlockQuantizedLinear(K, N_, True, bits=bits, tile_cols=tile).to(DEV)
    lin.scales.copy_(torch.from_numpy(s)); lin.zeros.copy_(torch.from_numpy(z))
    lin.quant_weight.copy_(torch.from_numpy(packed))
    bias = r.randn(N_).astype(np.float32)
    lin.bias.copy_(torch.from_numpy(bias))
    for M in (1, 7, 40):   # 40 rows -> dequant-GEMM branch
        x = r.randn(M, K).astype(np.float32)
        y = lin(torch.from_numpy(x).to(DEV)).cpu().numpy()
        ref = OQ.linear_forward(x, packed, s, z, bias, bits, tile, "fp32")
        np.testing.assert_allclose(y, ref, rtol=2e-4, atol=2e-5)


def test_non_integer_zero_falls_back_to_simt():
    packed, s, z = _bf16_layer(128, 256, 128, 9)
    z = z + 0.3
    z = OQ.bf16_round(z)
    lin = make_lin(128, 256, 4, 128, s, z, "bf16", packed)
    assert lin.prepare() is None
    x = OQ.bf16_round(np.random.RandomState(1).randn(1, 256).astype(np.float32))
    y = lin(torch.from_numpy(x).to(DEV, torch.bfloat16)).float().cpu().numpy()
    ulp_close(y, OQ.linear_forward(x, packed, s, z, None, 4, 128, "bf16"))


def test_errors_are_loud():
    from lit_llama.quantization import ColBlockQuantizedLinear

    lin = ColBlockQuantizedLinear(64, 32, False, bits=4, tile_cols=-1)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        lin(torch.zeros(1, 64))
    with pytest.raises(AssertionError):
        ColBlockQuantizedLinear(64, 32, None, bits=4, tile_cols=-1)
    with pytest.raises(AssertionError):
        ColBlockQuantizedLinear(64, 32, False, bits=3, tile_cols=-1)
