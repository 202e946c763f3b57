"""GPU parity: CUDA kernels (through the C-ABI) vs the numpy oracle and the reference's goldens."""
import os

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
        elif bits == 4 and lin.prepare(g[f"{name}.x"].reshape(-1, g[f"{name}.x"].shape[-1]).shape[0]) is not None:
            # streamed GEMV: exact dequant (no bf16 rounding of the weights) - bf16 tolerance vs the reference
            assert np.abs(y - ref).max() <= 0.02 * np.abs(ref).max(), name
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


def exact_close(y, x, packed, s, z, tile):
    """int4 GEMV contract: exact (q - zero) * scale weights, exact products, fp32 accumulation, one bf16
    rounding on store -> within 1 bf16 ulp of the exact result plus 2^-18 of sum |x||w| (fp32 summation)."""
    exact = OQ.linear_forward_exact(x, packed, s, z, 4, tile)
    mag = np.abs(x).astype(np.float64) @ np.abs(OQ.get_weight(packed, s, z, 4, tile)).astype(np.float64).T
    tol = np.abs(exact) * 2.0**-8 + mag * 2.0**-18 + 1e-30
    assert (np.abs(y - exact) <= tol).all(), float((np.abs(y - exact) / tol).max())


def stream_decode(buf, N, K, group):
    """Host-side inverse of ll_int4_prepare's fragment-ordered repack (include/llb200.h):
    stream bytes -> (codes [N][K] uint8, scale [N][K/KB] f32, zero [N][K/KB] f32)."""
    KB = 128 if group % 128 == 0 else (64 if group % 64 == 0 else 32)
    G, BLK = K // KB, KB * 8 + 64
    blocks = np.frombuffer(buf.tobytes(), dtype=np.uint8).reshape(N // 16, G, BLK)
    words = blocks[:, :, :KB * 8].copy().view(np.uint32).reshape(N // 16, G, KB * 2)
    tail = blocks[:, :, KB * 8:].copy().view(np.uint32).reshape(N // 16, G, 16)
    codes = np.zeros((N, K), np.uint8)
    r = np.arange(KB * 2)
    if KB == 32:
        lane, sp, e0 = r >> 1, np.zeros_like(r), (r & 1) * 4
    else:
        h, rr = r >> 7, r & 127
        lane, j = rr >> 2, rr & 3
        sp, e0 = 2 * h + (j >> 1), (j & 1) * 4
    g, tig = lane >> 2, lane & 3
    kin = sp * 32 + tig * 8 + e0                       # k offset inside the block of word r's first element
    rot = lambda w, sft: ((w >> np.uint32(sft)) | (w << np.uint32(32 - sft))) & np.uint32(15) if sft else w & np.uint32(15)
    for off, hi, dk in ((3, 0, 0), (19, 0, 1), (7, 1, 0), (23, 1, 1), (11, 0, 2), (27, 0, 3), (15, 1, 2), (31, 1, 3)):
        q = rot(words, off).astype(np.uint8)            # [tile][gk][r]
        for t in range(N // 16):
            for gk in range(G):
                codes[t * 16 + g + 8 * hi, gk * KB + kin + dk] = q[t, gk]
    e = np.arange(16)
    ncol = (e >> 1) + 8 * (e & 1)
    sc = np.zeros((N, G), np.float32)
    zr = np.zeros((N, G), np.float32)
    for t in range(N // 16):
        sc[t * 16 + ncol] = OQ.from_bf16_bits((tail[t] & 0xffff).astype(np.uint16)).T
        zr[t * 16 + ncol] = (-OQ.from_bf16_bits((tail[t] >> 16).astype(np.uint16)) - 16.0).T
    return codes, sc, zr


def stream_decode_u8(buf, N, K, group):
    """Host-side inverse of the kind-1 repack (LL_STREAM_U8, include/llb200.h): the lane's word 4 q + j of a block holds
    k = 32 tig + 4 (4 q + j) + i in byte i: column g in the low nibble, column g + 8 in the high nibble; tail (scale, -zero)."""
    KB = 128
    G, BLK = K // KB, KB * 8 + 64
    blocks = np.frombuffer(buf.tobytes(), dtype=np.uint8).reshape(N // 16, G, BLK)
    words = blocks[:, :, :KB * 8].copy().view(np.uint32).reshape(N // 16, G, KB * 2)
    tail = blocks[:, :, KB * 8:].copy().view(np.uint32).reshape(N // 16, G, 16)
    codes = np.zeros((N, K), np.uint8)
    r = np.arange(KB * 2)
    q, lane, j = r >> 7, (r & 127) >> 2, r & 3
    g, tig = lane >> 2, lane & 3
    kin = 32 * tig + 4 * (4 * q + j)
    for i in range(4):
        lo = ((words >> np.uint32(8 * i)) & np.uint32(15)).astype(np.uint8)
        hi = ((words >> np.uint32(8 * i + 4)) & np.uint32(15)).astype(np.uint8)
        for t in range(N // 16):
            for gk in range(G):
                codes[t * 16 + g, gk * KB + kin + i] = lo[t, gk]
                codes[t * 16 + g + 8, gk * KB + kin + i] = hi[t, gk]
    e = np.arange(16)
    ncol = (e >> 1) + 8 * (e & 1)
    sc = np.zeros((N, G), np.float32)
    zr = np.zeros((N, G), np.float32)
    for t in range(N // 16):
        sc[t * 16 + ncol] = OQ.from_bf16_bits((tail[t] & 0xffff).astype(np.uint16)).T
        zr[t * 16 + ncol] = (-OQ.from_bf16_bits((tail[t] >> 16).astype(np.uint16))).T
    return codes, sc, zr


@pytest.mark.parametrize("M", [1, 2])
@pytest.mark.parametrize("N,K,tile", [(48, 256, 128), (32, 128, 32), (64, 192, 64), (16, 512, -1), (4096, 4096, 128)])
def test_int4_stream_repack_is_bit_exact_reindexing(N, K, tile, M):
    """Every nibble of the checkpoint bytes lands at its documented stream position; scales exact; zero exact - for both
    stream kinds (batch 1 with 128-multiple groups takes the integer-MMA layout, everything else the bf16 one)."""
    from lit_llama import _native as NV

    packed, s, z = _bf16_layer(N, K, tile, N + K)
    lin = make_lin(N, K, 4, tile, s, z, "bf16", packed)
    st = lin.prepare(M)
    assert st is not None
    group = K if tile == -1 else tile
    kind = NV.stream_kind(N, K, group, M)
    assert kind == (NV.STREAM_U8 if (M == 1 and group % 128 == 0 and (K // 128) % 2 == 0) else NV.STREAM_BF16)
    codes, sc, zr = (stream_decode_u8 if kind == NV.STREAM_U8 else stream_decode)(st.cpu().numpy(), N, K, group)
    assert np.array_equal(codes, OQ.unpack_codes(packed, 4))
    KB = 128 if group % 128 == 0 else (64 if group % 64 == 0 else 32)
    rep = group // KB
    assert np.array_equal(sc, np.repeat(s, rep, axis=1))
    assert np.array_equal(zr, np.repeat(z, rep, axis=1))


@pytest.mark.parametrize("N,K,tile,M", [
    (384, 512, 128, 1), (4096, 4096, 128, 1), (1024, 11008, 128, 1), (128, 128, 32, 1), (512, 128, 32, 3),
    (256, 512, -1, 1), (2048, 4096, -1, 8), (1008, 256, 64, 2), (32000, 1024, 128, 1), (384, 2048, 128, 5),
    (12288, 4096, 128, 1), (4096, 11008, 128, 2), (16, 4096, 128, 1),
])
def test_int4_gemv_vs_oracle(N, K, tile, M):
    """Streamed int4 GEMV == exact dequant dot products (fp32 accumulation); and within the bf16 tolerance of
    the reference's fallback numerics (get_weight(bf16) + matmul)."""
    packed, s, z = _bf16_layer(N, K, tile, N + K + M)
    lin = make_lin(N, K, 4, tile, s, z, "bf16", packed)
    assert lin.prepare(M) is not None
    x = OQ.bf16_round(np.random.RandomState(1).randn(M, K).astype(np.float32))
    y = lin(torch.from_numpy(x).to(DEV, torch.bfloat16)).float().cpu().numpy()
    exact_close(y, x, packed, s, z, tile)
    ref = OQ.linear_forward(x, packed, s, z, None, 4, tile, "bf16")
    assert np.abs(y - ref).max() <= 0.02 * np.abs(ref).max()
    # deterministic: same bits on a second run (fixed-order cross-warp and cross-CTA combine)
    y2 = lin(torch.from_numpy(x).to(DEV, torch.bfloat16)).float().cpu().numpy()
    assert np.array_equal(y, y2)


@pytest.mark.parametrize("M", [1, 2])
def test_int4_gemv_forced_grid_matches(M):
    """Any CTA count gives the same bits: a tile is summed inside one CTA in a fixed order (both kernels)."""
    N_, K, tile = 512, 2048, 128
    packed, s, z = _bf16_layer(N_, K, tile, 5)
    lin = make_lin(N_, K, 4, tile, s, z, "bf16", packed)
    lin.prepare(M)
    xn = OQ.bf16_round(np.random.RandomState(2).randn(M, K).astype(np.float32))
    x = torch.from_numpy(xn).to(DEV, torch.bfloat16)
    outs = []
    for grid in (1, 2, 3, 7, 32, 100, 148, 512):
        y = torch.empty((M, N_), device=DEV, dtype=torch.bfloat16)
        lin._gemv(x, y, grid=grid)
        outs.append(y.float().cpu().numpy())
        exact_close(outs[-1], xn, packed, s, z, tile)
    assert all(np.array_equal(outs[0], o) for o in outs[1:])


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
    # the bf16 kernel needs -(16 + zero) exact in bf16: not so here -> reference-semantics SIMT kernel at batch 2
    assert lin.prepare(2) is None
    x = OQ.bf16_round(np.random.RandomState(1).randn(2, 256).astype(np.float32))
    y = lin(torch.from_numpy(x).to(DEV, torch.bfloat16)).float().cpu().numpy()
    ulp_close(y, OQ.linear_forward(x, packed, s, z, None, 4, 128, "bf16"))   # SIMT path: reference rounding steps
    # the batch-1 integer kernel carries any bf16 zero exactly (sum q x - zero sum x)
    assert lin.prepare(1) is not None
    y1 = lin(torch.from_numpy(x[:1]).to(DEV, torch.bfloat16)).float().cpu().numpy()
    exact_close(y1, x[:1], packed, s, z, 128)


def test_errors_are_loud():
    from lit_llama.quantization import ColBlockQuantizedLinear

    lin = ColBlockQuantizedLinear(64, 32, False, bits=4, tile_cols=-1)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        lin(torch.zeros(1, 64))
    with pytest.raises(AssertionError):
        ColBlockQuantizedLinear(64, 32, None, bits=4, tile_cols=-1)
    with pytest.raises(AssertionError):
        ColBlockQuantizedLinear(64, 32, False, bits=3, tile_cols=-1)


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (17, 384, 512), (512, 4096, 4096), (300, 1000, 328), (640, 11008, 4096),
                                   (4096, 12288, 1024), (130, 257, 72)])
def test_tcgen05_gemm_vs_fp32_reference(M, N, K):
    """tcgen05/TMEM GEMM == fp32 matmul of the same bf16 operands, up to fp32 summation order and one bf16 rounding."""
    from lit_llama import _native as NV

    g = torch.Generator(device=DEV).manual_seed(M + N + K)
    x = torch.randn((M, K), device=DEV, generator=g).bfloat16()
    w = (torch.randn((N, K), device=DEV, generator=g) * 0.05).bfloat16()
    b = torch.randn((N,), device=DEV, generator=g).bfloat16()
    for bias in (None, b):
        y = torch.empty((M, N), device=DEV, dtype=torch.bfloat16)
        NV.call("ll_gemm_bf16_tc", NV.ptr(x), NV.ptr(w), NV.ptr(bias), NV.ptr(y), M, N, K, NV.stream())
        ref = x.float() @ w.float().t()
        if bias is not None:
            ref = ref + bias.float()
        mag = x.float().abs() @ w.float().abs().t() + 1.0
        err = (y.float() - ref).abs()
        assert bool((err <= ref.abs() * 2.0**-8 + mag * 2.0**-16).all()), float((err / (ref.abs() * 2.0**-8 + mag * 2.0**-16)).max())


@pytest.mark.parametrize("N,K,tile,M", [(4096, 4096, 128, 512), (1024, 11008, 128, 40), (384, 512, -1, 100), (512, 256, 64, 33)])
def test_int4_large_m_linear_vs_oracle(N, K, tile, M):
    """Prefill path: bit-exact bf16 de-quantisation + tcgen05 GEMM == the reference's fallback branch
    F.linear(x, get_weight(bf16)) with fp32 accumulation."""
    packed, s, z = _bf16_layer(N, K, tile, N + K + M)
    lin = make_lin(N, K, 4, tile, s, z, "bf16", packed)
    w = lin.get_weight(dtype=torch.bfloat16).float().cpu().numpy()
    assert np.array_equal(w, OQ.get_weight(packed, s, z, 4, tile, "bf16"))   # fast unpack kernel: bit-exact
    x = OQ.bf16_round(np.random.RandomState(1).randn(M, K).astype(np.float32))
    y = lin(torch.from_numpy(x).to(DEV, torch.bfloat16)).float().cpu().numpy()
    ref = OQ.linear_forward(x, packed, s, z, None, 4, tile, "bf16")
    # same bf16 weights, fp32 accumulation in a different order, one bf16 rounding: 1 ulp (+ summation noise)
    mag = np.abs(x).astype(np.float64) @ np.abs(OQ.get_weight(packed, s, z, 4, tile, "bf16")).astype(np.float64).T
    assert (np.abs(y - ref) <= np.abs(ref) * 2.0**-7 + mag * 2.0**-16).all()
    assert (y != ref).mean() <= 0.05


@pytest.mark.parametrize("N,K,tile,M", [
    (4096, 4096, 128, 9), (4096, 4096, 128, 16), (4096, 4096, 128, 17), (4096, 4096, 128, 33), (11008, 4096, 128, 64),
    (4096, 11008, 128, 65), (4096, 11008, 128, 128), (12288, 4096, -1, 12), (32000, 4096, 128, 24), (128, 128, 128, 10),
    (256, 384, 128, 128),
])
def test_int4_fused_dequant_gemm_vs_oracle(N, K, tile, M):
    """9 <= M <= 128 rows: ll_int4_gemm_tc streams the packed weights once and de-quantises them in the main loop
    (tcgen05, weights in tensor memory) with the reference's bf16 rounding steps: == F.linear(x, get_weight(bf16)) of the
    reference's fallback branch up to the fp32 summation order.  Shapes: the 7B layers, per-row scales, one k-block, an odd
    number of k-blocks, every M padding class and its edges."""
    import ctypes as C
    from lit_llama import _native as NV

    packed, s, z = _bf16_layer(N, K, tile, N + K + M)
    lin = make_lin(N, K, 4, tile, s, z, "bf16", packed)
    x = OQ.bf16_round(np.random.RandomState(1).randn(M, K).astype(np.float32))
    xd = torch.from_numpy(x).to(DEV, torch.bfloat16)
    assert lin._fused_gemm_eligible(xd)
    NV.lib.ll_debug_gemv_counts  # (library loaded)
    y = lin(xd).float().cpu().numpy()
    assert NV.STREAM_U8 in lin._streams, "the fused path did not run (no kind-1 stream was built)"
    ref = OQ.linear_forward(x, packed, s, z, None, 4, tile, "bf16")
    mag = np.abs(x).astype(np.float64) @ np.abs(OQ.get_weight(packed, s, z, 4, tile, "bf16")).astype(np.float64).T
    assert (np.abs(y - ref) <= np.abs(ref) * 2.0**-7 + mag * 2.0**-16).all(), float(np.abs(y - ref).max())
    assert (y != ref).mean() <= 0.05
    y2 = lin(xd).float().cpu().numpy()
    assert np.array_equal(y, y2)            # fixed-order split-K sum: same bits run to run


def test_int4_fused_dequant_gemm_bias_and_fractional_zero():
    """bias is added after the split-K sum; a zero that is not an integer goes through bf16(q - zero) like get_weight."""
    N_, K, tile, M = 512, 1024, 128, 20
    packed, s, z = _bf16_layer(N_, K, tile, 77)
    z = OQ.bf16_round(z + 0.3)
    r = np.random.RandomState(5)
    bias = OQ.bf16_round(r.randn(N_).astype(np.float32))
    from lit_llama.quantization import ColBlockQuantizedLinear

    torch.set_default_dtype(torch.bfloat16)
    try:
        lin = ColBlockQuantizedLinear(K, N_, True, bits=4, tile_cols=tile)
    finally:
        torch.set_default_dtype(torch.float32)
    lin = lin.to(DEV)
    lin.scales.copy_(torch.from_numpy(s)); lin.zeros.copy_(torch.from_numpy(z)); lin.bias.copy_(torch.from_numpy(bias))
    lin.quant_weight.copy_(torch.from_numpy(packed))
    x = OQ.bf16_round(r.randn(M, K).astype(np.float32))
    y = lin(torch.from_numpy(x).to(DEV, torch.bfloat16)).float().cpu().numpy()
    from lit_llama import _native as NV
    assert NV.STREAM_U8 in lin._streams
    ref = OQ.linear_forward(x, packed, s, z, bias, 4, tile, "bf16")
    mag = np.abs(x).astype(np.float64) @ np.abs(OQ.get_weight(packed, s, z, 4, tile, "bf16")).astype(np.float64).T
    assert (np.abs(y - ref) <= np.abs(ref) * 2.0**-7 + mag * 2.0**-16 + 2.0**-8 * np.abs(bias)).all()
