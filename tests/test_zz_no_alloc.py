"""ll_int4_gemm_tc_static (include/llb200.h): the fused int4 de-quantise + GEMM with its split-K scratch in an array built
into the library, for callers that cannot call cudaMalloc.  No counterpart in the reference;
the checker is the existing entry point ll_int4_gemm_tc on the same inputs, whose own parity against the
oracle is tests/test_gpu_quant.py::test_int4_fused_dequant_gemm_vs_oracle.

This file is named to be collected LAST: a failure here hides nothing after it.
Both entry points are called through the C-ABI (ctypes), twice in a row with different inputs and row counts, so that the
second call finds the first call's partial sums in the shared array.
"""
import numpy as np
import pytest
import torch

from oracle import quant as OQ

pytestmark = pytest.mark.gpu
DEV = "cuda"
STATIC_BYTES = 32 << 20      # LL_INT4_GEMM_TC_STATIC_SCRATCH_BYTES


def _layer(N_, K, tile, seed):
    """RTN-quantised random layer with bf16 scales / zeros (the recipe of tests/test_gpu_quant.py::_bf16_layer)."""
    r = np.random.RandomState(seed)
    W = (r.randn(N_, K) * 0.02).astype(np.float32)
    Q, s, z = OQ.rtn_quantize(W, 4, tile)
    s, z = OQ.bf16_round(s), OQ.bf16_round(z)
    return OQ.pack_weight(Q, s, z, 4, tile, "bf16"), s, z


def _make_lin(N_, K, tile, packed, s, z):
    from lit_llama.quantization import ColBlockQuantizedLinear

    torch.set_default_dtype(torch.bfloat16)
    try:
        lin = ColBlockQuantizedLinear(K, N_, False, bits=4, tile_cols=tile)
    finally:
        torch.set_default_dtype(torch.float32)
    lin = lin.to(DEV)
    with torch.no_grad():
        # checkpoint layout of the packed bytes: [N][K/2] with strides (1, N) (ref quantization.py:373-377)
        lin.quant_weight.copy_(torch.from_numpy(packed))
        lin.scales.copy_(torch.from_numpy(np.ascontiguousarray(s)))
        lin.zeros.copy_(torch.from_numpy(np.ascontiguousarray(z)))
    return lin


def _both(NV, w, x, bias, M, N_, K, tile):
    """(existing entry point with a torch-allocated scratch, new entry point with none) on the same inputs."""
    nbytes = NV.call_ret("ll_int4_gemm_tc_scratch_bytes", M, N_, K)
    assert nbytes <= STATIC_BYTES
    scratch = torch.empty(max(1, nbytes // 4), dtype=torch.float32, device=DEV) if nbytes else None
    y_ref = torch.full((M, N_), float("nan"), dtype=torch.bfloat16, device=DEV)
    y_new = torch.full((M, N_), float("nan"), dtype=torch.bfloat16, device=DEV)
    NV.call("ll_int4_gemm_tc", NV.ptr(w), NV.ptr(x), NV.ptr(bias), NV.ptr(y_ref), NV.ptr(scratch), M, N_, K, tile, NV.stream())
    torch.cuda.synchronize()
    # nothing is allocated by the new entry point: neither through torch's allocator nor behind it (free device memory as
    # the driver reports it; the smallest split-K scratch these tests use is several MiB, the slack is 1 MiB)
    before, free_before = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0]
    NV.call("ll_int4_gemm_tc_static", NV.ptr(w), NV.ptr(x), NV.ptr(bias), NV.ptr(y_new), M, N_, K, tile, NV.stream())
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    assert torch.cuda.mem_get_info()[0] >= free_before - (1 << 20)
    return y_ref.float().cpu().numpy(), y_new.float().cpu().numpy(), nbytes


def test_static_scratch_entry_point_matches_ll_int4_gemm_tc_twice_in_a_row():
    from lit_llama import _native as NV

    N_, K, tile = 4096, 4096, 128
    packed, s, z = _layer(N_, K, tile, 11)
    lin = _make_lin(N_, K, tile, packed, s, z)
    w = lin._prepare_kind(NV.STREAM_U8)
    assert w is not None
    r = np.random.RandomState(2)
    bias = torch.from_numpy(OQ.bf16_round(r.randn(N_).astype(np.float32))).to(DEV, torch.bfloat16)
    wq = OQ.get_weight(packed, s, z, 4, tile, "bf16").astype(np.float64)
    # call 1: 128 rows (largest row class, 8 MiB of partial sums), with a bias; call 2: 33 rows of other data on the same
    # array (64-row class: a different layout over the same bytes), without a bias
    for M, b in ((128, bias), (33, None)):
        x = OQ.bf16_round(r.randn(M, K).astype(np.float32))
        xd = torch.from_numpy(x).to(DEV, torch.bfloat16)
        y_ref, y_new, nbytes = _both(NV, w, xd, b, M, N_, K, tile)
        assert nbytes > 0, "this shape was meant to split K (otherwise the scratch is not exercised)"
        assert np.isfinite(y_new).all()
        # same kernels, same fixed-order split-K sum: the same bits.  The tolerance the fused GEMM's own oracle test uses
        # (|ref| * 2^-7 + sum |x||w| * 2^-16) is checked as well, against the de-quantised weights.
        assert np.array_equal(y_new, y_ref), float(np.abs(y_new - y_ref).max())
        exact = x.astype(np.float64) @ wq.T + (0.0 if b is None else b.float().cpu().numpy().astype(np.float64))
        mag = np.abs(x).astype(np.float64) @ np.abs(wq).T
        assert (np.abs(y_new - exact) <= np.abs(exact) * 2.0**-7 + mag * 2.0**-16 + 2.0**-8).all()


def test_static_scratch_entry_point_without_split_and_rejections():
    from lit_llama import _native as NV

    fn = NV._fn["ll_int4_gemm_tc_static"]
    # one k-block: no split, no scratch touched
    N_, K, tile, M = 128, 128, 128, 10
    packed, s, z = _layer(N_, K, tile, 5)
    lin = _make_lin(N_, K, tile, packed, s, z)
    w = lin._prepare_kind(NV.STREAM_U8)
    x = torch.from_numpy(OQ.bf16_round(np.random.RandomState(3).randn(M, K).astype(np.float32))).to(DEV, torch.bfloat16)
    y_ref, y_new, nbytes = _both(NV, w, x, None, M, N_, K, tile)
    assert nbytes == 0 and np.array_equal(y_ref, y_new)

    # beyond the built-in maximum (the 32000-column head at 128 rows needs ~65 MB): a clear error and no launch
    assert NV.call_ret("ll_int4_gemm_tc_scratch_bytes", 128, 32000, 4096) > STATIC_BYTES
    y = torch.zeros((M, N_), dtype=torch.bfloat16, device=DEV)
    rc = fn(NV.ptr(w), NV.ptr(x), None, NV.ptr(y), 128, 32000, 4096, 128, NV.stream())
    assert rc == -3
    msg = NV.last_error()
    assert "built-in array" in msg and "ll_int4_gemm_tc" in msg
    # argument errors are reported like the existing entry point's
    assert fn(NV.ptr(w), NV.ptr(x), None, NV.ptr(y), 129, N_, K, tile, NV.stream()) == -1
    assert fn(NV.ptr(w), NV.ptr(x), None, NV.ptr(y), M, 100, K, tile, NV.stream()) == -3
    torch.cuda.synchronize()
    assert float(y.abs().max()) == 0.0
