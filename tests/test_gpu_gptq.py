"""GPTQ solver on the device (row Z2 / SURVEY 8f #1): ll_gptq_block + GPTQQuantizer against oracle/quant.py
(itself pinned to the reference's GPTQQuantizer outputs in tests/golden/gptq_solver.npz)."""
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent
for p_ in (str(ROOT), str(ROOT / "lit-llama-ja_b200")):
    if p_ not in sys.path:
        sys.path.insert(0, p_)

from oracle import quant as OQ  # noqa: E402

DEV = "cuda:0"


def _case(N_, K, seed, tokens=64):
    r = np.random.RandomState(seed)
    W = (r.randn(N_, K) * 0.05).astype(np.float32)
    X = [(r.randn(2, tokens, K) * (1 + 3 * (r.rand(K) < 0.05))).astype(np.float32) for _ in range(3)]
    H, _ = OQ.gptq_hessian(X)
    return W, X, H


@pytest.mark.parametrize("N_,K,blocksize,groupsize", [(24, 64, 32, -1), (40, 128, 128, -1), (70, 256, 128, 128), (33, 256, 64, 32),
                                                       (16, 96, 32, -1)])
def test_gptq_column_walk_is_bit_exact_given_the_same_hinv(N_, K, blocksize, groupsize):
    """Same Hinv and working weights in -> the kernel's column walk picks exactly the oracle's grid points inside a
    block (identical fp32 operation order); across blocks the cuBLAS update may round differently, so codes must
    agree on >= 99.5 % of the entries and the loss to 1e-3."""
    from lit_llama.quantization import gptq_solve

    W, X, H = _case(N_, K, N_ + K)
    Q, s, z, err, Hinv, W0 = OQ.gptq_quantize(W, H, 4, blocksize=blocksize, groupsize=groupsize, return_hinv=True)
    tile = K if groupsize == -1 else groupsize
    G = K // tile
    sc = torch.zeros((N_, G), device=DEV)
    zr = torch.zeros((N_, G), device=DEV)
    s0, z0 = OQ.find_params_weight(W, 15)
    sc[:] = torch.from_numpy(s0).to(DEV)
    zr[:] = torch.from_numpy(z0).to(DEV)
    Qd, loss = gptq_solve(torch.from_numpy(W0).to(DEV).contiguous(), torch.from_numpy(Hinv).to(DEV).contiguous(), sc, zr,
                          bits=4, blocksize=blocksize, groupsize=groupsize)
    Qd = Qd.cpu().numpy()
    first = slice(0, min(blocksize, K))
    assert np.array_equal(Qd[:, first], Q[:, first]), "first block must be bit-identical"
    if groupsize != -1:
        ng = max(1, blocksize // groupsize)
        assert np.array_equal(sc.cpu().numpy()[:, :ng], s[:, :ng]) and np.array_equal(zr.cpu().numpy()[:, :ng], z[:, :ng])
    same = (Qd == Q).mean()
    assert same >= 0.995, same
    assert abs(float(loss.sum()) - err) <= 1e-3 * abs(err) + 1e-9


def test_gptq_quantizer_matches_reference_golden():
    """Whole GPTQQuantizer (device Hessian, cuSOLVER factor, kernel walk, pack) on the reference's golden layer."""
    from lit_llama.quantization import GPTQQuantizer

    g = np.load(ROOT / "tests" / "golden" / "gptq_solver.npz")
    for name, actorder in (("plain", False), ("act", True)):
        W = g[f"{name}.W"]
        lin = torch.nn.Linear(W.shape[1], W.shape[0], bias=False).to(DEV)
        with torch.no_grad():
            lin.weight.copy_(torch.from_numpy(W))
        q = GPTQQuantizer(lin, bits=4, groupsize=-1, actorder=actorder, blocksize=32)
        for i in range(3):
            q.collect_input_stats(None, (torch.from_numpy(g[f"{name}.X{i}"]).to(DEV),), None)
        np.testing.assert_allclose(q.H.cpu().numpy(), g[f"{name}.H"], rtol=1e-4, atol=1e-5)
        qm, err = q.quantize()
        assert np.array_equal(qm.scales.cpu().numpy(), g[f"{name}.scales"]) and np.array_equal(qm.zeros.cpu().numpy(), g[f"{name}.zeros"])
        same = (OQ.unpack_codes(qm.quant_weight.cpu().numpy(), 4) == OQ.unpack_codes(g[f"{name}.packed"], 4)).mean()
        assert same >= 0.99, (name, same)          # cuSOLVER vs LAPACK rounding flips a rare borderline code
        assert abs(err - float(g[f"{name}.error"])) <= 0.03 * abs(float(g[f"{name}.error"])) + 1e-6
        assert tuple(qm.quant_weight.stride()) == (1, W.shape[0])


def test_gptq_beats_rtn_on_the_calibration_data_and_writes_grouped_layers():
    """The point of the solver: lower output error than round-to-nearest on the calibration inputs, for the per-row
    writer and for the group-128 layout the reference's own code cannot produce (its grouped branch raises)."""
    from lit_llama.quantization import ColBlockQuantizedLinear, GPTQQuantizer, rtn_quantize_

    N_, K = 256, 512
    W, X, H = _case(N_, K, 3, tokens=256)
    x = torch.from_numpy(np.concatenate([xi.reshape(-1, K) for xi in X])).to(DEV)
    lin = torch.nn.Linear(K, N_, bias=False).to(DEV)
    with torch.no_grad():
        lin.weight.copy_(torch.from_numpy(W))
    y = x @ lin.weight.t()
    for groupsize in (-1, 128):
        q = GPTQQuantizer(lin, bits=4, groupsize=groupsize)
        for xi in X:
            q.collect_input_stats(None, (torch.from_numpy(xi).to(DEV),), None)
        qm, err = q.quantize()
        assert qm.scales.shape == (N_, 1 if groupsize == -1 else K // 128) and err > 0
        rtn = rtn_quantize_(ColBlockQuantizedLinear(K, N_, False, bits=4, tile_cols=groupsize).to(DEV), lin.weight)
        e_gptq = float((x @ qm.get_weight(torch.float32).t() - y).pow(2).mean())
        e_rtn = float((x @ rtn.get_weight(torch.float32).t() - y).pow(2).mean())
        assert e_gptq < 0.95 * e_rtn, (groupsize, e_gptq, e_rtn)   # near-isotropic random inputs: a modest but certain gain
        # the packed module is on the reference's grid: dequantised weights are scale * (code - zero)
        codes = OQ.unpack_codes(qm.quant_weight.cpu().numpy(), 4)
        tile = K if groupsize == -1 else groupsize
        s_ = np.repeat(qm.scales.cpu().numpy(), tile, axis=1)
        z_ = np.repeat(qm.zeros.cpu().numpy(), tile, axis=1)
        np.testing.assert_array_equal(qm.get_weight(torch.float32).cpu().numpy(), ((codes - z_) * s_).astype(np.float32))


def test_blockwise_writer_on_the_tiny_model(tmp_path):
    """quantize/gptq.py::llama_blockwise_quantization on the tiny golden LLaMA with the golden calibration ids:
    logits close to the reference writer's own quantised model (golden), checkpoint reloads through the class-swap
    door in the reference layout, and group-128 (which the reference cannot write) loads and runs too."""
    import importlib.util

    import tests.test_gpu_model as TM
    from lit_llama import LLaMA, LLaMAConfig
    from lit_llama.quantization import ColBlockQuantizedLinear
    from lit_llama.utils import quantization

    spec = importlib.util.spec_from_file_location("b200_gptq", ROOT / "lit-llama-ja_b200" / "quantize" / "gptq.py")
    gq = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gq)
    g = np.load(ROOT / "tests" / "golden" / "model_tiny.npz")
    idx = torch.from_numpy(g["idx"]).to(DEV)
    samples = torch.from_numpy(np.random.RandomState(3).randint(0, 100, size=(8, 64)).astype(np.int64)).to(DEV)   # make_golden.py
    m = TM.build_fp_model()
    with torch.no_grad():
        fp_logits = m(idx).float().cpu().numpy()
    gq.llama_blockwise_quantization(m, samples, DEV, bits=4, verbose=False)
    assert all(isinstance(m.get_submodule(n), ColBlockQuantizedLinear) for n in TM.linear_names())
    with torch.no_grad():
        got = m(idx).float().cpu().numpy()
    ref = g["gptq.logits"]                                   # the reference writer + reference forward, same inputs
    # two independently rounded GPTQ runs (cuSOLVER vs LAPACK factor, different error cascades) are not the same
    # codes; the bar is: as accurate against the fp32 model as the reference's own quantised model, and no further
    # from it than twice its quantisation noise
    noise = np.abs(ref - fp_logits).mean()
    assert np.abs(got - fp_logits).max() <= 1.5 * np.abs(ref - fp_logits).max() + 1e-3
    assert np.abs(got - fp_logits).mean() <= 1.25 * noise
    assert np.abs(got - ref).mean() <= 2.0 * noise
    top_fp = fp_logits.argmax(-1)
    assert (got.argmax(-1) == top_fp).mean() >= (ref.argmax(-1) == top_fp).mean() - 0.1
    # reference layout on disk, and it loads through the quantization("gptq.int4") door
    path = tmp_path / "tiny-gptq.pth"
    torch.save(m.state_dict(), path)
    sd = torch.load(path)
    qw = sd["transformer.h.0.attn.c_attn.quant_weight"]
    assert qw.dtype == torch.uint8 and tuple(qw.shape) == (384, 64) and tuple(qw.stride()) == (1, 384)
    assert set(g[k].shape for k in g.files if k.startswith("gptq.sd.") and k.endswith("c_attn.scales")) == {tuple(sd["transformer.h.0.attn.c_attn.scales"].shape)}
    with quantization("gptq.int4"):
        m2 = LLaMA(LLaMAConfig(**TM.TINY))
    m2.load_state_dict(sd)
    m2 = m2.to(DEV).eval()
    with torch.no_grad():
        assert np.array_equal(m2(idx).float().cpu().numpy(), got)
    # grouped writer (the reference's grouped branch raises): scales per 32-column group
    m3 = TM.build_fp_model()
    gq.llama_blockwise_quantization(m3, samples, DEV, bits=4, groupsize=32, verbose=False)
    assert m3.transformer.h[0].attn.c_attn.scales.shape == (384, 4)
    with torch.no_grad():
        got3 = m3(idx).float().cpu().numpy()
    assert np.abs(got3 - fp_logits).max() <= np.abs(got - fp_logits).max() * 1.2 + 1e-3    # finer groups: no worse
