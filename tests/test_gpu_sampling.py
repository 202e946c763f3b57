"""Fused top-k sampling kernel (csrc/sampling.cu) against oracle/sampling.py, and generate()'s device-side loop."""
import ctypes as C
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
from oracle import sampling as OS  # noqa: E402

DEV = "cuda:0"


def _draws(logits, T, k, us, pos=3):
    from lit_llama import _native as N

    V = logits.size
    lg = torch.from_numpy(logits).to(DEV, torch.bfloat16)
    u = torch.tensor(us, dtype=torch.float32, device=DEV)
    step = torch.zeros(1, dtype=torch.int32, device=DEV)
    posd = torch.tensor([pos, pos, pos + 1], dtype=torch.int32, device=DEV)
    out = torch.full((len(us) + pos + 2,), -1, dtype=torch.int64, device=DEV)
    tok = torch.zeros(1, dtype=torch.int64, device=DEV)
    got = []
    for i in range(len(us)):
        posd[0] = pos + i
        N.call("ll_sample_topk", N.ptr(lg), V, float(T), 0 if k is None else k, N.ptr(u), N.ptr(step), N.ptr(posd), N.ptr(out),
               N.ptr(tok), N.stream())
        got.append(int(tok.item()))
    assert int(step.item()) == len(us)
    assert out[pos + 1: pos + 1 + len(us)].tolist() == got and int(out[pos]) == -1
    return got


@pytest.mark.parametrize("V,k,T", [(32000, 200, 0.8), (32000, 1, 1.0), (32000, None, 1.0), (1000, 50, 0.5), (40000, 7, 2.0),
                                   (31, 5, 1.0), (65536, 300, 0.9)])
def test_sample_topk_matches_oracle(V, k, T):
    r = np.random.RandomState(V + (k or 0))
    logits = OQ.bf16_round(r.randn(V).astype(np.float32) * 4.0)
    if V > 100:
        logits[r.randint(0, V, 3)] = np.sort(logits)[-(k or 10)]     # ties at the threshold are kept
    us = np.concatenate([r.rand(60), [0.0, 1.0 - 2.0**-24, 0.5]]).astype(np.float32)
    got = _draws(logits, T, k, us)
    p = OS.probabilities(logits, T, k, "bf16")
    kept = p > 0
    n_close = 0
    for g, u in zip(got, us):
        assert kept[g] or p[g] == 0 and OS.probabilities(logits, T, k)[g] >= 0, "drew a masked-out token"
        want = OS.draw(p, float(u))
        if g != want:
            # only a draw within fp32 summation noise of a CDF boundary may land on the neighbouring kept entry
            assert OS.draw_margin(p, float(u)) < 2e-5, (g, want, u)
            n_close += 1
    assert n_close <= 3
    if k == 1:
        assert all(logits[g] == logits.max() for g in got)


def test_sample_topk_distribution():
    """Empirical frequencies over a uniform grid of u follow the oracle's probabilities."""
    r = np.random.RandomState(5)
    logits = OQ.bf16_round(r.randn(512).astype(np.float32) * 2.0)
    n = 3000
    us = ((np.arange(n) + 0.5) / n).astype(np.float32)
    got = _draws(logits, 0.7, 20, us)
    p = OS.probabilities(logits, 0.7, 20)
    freq = np.bincount(got, minlength=512) / n
    assert np.abs(freq - p / p.sum()).max() < 2e-3
    assert set(np.nonzero(freq)[0]) <= set(np.nonzero(p)[0])


def test_generate_on_device_is_greedy_consistent_for_top1():
    """top_k=1 keeps only the maximal logits (ties included, as the reference's ``logits < kth`` mask does): every
    token the graph-replayed device loop emits must attain the maximum of the logits that the per-module path
    computes when it is fed the same prefix."""
    import generate as G
    import tests.test_gpu_model as TM

    g = np.load(ROOT / "tests" / "golden" / "model_tiny.npz")
    m = TM.build_q_model(g, "rtn32bf16.", torch.bfloat16)
    prompt = torch.from_numpy(g["idx"][0, :6]).to(DEV)
    n = prompt.numel()
    for msl in (None, 12):
        m.reset_cache()
        fast = G.generate(m, prompt, 10, max_seq_length=msl, top_k=1)
        assert m._engine is not None and m._engine.sample_graphs, "device-side loop was not used"
        assert fast.dtype == prompt.dtype and fast.shape == (n + 10,) and torch.equal(fast[:n], prompt)
        msl_ = msl or min(n + 10, m.config.block_size)
        m.reset_cache()
        m.use_fused_decode = False
        try:
            logits = m(prompt.view(1, -1), msl_, torch.arange(n, device=DEV))[0, -1].float()
            for i in range(10):
                assert logits[fast[n + i]] == logits.max(), (i, fast.tolist())
                logits = m(fast[n + i].view(1, 1), msl_, torch.tensor([n + i], device=DEV))[0, -1].float()
        finally:
            m.use_fused_decode = True
    # sampled run: valid ids, reproducible under the same seed, eos cut like the reference
    m.reset_cache(); torch.manual_seed(7)
    a = G.generate(m, prompt, 10, temperature=0.8, top_k=5)
    m.reset_cache(); torch.manual_seed(7)
    b = G.generate(m, prompt, 10, temperature=0.8, top_k=5)
    assert torch.equal(a, b) and int(a.max()) < m.config.padded_vocab_size and torch.equal(a[:6], prompt)
    eos = int(a[8])
    m.reset_cache(); torch.manual_seed(7)
    c = G.generate(m, prompt, 10, temperature=0.8, top_k=5, eos_id=eos)
    first = 6 + int((a[6:] == eos).nonzero()[0])
    assert torch.equal(c, a[:first])


@pytest.mark.parametrize("misaligned", [False, True])
def test_sample_advance_embed_equals_the_three_kernels(misaligned):
    """ll_sample_advance_embed = ll_sample_topk, ll_advance_pos, ll_embed_ssq in one launch: same token, same out_tokens
    entry, same position triple (including the ring wrap) and the same residual row, over a run of draws; the 16-byte load
    path of the logits and the scalar one (row at an odd offset) agree."""
    from lit_llama import _native as N

    V, dim, S, max_pos = 32000, 512, 7, 64
    r = np.random.RandomState(11)
    base = torch.from_numpy(OQ.bf16_round(r.randn(V + 8).astype(np.float32) * 3.0)).to(DEV, torch.bfloat16)
    lg = base[1:1 + V] if misaligned else base[:V].clone()
    wte = torch.from_numpy(OQ.bf16_round(r.randn(V, dim).astype(np.float32))).to(DEV, torch.bfloat16)
    us = r.rand(12).astype(np.float32)
    u = torch.tensor(us, device=DEV)

    def state():
        return dict(step=torch.zeros(1, dtype=torch.int32, device=DEV), pos=torch.tensor([3, 3, 4], dtype=torch.int32, device=DEV),
                    out=torch.full((32,), -1, dtype=torch.int64, device=DEV), tok=torch.zeros(1, dtype=torch.int64, device=DEV),
                    x=torch.zeros((1, dim), dtype=torch.bfloat16, device=DEV))

    a, b = state(), state()
    for _ in range(len(us)):
        N.call("ll_sample_topk", N.ptr(lg), V, 0.8, 200, N.ptr(u), N.ptr(a["step"]), N.ptr(a["pos"]), N.ptr(a["out"]), N.ptr(a["tok"]),
               N.stream())
        N.call("ll_advance_pos", N.ptr(a["pos"]), S, max_pos, N.stream())
        N.call("ll_embed_ssq", N.ptr(wte), N.ptr(a["tok"]), N.ptr(a["x"]), None, 1, dim, N.stream())
        N.call("ll_sample_advance_embed", N.ptr(lg), V, 0.8, 200, N.ptr(u), N.ptr(b["step"]), N.ptr(b["pos"]), N.ptr(b["out"]),
               N.ptr(b["tok"]), S, max_pos, N.ptr(wte), N.ptr(b["x"]), dim, N.stream())
        torch.cuda.synchronize()
        for k in a:
            assert torch.equal(a[k], b[k]), k
        assert torch.equal(b["x"][0], wte[int(b["tok"].item())])
    assert b["pos"].tolist() == [15, 15 % S, S]
