"""Pin the numpy oracle against vectors produced by the unmodified reference
(tests/golden/make_golden.py).  CPU only."""
import numpy as np
import pytest

from oracle import model as OM
from oracle import quant as OQ
from tests.util_tiny import TINY, oracle_fp_model, oracle_q_model, tiny_weights


def _cases(g):
    for c in g["cases"]:
        name, N, K, bits, tile = str(c).split(",")
        yield name, int(N), int(K), int(bits), int(tile)


def test_bf16_round_matches_torch():
    import torch

    x = np.random.RandomState(0).randn(4096).astype(np.float32) * 37.0
    x[:4] = [0.0, -0.0, 1e-40, 3.3895314e38]
    ref = torch.from_numpy(x).bfloat16().float().numpy()
    assert np.array_equal(OQ.bf16_round(x), ref)
    assert np.array_equal(OQ.from_bf16_bits(OQ.to_bf16_bits(x)), ref)


def test_rtn_params_and_grid(golden):
    g = golden["quant"]
    for name, N, K, bits, tile in _cases(g):
        Q, s, z = OQ.rtn_quantize(g[f"{name}.W"], bits, tile)
        for j in range(s.shape[1]):
            assert np.array_equal(s[:, j : j + 1], g[f"{name}.scale_f32.{j}"]), (name, j)
            assert np.array_equal(z[:, j : j + 1], g[f"{name}.zero_f32.{j}"]), (name, j)
        assert np.array_equal(Q, g[f"{name}.Q"]), name


@pytest.mark.parametrize("dt", ["fp32", "bf16"])
def test_pack_unpack_forward_bit_exact(golden, dt):
    g = golden["quant"]
    for name, N, K, bits, tile in _cases(g):
        s, z = g[f"{name}.{dt}.scales"], g[f"{name}.{dt}.zeros"]
        packed = OQ.pack_weight(g[f"{name}.Q"], s, z, bits, tile, dt)
        assert np.array_equal(packed, g[f"{name}.{dt}.packed"]), (name, dt)
        w = OQ.get_weight(packed, s, z, bits, tile, dt)
        assert np.array_equal(w, g[f"{name}.{dt}.get_weight"]), (name, dt)
        y = OQ.linear_forward(g[f"{name}.x"], packed, s, z, None, bits, tile, dt)
        ref = g[f"{name}.{dt}.forward"]
        if dt == "fp32":
            np.testing.assert_allclose(y, ref, rtol=1e-5, atol=1e-6)
        else:  # only the fp32 summation order differs -> at most one bf16 ulp, rarely
            ulp = np.abs(ref) * 2.0**-7 + 1e-30
            assert (np.abs(y - ref) <= ulp).all()
            assert (y != ref).mean() < 0.02


def test_truncation_quirk_is_reproduced(golden):
    """pack_weight truncates x/scale+zero: some codes land one below the
    RTN code; the oracle must reproduce the reference's bytes, not 'fix' them."""
    g = golden["quant"]
    n_off = 0
    for name, N, K, bits, tile in _cases(g):
        s, z = g[f"{name}.fp32.scales"], g[f"{name}.fp32.zeros"]
        t = K if tile == -1 else tile
        codes = OQ.unpack_codes(g[f"{name}.fp32.packed"], bits).astype(np.float32)
        W = g[f"{name}.W"]
        rtn = np.clip(np.round(W / np.repeat(s, t, 1)[:, :K]) + np.repeat(z, t, 1)[:, :K], 0, 2**bits - 1)
        n_off += int((codes != rtn).sum())
    assert n_off > 0


def test_gptq_solver(golden):
    g = golden["gptq_solver"]
    for name, actorder in (("plain", False), ("act", True)):
        H, n = OQ.gptq_hessian([g[f"{name}.X{i}"] for i in range(3)])
        np.testing.assert_allclose(H, g[f"{name}.H"], rtol=2e-5, atol=1e-6)
        Q, s, z, err = OQ.gptq_quantize(g[f"{name}.W"], g[f"{name}.H"], 4, blocksize=32, actorder=actorder)
        assert np.array_equal(s, g[f"{name}.scales"]) and np.array_equal(z, g[f"{name}.zeros"])
        packed = OQ.pack_weight(Q, s, z, 4, -1)
        ref = g[f"{name}.packed"]
        # LAPACK rounding may flip a rare borderline code; demand >= 99.5 % identical nibbles
        same = (OQ.unpack_codes(packed, 4) == OQ.unpack_codes(ref, 4)).mean()
        assert same >= 0.995, (name, same)
        assert abs(err - float(g[f"{name}.error"])) <= 0.02 * abs(float(g[f"{name}.error"])) + 1e-6


def test_ops(golden):
    g = golden["ops"]
    np.testing.assert_allclose(OM.rmsnorm(g["rms.x"], g["rms.scale"], dtype="fp32"), g["rms.fp32"], rtol=2e-6, atol=1e-6)
    # goldens were produced on the CPU: torch's CPU bf16 rsqrt double-rounds (see rmsnorm docstring)
    assert np.array_equal(OM.rmsnorm(g["rms.x"], g["rms.scale"], dtype="bf16", rsqrt_mode="cpu"), g["rms.bf16"])
    cache = OM.build_rope_cache(96, 32)
    np.testing.assert_allclose(cache, g["rope.cache"], rtol=0, atol=2e-6)
    rope = g["rope.cache"][g["rope.pos"]]
    np.testing.assert_allclose(OM.apply_rope(g["rope.x"], rope, "fp32"), g["rope.fp32"], rtol=1e-6, atol=1e-6)
    assert np.array_equal(OM.apply_rope(OQ.bf16_round(g["rope.x"]), rope, "bf16"), g["rope.bf16"])


def test_tiny_model_fp32(golden):
    g = golden["model_tiny"]
    m = oracle_fp_model()
    logits = m.forward(g["idx"])
    np.testing.assert_allclose(logits, g["fp32.logits"], rtol=1e-4, atol=2e-5)
    # KV-cache decode incl. the roll branch (max_seq_length 12 < 16 tokens)
    msl = 12
    steps = [m.forward(g["idx"][:1, :10], msl, np.arange(10))[:, -1]]
    for t in range(10, 16):
        steps.append(m.forward(g["idx"][:1, t : t + 1], msl, np.array([t]))[:, -1])
    np.testing.assert_allclose(np.stack(steps), g["fp32.kv_steps"], rtol=1e-4, atol=2e-5)
    np.testing.assert_allclose(m.kv_caches[0][0], g["fp32.kv_k0"], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(m.kv_caches[1][1], g["fp32.kv_v1"], rtol=1e-4, atol=1e-5)
    m.reset_cache()
    prompt = g["idx"][0, :8]
    assert np.array_equal(OM.generate(m, prompt, 12, top_k=1), g["fp32.generate_top1"])
    m.reset_cache()
    assert np.array_equal(OM.generate(m, prompt, 12, max_seq_length=10, top_k=1), g["fp32.generate_roll"])


def test_tiny_model_gptq_checkpoint(golden):
    """The state dict written by the reference's quantize/gptq.py, run through the oracle."""
    g = golden["model_tiny"]
    m = oracle_q_model(g, "gptq.sd.", tile_cols=-1, dtype="fp32")
    np.testing.assert_allclose(m.forward(g["idx"]), g["gptq.logits"], rtol=1e-4, atol=3e-5)
    assert np.array_equal(OM.generate(m, g["idx"][0, :8], 12, top_k=1), g["gptq.generate_top1"])


@pytest.mark.parametrize("dt", ["fp32", "bf16"])
def test_tiny_model_rtn_group32(golden, dt):
    g = golden["model_tiny"]
    prefix = "rtn32." if dt == "fp32" else "rtn32bf16."
    m = oracle_q_model(g, prefix, tile_cols=32, dtype=dt)
    m.rsqrt_mode = "cpu"   # goldens come from a CPU run of the reference
    logits = m.forward(g["idx"])
    ref = g[f"rtn32.{dt}.logits"]
    if dt == "fp32":
        np.testing.assert_allclose(logits, ref, rtol=1e-4, atol=3e-5)
    else:
        # bf16 end to end: torch's CPU SDPA / summation order differ from numpy's;
        # tolerance = the reference's own stated bf16 bar (tests/test_model.py:131), scaled
        assert np.abs(logits - ref).max() <= 0.05 * np.abs(ref).max()
        assert (logits.argmax(-1) == ref.argmax(-1)).mean() >= 0.9
    msl = 16
    steps = [m.forward(g["idx"][:1, :10], msl, np.arange(10))[:, -1]]
    for t in range(10, 16):
        steps.append(m.forward(g["idx"][:1, t : t + 1], msl, np.array([t]))[:, -1])
    ref = g[f"rtn32.{dt}.kv_steps"]
    tol = dict(rtol=1e-4, atol=3e-5) if dt == "fp32" else dict(rtol=0, atol=0.05 * np.abs(ref).max())
    np.testing.assert_allclose(np.stack(steps), ref, **tol)
    if dt == "fp32":
        m.reset_cache()
        assert np.array_equal(OM.generate(m, g["idx"][0, :8], 12, top_k=1), g["rtn32.fp32.generate_top1"])


def test_sampling_oracle_matches_torch_ops_of_the_reference():
    """oracle/sampling.probabilities == the reference's own op sequence (generate.py:62-69) run with torch on CPU
    in bf16 and fp32: same kept set, probabilities within one bf16 ulp (the softmax's exp differs in the last bit)."""
    import torch
    from oracle import sampling as OS
    from oracle import quant as OQ

    r = np.random.RandomState(11)
    for V, k, T in ((32000, 200, 0.8), (1000, 1, 1.0), (4096, None, 1.3), (777, 50, 0.5)):
        logits = OQ.bf16_round(r.randn(V).astype(np.float32) * 3.0)
        logits[r.randint(0, V, 5)] = logits.max()          # ties at the top
        t = torch.from_numpy(logits).to(torch.bfloat16) / T
        if k is not None:
            v, _ = torch.topk(t, min(k, t.size(-1)))
            t = torch.where(t < v[[-1]], -float("Inf"), t)
        want = torch.nn.functional.softmax(t, dim=-1).float().numpy()
        got = OS.probabilities(logits, T, k, "bf16")
        assert np.array_equal(got > 0, want > 0)
        assert np.all(np.abs(got - want) <= np.maximum(want, got) * 2.0**-7)
        assert (got != want).mean() < 0.05
    p = OS.probabilities(logits, 0.5, 50)
    counts = np.bincount([OS.draw(p, u) for u in (np.arange(4000) + 0.5) / 4000], minlength=p.size)
    assert np.abs(counts / 4000 - p / p.sum()).max() < 1e-3     # inverse CDF reproduces the distribution
