"""Parity at the HEADLINE configuration's layer shapes (BASELINE.json configs[1], SURVEY 8d "Correctness gates"):
LLaMA-7B dimensions (n_embd 4096, 32 heads of 128, n_hidden 11008, vocab 32000), GPTQ int4 group 128 (and per-row), bf16,
a 512-token prompt then teacher-forced single-token steps - the shapes that select the SHIPPED decode kernels
(`gemv_w4a8_kernel<{ROPE_KV,RESIDUAL,SWIGLU,STORE},{norm}>` at batch 1, `gemv_int4_kernel<8,4,2,...>` at batch 2,
`decode_attn_kernel<bf16,128,...>` with 8 splits).

Three references on identical weights and tokens:
  (a) oracle/model.py - the numpy restatement with the reference's rounding rules (get_weight -> bf16 weights, F.linear);
  (b) this repo's per-module path (one kernel per op, no fusion, no CUDA graph);
  (c) the UNMODIFIED reference's own modules on the same GPU (baseline/_ref, built by baseline/install_ref.sh):
      LLaMA.forward over ColBlockQuantizedLinear.forward - for group 128 its get_weight + F.linear branch
      (ref quantization.py:390-421), for per-row its Triton kernel (ref quantization.py:185-331; opt-in, see below).

Layer count is 2 here; the full depth (32 layers, 512 + 256 steps on the reference's greedy stream, and the reference's own
fp32 result on the same tokens) is run by profiles/parity_7b.py -> profiles/r02_parity_config2_*.json.  The stated
tolerance is in TOL below and in DESIGN.md section 1.
"""
import ctypes as C
import os

import numpy as np
import pytest
import torch

from oracle import model as OM

pytestmark = pytest.mark.gpu
DEV = "cuda"

CFG2 = dict(block_size=2048, vocab_size=32000, n_layer=2, n_head=32, n_embd=4096)
PROMPT = 512

# Stated bf16 tolerance.  The reference's one stated tolerance is atol 5e-3 / rtol 1e-3 on logits of
# an fp32 16-layer E=32 model (ref tests/test_model.py:131).  Under bf16-true every op rounds to 2^-9 relative and the two
# numerics the reference itself has for a quantized Linear (bf16-rounded de-quantised weights in the fallback branch, exact
# (q - zero) * scale in the Triton branch) already differ by that much per layer, so the bar is relative to the logit
# scale: max |diff| <= 4 % of max |logit|, relative Frobenius error <= 3 %, measured per decode step
# (measured on B200, 2 layers: 1.7 % and 1.8-2.0 % against the oracle with the fallback branch's bf16-rounded weights).
TOL = dict(max_rel=0.04, fro=0.03)


def _lin_names(n_layer):
    names = ["lm_head"]
    for i in range(n_layer):
        p = f"transformer.h.{i}."
        names += [p + "attn.c_attn", p + "attn.c_proj", p + "mlp.c_fc1", p + "mlp.c_fc2", p + "mlp.c_proj"]
    return names


def build_cfg2_model(group, n_layer=2, seed=0, mode="gptq.int4"):
    """Random-init model of 7B layer shapes, RTN-quantised on-grid per group, norm scales 1 +- 0.1."""
    from lit_llama import LLaMA, LLaMAConfig
    from lit_llama.quantization import ColBlockQuantizedLinear, rtn_quantize_
    from lit_llama.utils import EmptyInitOnDevice, quantization

    cfg = LLaMAConfig(**{**CFG2, "n_layer": n_layer})
    with EmptyInitOnDevice(device=DEV, dtype=torch.bfloat16), quantization(mode, groupsize=group):
        m = LLaMA(cfg)
    gen = torch.Generator(device=DEV).manual_seed(seed)
    with torch.no_grad():
        for mod in m.modules():
            if isinstance(mod, ColBlockQuantizedLinear):
                w = torch.randn((mod.out_features, mod.in_features), device=DEV, dtype=torch.float32, generator=gen) * 0.02
                rtn_quantize_(mod, w)
        m.transformer.wte.weight.copy_(torch.randn(m.transformer.wte.weight.shape, device=DEV, generator=gen) * 0.02)
        for n, p in m.named_parameters():
            if n.endswith(".scale"):
                p.copy_(1.0 + 0.1 * torch.randn(p.shape, device=DEV, generator=gen))
    return m.eval()


def oracle_of(m, group, bits=4):
    """oracle/model.py over the SAME quantized state (checkpoint-layout buffers read back from the device)."""
    cfg = m.config
    sd = m.state_dict()
    lin = {}
    for n in _lin_names(cfg.n_layer):
        lin[n] = OM.QLinear(sd[n + ".quant_weight"].cpu().numpy(), sd[n + ".scales"].float().cpu().numpy(),
                            sd[n + ".zeros"].float().cpu().numpy(), bits, group, cache=True)
    norms = {k: v.float().cpu().numpy() for k, v in sd.items() if k.endswith(".scale")}
    ocfg = OM.Config(block_size=cfg.block_size, vocab_size=cfg.vocab_size, n_layer=cfg.n_layer, n_head=cfg.n_head, n_embd=cfg.n_embd)
    return OM.LLaMA(ocfg, lin, sd["transformer.wte.weight"].float().cpu().numpy(), norms, "bf16")


def tokens(n, seed=0):
    return torch.randint(0, 32000, (1, n), dtype=torch.int64, generator=torch.Generator().manual_seed(seed))


def run_ours(m, idx, msl, n_prefill, fused):
    m.use_fused_decode = fused
    m.reset_cache()
    m._engine = None if not fused else m._engine
    out = []
    idx = idx.to(DEV)
    with torch.no_grad():
        out.append(m(idx[:, :n_prefill], msl, torch.arange(n_prefill, device=DEV))[:, -1].float().cpu().numpy())
        for t in range(n_prefill, idx.shape[1]):
            out.append(m(idx[:, t:t + 1], msl, torch.tensor([t], device=DEV))[:, -1].float().cpu().numpy())
    return np.concatenate(out)


def run_oracle(o, idx, msl, n_prefill):
    o.reset_cache()
    idx = idx.numpy()
    out = [o.forward(idx[:, :n_prefill], msl, np.arange(n_prefill))[:, -1]]
    for t in range(n_prefill, idx.shape[1]):
        out.append(o.forward(idx[:, t:t + 1], msl, np.array([t]))[:, -1])
    return np.concatenate(out)


def run_reference_gpu(rm, idx, msl, n_prefill):
    """The reference's own LLaMA.forward with its KV cache (ref model.py:84-128), teacher-forced."""
    rm.kv_caches.clear()        # (the reference's reset_cache() dereferences mask_cache, which is None before the first forward)
    idx = idx.to(DEV)
    out = []
    with torch.no_grad():
        out.append(rm(idx[:, :n_prefill], msl, torch.arange(n_prefill, device=DEV))[:, -1].float().cpu().numpy())
        for t in range(n_prefill, idx.shape[1]):
            out.append(rm(idx[:, t:t + 1], msl, torch.tensor([t], device=DEV))[:, -1].float().cpu().numpy())
    return np.concatenate(out)


def metrics(a, ref):
    """per-run: max |diff| / max |ref|, worst per-step relative Frobenius error, top-1 agreement"""
    d = np.abs(a - ref)
    fro = np.sqrt((d.astype(np.float64) ** 2).sum(-1)) / np.sqrt((ref.astype(np.float64) ** 2).sum(-1))
    return dict(max_rel=float(d.max() / np.abs(ref).max()), max_abs=float(d.max()), fro=float(fro.max()),
                top1=float((a.argmax(-1) == ref.argmax(-1)).mean()))


def check(name, a, ref, record):
    mt = metrics(a, ref)
    record[name] = mt
    print(f"[config2 parity] {name}: max|diff| {mt['max_abs']:.4g} = {100 * mt['max_rel']:.2f} % of max|logit|, "
          f"rel. Frobenius {100 * mt['fro']:.2f} %, top-1 agreement {100 * mt['top1']:.1f} %")
    assert mt["max_rel"] <= TOL["max_rel"], (name, mt)
    assert mt["fro"] <= TOL["fro"], (name, mt)
    return mt


def gemv_counts(reset=False):
    from lit_llama import _native as NV

    out = (C.c_int32 * 21)()
    NV.lib.ll_debug_gemv_counts.argtypes = [C.c_void_p, C.c_int]
    NV.lib.ll_debug_gemv_counts.restype = None
    NV.lib.ll_debug_gemv_counts(out, 1 if reset else 0)
    a = np.array(out[:20]).reshape(5, 2, 2)
    return a, out[20]


@pytest.fixture(scope="module")
def g128():
    m = build_cfg2_model(128)
    return m, oracle_of(m, 128)


@pytest.mark.parametrize("msl,n_prefill,n_steps", [(768, PROMPT, 64), (768, 700, 60), (520, PROMPT, 24)])
def test_fused_decode_chain_at_7b_shapes_vs_oracle_and_module_path(g128, msl, n_prefill, n_steps):
    """The shipped chain: specialised GEMVs with fused RMSNorm / RoPE+KV / SwiGLU / residual, hs=128 attention with 8 splits;
    positions 512..575 and 700..759 of a 768-slot cache, and 512..535 through the wrap of a 520-slot ring."""
    from lit_llama import _native as NV

    m, o = g128
    idx = tokens(n_prefill + n_steps, seed=msl + n_prefill)
    m._engine = None            # a fresh engine: its capture goes through the launcher, which counts the kernels it picks
    gemv_counts(reset=True)
    fused = run_ours(m, idx, msl, n_prefill, fused=True)
    eng = m._engine
    assert eng is not None and eng.n_replays > 0, "the fused engine did not run"
    assert eng.splits == 8 and eng.layers[0][0].epilogue == NV.EPI_ROPE_KV
    spec, generic = gemv_counts()
    # the batch-1 specialisations of the decode chain were launched: (ROPE_KV, norm), (RESIDUAL, no norm), (SWIGLU, norm),
    # (STORE, norm) [lm_head]
    for e, n in ((NV.EPI_ROPE_KV, 1), (NV.EPI_RESIDUAL, 0), (NV.EPI_SWIGLU, 1), (NV.EPI_STORE, 1)):
        assert spec[e, n, 1] > 0, f"specialised kernel (epilogue {e}, norm {n}, batch 1) never launched: {spec.tolist()}"
    module = run_ours(m, idx, msl, n_prefill, fused=False)
    want = run_oracle(o, idx, msl, n_prefill)
    rec = {}
    check(f"fused engine vs oracle (S={msl}, prefill {n_prefill}, {n_steps} steps)", fused, want, rec)
    check("per-module path vs oracle", module, want, rec)
    check("fused engine vs per-module path", fused, module, rec)
    # deterministic run to run
    again = run_ours(m, idx, msl, n_prefill, fused=True)
    assert np.array_equal(fused, again)


def test_attention_emitted_digit_planes_give_identical_bits(g128, monkeypatch):
    """attn.c_proj's input as digit planes written by the attention kernel's leader CTAs (ll_decode_attn_digits ->
    ll_gemv_args.x_planes / x_rec) against the GEMV converting the bf16 row itself: the conversion is restated with the
    same operations in the same order, so the logits must be identical bit for bit (positions through a ring wrap)."""
    m, _ = g128
    msl, n_prefill, n_steps = 520, PROMPT, 16
    idx = tokens(n_prefill + n_steps, seed=77)
    m._engine = None
    with_digits = run_ours(m, idx, msl, n_prefill, fused=True)
    assert m._engine is not None and m._engine.attn_digits, "the engine did not take the digit-emitting attention"
    monkeypatch.setenv("LLB200_ATTN_DIGITS", "0")
    m._engine = None
    m.kv_caches.clear()
    without = run_ours(m, idx, msl, n_prefill, fused=True)
    assert m._engine is not None and not m._engine.attn_digits
    m._engine = None
    assert np.array_equal(with_digits, without)


def test_gptq_int8_fused_decode_chain_at_7b_shapes_vs_oracle_and_module_path():
    """gptq.int8 (8-bit GPTQ weights, group 128) through the same fused decode chain: every Linear takes stream kind 2 of
    the batch-1 GEMV; against the oracle with the reference's rounding rules at 8 bits and the per-module path."""
    from lit_llama import _native as NV

    m = build_cfg2_model(128, mode="gptq.int8")
    assert m.lm_head.bits == 8
    o = oracle_of(m, 128, bits=8)
    msl, n_prefill, n_steps = 768, PROMPT, 24
    idx = tokens(n_prefill + n_steps, seed=8)
    m._engine = None
    gemv_counts(reset=True)
    fused = run_ours(m, idx, msl, n_prefill, fused=True)
    eng = m._engine
    assert eng is not None and eng.n_replays > 0, "the fused engine did not run for gptq.int8"
    assert eng.layers[0][0].stream_kind == NV.STREAM_W8 and eng.parity
    spec, generic = gemv_counts()
    for e, n in ((NV.EPI_ROPE_KV, 1), (NV.EPI_RESIDUAL, 0), (NV.EPI_SWIGLU, 1), (NV.EPI_STORE, 1)):
        assert spec[e, n, 1] > 0, spec.tolist()
    module = run_ours(m, idx, msl, n_prefill, fused=False)
    want = run_oracle(o, idx, msl, n_prefill)
    rec = {}
    check("gptq.int8 fused engine vs oracle", fused, want, rec)
    check("gptq.int8 per-module path vs oracle", module, want, rec)
    check("gptq.int8 fused engine vs per-module path", fused, module, rec)
    assert np.array_equal(fused, run_ours(m, idx, msl, n_prefill, fused=True))


@pytest.mark.parametrize("group", [128])
def test_fused_decode_chain_vs_the_reference_modules_on_the_gpu(g128, group):
    """(c): the unmodified reference's LLaMA + ColBlockQuantizedLinear (group 128 -> get_weight + F.linear on CUDA) loaded
    from OUR state dict (strict: the checkpoint layout is the contract), same tokens, teacher-forced."""
    from baseline import ref_loader

    if not ref_loader.available():
        pytest.skip("baseline/_ref not installed (baseline/install_ref.sh)")
    ref = ref_loader.load()
    m, _ = g128
    rm = ref_loader.build_quantized_model(ref, ref.model.LLaMAConfig(**CFG2), tile_cols=group, device=DEV, dtype=torch.bfloat16)
    rm.load_state_dict(m.state_dict(), strict=True)
    assert type(rm.lm_head).__module__ == "lit_llama.quantization" and "baseline/_ref" in ref.quantization.__file__
    idx = tokens(PROMPT + 48, seed=7)
    want = run_reference_gpu(rm, idx, 768, PROMPT)
    fused = run_ours(m, idx, 768, PROMPT, fused=True)
    assert m._engine is not None and m._engine.n_replays > 0
    check("fused engine vs the reference's own modules on this GPU (group 128)", fused, want, {})


@pytest.mark.skipif(os.environ.get("LLB200_TEST_TRITON_REF") != "1",
                    reason="opt-in (Triton autotunes 11 configurations per layer shape on first use: minutes): LLB200_TEST_TRITON_REF=1")
def test_per_row_chain_vs_the_reference_triton_kernel():
    """Per-row scales (tile_cols = -1, what the reference's writer ships): the reference's forward goes through its Triton
    kernel (ref quantization.py:282-331, M padded to 256); ours through the same fused chain with group = K."""
    from baseline import ref_loader

    if not ref_loader.available():
        pytest.skip("baseline/_ref not installed")
    ref = ref_loader.load()
    assert ref.quantization.triton is not None and ref.quantization.qlinear_4bit_weight is not None
    m = build_cfg2_model(-1)
    assert m.lm_head.scales.shape[1] == 1
    rm = ref_loader.build_quantized_model(ref, ref.model.LLaMAConfig(**CFG2), tile_cols=-1, device=DEV, dtype=torch.bfloat16)
    rm.load_state_dict(m.state_dict(), strict=True)
    idx = tokens(256 + 24, seed=9)
    want = run_reference_gpu(rm, idx, 512, 256)
    fused = run_ours(m, idx, 512, 256, fused=True)
    assert m._engine is not None and m._engine.n_replays > 0
    check("fused engine (per-row) vs the reference's Triton path on this GPU", fused, want, {})


# ----------------------------------------------------------------------------- direct kernel tests at the shipped shapes
def _attn_ref(q, k, v, n_valid):
    """float64 softmax attention over the first n_valid slots (slot order is irrelevant to the result)."""
    B, nh, hs = q.shape
    out = np.zeros((B, nh, hs))
    for b in range(B):
        for h in range(nh):
            s = (k[b, h, :n_valid].astype(np.float64) @ q[b, h].astype(np.float64)) / np.sqrt(hs)
            p = np.exp(s - s.max())
            out[b, h] = (p / p.sum()) @ v[b, h, :n_valid].astype(np.float64)
    return out


@pytest.mark.parametrize("B,S,pos", [(1, 768, 0), (1, 768, 511), (1, 768, 640), (1, 768, 767), (1, 768, 1000), (8, 768, 700),
                                     (1, 2048, 2047), (8, 2048, 1500)])
@pytest.mark.parametrize("splits", [1, 2, 3, 5, 8])
def test_decode_attention_hs128_vs_float64_softmax(B, S, pos, splits):
    """ll_decode_attn at head_size 128, 32 heads: every split count the engine can pick (1..8: one cluster per head), a cold
    cache (pos 0), a position beyond the ring (all S slots valid), batch 1 and 8."""
    from lit_llama import _native as NV

    nh, hs = 32, 128
    g = torch.Generator(device=DEV).manual_seed(B * 100000 + S + pos)
    q = (torch.randn((B, nh * hs), device=DEV, generator=g)).bfloat16()
    k = (torch.randn((B, nh, S, hs), device=DEV, generator=g)).bfloat16()
    v = (torch.randn((B, nh, S, hs), device=DEV, generator=g)).bfloat16()
    n_valid = min(pos + 1, S)
    pos3 = torch.tensor([pos, pos % S, n_valid], dtype=torch.int32, device=DEV)
    y = torch.empty((B, nh * hs), device=DEV, dtype=torch.bfloat16)
    scratch = torch.empty(NV.attn_scratch_bytes(B, nh, hs, splits) // 4 + 16, dtype=torch.float32, device=DEV)
    counters = torch.zeros(B * nh, dtype=torch.int32, device=DEV)
    for _ in range(2):   # twice: the counters must come back to zero
        NV.call("ll_decode_attn", NV.ptr(q), NV.ptr(k), NV.ptr(v), NV.ptr(pos3), NV.ptr(y), NV.LL_BF16, B, nh, hs, S, splits,
                NV.ptr(scratch), NV.ptr(counters), 0, NV.stream())
    torch.cuda.synchronize()
    assert int(counters.abs().sum()) == 0
    want = _attn_ref(q.float().cpu().numpy().reshape(B, nh, hs), k.float().cpu().numpy(), v.float().cpu().numpy(), n_valid)
    got = y.float().cpu().numpy().reshape(B, nh, hs)
    # fp32 online softmax, one bf16 rounding of the output
    assert np.abs(got - want).max() <= 2.0 ** -8 * np.abs(want).max() + 1e-6, np.abs(got - want).max()


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("B,S,pos,splits", [(1, 768, 0, 8), (1, 768, 5, 8), (1, 768, 95, 8), (1, 768, 96, 8), (1, 768, 97, 8),
                                            (1, 768, 640, 8), (1, 768, 767, 8), (1, 768, 1000, 8), (8, 768, 700, 8),
                                            (1, 520, 519, 8), (1, 520, 777, 8), (2, 384, 383, 4), (1, 192, 100, 2)])
def test_decode_attention_staged_kv_paths_agree(B, S, pos, splits, dtype):
    """The K/V-staging decode attention (decode_attn_pf_kernel: this split's rows copied to shared memory by bulk TMA ahead
    of the wait, the producer's row patched in afterwards) against a float64 softmax, and its three routes against each
    other BIT FOR BIT: staged rows, staged rows with the patched row poisoned first (the patch must repair it), and the
    global-memory route taken when the position guess was stale (forced by the test flag)."""
    from lit_llama import _native as NV

    if dtype == torch.float32 and 2 * ((S + splits - 1) // splits) * 128 * 4 > 50 * 1024:
        pytest.skip("fp32 rows of this split do not fit the staging budget (the direct kernel serves it)")
    nh, hs = 32, 128
    g = torch.Generator(device=DEV).manual_seed(B * 100000 + S + pos)
    q = (torch.randn((B, nh * hs), device=DEV, generator=g)).to(dtype)
    k = (torch.randn((B, nh, S, hs), device=DEV, generator=g)).to(dtype)
    v = (torch.randn((B, nh, S, hs), device=DEV, generator=g)).to(dtype)
    n_valid = min(pos + 1, S)
    pos3 = torch.tensor([pos, pos % S, n_valid], dtype=torch.int32, device=DEV)
    scratch = torch.empty(NV.attn_scratch_bytes(B, nh, hs, splits) // 4 + 16, dtype=torch.float32, device=DEV)
    counters = torch.zeros(B * nh, dtype=torch.int32, device=DEV)
    outs = []
    for flags in (0, 0x200, 0x100, 0x300):
        y = torch.full((B + 2, nh * hs), 7.0, device=DEV, dtype=dtype)
        NV.call("ll_decode_attn", NV.ptr(q), NV.ptr(k), NV.ptr(v), NV.ptr(pos3), NV.ptr(y[1:]),
                NV.LL_BF16 if dtype == torch.bfloat16 else NV.LL_F32, B, nh, hs, S, splits,
                NV.ptr(scratch), NV.ptr(counters), flags, NV.stream())
        torch.cuda.synchronize()
        assert (y[0] == 7.0).all() and (y[-1] == 7.0).all()
        outs.append(y[1:-1].float().cpu().numpy().reshape(B, nh, hs))
    want = _attn_ref(q.float().cpu().numpy().reshape(B, nh, hs), k.float().cpu().numpy(), v.float().cpu().numpy(), n_valid)
    tol = (2.0 ** -8 if dtype == torch.bfloat16 else 2.0 ** -18) * np.abs(want).max() + 1e-6
    assert np.abs(outs[0] - want).max() <= tol, np.abs(outs[0] - want).max()
    for o in outs[1:]:
        assert np.array_equal(outs[0], o)


@pytest.mark.parametrize("S,pos", [(768, 640), (768, 767), (520, 777)])
def test_staged_attention_behind_c_attn_as_programmatic_dependent(S, pos):
    """The race the staging kernel is built around: c_attn (fused RMSNorm, RoPE, K/V row written at the ring slot) followed
    by the attention kernel as its PROGRAMMATIC DEPENDENT, whose bulk copies of K and V run while c_attn is still computing.
    The slot's row holds large garbage beforehand, so a result from the stale copy would be far off.  Same bits as the two
    kernels launched plainly one after the other, over several rounds; the result also matches the float64 softmax."""
    from lit_llama import _native as NV
    from oracle import quant as OQ

    E, nh, hs, group, splits = 4096, 32, 128, 128, 8
    r = np.random.RandomState(S + pos)
    pk, s, z = _bf16_layer(3 * E, E, group, 26)
    lin = _stream_of(pk, s, z, 3 * E, E, group, 1)
    rope_d = torch.from_numpy(OM.build_rope_cache(1024, hs)).to(DEV)
    nw = torch.from_numpy(OQ.bf16_round((1 + 0.1 * r.randn(E)).astype(np.float32))).to(DEV, torch.bfloat16)
    x = torch.from_numpy(OQ.bf16_round(r.randn(1, E).astype(np.float32))).to(DEV, torch.bfloat16)
    g = torch.Generator(device=DEV).manual_seed(pos)
    k0 = torch.randn((1, nh, S, hs), device=DEV, generator=g).bfloat16()
    v0 = torch.randn((1, nh, S, hs), device=DEV, generator=g).bfloat16()
    slot, n_valid = pos % S, min(pos + 1, S)
    k0[:, :, slot] = 300.0
    v0[:, :, slot] = -300.0
    pos3 = torch.tensor([pos, slot, n_valid], dtype=torch.int32, device=DEV)
    scratch = torch.empty(NV.attn_scratch_bytes(1, nh, hs, splits) // 4 + 16, dtype=torch.float32, device=DEV)
    counters = torch.zeros(nh, dtype=torch.int32, device=DEV)

    def chain(pdl):
        kc, vc = k0.clone(), v0.clone()
        q = torch.zeros((1, E), device=DEV, dtype=torch.bfloat16)
        y = torch.zeros((1, E), device=DEV, dtype=torch.bfloat16)
        a = NV.GemvArgs()
        w, a.stream_kind = lin.stream_for(1)
        a.w, a.N, a.K, a.group, a.M = w.data_ptr(), 3 * E, E, group, 1
        a.x, a.y, a.norm_weight, a.eps = x.data_ptr(), q.data_ptr(), nw.data_ptr(), 1e-5
        a.epilogue, a.flags, a.grid = NV.EPI_ROPE_KV, (NV.GEMV_PDL | NV.GEMV_EVEN) if pdl else 0, 0
        a.rope, a.pos, a.k_cache, a.v_cache = rope_d.data_ptr(), pos3.data_ptr(), kc.data_ptr(), vc.data_ptr()
        a.n_head, a.head_size, a.cache_len = nh, hs, S
        torch.cuda.synchronize()
        NV.gemv(a)
        NV.call("ll_decode_attn", NV.ptr(q), NV.ptr(kc), NV.ptr(vc), NV.ptr(pos3), NV.ptr(y), NV.LL_BF16, 1, nh, hs, S, splits,
                NV.ptr(scratch), NV.ptr(counters), NV.GEMV_PDL if pdl else 0, NV.stream())
        torch.cuda.synchronize()
        return y.float().cpu().numpy(), q, kc, vc

    plain, q, kc, vc = chain(False)
    assert float(kc[:, :, slot].abs().max()) < 100.0, "c_attn did not write the K row"
    want = _attn_ref(q.float().cpu().numpy().reshape(1, nh, hs), kc.float().cpu().numpy(), vc.float().cpu().numpy(), n_valid)
    assert np.abs(plain.reshape(1, nh, hs) - want).max() <= 2.0 ** -8 * np.abs(want).max() + 1e-6
    for _ in range(5):
        assert np.array_equal(chain(True)[0], plain)


BITS = 4   # weight width the helpers below build layers with (the gptq.int8 test switches it to 8)


def _bf16_layer(N_, K, group, seed):
    from oracle import quant as OQ

    r = np.random.RandomState(seed)
    Q, s, z = OQ.rtn_quantize((r.randn(N_, K) * 0.02).astype(np.float32), BITS, group)
    s, z = OQ.bf16_round(s), OQ.bf16_round(z)
    return OQ.pack_weight(Q, s, z, BITS, group, "bf16"), s, z


def _stream_of(packed, s, z, N_, K, group, M=1):
    from lit_llama.quantization import ColBlockQuantizedLinear

    torch.set_default_dtype(torch.bfloat16)
    try:
        lin = ColBlockQuantizedLinear(K, N_, False, bits=BITS, tile_cols=group)
    finally:
        torch.set_default_dtype(torch.float32)
    lin = lin.to(DEV)
    lin.scales.copy_(torch.from_numpy(s)); lin.zeros.copy_(torch.from_numpy(z))
    lin.quant_weight.copy_(torch.from_numpy(packed))
    assert lin.prepare(M) is not None
    return lin


def _exact(x, packed, s, z, group):
    from oracle import quant as OQ

    return OQ.linear_forward_exact(x, packed, s, z, BITS, group)


def _mag(x, packed, s, z, group):
    from oracle import quant as OQ

    return np.abs(x).astype(np.float64) @ np.abs(OQ.get_weight(packed, s, z, BITS, group)).astype(np.float64).T


@pytest.mark.parametrize("M", [1, 2])
def test_specialised_gemv_epilogues_with_fused_norm_at_7b_shapes(M):
    """ll_int4_gemv called directly, one launch per layer kind of the decode chain, against the exact oracle:
    c_attn (12288 x 4096, fused RMSNorm, RoPE + KV write), attn.c_proj (4096 x 4096, residual), fc1|fc2 (2 x 11008 x 4096,
    fused RMSNorm, SwiGLU), mlp.c_proj (4096 x 11008, residual), lm_head (32000 x 4096, fused RMSNorm, store)."""
    from lit_llama import _native as NV
    from oracle import quant as OQ

    E, H, V, nh, hs, S, group = 4096, 11008, 32000, 32, 128, 64, 128
    r = np.random.RandomState(5 + M)
    eps = 1e-5

    def bf(a):
        return OQ.bf16_round(np.asarray(a, np.float32))

    def dev(a):
        return torch.from_numpy(np.ascontiguousarray(a)).to(DEV, torch.bfloat16)

    def gemv(lin, x, y, epi, norm_w=None, lin2=None, **kw):
        a = NV.GemvArgs()
        w, a.stream_kind = lin.stream_for(M)
        a.w = w.data_ptr()
        if lin2 is not None:
            a.w2 = lin2.stream_for(M)[0].data_ptr()
        a.N, a.K, a.group, a.M = lin.out_features, lin.in_features, group, M
        a.x, a.y = x.data_ptr(), y.data_ptr()
        if norm_w is not None:
            a.norm_weight, a.eps = norm_w.data_ptr(), eps
        a.epilogue, a.flags, a.grid = epi, 0, 0
        for k, v in kw.items():
            setattr(a, k, v)
        NV.gemv(a)
        torch.cuda.synchronize()

    def close(y, exact, mag, extra=0.0):
        assert (np.abs(y - exact) <= np.abs(exact) * 2.0 ** -7 + mag * 2.0 ** -17 + extra).all(), float(np.abs(y - exact).max())

    gemv_counts(reset=True)
    xraw = bf(r.randn(M, E))
    nw = bf(1 + 0.1 * r.randn(E))
    xn = OM.rmsnorm(xraw, nw, eps, "bf16", "cuda")

    # ---- lm_head: fused norm + store
    pk, s, z = _bf16_layer(V, E, group, 21)
    lin = _stream_of(pk, s, z, V, E, group, M)
    y = torch.empty((M, V), device=DEV, dtype=torch.bfloat16)
    gemv(lin, dev(xraw), y, NV.EPI_STORE, norm_w=dev(nw))
    close(y.float().cpu().numpy(), _exact(xn, pk, s, z, group), _mag(xn, pk, s, z, group))

    # ---- fc1|fc2: fused norm + SwiGLU
    pk1, s1, z1 = _bf16_layer(H, E, group, 22)
    pk2, s2, z2 = _bf16_layer(H, E, group, 23)
    l1, l2 = _stream_of(pk1, s1, z1, H, E, group, M), _stream_of(pk2, s2, z2, H, E, group, M)
    y = torch.empty((M, H), device=DEV, dtype=torch.bfloat16)
    gemv(l1, dev(xraw), y, NV.EPI_SWIGLU, norm_w=dev(nw), lin2=l2)
    a, b = bf(_exact(xn, pk1, s1, z1, group)), bf(_exact(xn, pk2, s2, z2, group))
    want = bf(bf(a / (1 + np.exp(-a))) * b)
    got = y.float().cpu().numpy()
    # two rounded factors: a flip of either input's last bf16 bit moves the product by ~2^-8 relative
    assert (np.abs(got - want) <= np.abs(want) * 2.0 ** -6 + 1e-6).mean() >= 0.999
    assert np.abs(got - want).max() <= 2.0 ** -5 * np.abs(want).max()

    # ---- mlp.c_proj: K = 11008, residual add in place
    pk, s, z = _bf16_layer(E, H, group, 24)
    lin = _stream_of(pk, s, z, E, H, group, M)
    h = bf(r.randn(M, H))
    res = bf(r.randn(M, E))
    y = dev(res)
    gemv(lin, dev(h), y, NV.EPI_RESIDUAL)
    ex = _exact(h, pk, s, z, group)
    want = bf(res + bf(ex))
    got = y.float().cpu().numpy()
    assert (np.abs(got - want) <= 2.0 ** -7 * np.maximum(np.abs(want), np.abs(ex)) + _mag(h, pk, s, z, group) * 2.0 ** -17).all()

    # ---- attn.c_proj: K = 4096, residual
    pk, s, z = _bf16_layer(E, E, group, 25)
    lin = _stream_of(pk, s, z, E, E, group, M)
    att = bf(r.randn(M, E))
    y = dev(res)
    gemv(lin, dev(att), y, NV.EPI_RESIDUAL)
    ex = _exact(att, pk, s, z, group)
    want = bf(res + bf(ex))
    got = y.float().cpu().numpy()
    assert (np.abs(got - want) <= 2.0 ** -7 * np.maximum(np.abs(want), np.abs(ex)) + _mag(att, pk, s, z, group) * 2.0 ** -17).all()

    # ---- c_attn: fused norm + RoPE (interleaved pairs, ref model.py:312-329) + KV write at the ring slot
    pk, s, z = _bf16_layer(3 * E, E, group, 26)
    lin = _stream_of(pk, s, z, 3 * E, E, group, M)
    pos, slot = 37, 5
    rope = OM.build_rope_cache(64, hs)
    q = torch.zeros((M, E), device=DEV, dtype=torch.bfloat16)
    kc = torch.zeros((M, nh, S, hs), device=DEV, dtype=torch.bfloat16)
    vc = torch.zeros((M, nh, S, hs), device=DEV, dtype=torch.bfloat16)
    pos3 = torch.tensor([pos, slot, min(pos + 1, S)], dtype=torch.int32, device=DEV)
    rope_d = torch.from_numpy(rope).to(DEV)
    gemv(lin, dev(xraw), q, NV.EPI_ROPE_KV, norm_w=dev(nw), rope=rope_d.data_ptr(), pos=pos3.data_ptr(),
         k_cache=kc.data_ptr(), v_cache=vc.data_ptr(), n_head=nh, head_size=hs, cache_len=S)
    qkv = bf(_exact(xn, pk, s, z, group))
    qr = OM.apply_rope(qkv[:, :E].reshape(M, 1, nh, hs), rope[pos:pos + 1], "bf16").reshape(M, E)
    kr = OM.apply_rope(qkv[:, E:2 * E].reshape(M, 1, nh, hs), rope[pos:pos + 1], "bf16").reshape(M, nh, hs)
    vr = qkv[:, 2 * E:].reshape(M, nh, hs)
    scale = np.abs(qkv).max()
    assert np.abs(q.float().cpu().numpy() - qr).max() <= 2.0 ** -6 * scale
    assert np.abs(kc[:, :, slot].float().cpu().numpy() - kr).max() <= 2.0 ** -6 * scale
    assert np.abs(vc[:, :, slot].float().cpu().numpy() - vr).max() <= 2.0 ** -6 * scale
    kc[:, :, slot] = 0; vc[:, :, slot] = 0
    assert float(kc.abs().sum()) == 0 and float(vc.abs().sum()) == 0, "the KV epilogue wrote outside its slot"

    spec, _ = gemv_counts()
    m1 = 1 if M == 1 else 0
    for e, n in ((NV.EPI_ROPE_KV, 1), (NV.EPI_RESIDUAL, 0), (NV.EPI_SWIGLU, 1), (NV.EPI_STORE, 1)):
        assert spec[e, n, m1] > 0, spec.tolist()


def test_q8_gemv_epilogues_at_7b_shapes(monkeypatch):
    """gptq.int8 (ColBlockQuantizedLinear bits = 8, ref utils.py:182-184): stream kind 2 of the batch-1 GEMV - the same
    five layer kinds of the decode chain with every fused prologue / epilogue, against the exact oracle at 8 bits; then a
    row with an ODD number of 128-element k-blocks (43), a ragged last CTA share and canary rows around the output."""
    import sys
    from lit_llama import _native as NV
    from oracle import quant as OQ

    monkeypatch.setattr(sys.modules[__name__], "BITS", 8)
    test_specialised_gemv_epilogues_with_fused_norm_at_7b_shapes(1)
    N_, K, group = 1040, 5504, 128
    pk, s, z = _bf16_layer(N_, K, group, 33)
    lin = _stream_of(pk, s, z, N_, K, group, 1)
    x = OQ.bf16_round(np.random.RandomState(6).randn(1, K).astype(np.float32))
    buf = torch.full((3, N_), 1024.0, device=DEV, dtype=torch.bfloat16)
    a = NV.GemvArgs()
    w, a.stream_kind = lin.stream_for(1)
    assert a.stream_kind == NV.STREAM_W8 and w is not None
    a.w, a.N, a.K, a.group, a.M = w.data_ptr(), N_, K, group, 1
    xd = torch.from_numpy(x).to(DEV, torch.bfloat16)
    a.x, a.y = xd.data_ptr(), buf[1:].data_ptr()
    a.epilogue, a.flags, a.grid = NV.EPI_STORE, 0, 0
    NV.gemv(a)
    torch.cuda.synchronize()
    got = buf.float().cpu().numpy()
    assert (got[0] == 1024.0).all() and (got[2] == 1024.0).all()
    exact = _exact(x, pk, s, z, group)
    assert (np.abs(got[1:2] - exact) <= np.abs(exact) * 2.0 ** -8 + _mag(x, pk, s, z, group) * 2.0 ** -18 + 1e-30).all()
    # the module's forward takes the same route at batch 1 and the reference-semantics kernel at batch 2
    y1 = lin(xd).float().cpu().numpy()
    assert np.array_equal(y1, got[1:2])
    x2 = torch.cat([xd, xd])
    y2 = lin(x2).float().cpu().numpy()
    ref = OQ.linear_forward(OQ.bf16_round(np.concatenate([x, x])), pk, s, z, None, 8, group, "bf16")
    assert np.abs(y2 - ref).max() <= 2.0 ** -6 * np.abs(ref).max()


@pytest.mark.parametrize("M", [1, 2])
@pytest.mark.parametrize("N_,K", [(4096, 4096), (1040, 11008)])
def test_gemv_writes_nothing_outside_its_output(M, N_, K):
    """compute-sanitizer is closed on this pool (profiles/r02_sanitizer_closed.txt): stray stores are looked for directly -
    the output rows sit between canary rows that must come back untouched (both kernels, a ragged last CTA share)."""
    from lit_llama import _native as NV
    from oracle import quant as OQ

    group = 128
    pk, s, z = _bf16_layer(N_, K, group, 31)
    lin = _stream_of(pk, s, z, N_, K, group, M)
    x = OQ.bf16_round(np.random.RandomState(3).randn(M, K).astype(np.float32))
    buf = torch.full((M + 2, N_), 1024.0, device=DEV, dtype=torch.bfloat16)
    a = NV.GemvArgs()
    w, a.stream_kind = lin.stream_for(M)
    a.w = w.data_ptr()
    a.N, a.K, a.group, a.M = N_, K, group, M
    xd = torch.from_numpy(x).to(DEV, torch.bfloat16)
    a.x, a.y = xd.data_ptr(), buf[1:].data_ptr()
    a.epilogue, a.flags, a.grid = NV.EPI_STORE, 0, 0
    NV.gemv(a)
    torch.cuda.synchronize()
    got = buf.float().cpu().numpy()
    assert (got[0] == 1024.0).all() and (got[-1] == 1024.0).all(), "the GEMV wrote outside its output rows"
    exact = _exact(x, pk, s, z, group)
    assert (np.abs(got[1:-1] - exact) <= np.abs(exact) * 2.0 ** -8 + _mag(x, pk, s, z, group) * 2.0 ** -18 + 1e-30).all()
