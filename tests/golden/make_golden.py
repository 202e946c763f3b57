"""Generate golden input/output vectors by running the UNMODIFIED reference
(/root/reference) in the build container.  Not runnable on the GPU box (the reference is
absent there); the ``*.npz`` files it writes are committed beside it.

    python tests/golden/make_golden.py

The reference imports once ``lightning`` is stubbed (it is needed only for two class
names at lit_llama/utils.py:13) - SURVEY.md appendix A.
"""
import importlib.util
import io
import sys
import types
from contextlib import redirect_stdout
from pathlib import Path

import numpy as np
import torch

REF = Path("/root/reference")
OUT = Path(__file__).parent


def _import_reference():
    L = types.ModuleType("lightning")
    F = types.ModuleType("lightning.fabric")
    S = types.ModuleType("lightning.fabric.strategies")
    S.DeepSpeedStrategy = type("DeepSpeedStrategy", (), {})
    S.FSDPStrategy = type("FSDPStrategy", (), {})
    F.strategies = S
    L.fabric = F
    L.Fabric = object
    L.seed_everything = lambda s: torch.manual_seed(s)
    sys.modules.update({"lightning": L, "lightning.fabric": F, "lightning.fabric.strategies": S})
    sys.path.insert(0, str(REF))
    import lit_llama  # noqa
    from lit_llama import quantization, utils, model  # noqa

    def load(name, path):
        spec = importlib.util.spec_from_file_location(name, path)
        m = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(m)
        return m

    gen = load("ref_generate", REF / "generate.py")
    try:
        gptq = load("ref_gptq", REF / "quantize" / "gptq.py")
    except Exception as e:  # datasets import etc.
        print("quantize/gptq.py import failed:", e)
        gptq = None
    return quantization, utils, model, gen, gptq


def f32(t):
    return t.detach().float().cpu().numpy().copy()


def rs(seed):
    return np.random.RandomState(seed)


def golden_quant(qz):
    """pack / unpack / fallback-forward vectors for ColBlockQuantizedLinear (rows C,P,G,F,Z)."""
    out = {}
    cases = [  # name, N, K, bits, tile_cols
        ("a", 48, 128, 4, -1), ("b", 32, 256, 4, 128), ("c", 16, 192, 4, 64),
        ("d", 24, 64, 8, -1), ("e", 8, 96, 8, 32), ("f", 40, 80, 4, 16), ("g", 8, 70, 4, 32),
    ]
    meta = []
    for name, N, K, bits, tile in cases:
        r = rs(hash((N, K, bits)) % 2**31)
        W = (r.randn(N, K) * 0.02).astype(np.float32)
        W[1] = 0.0                      # all-zero row -> (-1,+1) branch
        W[2] = np.abs(W[2])             # xmin clamps to 0
        x = r.randn(3, 5, K).astype(np.float32)
        for wdt_name, wdt in (("fp32", torch.float32), ("bf16", torch.bfloat16)):
            torch.set_default_dtype(wdt)
            lin = qz.ColBlockQuantizedLinear(K, N, False, bits=bits, tile_cols=tile)
            torch.set_default_dtype(torch.float32)
            dummy = torch.nn.Linear(K, N, bias=False)
            g = qz.GPTQQuantizer(dummy, bits=bits, groupsize=tile)
            Wt = torch.from_numpy(W)
            tc = lin.tile_cols
            Q = torch.empty_like(Wt)
            for j in range(lin.scales.shape[1]):
                sl = slice(j * tc, (j + 1) * tc)
                s, z = g.find_params_weight(Wt[:, sl])
                lin.scales[:, j : j + 1] = s
                lin.zeros[:, j : j + 1] = z
                Q[:, sl] = g.quantize_weight(Wt[:, sl], s, z, g.maxq)
                if wdt_name == "fp32":
                    out[f"{name}.scale_f32.{j}"] = f32(s)
                    out[f"{name}.zero_f32.{j}"] = f32(z)
            if wdt_name == "fp32":
                out[f"{name}.Q"] = f32(Q)
            lin.pack_weight(Q.to(wdt))
            assert lin.quant_weight.stride() == (1, N)
            out[f"{name}.{wdt_name}.packed"] = lin.quant_weight.numpy().copy()
            out[f"{name}.{wdt_name}.scales"] = f32(lin.scales)
            out[f"{name}.{wdt_name}.zeros"] = f32(lin.zeros)
            out[f"{name}.{wdt_name}.get_weight"] = f32(lin.get_weight(dtype=wdt))
            out[f"{name}.{wdt_name}.forward"] = f32(lin(torch.from_numpy(x).to(wdt)))
        out[f"{name}.W"] = W
        out[f"{name}.x"] = x
        meta.append((name, N, K, bits, tile))
    out["cases"] = np.array([f"{n},{N},{K},{b},{t}" for n, N, K, b, t in meta])
    np.savez_compressed(OUT / "quant.npz", **out)
    print("quant.npz", len(out))


def golden_gptq_solver(qz):
    """GPTQQuantizer.collect_input_stats + quantize on one small layer (row Z2)."""
    out = {}
    for name, N, K, actorder in (("plain", 24, 64, False), ("act", 16, 96, True)):
        r = rs(77 + N)
        W = (r.randn(N, K) * 0.05).astype(np.float32)
        X = [r.randn(2, 7, K).astype(np.float32) for _ in range(3)]
        lin = torch.nn.Linear(K, N, bias=False)
        with torch.no_grad():
            lin.weight.copy_(torch.from_numpy(W))
        g = qz.GPTQQuantizer(lin, bits=4, groupsize=-1, actorder=actorder, blocksize=32)
        for xi in X:
            g.collect_input_stats(None, (torch.from_numpy(xi),), None)
        out[f"{name}.H"] = f32(g.H)
        qm, err = g.quantize()
        out[f"{name}.W"] = W
        for i, xi in enumerate(X):
            out[f"{name}.X{i}"] = xi
        out[f"{name}.packed"] = qm.quant_weight.numpy().copy()
        out[f"{name}.scales"] = f32(qm.scales)
        out[f"{name}.zeros"] = f32(qm.zeros)
        out[f"{name}.error"] = np.float64(err)
        out[f"{name}.deq"] = f32(qm.get_weight())
    np.savez_compressed(OUT / "gptq_solver.npz", **out)
    print("gptq_solver.npz")


def golden_ops(model):
    """RMSNorm / RoPE / attention-with-cache vectors (rows N, R, A)."""
    out = {}
    r = rs(5)
    x = r.randn(2, 6, 64).astype(np.float32) * 3
    scale = (1 + 0.1 * r.randn(64)).astype(np.float32)
    norm = model.RMSNorm(64)
    with torch.no_grad():
        norm.scale.copy_(torch.from_numpy(scale))
    out["rms.x"], out["rms.scale"] = x, scale
    out["rms.fp32"] = f32(norm(torch.from_numpy(x)))
    out["rms.bf16"] = f32(norm.to(torch.bfloat16)(torch.from_numpy(x).bfloat16()))
    cache = model.build_rope_cache(seq_len=96, n_elem=32, dtype=torch.int64, device="cpu")
    assert cache.dtype == torch.float32
    out["rope.cache"] = f32(cache)
    xq = r.randn(2, 7, 4, 32).astype(np.float32)
    pos = torch.tensor([3, 4, 5, 6, 7, 50, 95])
    out["rope.x"] = xq
    out["rope.pos"] = pos.numpy()
    out["rope.fp32"] = f32(model.apply_rope(torch.from_numpy(xq), cache.index_select(0, pos)))
    out["rope.bf16"] = f32(model.apply_rope(torch.from_numpy(xq).bfloat16(), cache.index_select(0, pos)))
    np.savez_compressed(OUT / "ops.npz", **out)
    print("ops.npz")


TINY = dict(block_size=64, vocab_size=100, n_layer=2, n_head=4, n_embd=128)


def tiny_weights(cfg):
    """Deterministic weights keyed by the reference's state-dict names (numpy legacy RNG)."""
    r = rs(1234)
    E, L = cfg["n_embd"], cfg["n_layer"]
    V = cfg["vocab_size"] + (-cfg["vocab_size"]) % 64
    H = int(2 * 4 * E / 3)
    H = H + (-H) % 256
    sd = {"lm_head.weight": (V, E), "transformer.wte.weight": (V, E)}
    for i in range(L):
        p = f"transformer.h.{i}."
        sd[p + "attn.c_attn.weight"] = (3 * E, E)
        sd[p + "attn.c_proj.weight"] = (E, E)
        sd[p + "mlp.c_fc1.weight"] = (H, E)
        sd[p + "mlp.c_fc2.weight"] = (H, E)
        sd[p + "mlp.c_proj.weight"] = (E, H)
    out = {k: (r.randn(*s) * (0.5 if "wte" in k else 0.08)).astype(np.float32) for k, s in sd.items()}
    for i in range(L):
        for n in ("rms_1", "rms_2"):
            out[f"transformer.h.{i}.{n}.scale"] = (1 + 0.1 * r.randn(E)).astype(np.float32)
    out["transformer.ln_f.scale"] = (1 + 0.1 * r.randn(E)).astype(np.float32)
    return out


def golden_model(qz, utils, model, gen, gptq):
    """Tiny LLaMA (BASELINE.json configs[0]): fp32 forward / KV-cache decode / generate,
    the reference GPTQ writer's int4 checkpoint and its forward + generate, and a
    group-32 RTN int4 model in fp32 and bf16 (rows M, A, K, S, Q, L)."""
    out = {}
    cfg = model.LLaMAConfig(**TINY)
    W = tiny_weights(TINY)
    m = model.LLaMA(cfg)
    m.load_state_dict({k: torch.from_numpy(v) for k, v in W.items()})
    m.eval()
    r = rs(99)
    idx = r.randint(0, 100, size=(2, 16)).astype(np.int64)
    out["idx"] = idx
    with torch.no_grad():
        out["fp32.logits"] = f32(m(torch.from_numpy(idx)))
        # KV-cache path: prefill 10 then 6 single-token steps, max_seq_length 12 -> rolls
        msl = 12
        steps = []
        pos = torch.arange(0, 10)
        lg = m(torch.from_numpy(idx[:1, :10]), msl, pos)
        steps.append(f32(lg[:, -1]))
        for t in range(10, 16):
            pos = torch.tensor([t])
            lg = m(torch.from_numpy(idx[:1, t : t + 1]), msl, pos)
            steps.append(f32(lg[:, -1]))
        out["fp32.kv_steps"] = np.stack(steps)
        out["fp32.kv_k0"] = f32(m.kv_caches[0][0])
        out["fp32.kv_v1"] = f32(m.kv_caches[1][1])
        m.reset_cache()
        prompt = torch.from_numpy(idx[0, :8]).to(torch.int64)
        out["fp32.generate_top1"] = gen.generate(m, prompt, 12, top_k=1).numpy()
        m.reset_cache()
        out["fp32.generate_roll"] = gen.generate(m, prompt, 12, max_seq_length=10, top_k=1).numpy()
        m.reset_cache()

    # ---- reference GPTQ writer (per-row scales, actorder) -> state dict = checkpoint layout
    if gptq is not None:
        torch.manual_seed(0)
        samples = torch.from_numpy(rs(3).randint(0, 100, size=(8, 64)).astype(np.int64))
        with redirect_stdout(io.StringIO()):
            gptq.llama_blockwise_quantization(m, samples, "cpu", bits=4)
        sd = m.state_dict()
        path = OUT / "_tiny_gptq.pth"
        torch.save(sd, path)
        with utils.lazy_load(path) as ck:
            assert utils.llama_model_lookup is not None
            with utils.quantization("gptq.int4"):
                mq = model.LLaMA(cfg)
            mq.load_state_dict(ck)
        path.unlink()
        mq.eval()
        for k, v in sd.items():
            if v.dtype == torch.uint8:
                assert v.stride() == (1, v.shape[0]), (k, v.stride())
                out["gptq.sd." + k] = v.numpy().copy()
            elif "scales" in k or "zeros" in k:
                out["gptq.sd." + k] = f32(v)
        with torch.no_grad():
            out["gptq.logits"] = f32(mq(torch.from_numpy(idx)))
            out["gptq.generate_top1"] = gen.generate(mq, prompt, 12, top_k=1).numpy()
            mq.reset_cache()

    # ---- RTN group-32 int4, fp32 and bf16 activations (fallback forward path)
    for dt_name, dt in (("fp32", torch.float32), ("bf16", torch.bfloat16)):
        torch.set_default_dtype(dt)
        holder = {}
        for k, v in W.items():
            if not k.endswith(".weight") or "wte" in k:
                continue
            N, K = v.shape
            lin = qz.ColBlockQuantizedLinear(K, N, False, bits=4, tile_cols=32)
            g = qz.GPTQQuantizer(torch.nn.Linear(K, N, bias=False), bits=4, groupsize=32)
            Wt = torch.from_numpy(v)
            Qw = torch.empty_like(Wt)
            for j in range(lin.scales.shape[1]):
                sl = slice(j * 32, (j + 1) * 32)
                s, z = g.find_params_weight(Wt[:, sl])
                lin.scales[:, j : j + 1] = s
                lin.zeros[:, j : j + 1] = z
                Qw[:, sl] = g.quantize_weight(Wt[:, sl], s, z, 15)
            lin.pack_weight(Qw.to(dt))
            holder[k[: -len(".weight")]] = lin
        mq = model.LLaMA(cfg)
        torch.set_default_dtype(torch.float32)
        mq.load_state_dict({k: torch.from_numpy(v) for k, v in W.items()})
        mq = mq.to(dt)
        for name, lin in holder.items():
            parent, attr = name.rsplit(".", 1) if "." in name else ("", name)
            setattr(mq.get_submodule(parent) if parent else mq, attr, lin)
            if dt_name == "fp32":
                out[f"rtn32.{name}.packed"] = lin.quant_weight.numpy().copy()
                out[f"rtn32.{name}.scales"] = f32(lin.scales)
                out[f"rtn32.{name}.zeros"] = f32(lin.zeros)
            else:
                out[f"rtn32bf16.{name}.packed"] = lin.quant_weight.numpy().copy()
                out[f"rtn32bf16.{name}.scales"] = f32(lin.scales)
                out[f"rtn32bf16.{name}.zeros"] = f32(lin.zeros)
        mq.eval()
        with torch.no_grad():
            out[f"rtn32.{dt_name}.logits"] = f32(mq(torch.from_numpy(idx)))
            steps = []
            msl = 16
            lg = mq(torch.from_numpy(idx[:1, :10]), msl, torch.arange(0, 10))
            steps.append(f32(lg[:, -1]))
            for t in range(10, 16):
                lg = mq(torch.from_numpy(idx[:1, t : t + 1]), msl, torch.tensor([t]))
                steps.append(f32(lg[:, -1]))
            out[f"rtn32.{dt_name}.kv_steps"] = np.stack(steps)
            mq.reset_cache()
            out[f"rtn32.{dt_name}.generate_top1"] = gen.generate(mq, prompt, 12, top_k=1).numpy()
            mq.reset_cache()
    np.savez_compressed(OUT / "model_tiny.npz", **out)
    print("model_tiny.npz", sum(v.nbytes for v in out.values()) // 1024, "KiB")


def main():
    torch.manual_seed(0)
    torch.set_num_threads(8)
    qz, utils, model, gen, gptq = _import_reference()
    golden_quant(qz)
    golden_gptq_solver(qz)
    golden_ops(model)
    golden_model(qz, utils, model, gen, gptq)


if __name__ == "__main__":
    main()
