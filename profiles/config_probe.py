"""Measurements of BASELINE.json configs 3 and 4 through the public model API.

    python profiles/config_probe.py int8      # config 3: LLaMA-7B llm.int8, batch 32 single-token steps, 2048-slot KV cache
    python profiles/config_probe.py prefill   # config 4: LLaMA-13B GPTQ int4 g128 prefill, batch 8 x 2048 tokens

Device time with CUDA events after warm-up; one JSON line per config (algorithmic bytes / flops, achieved rate,
fraction of the measured peak in MEASURED_PEAKS.json).  Random-init weights, synthetic ids.
"""
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "lit-llama-ja_b200")]
import bench  # noqa: E402
from lit_llama import LLaMA  # noqa: E402
from lit_llama.quantization import Linear8bitLt  # noqa: E402
from lit_llama.utils import EmptyInitOnDevice  # noqa: E402

dev = torch.device("cuda:0")
try:
    PEAKS = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
except Exception:
    PEAKS = {}


def timed(fn, warm=2, reps=5):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def quantized_nk(model, cls):
    return sum(m.in_features * m.out_features for m in model.modules() if isinstance(m, cls))


@torch.no_grad()
def int8_batched_decode(B=32, S=2048, outlier_frac=0.001):
    with EmptyInitOnDevice(device=dev, dtype=torch.bfloat16, quantization_mode="llm.int8"):
        model = LLaMA.from_name("7B")
    gen = torch.Generator(device=dev).manual_seed(0)
    for mod in model.modules():
        if isinstance(mod, Linear8bitLt):
            mod._quantize_weight(torch.randn((mod.out_features, mod.in_features), device=dev, dtype=torch.bfloat16, generator=gen) * 0.02)
    wte = model.transformer.wte.weight
    wte.copy_(torch.randn(wte.shape, device=dev, generator=gen) * 0.02)
    # ~0.1 % of the embedding features are scaled x20 so that the outlier path (|x| >= 6) is exercised
    cols = torch.randperm(wte.shape[1], device=dev, generator=gen)[: max(1, int(outlier_frac * wte.shape[1]))]
    wte[:, cols] *= 400.0
    for n, p in model.named_parameters():
        if n.endswith(".scale"):
            p.fill_(1.0)
    model.eval()
    idx = torch.randint(0, 32000, (B, 1), device=dev, generator=gen)
    pos = torch.tensor([S - 1], device=dev)
    model(idx, S, torch.tensor([0], device=dev))          # allocates the KV caches
    for kc, vc in model.kv_caches:                        # a full cache of plausible values
        kc.normal_(0, 0.5, generator=gen)
        vc.normal_(0, 0.5, generator=gen)
    ms = timed(lambda: model(idx, S, pos), warm=2, reps=5)
    import os
    if os.environ.get("PROBE_OUTLIERS") == "1":   # outlier columns (any |x| >= 6 over the batch) per Linear8bitLt input
        counts = {}
        hooks = []
        for name, mod in model.named_modules():
            if isinstance(mod, Linear8bitLt):
                hooks.append(mod.register_forward_pre_hook(
                    lambda m, a, name=name: counts.__setitem__(name, int((a[0].reshape(-1, a[0].shape[-1]).half().abs() >= 6.0).any(0).sum()))))
        model(idx, S, pos)
        for h in hooks:
            h.remove()
        kinds = {}
        for k, v in counts.items():
            kinds.setdefault(k.split(".")[-1] if "lm_head" not in k else "lm_head", []).append(v)
        print("outlier columns per layer kind (min / median / max):",
              {k: (min(v), sorted(v)[len(v) // 2], max(v)) for k, v in kinds.items()}, flush=True)
    if os.environ.get("PROBE_NCU") == "1":   # one step inside the profiler range (ncu --profile-from-start off)
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        model(idx, S, pos)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
    # the same step replayed from a CUDA graph: the per-module path issues ~30 launches per layer from Python and is
    # host-bound at this size; the replay shows the device time of the same kernels
    ms_graph = None
    try:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2):
                model(idx, S, pos)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            model(idx, S, pos)
        ms_graph = timed(g.replay, warm=2, reps=10)
    except Exception as e:  # noqa: BLE001
        ms_graph = f"capture failed: {type(e).__name__}: {e}"[:300]
    cfg = model.config
    wbytes = quantized_nk(model, Linear8bitLt)            # int8: one byte per weight
    kvbytes = 2 * cfg.n_layer * B * S * cfg.n_embd * 2
    peak = PEAKS.get("hbm_gbs")
    gbs = (wbytes + kvbytes) / ms / 1e6
    return {"config": "3: LLaMA-7B llm.int8, batch 32 single-token step, 2048-slot KV cache (full)", "ms_per_step": ms,
            "tokens_per_s": B / ms * 1e3, "weight_bytes": wbytes, "kv_bytes": kvbytes, "achieved_gbs": gbs,
            "hbm_peak_gbs": peak, "frac": gbs / peak if peak else None, "path": "per-module (Linear8bitLt + ll_decode_attn)",
            "ms_per_step_graph_replay": ms_graph,
            "graph_replay_gbs": (wbytes + kvbytes) / ms_graph / 1e6 if isinstance(ms_graph, float) else None}


@torch.no_grad()
def int4_prefill(B=8, T=2048):
    from lit_llama.quantization import ColBlockQuantizedLinear

    model = bench.build_model(dev, name="13B")
    gen = torch.Generator(device=dev).manual_seed(0)
    idx = torch.randint(0, 32000, (B, T), device=dev, generator=gen)

    def run():
        model(idx)                                        # no-cache forward of the whole batch (evaluate/full.py style)
    ms = timed(run, warm=1, reps=3)
    nk = quantized_nk(model, ColBlockQuantizedLinear)
    flops = 2.0 * B * T * nk
    peak = PEAKS.get("bf16_tflops_sustained")
    tf = flops / ms / 1e9
    return {"config": "4: LLaMA-13B GPTQ int4 g128 prefill, batch 8 x 2048 tokens", "ms": ms, "tokens_per_s": B * T / ms * 1e3,
            "linear_flops": flops, "achieved_tflops": tf, "bf16_peak_tflops_sustained": peak,
            "frac": tf / peak if peak else None,
            "path": "get_weight_int4_bf16 unpack + tcgen05 GEMM per Linear; SDPA, norms, RoPE included in the time"}


@torch.no_grad()
def int8_parts(B=32, S=2048):
    """Where a config-3 step goes: each Linear8bitLt shape at M = 32 and the decode attention at B = 32, S = 2048."""
    from lit_llama import _native as NV

    E, H, nh, hs = 4096, 11008, 32, 128
    out = {}
    for name, (n, k) in {"c_attn": (3 * E, E), "c_proj": (E, E), "fc": (H, E), "down": (E, H), "lm_head": (32000, E)}.items():
        lin = Linear8bitLt(k, n, bias=False, device=dev)
        lin._quantize_weight(torch.randn((n, k), device=dev, dtype=torch.bfloat16) * 0.02)
        x = torch.randn((B, k), device=dev, dtype=torch.bfloat16)
        ms = timed(lambda: lin(x), warm=3, reps=20)
        out[name] = {"us": ms * 1e3, "gbs": n * k / ms / 1e6}
        del lin
    q = torch.randn((B, nh, 1, hs), device=dev, dtype=torch.bfloat16)
    kc = torch.randn((B, nh, S, hs), device=dev, dtype=torch.bfloat16)
    vc = torch.randn((B, nh, S, hs), device=dev, dtype=torch.bfloat16)
    y = torch.empty((B, 1, nh * hs), device=dev, dtype=torch.bfloat16)
    pos3 = torch.tensor([S - 1, S - 1, S], dtype=torch.int32, device=dev)
    for splits in (1, 2, 4, 8):
        ws = NV.Workspace.get(dev).ensure(attn_bytes=NV.attn_scratch_bytes(B, nh, hs, splits), attn_counters=B * nh)
        ms = timed(lambda: NV.call("ll_decode_attn", NV.ptr(q), NV.ptr(kc), NV.ptr(vc), NV.ptr(pos3), NV.ptr(y), NV.LL_BF16,
                                   B, nh, hs, S, splits, NV.ptr(ws.attn_scratch), NV.ptr(ws.attn_counters), 0, NV.stream()), warm=3, reps=20)
        out[f"attn_splits{splits}"] = {"us": ms * 1e3, "gbs": 2 * kc.numel() * 2 / ms / 1e6}
    return out


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "both"
    if which in ("int8", "both"):
        print(json.dumps(int8_batched_decode()), flush=True)
        torch.cuda.empty_cache()
    if which == "int8parts":
        print(json.dumps(int8_parts()), flush=True)
    if which in ("prefill", "both"):
        print(json.dumps(int4_prefill()), flush=True)
