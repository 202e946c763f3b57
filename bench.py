"""Headline benchmark: LLaMA-7B GPTQ-int4 (group 128, bf16) batch-1 decode tokens/s on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A "step" is one pass of the hot path over one synthetic input: ``generate()`` of NEW_TOKENS tokens
after a PROMPT_LEN-token prompt (BASELINE.json configs[1]; the reference prints the same
tokens/sec, generate.py:146-153, prefill included).  Prints ONE JSON line (rank 0).

* ``value``: tokens/s with the prompt already resident in HBM, device-timed (CUDA events).
* ``e2e``:   the same through the public API with a HOST prompt (pinned H2D copy inside the timed
             region) and a D2H read of the generated ids.
* ``roofline``: the int4 GEMV kernel - algorithmic bytes of one token's GEMVs / their device time.
* ``cpu_baseline``: the reference's CPU path (dequantise-then-matmul), restated in oracle/, timed on
             the host cores on a bounded sample.
``--impl reference`` times the UNMODIFIED reference (baseline/_ref, installed by baseline/install_ref.sh) on the host
cores: its own LLaMA.forward over its own ColBlockQuantizedLinear (get_weight + F.linear on a CPU), all host threads.
It imports nothing of this repo's package and loads none of its libraries.

``--gpus N`` with N > 1 runs the SHARDED path the north star names: the same 7B sequence decoded by N tensor-parallel
ranks (column / row sharded Linears, one peer all-reduce per row-parallel layer) - ``scaling: strong``; the LLaMA-65B
tensor-parallel decode step and the N-replica throughput are reported beside it under ``extra``.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
for p in (str(ROOT), str(ROOT / "lit-llama-ja_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

MODEL = os.environ.get("LLB200_BENCH_MODEL", "7B")
PROMPT_LEN = int(os.environ.get("LLB200_BENCH_PROMPT", 512))
NEW_TOKENS = int(os.environ.get("LLB200_BENCH_NEW", 256))
GROUP = 128
METRIC = "decode_tokens_per_sec"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    return ap.parse_args()


def config_dict(n_gpus):
    return {
        "workload": f"LLaMA-{MODEL} GPTQ int4 group{GROUP} bf16, batch-1: {PROMPT_LEN}-token prompt + {NEW_TOKENS} new tokens "
                    f"(generate(), top_k=200, temperature=0.8), random-init weights RTN-quantised",
        "prompt_len": PROMPT_LEN, "new_tokens": NEW_TOKENS, "batch": 1,
        "parallelism": "single GPU" if n_gpus == 1 else
        f"tensor-parallel TP{n_gpus}: one sequence over {n_gpus} GPUs (column-parallel c_attn/c_fc1/c_fc2, row-parallel c_proj "
        f"+ one-traversal push exchange tp_ll_allreduce_add_kernel over NVLink, KV cache sharded by head)",
        "l2": "inputs larger than L2 (3.5 GB of weights streamed per token vs 126 MB L2)",
    }


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1] or [r for _, r in self.rows[-3:]]
        sm, mx, reasons = [], None, set()
        for r in rows:
            try:
                sm.append(float(r[0])); mx = float(r[1])
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------- synthetic model
def build_model(device, name=MODEL, group=GROUP, seed=0, mode="gptq.int4"):
    """Random-init LLaMA of the named size, RTN-quantised per group on the device."""
    import torch
    from lit_llama import LLaMA
    from lit_llama.quantization import ColBlockQuantizedLinear, rtn_quantize_
    from lit_llama.utils import EmptyInitOnDevice, quantization

    with EmptyInitOnDevice(device=device, dtype=torch.bfloat16), quantization(mode, groupsize=group):
        model = LLaMA.from_name(name)
    gen = torch.Generator(device=device).manual_seed(seed)
    with torch.no_grad():
        for mod in model.modules():
            if isinstance(mod, ColBlockQuantizedLinear):
                w = torch.randn((mod.out_features, mod.in_features), device=device, dtype=torch.float32, generator=gen) * 0.02
                rtn_quantize_(mod, w)
                del w
        model.transformer.wte.weight.copy_(torch.randn(model.transformer.wte.weight.shape, device=device, generator=gen) * 0.02)
        for n, p in model.named_parameters():
            if n.endswith(".scale"):
                p.fill_(1.0)
    return model.eval()


# ----------------------------------------------------------------------------- CPU baseline / reference arm
# Model dimensions by name (the reference's llama_configs, lit_llama/model.py:49-56): kept here so that the reference arm
# imports nothing of this repo's package (which would load libllb200.so into the reference's process).
MODEL_DIMS = {"tiny": dict(n_layer=2, n_head=4, n_embd=128),   # tests only (tests/test_host_logic.py)
              "7B": dict(n_layer=32, n_head=32, n_embd=4096), "13B": dict(n_layer=40, n_head=40, n_embd=5120),
              "30B": dict(n_layer=60, n_head=52, n_embd=6656), "65B": dict(n_layer=80, n_head=64, n_embd=8192)}


def _n_hidden(n_embd):
    h = int(2 * (4 * n_embd) / 3)
    return h if h % 256 == 0 else h + 256 - h % 256


def _host_threads():
    import torch

    n = os.cpu_count() or 1
    try:
        n = len(os.sched_getaffinity(0))
    except Exception:
        pass
    torch.set_num_threads(n)        # torchrun exports OMP_NUM_THREADS=1: the CPU arm uses every core it may
    return n


_REF_CPU = {}


def _reference_cpu_model():
    """The unmodified reference's LLaMA (baseline/_ref) at the benchmark's dimensions on the CPU, fp32 (what its
    generate.py picks without CUDA), every Linear its ColBlockQuantizedLinear(bits=4, tile_cols=GROUP) with synthetic
    packed bytes / scales / zeros.  Built once per process."""
    if "m" in _REF_CPU:
        return _REF_CPU["m"]
    import torch
    from baseline import ref_loader

    ref = ref_loader.load()
    d = MODEL_DIMS[MODEL]
    cfg = ref.model.LLaMAConfig(block_size=2048, vocab_size=32000, **d)
    m = ref_loader.build_quantized_model(ref, cfg, tile_cols=GROUP, device="cpu", dtype=torch.float32)
    g = torch.Generator().manual_seed(0)
    with torch.no_grad():
        for name, buf in m.named_buffers():
            if name.endswith("quant_weight"):
                buf.copy_(torch.randint(0, 256, buf.shape, dtype=torch.uint8, generator=g))
            elif name.endswith("scales"):
                buf.copy_(torch.rand(buf.shape, generator=g) * 0.004 + 0.001)
            elif name.endswith("zeros"):
                buf.copy_(torch.randint(0, 16, buf.shape, generator=g).float())
        for name, p_ in m.named_parameters():
            p_.copy_(torch.randn(p_.shape, generator=g) * 0.02 if name.endswith("wte.weight") else torch.ones(p_.shape))
    _REF_CPU["m"] = m
    return m


def cpu_baseline(prefill_sample=None, decode_tokens=2):
    """Tokens/s of the benchmark's workload on the host cores, from a bounded sample.  kind "reference": the unmodified
    reference's LLaMA.forward (baseline/_ref) - a prefill of `prefill_sample` prompt tokens (default: the whole prompt; a
    shorter one is scaled to PROMPT_LEN) and `decode_tokens` single-token steps against its KV cache, timed;
    tokens/s = NEW_TOKENS / (t_prefill + NEW_TOKENS * t_token).  kind "port" (only if baseline/_ref is absent): the
    oracle's restatement of its quantized Linears alone."""
    repeats = 1
    prefill_sample = PROMPT_LEN if prefill_sample is None else min(prefill_sample, PROMPT_LEN)
    import torch

    cores = _host_threads()
    try:
        from baseline import ref_loader
        have_ref = ref_loader.available()
    except Exception:
        have_ref = False
    if not have_ref:
        from oracle import torch_cpu

        d = MODEL_DIMS[MODEL]
        sec, sample = torch_cpu.decode_token_seconds(d["n_embd"], _n_hidden(d["n_embd"]), 32000, d["n_layer"], GROUP, 1)
        return {"value": 1.0 / sec, "unit": "tokens/s", "cores": cores, "kind": "port", "sample": sample + " (decode tokens only)"}
    m = _reference_cpu_model()
    msl = PROMPT_LEN + NEW_TOKENS
    P = prefill_sample
    idx = torch.randint(0, 32000, (1, P + decode_tokens), generator=torch.Generator().manual_seed(0))
    with torch.no_grad():
        m.kv_caches.clear()
        t0 = time.perf_counter()
        m(idx[:, :P], msl, torch.arange(P))
        t1 = time.perf_counter()
        for t in range(P, P + decode_tokens):
            m(idx[:, t:t + 1], msl, torch.tensor([t]))
        t2 = time.perf_counter()
    m.kv_caches.clear()
    t_pre, t_tok = (t1 - t0) * PROMPT_LEN / P, (t2 - t1) / decode_tokens
    sec = t_pre + NEW_TOKENS * t_tok
    scaled = "" if P == PROMPT_LEN else f" (a {P}-token prefill took {t1 - t0:.1f} s, scaled x{PROMPT_LEN / P:g})"
    return {"value": NEW_TOKENS / sec, "unit": "tokens/s", "cores": cores, "kind": "reference",
            "sample": f"unmodified reference LLaMA-{MODEL}.forward (baseline/_ref, fp32, ColBlockQuantizedLinear group {GROUP}: "
                      f"get_weight + F.linear) on the host cores: {PROMPT_LEN}-token prefill {t_pre:.1f} s{scaled} + {decode_tokens} "
                      f"single-token step(s) at {t_tok:.2f} s each; tokens/s = {NEW_TOKENS} / (t_prefill + {NEW_TOKENS} * t_token), "
                      f"i.e. the other {NEW_TOKENS - decode_tokens} tokens are extrapolated from the timed one(s)"}


def run_reference(args, rank, world):
    """--impl reference: the reference's own implementation on the host cores (rank 0 only; nothing of this repo's
    package is imported, no CUDA library of it is loaded)."""
    if rank != 0:
        return
    t0 = time.perf_counter()
    best = cpu_baseline(prefill_sample=None, decode_tokens=2)
    ms = 1000.0 * NEW_TOKENS / best["value"]
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": best["value"], "unit": "tokens/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak" if args.gpus == 1 else "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config_dict(args.gpus),
        "cpu_baseline": best, "gpu_launches": 0,
        "e2e": {"value": best["value"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "extrapolated": True, "wall_s": time.perf_counter() - t0,
    }))


# ----------------------------------------------------------------------------- B200 arm
def algorithmic_bytes(model, n_positions_avg):
    """bytes one decoded token must move - packed weights + scale/zero tables (+ KV)."""
    from lit_llama.quantization import ColBlockQuantizedLinear

    wbytes = 0
    for mod in model.modules():
        if isinstance(mod, ColBlockQuantizedLinear):
            n, k = mod.out_features, mod.in_features
            wbytes += n * k // 2 + 2 * 2 * n * (k // mod.tile_cols) + 2 * (n + k)
    cfg = model.config
    kv = 2 * cfg.n_layer * cfg.n_embd * 2 * n_positions_avg
    return wbytes, kv


def main():
    args = parse()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch
    import torch.distributed as dist

    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)

    from __graft_entry__ import build
    build()
    import generate as G
    from lit_llama import _native as NV

    model = build_model(device)
    replica = None
    if world > 1:
        # (a) what N independent replicas deliver (no collective), measured first on the unsharded model: an extra
        replica = replica_tokens_per_s(model, device, world, dist)
        # (b) the sharded path: every rank holds 1/N of each Linear and of the KV cache
        from lit_llama.tensor_parallel import tensor_parallelize_
        tensor_parallelize_(model, rank, world)
        torch.cuda.empty_cache()
    prompt_host = torch.randint(0, 32000, (PROMPT_LEN,), dtype=torch.int32, generator=torch.Generator().manual_seed(0)).pin_memory()
    prompt_dev = prompt_host.to(device)
    msl = PROMPT_LEN + NEW_TOKENS
    torch.manual_seed(1234)

    def step_resident():
        model.reset_cache()
        return G.generate(model, prompt_dev, NEW_TOKENS, max_seq_length=msl, temperature=0.8, top_k=200)

    def step_e2e():
        model.reset_cache()
        ids = prompt_host.to(device, non_blocking=True)
        return G.generate(model, ids, NEW_TOKENS, max_seq_length=msl, temperature=0.8, top_k=200).cpu()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=device)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), t0, time.perf_counter()

    for _ in range(args.warmup):
        step_resident()
    sampler = ClockSampler(local) if rank == 0 else None
    NV.LAUNCHES = 0
    replays0 = model._engine.n_replays if model._engine is not None else 0
    torch.cuda.profiler.start()        # `ncu --profile-from-start off` sees exactly the timed region
    ms_res, t0, t1 = timed(step_resident, args.steps)
    torch.cuda.profiler.stop()
    launches = NV.LAUNCHES + (model._engine.n_replays - replays0) * model._engine.n_launches
    clocks = sampler.stop(t0, t1) if sampler else None
    for _ in range(min(args.warmup, 1)):
        step_e2e()
    ms_e2e, _, _ = timed(step_e2e, args.steps)

    # tensor parallel: ONE sequence is decoded by all ranks together, so the job's tokens are NEW_TOKENS per step
    value = NEW_TOKENS * args.steps / (ms_res / 1e3)
    e2e_value = NEW_TOKENS * args.steps / (ms_e2e / 1e3)

    out = {
        "metric": METRIC, "value": value, "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_res / args.steps, "higher_is_better": True, "scaling": "weak" if world == 1 else "strong",
        "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic", "config": config_dict(world), "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": "tokens/s", "h2d_bytes_per_step": PROMPT_LEN * 4 * world,
                "d2h_bytes_per_step": (PROMPT_LEN + NEW_TOKENS) * 4 * world},
        "gpu_launches": int(launches) * world,
    }
    if world == 1:
        if rank == 0:
            out.update(kernel_metrics(model, device))
            if os.environ.get("LLB200_BENCH_REFGPU", "1") != "0":
                out["reference_gpu"] = reference_gpu(model, device)
            if os.environ.get("LLB200_BENCH_CPU", "1") != "0":     # (profiling runs skip the host leg)
                out["cpu_baseline"] = cpu_baseline(prefill_sample=64, decode_tokens=1)
    else:
        out["comm"] = {"collective": "tp_ll_allreduce_add_kernel (one NVLink traversal: {word, epoch} units pushed into every peer's "
                                     "receive buffer, summed from local memory; a programmatic-launch link inside the step's CUDA graph)",
                       "exchanges_per_token": 2 * model.config.n_layer, "bytes_per_exchange": model.config.n_embd * 2,
                       "prefill": "NCCL all_reduce (module path)"}
        extra = {"replicas_tokens_per_s": replica}
        extra.update(tp_decode_step_metrics(model, device, world, dist, "7B"))
        del model
        torch.cuda.empty_cache()
        if os.environ.get("LLB200_BENCH_65B", "1") != "0":
            extra.update(tp_65b_probe(device, rank, world, dist))
        out["extra"] = extra
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def replica_tokens_per_s(model, device, world, dist):
    """N independent replicas (no collective): tokens/s summed over the ranks, from one short generate() each."""
    import torch
    import generate as G

    ids = torch.randint(0, 32000, (PROMPT_LEN,), dtype=torch.int32, generator=torch.Generator().manual_seed(0)).to(device)
    msl = PROMPT_LEN + NEW_TOKENS
    for _ in range(2):
        model.reset_cache()
        G.generate(model, ids, NEW_TOKENS, max_seq_length=msl, temperature=0.8, top_k=200)
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        model.reset_cache()
        G.generate(model, ids, NEW_TOKENS, max_seq_length=msl, temperature=0.8, top_k=200)
    e1.record(); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=device)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    model.reset_cache()
    model._engine = None
    return world * 3 * NEW_TOKENS / (float(ms.item()) / 1e3)


def tp_decode_step_metrics(model, device, world, dist, name):
    """Device time of one tensor-parallel decode step (graph replay, position PROMPT_LEN + NEW_TOKENS / 2), max over ranks,
    and the fraction of the HBM rate each rank reaches on ITS share of the bytes."""
    import torch

    eng = model._engine
    if eng is None:
        with torch.no_grad():
            model(torch.zeros((1, 4), dtype=torch.int64, device=device), PROMPT_LEN + NEW_TOKENS, torch.arange(4, device=device))
        eng = model._fused_engine(1, PROMPT_LEN + NEW_TOKENS)
    pos = PROMPT_LEN + NEW_TOKENS // 2
    eng.pos.copy_(torch.tensor([pos, pos % eng.S, min(pos + 1, eng.S)], dtype=torch.int32))
    for _ in range(5):
        eng.step()
    reps = 30
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dist.barrier(); torch.cuda.synchronize(); e0.record()
    for _ in range(reps):
        eng.step()
    e1.record(); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / reps], device=device)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    wbytes, kv = algorithmic_bytes(model, pos)
    peaks_path = ROOT / "MEASURED_PEAKS.json"
    peak = float(json.loads(peaks_path.read_text())["hbm_gbs"]) if peaks_path.exists() else 6650.0
    gbs = (wbytes + kv / world) / (ms * 1e-3) / 1e9
    return {f"tp_decode_step_{name}": {"ms_per_token": ms, "tokens_per_s": 1e3 / ms, "position": pos,
                                       "bytes_per_rank": wbytes + kv // world, "achieved_gbs_per_rank": gbs,
                                       "frac_of_hbm_peak_per_rank": gbs / peak, "kernels_per_token": eng.n_launches}}


def tp_65b_probe(device, rank, world, dist):
    """BASELINE.json configs[4]: LLaMA-65B int4 group 128, batch-1 tensor-parallel decode step on `world` GPUs."""
    import torch
    from lit_llama.tensor_parallel import tensor_parallelize_

    try:
        model = build_model(device, name="65B")
        tensor_parallelize_(model, rank, world)
        torch.cuda.empty_cache()
        out = tp_decode_step_metrics(model, device, world, dist, "65B")
        del model
        torch.cuda.empty_cache()
        return out
    except Exception as e:  # an extra must not take the headline line down
        return {"tp_decode_step_65B": {"error": repr(e)[:300]}}


def reference_gpu(model, device):
    """SURVEY section 0's bar: the reference's own GPU path on this B200, identical weights (the model's state dict loaded
    into the unmodified reference's LLaMA from baseline/_ref): decode tokens/s of its LLaMA.forward with KV cache, batch 1.
    Group 128 takes its get_weight + F.linear branch (ref quantization.py:420-421); its Triton kernel only serves per-row
    scales (measured by profiles/parity_7b.py --per-row; numbers in BASELINE.md)."""
    import torch

    try:
        from baseline import ref_loader
        if not ref_loader.available():
            return {"unavailable": "baseline/_ref not installed"}
        ref = ref_loader.load()
        cfg = model.config
        rcfg = ref.model.LLaMAConfig(block_size=cfg.block_size, vocab_size=cfg.vocab_size, n_layer=cfg.n_layer,
                                     n_head=cfg.n_head, n_embd=cfg.n_embd)
        rm = ref_loader.build_quantized_model(ref, rcfg, tile_cols=GROUP, device=device, dtype=torch.bfloat16)
        rm.load_state_dict(model.state_dict(), strict=True)
        msl = PROMPT_LEN + NEW_TOKENS
        idx = torch.randint(0, 32000, (1, PROMPT_LEN + 12), generator=torch.Generator().manual_seed(0)).to(device)
        with torch.no_grad():
            rm(idx[:, :PROMPT_LEN], msl, torch.arange(PROMPT_LEN, device=device))
            for t in range(PROMPT_LEN, PROMPT_LEN + 4):
                rm(idx[:, t:t + 1], msl, torch.tensor([t], device=device))
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(); e0.record()
            for t in range(PROMPT_LEN + 4, PROMPT_LEN + 12):
                rm(idx[:, t:t + 1], msl, torch.tensor([t], device=device))
            e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 8
        del rm
        torch.cuda.empty_cache()
        return {"fallback_g128_decode_tokens_per_s": 1e3 / ms, "ms_per_token": ms,
                "what": "unmodified reference LLaMA.forward (baseline/_ref), bf16, ColBlockQuantizedLinear group 128 -> "
                        "get_weight + F.linear on CUDA, batch 1, 8 single-token steps after the 512-token prefill, same weights"}
    except Exception as e:
        return {"unavailable": repr(e)[:300]}


def kernel_metrics(model, device):
    """Roofline of the dominant kernel (int4 GEMV) and decode-only token time, both device-timed with
    CUDA events on the launching stream, in CUDA graphs (back-to-back launches as in production)."""
    import torch
    from lit_llama import _native as NV

    eng = model._engine
    peaks_path = ROOT / "MEASURED_PEAKS.json"
    if peaks_path.exists():
        peak, peak_src = float(json.loads(peaks_path.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    wbytes, _ = algorithmic_bytes(model, 0)

    # (a) one token's GEMV launches only (weights 3.5 GB >> L2, so every launch streams from HBM)
    def gemvs():
        n = 0
        for a_attn, kc, vc, a_proj, a_fc, a_down in eng.layers:
            for a in (a_attn, a_proj, a_fc, a_down):
                NV.gemv(a); n += 1
        NV.gemv(eng.a_head)
        return n + 1

    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        n_gemv = gemvs()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        gemvs()
    reps = 20
    for _ in range(3):
        g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(reps):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    gemv_ms = e0.elapsed_time(e1) / reps
    achieved = wbytes / (gemv_ms * 1e-3) / 1e9

    # (b) whole fused decode step (graph replay) at the run's average position
    pos_avg = PROMPT_LEN + NEW_TOKENS // 2
    eng.pos.copy_(torch.tensor([pos_avg, pos_avg % eng.S, min(pos_avg + 1, eng.S)], dtype=torch.int32))
    for _ in range(3):
        eng.step()
    torch.cuda.synchronize(); e0.record()
    for _ in range(reps):
        eng.step()
    e1.record(); torch.cuda.synchronize()
    tok_ms = e0.elapsed_time(e1) / reps
    _, kv = algorithmic_bytes(model, pos_avg)
    # DRAM traffic per launch from the committed `ncu --set full` capture of this kernel (profiles/): the four
    # per-layer launches measured, the lm_head taken at its algorithmic bytes; None if the capture is not there
    traffic = None
    tpath = ROOT / "profiles" / "r02_gemv_traffic.json"
    if not tpath.exists():
        tpath = ROOT / "profiles" / "r01_gemv_traffic.json"
    if tpath.exists() and model.config.n_layer == 32 and model.config.n_embd == 4096:
        t = json.loads(tpath.read_text())
        per_layer = sum(t["dram_read_bytes"].values())
        head_mod = model.lm_head
        head = (head_mod.out_features * head_mod.in_features // 2
                + 4 * head_mod.out_features * (head_mod.in_features // head_mod.tile_cols)
                + 2 * (head_mod.out_features + head_mod.in_features))
        traffic = (32 * per_layer + head) / n_gemv
    return {
        "roofline": {"bound": "hbm", "kernel": "gemv_int4_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": wbytes / n_gemv, "avg_launch_us": 1e3 * gemv_ms / n_gemv,
                     "launches_per_token": n_gemv},
        "decode_step": {"ms_per_token": tok_ms, "tokens_per_s": 1e3 / tok_ms, "position": pos_avg,
                        "algorithmic_bytes": wbytes + kv, "achieved_gbs": (wbytes + kv) / (tok_ms * 1e-3) / 1e9,
                        "frac_of_peak": (wbytes + kv) / (tok_ms * 1e-3) / 1e9 / peak, "kernels_per_token": eng.n_launches,
                        "pdl": eng.use_pdl, "cuda_graph": eng.use_graph},
    }


if __name__ == "__main__":
    main()
