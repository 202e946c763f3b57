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
``--impl reference`` times that CPU path alone (the reference has no other implementation that
can run here: its GPU path is Triton/bitsandbytes - see DESIGN.md).
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
        "parallelism": "single GPU" if n_gpus == 1 else f"{n_gpus} independent replicas (one sequence per GPU)",
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
def build_model(device, name=MODEL, group=GROUP, seed=0):
    """Random-init LLaMA of the named size, RTN-quantised per group on the device."""
    import torch
    from lit_llama import LLaMA
    from lit_llama.quantization import ColBlockQuantizedLinear, rtn_quantize_
    from lit_llama.utils import EmptyInitOnDevice, quantization

    with EmptyInitOnDevice(device=device, dtype=torch.bfloat16), quantization("gptq.int4", groupsize=group):
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


# ----------------------------------------------------------------------------- CPU baseline
def cpu_baseline(repeats=2):
    import torch
    from lit_llama.model import LLaMAConfig
    from oracle import torch_cpu

    cfg = LLaMAConfig.from_name(MODEL)
    sec, sample = torch_cpu.decode_token_seconds(cfg.n_embd, cfg.n_hidden, cfg.padded_vocab_size, cfg.n_layer, GROUP, repeats)
    return {"value": 1.0 / sec, "unit": "tokens/s", "cores": torch.get_num_threads(), "kind": "port", "sample": sample}


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU path (oracle port; /root/reference is a Python tree that
    cannot travel and whose GPU path needs Triton/bitsandbytes).  Rank 0 only."""
    if rank != 0:
        return
    t0 = time.perf_counter()
    for _ in range(max(0, min(args.warmup, 1))):
        cpu_baseline(1)
    vals = [cpu_baseline(1) for _ in range(max(1, min(args.steps, 3)))]
    best = max(vals, key=lambda v: v["value"])
    ms = 1000.0 * NEW_TOKENS / best["value"]
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": best["value"], "unit": "tokens/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config_dict(args.gpus),
        "cpu_baseline": best, "gpu_launches": 0,
        "e2e": {"value": best["value"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "wall_s": time.perf_counter() - t0,
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

    value = world * NEW_TOKENS * args.steps / (ms_res / 1e3)
    e2e_value = world * NEW_TOKENS * args.steps / (ms_e2e / 1e3)

    out = {
        "metric": METRIC, "value": value, "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_res / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic", "config": config_dict(world), "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": "tokens/s", "h2d_bytes_per_step": PROMPT_LEN * 4,
                "d2h_bytes_per_step": (PROMPT_LEN + NEW_TOKENS) * 4},
        "gpu_launches": int(launches),
    }
    if rank == 0:
        out.update(kernel_metrics(model, device))
        if world == 1:
            out["cpu_baseline"] = cpu_baseline(2)
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


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
