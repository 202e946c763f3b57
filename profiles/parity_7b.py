"""Config-2 parity at FULL depth (SURVEY 8d "Correctness gates"): LLaMA-7B int4, random-init weights of bench.py, a
512-token prompt, then 256 decode steps.  The UNMODIFIED reference (baseline/_ref) runs greedily on the same B200 from the
same state dict; this repo's fused engine is teacher-forced on the reference's token stream; per step: max |diff| of the
logits, relative Frobenius error, top-1 agreement.  Also times the reference's decode on this GPU (SURVEY section 0's bar).

    python profiles/parity_7b.py [--per-row] [--steps 256] [--out profiles/r02_parity_config2.json]

--per-row: per-row scales (tile_cols = -1, what the reference's writer ships) -> the reference takes its Triton kernel
(autotunes on first use: minutes).
"""
import argparse
import json
import sys
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "lit-llama-ja_b200")]
import bench  # noqa: E402
from baseline import ref_loader  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--per-row", action="store_true")
ap.add_argument("--steps", type=int, default=256)
ap.add_argument("--layers", type=int, default=0)
ap.add_argument("--out", default="")
ap.add_argument("--no-fp32", dest="no_fp32", action="store_true")
a = ap.parse_args()

dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
group = -1 if a.per_row else 128
PROMPT, S = 512, 512 + a.steps
model = bench.build_model(dev, group=group)
if a.layers:
    del model.transformer.h[a.layers:]
    model.config.n_layer = a.layers
with torch.no_grad():
    g = torch.Generator(device=dev).manual_seed(3)
    for n, p in model.named_parameters():
        if n.endswith(".scale"):
            p.copy_(1.0 + 0.1 * torch.randn(p.shape, device=dev, generator=g))
cfg = model.config
ref = ref_loader.load()
rcfg = ref.model.LLaMAConfig(block_size=cfg.block_size, vocab_size=cfg.vocab_size, n_layer=cfg.n_layer, n_head=cfg.n_head, n_embd=cfg.n_embd)
rm = ref_loader.build_quantized_model(ref, rcfg, tile_cols=group, device=dev, dtype=torch.bfloat16)
rm.load_state_dict(model.state_dict(), strict=True)
path = "Triton kernel (quantization.py:185-331)" if a.per_row else "get_weight + F.linear (quantization.py:390-421)"

prompt = torch.randint(0, 32000, (1, PROMPT), generator=torch.Generator().manual_seed(0)).to(dev)
ref_logits, toks = [], []
with torch.no_grad():
    t0 = time.perf_counter()
    lg = rm(prompt, S, torch.arange(PROMPT, device=dev))[:, -1]
    torch.cuda.synchronize()
    t_prefill_ref = time.perf_counter() - t0
    ref_logits.append(lg.float().cpu().numpy())
    tok = lg.argmax(-1, keepdim=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n_timed = 0
    for t in range(PROMPT, PROMPT + a.steps - 1):
        toks.append(tok)
        if t == PROMPT + 8:
            e0.record()
        lg = rm(tok, S, torch.tensor([t], device=dev))[:, -1]
        if t >= PROMPT + 8:
            n_timed += 1
        ref_logits.append(lg.float().cpu().numpy())
        tok = lg.argmax(-1, keepdim=True)
    e1.record(); torch.cuda.synchronize()
ref_ms = e0.elapsed_time(e1) / max(1, n_timed)
ref_logits = np.concatenate(ref_logits)
del rm
torch.cuda.empty_cache()

# the same reference modules in fp32 on the same tokens: how far the reference's OWN bf16 path is from its fp32 result
truth = None
if not a.no_fp32:
    rm32 = ref_loader.build_quantized_model(ref, rcfg, tile_cols=group, device=dev, dtype=torch.float32)
    rm32.load_state_dict(model.state_dict(), strict=True)
    truth = []
    with torch.no_grad():
        truth.append(rm32(prompt, S, torch.arange(PROMPT, device=dev))[:, -1].float().cpu().numpy())
        for i, t in enumerate(range(PROMPT, PROMPT + a.steps - 1)):
            truth.append(rm32(toks[i], S, torch.tensor([t], device=dev))[:, -1].float().cpu().numpy())
    truth = np.concatenate(truth)
    del rm32
    torch.cuda.empty_cache()

ours = []
with torch.no_grad():
    model.reset_cache()
    ours.append(model(prompt, S, torch.arange(PROMPT, device=dev))[:, -1].float().cpu().numpy())
    for i, t in enumerate(range(PROMPT, PROMPT + a.steps - 1)):
        ours.append(model(toks[i], S, torch.tensor([t], device=dev))[:, -1].float().cpu().numpy())
assert model._engine is not None and model._engine.n_replays > 0
ours = np.concatenate(ours)
eng = model._engine
for _ in range(3):
    eng.step()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); e0.record()
for _ in range(20):
    eng.step()
e1.record(); torch.cuda.synchronize()
our_ms = e0.elapsed_time(e1) / 20

d = np.abs(ours - ref_logits)
scale = np.abs(ref_logits).max()
fro = np.sqrt((d.astype(np.float64) ** 2).sum(-1)) / np.sqrt((ref_logits.astype(np.float64) ** 2).sum(-1))
top1 = ours.argmax(-1) == ref_logits.argmax(-1)
# how decisive the reference's own top-1 was where the two disagree: its gap to the runner-up, in units of max |diff|
srt = np.sort(ref_logits, -1)
gap = srt[:, -1] - srt[:, -2]
res = {
    "config": f"LLaMA-7B x {cfg.n_layer} layers, int4 {'per-row' if a.per_row else 'group 128'}, bf16, prompt {PROMPT}, {a.steps} steps "
              f"teacher-forced on the reference's greedy stream",
    "reference_path": path,
    "max_abs_diff": float(d.max()), "max_logit": float(scale), "max_rel": float(d.max() / scale),
    "rel_frobenius_max": float(fro.max()), "rel_frobenius_mean": float(fro.mean()),
    "top1_agreement": float(top1.mean()), "steps": int(len(top1)),
    "top1_disagreements_gap_over_maxdiff": [float(x) for x in (gap[~top1] / max(d.max(), 1e-30))[:16]],
    "reference_gpu_decode_ms_per_token": ref_ms, "reference_gpu_decode_tokens_per_s": 1e3 / ref_ms,
    "reference_gpu_prefill_s": t_prefill_ref,
    "ours_decode_step_ms": our_ms, "ours_decode_tokens_per_s": 1e3 / our_ms,
}
if truth is not None:
    def dist(x):
        dd = np.abs(x - truth)
        fr = np.sqrt((dd.astype(np.float64) ** 2).sum(-1)) / np.sqrt((truth.astype(np.float64) ** 2).sum(-1))
        return {"max_rel": float(dd.max() / np.abs(truth).max()), "rel_frobenius_mean": float(fr.mean()),
                "rel_frobenius_max": float(fr.max()), "top1_agreement": float((x.argmax(-1) == truth.argmax(-1)).mean())}
    res["vs_reference_fp32_on_the_same_tokens"] = {"reference_bf16": dist(ref_logits), "ours_bf16": dist(ours)}
print(json.dumps(res, indent=1))
if a.out:
    Path(a.out).write_text(json.dumps(res, indent=1))
