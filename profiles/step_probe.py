"""Decode-step probe: device time of one fused decode step of the 7B int4 model (CUDA graph), with
PDL on/off, and of its parts (GEMVs only, attention only) - to see where a token's time goes."""
import os
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "lit-llama-ja_b200")]
import bench  # noqa: E402
from lit_llama import _native as NV  # noqa: E402
from lit_llama.engine import DecodeEngine  # noqa: E402

dev = torch.device("cuda:0")
model = bench.build_model(dev)
S = 768
with torch.no_grad():
    model(torch.zeros((1, 4), dtype=torch.int64, device=dev), S, torch.arange(4, device=dev))   # builds rope / kv caches
POS = int(os.environ.get("PROBE_POS", 640))


def timed_graph(fn, reps=20):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    for _ in range(3):
        g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(reps):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    return 1e3 * e0.elapsed_time(e1) / reps


for pdl in (True,):
    os.environ["LLB200_PDL"] = "1" if pdl else "0"
    if "PROBE_APDL" in os.environ:          # otherwise the engine's own default (on with the EVEN/ODD footprints)
        os.environ["LLB200_PDL_ATTN"] = os.environ["PROBE_APDL"]
    model.kv_caches.clear(); model._engine = None
    eng = model._fused_engine(1, S)
    apdl = eng.attn_pdl
    eng.pos.copy_(torch.tensor([POS, POS % S, min(POS + 1, S)], dtype=torch.int32))
    flags = NV.GEMV_PDL if (pdl and apdl) else 0
    cfg = model.config

    def attn_only():
        for a_attn, kc, vc, a_proj, a_fc, a_down in eng.layers:
            NV.call("ll_decode_attn", NV.ptr(eng.q), NV.ptr(kc), NV.ptr(vc), NV.ptr(eng.pos), NV.ptr(eng.att), NV.LL_BF16,
                    1, cfg.n_head, cfg.head_size, S, eng.splits, NV.ptr(eng.attn_scratch), NV.ptr(eng.attn_counters), flags, NV.stream())

    def gemvs(which):
        def f():
            for layer in eng.layers:
                for i in which:
                    NV.gemv(layer[i])
        return f

    def seq(spec):
        def f():
            for layer in eng.layers:
                for c in spec:
                    if c == "A":
                        NV.call("ll_decode_attn", NV.ptr(eng.q), NV.ptr(layer[1]), NV.ptr(layer[2]), NV.ptr(eng.pos), NV.ptr(eng.att), NV.LL_BF16,
                                1, cfg.n_head, cfg.head_size, S, eng.splits, NV.ptr(eng.attn_scratch), NV.ptr(eng.attn_counters), flags, NV.stream())
                    else:
                        NV.gemv(layer[{"q": 0, "p": 3, "f": 4, "d": 5}[c]])
        return f

    parts = {sp: timed_graph(seq(sp)) for sp in os.environ.get("PROBE_SEQS", "q,A,p,f,d,qpfd,qApfd").split(",")}
    import subprocess
    clk = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm,power.draw", "--format=csv,noheader", "-i", "0"],
                         capture_output=True, text=True).stdout.strip()
    print(f"[{clk}] cfg={os.environ.get('LLB200_GEMV_CFG', 'default')} PDL={pdl} parity={eng.parity} attnPDL={apdl}: full {timed_graph(eng._run_step):7.1f} | " + " ".join(f"{k}={v:.0f}" for k, v in parts.items()), flush=True)
