"""Is a GEMV faster when its weights are L2 hits?  32 launches of one layer kind over the 32 blocks (weights from HBM)
against 32 launches of block 0's layer (26-48 MB: L2 hits after the first), same CUDA-graph timing as step_probe.py."""
import os
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "lit-llama-ja_b200")]
os.environ["LLB200_L2_PREFETCH"] = "0"
import bench  # noqa: E402
from lit_llama import _native as NV  # noqa: E402

dev = torch.device("cuda:0")
model = bench.build_model(dev)
S = 768
with torch.no_grad():
    model(torch.zeros((1, 4), dtype=torch.int64, device=dev), S, torch.arange(4, device=dev))
model.kv_caches.clear(); model._engine = None
eng = model._fused_engine(1, S)
eng.pos.copy_(torch.tensor([640, 640, 641], dtype=torch.int32))


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


IDX = {"q": 0, "p": 3, "f": 4, "d": 5}
for kind in "qpfd":
    def hbm():
        for layer in eng.layers:
            NV.gemv(layer[IDX[kind]])

    def l2():
        for _ in eng.layers:
            NV.gemv(eng.layers[0][IDX[kind]])

    def two():   # alternate two blocks' layers: twice the footprint
        for i in range(len(eng.layers)):
            NV.gemv(eng.layers[i & 1][IDX[kind]])

    print(f"{kind}: 32 launches over 32 blocks {timed_graph(hbm):7.1f} us | block 0 only {timed_graph(l2):7.1f} us | blocks 0,1 alternating {timed_graph(two):7.1f} us", flush=True)
