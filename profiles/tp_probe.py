"""Tensor-parallel decode probe (BASELINE.json config 5): LLaMA-<MODEL> int4 g128, batch-1 decode step time with
TP = WORLD_SIZE ranks (NCCL all-reduce after the row-parallel GEMVs, all in one CUDA graph).
Launch: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 profiles/tp_probe.py [MODEL]"""
import json
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "lit-llama-ja_b200")]
import bench  # noqa: E402
from lit_llama.tensor_parallel import tensor_parallelize_  # noqa: E402

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
name = sys.argv[1] if len(sys.argv) > 1 else "7B"
S, POS = 768, 640
model = bench.build_model(dev, name=name)
wbytes_full, _ = bench.algorithmic_bytes(model, 0)
if world > 1:
    tensor_parallelize_(model, rank, world)
    torch.cuda.empty_cache()
with torch.no_grad():
    model(torch.zeros((1, 4), dtype=torch.int64, device=dev), S, torch.arange(4, device=dev))
eng = model._fused_engine(1, S)
assert eng is not None
eng.pos.copy_(torch.tensor([POS, POS % S, min(POS + 1, S)], dtype=torch.int32))
for _ in range(5):
    eng.step()
reps = 30
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
if world > 1:
    dist.barrier()
torch.cuda.synchronize(); e0.record()
for _ in range(reps):
    eng.step()
e1.record(); torch.cuda.synchronize()
ms = torch.tensor([e0.elapsed_time(e1) / reps], device=dev)
if world > 1:
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
wbytes, kv = bench.algorithmic_bytes(model, POS)
if rank == 0:
    print(json.dumps({"model": name, "tp": world, "ms_per_token": float(ms), "tokens_per_s": 1e3 / float(ms),
                      "weight_bytes_per_rank": wbytes, "weight_bytes_full": wbytes_full,
                      "achieved_gbs_per_rank": (wbytes + kv / world) / (float(ms) * 1e-3) / 1e9, "kernels_per_token": eng.n_launches}))
if world > 1:
    dist.barrier(); dist.destroy_process_group()
