"""Multi-GPU tensor-parallel check (torchrun, see run_tp.sh): prints progress lines so a stall can be located.
  python -m torch.distributed.run ... profiles/tp_check.py [push|peer|nccl]
1. the peer-memory all-reduce kernel alone against a host sum;  2. module path and fused engine on the tiny golden
model against the single-GPU result;  3. logits identical on every rank and from run to run."""
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "lit-llama-ja_b200")]
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
mode = sys.argv[1] if len(sys.argv) > 1 else "peer"
os.environ["LLB200_TP_EXCHANGE"] = mode


def log(*a):
    print(f"[rank {rank}]", *a, flush=True)


import tests.test_gpu_model as TM  # noqa: E402
from lit_llama.tensor_parallel import tensor_parallelize_  # noqa: E402

TM.DEV = dev = f"cuda:{local}"

if mode == "peer":
    import ctypes as C
    import torch.distributed._symmetric_memory as symm
    from lit_llama import _native as N

    n = 8192
    buf = symm.empty((2, n), dtype=torch.bfloat16, device=dev)
    flg = symm.empty((64,), dtype=torch.int32, device=dev)
    buf.zero_(), flg.zero_()
    hb, hf = symm.rendezvous(buf, dist.group.WORLD), symm.rendezvous(flg, dist.group.WORLD)
    torch.cuda.synchronize()
    dist.barrier()
    log("symmetric memory ready")
    epoch = torch.zeros(8, dtype=torch.int32, device=dev)
    FL = 1 if os.environ.get("TP_CHECK_PDL") == "1" else 0    # the exchange kernel as a programmatic-launch link (4 CTAs at n = 8192)
    flags = (C.c_uint64 * world)(*[int(p) for p in hf.buffer_ptrs])
    parts = [(C.c_uint64 * world)(*[int(p) + j * n * 2 for p in hb.buffer_ptrs]) for j in range(2)]
    gen = torch.Generator().manual_seed(5)
    x0 = torch.randn(n, generator=gen).bfloat16()
    allp = torch.randn(6, world, n, generator=gen).bfloat16()          # same on every rank
    x = x0.to(dev)
    want = x0.clone()
    for it in range(6):
        buf[it & 1].copy_(allp[it, rank])
        N.call("ll_tp_allreduce_add", N.ptr(x), C.cast(parts[it & 1], C.c_void_p), C.cast(flags, C.c_void_p), N.ptr(epoch),
               rank, world, n, FL, N.stream())
        s = torch.zeros(n)
        for r in range(world):
            s += allp[it, r].float()
        want = (want.float() + s.bfloat16().float()).bfloat16()
    torch.cuda.synchronize()
    log("all-reduce kernel bit-exact vs host sum:", bool(torch.equal(x.cpu(), want)), "epoch", int(epoch[0].item()))
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    dist.barrier()
    ev[0].record()
    for it in range(200):
        N.call("ll_tp_allreduce_add", N.ptr(x), C.cast(parts[it & 1], C.c_void_p), C.cast(flags, C.c_void_p), N.ptr(epoch),
               rank, world, n, FL, N.stream())
    ev[1].record()
    torch.cuda.synchronize()
    log(f"all-reduce+add of {n} bf16: {ev[0].elapsed_time(ev[1]) / 200 * 1e3:.2f} us per call (back-to-back launches)")

if mode in ("peer", "ll"):
    # the one-traversal exchange (tp_ll_allreduce_add_kernel) alone: bit-exact against the host's rank-order sum, then timed
    import ctypes as C
    import torch.distributed._symmetric_memory as symm
    from lit_llama import _native as N

    for n in (4096, 8192):
        nbytes = N.call_ret("ll_tp_ll_recv_bytes", world, n)
        recv = symm.empty((nbytes,), dtype=torch.uint8, device=dev)
        recv.zero_()
        hr = symm.rendezvous(recv, dist.group.WORLD)
        torch.cuda.synchronize()
        dist.barrier()
        peers = (C.c_uint64 * world)(*[int(p) for p in hr.buffer_ptrs])
        ep = torch.zeros(8, dtype=torch.int32, device=dev)
        FL2 = 1 if os.environ.get("TP_CHECK_PDL") == "1" else 0
        gen = torch.Generator().manual_seed(7 + n)
        x0 = torch.randn(n, generator=gen).bfloat16()
        allp = torch.randn(7, world, n, generator=gen).bfloat16()
        x, want = x0.to(dev), x0.clone()
        part = torch.zeros(n, dtype=torch.bfloat16, device=dev)
        for it in range(7):
            part.copy_(allp[it, rank])
            N.call("ll_tp_ll_allreduce_add", N.ptr(x), N.ptr(part), C.cast(peers, C.c_void_p), N.ptr(ep), rank, world, n, FL2, N.stream())
            ssum = torch.zeros(n)
            for r in range(world):
                ssum += allp[it, r].float()
            want = (want.float() + ssum.bfloat16().float()).bfloat16()
        torch.cuda.synchronize()
        log(f"LL exchange n={n}: bit-exact vs host sum:", bool(torch.equal(x.cpu(), want)))
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        dist.barrier()
        ev[0].record()
        for it in range(200):
            N.call("ll_tp_ll_allreduce_add", N.ptr(x), N.ptr(part), C.cast(peers, C.c_void_p), N.ptr(ep), rank, world, n, FL2, N.stream())
        ev[1].record()
        torch.cuda.synchronize()
        log(f"LL exchange of {n} bf16: {ev[0].elapsed_time(ev[1]) / 200 * 1e3:.2f} us per call (back-to-back launches)")

g = np.load(ROOT / "tests" / "golden" / "model_tiny.npz")
idx = torch.from_numpy(g["idx"][:1]).to(dev)
ref_model = TM.build_q_model(g, "rtn32bf16.", torch.bfloat16)
ref = TM._decode_run(ref_model, idx, 16, 10, fused=True)
log("reference done")
m = TM.build_q_model(g, "rtn32bf16.", torch.bfloat16)
tensor_parallelize_(m, rank, world)
a = TM._decode_run(m, idx, 16, 10, fused=False)
log("module path err", float(np.abs(a - ref).max() / np.abs(ref).max()))
b = TM._decode_run(m, idx, 16, 10, fused=True)
e = m._engine
log(f"engine err {float(np.abs(b - ref).max() / np.abs(ref).max()):.5f} tp {e.tp} exchange {e.exchange} graph {e.graph is not None}")
b2 = TM._decode_run(m, idx, 16, 10, fused=True)
t = torch.from_numpy(np.ascontiguousarray(b)).to(dev)
gath = [torch.empty_like(t) for _ in range(world)]
dist.all_gather(gath, t)
log("run-to-run identical", bool(np.array_equal(b, b2)), "identical across ranks", all(bool(torch.equal(gath[0], x_)) for x_ in gath))
dist.barrier()
dist.destroy_process_group()
log("done")
