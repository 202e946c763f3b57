"""Tensor-parallel sharding maths on CPU: world-size-2 gloo processes, oracle numerics (no CUDA compute).
Column shards gathered == the full layer; row shards all-reduced == the full layer (fp32 summation order);
the fused-qkv head split maps to the right rows."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent
for p in (str(ROOT), str(ROOT / "lit-llama-ja_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _layer(N, K, tile, seed):
    from oracle import quant as OQ

    r = np.random.RandomState(seed)
    W = (r.randn(N, K) * 0.05).astype(np.float32)
    Q, s, z = OQ.rtn_quantize(W, 4, tile)
    return OQ.pack_weight(Q, s, z, 4, tile), s, z


def _worker(rank, world, port, results):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from lit_llama import tensor_parallel as TP
        from oracle import quant as OQ

        E, nh, H, tile = 128, 4, 512, 32
        x = np.random.RandomState(7).randn(3, E).astype(np.float32)
        # ---- column parallel: fused qkv split by head inside each third
        qw, s, z = _layer(3 * E, E, tile, 1)
        full = OQ.linear_forward(x, qw, s, z, None, 4, tile)
        rows = TP.qkv_rows(E, nh, rank, world).numpy()
        mine = OQ.linear_forward(x, *TP.shard_state_oracle(qw, s, z, rows=rows, tile_cols=tile), None, 4, tile)
        h0, h1 = TP.head_range(nh, rank, world)
        hs = E // nh
        want = np.concatenate([full[:, t * E + h0 * hs:t * E + h1 * hs] for t in range(3)], axis=1)
        assert np.array_equal(mine, want)
        gathered = [torch.zeros_like(torch.from_numpy(mine)) for _ in range(world)]
        dist.all_gather(gathered, torch.from_numpy(mine))
        q_all = torch.cat([g[:, :g.shape[1] // 3] for g in gathered], dim=1).numpy()
        assert np.array_equal(q_all, full[:, :E])
        # ---- row parallel: split K in whole groups, all-reduce the partial sums
        for per_row in (False, True):
            t = -1 if per_row else tile
            qw, s, z = _layer(E, H, t, 2)
            xh = np.random.RandomState(8).randn(3, H).astype(np.float32)
            full = OQ.linear_forward(xh, qw, s, z, None, 4, t)
            a, b = TP.split_even(H, tile, rank, world)
            sq, ss, sz = TP.shard_state_oracle(qw, s, z, k_range=(a, b), tile_cols=t)
            part = OQ.linear_forward(xh[:, a:b], sq, ss, sz, None, 4, -1 if per_row else tile)
            tot = torch.from_numpy(part.copy())
            dist.all_reduce(tot)
            np.testing.assert_allclose(tot.numpy(), full, rtol=1e-5, atol=1e-6)
        # ---- uneven split (65B at TP=8: 172 groups over 8 ranks)
        spans = [TP.split_even(22016, 128, r, 8) for r in range(8)]
        assert spans[0][0] == 0 and spans[-1][1] == 22016 and all(spans[i][1] == spans[i + 1][0] for i in range(7))
        assert sorted({(b - a) // 128 for a, b in spans}) == [21, 22]
        results[rank] = "ok"
    except Exception as e:  # noqa: BLE001
        results[rank] = f"{type(e).__name__}: {e}"
    finally:
        dist.destroy_process_group()


def test_tensor_parallel_sharding_world2_gloo():
    world = 2
    mgr = mp.Manager()
    results = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), results), nprocs=world, join=True)
    assert dict(results) == {0: "ok", 1: "ok"}, dict(results)


def test_shard_modules_roundtrip_cpu():
    """shard_column / shard_row on CPU tensors keep the checkpoint strides and the right bytes."""
    from lit_llama import tensor_parallel as TP
    from lit_llama.quantization import ColBlockQuantizedLinear
    from oracle import quant as OQ

    N_, K, tile = 96, 256, 64
    qw, s, z = _layer(N_, K, tile, 3)
    lin = ColBlockQuantizedLinear(K, N_, False, bits=4, tile_cols=tile)
    lin.quant_weight.copy_(torch.from_numpy(qw)); lin.scales.copy_(torch.from_numpy(s)); lin.zeros.copy_(torch.from_numpy(z))
    rows = torch.arange(16, 48)
    c = TP.shard_column(lin, rows)
    assert c.quant_weight.stride() == (1, 32) and c.out_features == 32
    assert np.array_equal(c.quant_weight.numpy(), qw[16:48]) and np.array_equal(c.scales.numpy(), s[16:48])
    r = TP.shard_row(lin, 64, 192, keep_bias=True)
    assert r.in_features == 128 and r.quant_weight.stride() == (1, N_)
    assert np.array_equal(r.quant_weight.numpy(), qw[:, 32:96]) and np.array_equal(r.zeros.numpy(), z[:, 1:3])
    with pytest.raises(AssertionError):
        TP.shard_row(lin, 32, 192, keep_bias=True)     # not on a group boundary


@pytest.mark.parametrize("H", [11008, 22016, 13824])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_hidden_split_gives_even_k_block_counts(H, world):
    """The row-parallel split of mlp.c_proj (7B / 65B / 13B hidden sizes, group 128 and per-row): contiguous, covers H, whole
    scale groups, and every shard holds an EVEN number of 128-element k-blocks (the batch-1 integer-MMA GEMV's shape rule)."""
    from lit_llama.tensor_parallel import hidden_split_unit, split_even

    for tile_cols in (128, H):
        unit = hidden_split_unit(H, tile_cols, H)
        assert unit == 256 and H % unit == 0
        edges = [split_even(H, unit, r, world) for r in range(world)]
        assert edges[0][0] == 0 and edges[-1][1] == H
        for (a, b), (c, _) in zip(edges, edges[1:] + [(H, H)]):
            assert b == c and (b - a) % 256 == 0 and b > a
        sizes = [b - a for a, b in edges]
        assert max(sizes) - min(sizes) <= 256
    # a hidden size that is no multiple of 256 keeps the plain group split
    assert hidden_split_unit(11136, 128, 11136) == 128
