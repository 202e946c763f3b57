"""Tensor-parallel quantized LLaMA over one NVSwitch node (BASELINE.json config 5: LLaMA-65B int4 at 2/4/8 GPUs).

The reference is single-device (generate.py:122 ``devices=1``): sharding is new here and is validated
against the unsharded result on identical weights.  Megatron-style, one exchange per row-parallel layer:

* column-parallel (split the output rows N of ``quant_weight`` / ``scales`` / ``zeros``): ``attn.c_attn`` (by HEAD
  inside each of the q / k / v thirds - rows are ordered [q; k; v], ref model.py:204), ``mlp.c_fc1``, ``mlp.c_fc2``;
* row-parallel (split the input features K in whole scale groups; per-row scales are simply reused):
  ``attn.c_proj``, ``mlp.c_proj`` - followed by ONE all-reduce (sum) of the ``[B, T, n_embd]`` partial result
  (NCCL over NVLink / NVSwitch; 2 per block, 160 per token at 65B, 16 KB each at batch 1);
* KV cache and attention are sharded by head; embedding, RMSNorm, residual stream and ``lm_head`` are replicated
  (``lm_head`` column-parallel would need a gather of the logits for sampling; at batch 1 its 1 % of the bytes
  is not worth a second collective).

The pure index helpers below are exercised on CPU with world-size-2 ``gloo`` (tests/test_tp_gloo.py) against the
oracle; the CUDA path is exercised with ``gpurun --gpus 2`` (tests/test_gpu_tp.py, launched by torchrun).
"""
from __future__ import annotations

from typing import List, Tuple

import torch
import torch.distributed as dist

from .quantization import ColBlockQuantizedLinear


# ----------------------------------------------------------------------------- index helpers (pure)
def head_range(n_head: int, rank: int, world: int) -> Tuple[int, int]:
    """Heads [h0, h1) of ``rank`` - as even as possible, earlier ranks take the remainder."""
    base, rem = divmod(n_head, world)
    h0 = rank * base + min(rank, rem)
    return h0, h0 + base + (1 if rank < rem else 0)


def qkv_rows(n_embd: int, n_head: int, rank: int, world: int) -> torch.Tensor:
    """Row indices of the fused ``c_attn`` weight ([q; k; v], each ``n_embd`` rows, head-major) held by ``rank``."""
    hs = n_embd // n_head
    h0, h1 = head_range(n_head, rank, world)
    one = torch.arange(h0 * hs, h1 * hs)
    return torch.cat([one, one + n_embd, one + 2 * n_embd])


def split_even(n: int, unit: int, rank: int, world: int) -> Tuple[int, int]:
    """[a, b) of ``n`` split over ``world`` ranks in whole ``unit``s (n % unit == 0); uneven when units % world != 0
    (65B at TP=8: n_hidden = 22016 = 172 groups of 128 -> 22/21 groups per rank)."""
    assert n % unit == 0
    units = n // unit
    base, rem = divmod(units, world)
    a = rank * base + min(rank, rem)
    return a * unit, (a + base + (1 if rank < rem else 0)) * unit


def hidden_split_unit(H: int, tile_cols: int, in_features: int) -> int:
    """Granularity of the row-parallel split of ``mlp.c_proj`` (and of the matching column split of c_fc1 / c_fc2): whole
    scale groups (128 for per-row scales), doubled to 256 where H allows it - an EVEN number of 128-element k-blocks per
    shard is what the batch-1 integer-MMA GEMV takes (csrc/gemv_int4.cu i8_shape_ok); an odd count falls back to the
    slower bf16 kernel (7B: 86 blocks -> 44 / 42 at two ranks instead of 43 / 43; 65B at four ranks: 44, 44, 42, 42)."""
    unit = tile_cols if tile_cols != in_features else 128 if H % 128 == 0 else 2
    if (2 * unit) % 256 == 0 and H % (2 * unit) == 0 and unit % 256 != 0:
        unit *= 2
    return unit


# ----------------------------------------------------------------------------- sharding one layer
def _new_like(lin: ColBlockQuantizedLinear, in_features: int, out_features: int) -> ColBlockQuantizedLinear:
    dt = lin.scales.dtype
    prev = torch.get_default_dtype()
    torch.set_default_dtype(dt if dt.is_floating_point else prev)
    try:
        tile = -1 if lin.tile_cols == lin.in_features else lin.tile_cols
        new = ColBlockQuantizedLinear(in_features, out_features, lin.bias is not None, bits=lin.bits, tile_cols=tile)
    finally:
        torch.set_default_dtype(prev)
    return new.to(lin.quant_weight.device)


def shard_column(lin: ColBlockQuantizedLinear, rows: torch.Tensor) -> ColBlockQuantizedLinear:
    """Keep output rows ``rows`` (logical dim 0 of quant_weight / scales / zeros / bias)."""
    rows = rows.to(lin.quant_weight.device)
    new = _new_like(lin, lin.in_features, int(rows.numel()))
    new.quant_weight = lin.quant_weight.index_select(0, rows).t().contiguous().t()   # keep the (1, N) strides
    new.scales = lin.scales.index_select(0, rows).contiguous()
    new.zeros = lin.zeros.index_select(0, rows).contiguous()
    if lin.bias is not None:
        new.bias = lin.bias.index_select(0, rows).contiguous()
    return new


def shard_row(lin: ColBlockQuantizedLinear, k0: int, k1: int, keep_bias: bool) -> ColBlockQuantizedLinear:
    """Keep input features [k0, k1) (whole scale groups, or any multiple of the entries per byte for per-row scales)."""
    epb = lin.entries_per_byte
    assert k0 % epb == 0 and k1 % epb == 0
    per_row = lin.tile_cols == lin.in_features
    if not per_row:
        assert k0 % lin.tile_cols == 0 and k1 % lin.tile_cols == 0, "row-parallel split must fall on scale-group boundaries"
    new = _new_like(lin, k1 - k0, lin.out_features)
    new.quant_weight = lin.quant_weight[:, k0 // epb:k1 // epb].t().contiguous().t()
    if per_row:
        new.scales, new.zeros = lin.scales.clone(), lin.zeros.clone()
    else:
        g0, g1 = k0 // lin.tile_cols, k1 // lin.tile_cols
        new.scales, new.zeros = lin.scales[:, g0:g1].contiguous(), lin.zeros[:, g0:g1].contiguous()
    if lin.bias is not None:
        new.bias = lin.bias.clone() if keep_bias else torch.zeros_like(lin.bias)   # the bias is added once (rank 0)
    return new


class RowParallelLinear(torch.nn.Module):
    """A row-parallel shard + the all-reduce that completes it (the path's only exchange step)."""

    def __init__(self, inner: torch.nn.Module, group=None):
        super().__init__()
        self.inner = inner
        self.group = group

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        y = self.inner(x)
        dist.all_reduce(y, op=dist.ReduceOp.SUM, group=self.group)
        return y


# ----------------------------------------------------------------------------- sharding the model
def tensor_parallelize_(model, rank: int, world: int, group=None):
    """Shard ``model`` (a quantized ``LLaMA`` holding the FULL weights on this rank's device) in place."""
    if world == 1:
        return model
    cfg = model.config
    E, H, nh = cfg.n_embd, cfg.n_hidden, cfg.n_head
    assert nh >= world, "more ranks than attention heads"
    hs = E // nh
    h0, h1 = head_range(nh, rank, world)
    for blk in model.transformer.h:
        attn, mlp = blk.attn, blk.mlp
        for lin in (attn.c_attn, attn.c_proj, mlp.c_fc1, mlp.c_fc2, mlp.c_proj):
            assert isinstance(lin, ColBlockQuantizedLinear), "tensor parallelism is implemented for the GPTQ modules"
        attn.c_attn = shard_column(attn.c_attn, qkv_rows(E, nh, rank, world))
        attn.c_proj = RowParallelLinear(shard_row(attn.c_proj, h0 * hs, h1 * hs, keep_bias=rank == 0), group)
        attn.n_head = h1 - h0
        a, b = split_even(H, hidden_split_unit(H, mlp.c_proj.tile_cols, mlp.c_proj.in_features), rank, world)
        rows = torch.arange(a, b)
        mlp.c_fc1 = shard_column(mlp.c_fc1, rows)
        mlp.c_fc2 = shard_column(mlp.c_fc2, rows)
        mlp.c_proj = RowParallelLinear(shard_row(mlp.c_proj, a, b, keep_bias=rank == 0), group)
    model._tp = (rank, world, group)
    model.kv_caches.clear()
    model._engine = None
    return model


def shard_state_oracle(qw, scales, zeros, rows=None, k_range=None, bits: int = 4, tile_cols: int = -1):
    """numpy twin of shard_column / shard_row for the CPU (gloo) tests: arrays in, arrays out."""
    epb = 8 // bits
    if rows is not None:
        return qw[rows], scales[rows], zeros[rows]
    k0, k1 = k_range
    K = qw.shape[1] * epb
    if tile_cols in (-1, K):
        return qw[:, k0 // epb:k1 // epb], scales, zeros
    return qw[:, k0 // epb:k1 // epb], scales[:, k0 // tile_cols:k1 // tile_cols], zeros[:, k0 // tile_cols:k1 // tile_cols]
