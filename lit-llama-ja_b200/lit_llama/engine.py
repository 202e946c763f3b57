"""Fused single-token decode step for an int4 (GPTQ) / bf16 LLaMA, captured in a CUDA graph.

One decode step = 1 embedding kernel + 5 kernels per block + 1 lm_head kernel, all launched with
programmatic dependent launch so each GEMV's TMA weight stream starts while its predecessor drains:

    embed                                        wte row
    per block:
      gemv c_attn   [RMSNorm prologue | RoPE + KV-cache write epilogue]      (ref model.py:204-228)
      decode_attn   split-KV, deterministic combine                          (ref model.py:237)
      gemv c_proj   [residual-add epilogue]                                  (ref model.py:172, 242)
      gemv c_fc1|c_fc2 [RMSNorm prologue | SwiGLU epilogue]                  (ref model.py:258)
      gemv mlp.c_proj  [residual-add epilogue]                               (ref model.py:173, 259)
    gemv lm_head    [RMSNorm (ln_f) prologue]                                (ref model.py:125-127)

Positions live in a device int32[3] (pos, ring slot, #valid slots) that ``ll_set_pos`` /
``ll_advance_pos`` update, so the captured graph is replayed unchanged for every token.
"""
from __future__ import annotations

import os

import torch

import torch.distributed as dist

from . import _native as N
from .quantization import ColBlockQuantizedLinear


def _unwrap(lin):
    """(quantized Linear, process group or None) - row-parallel shards come wrapped (tensor_parallel.py)."""
    inner = getattr(lin, "inner", None)
    return (inner, getattr(lin, "group", None)) if inner is not None else (lin, None)


def _env_flag(name: str, default: bool) -> bool:
    v = os.environ.get(name)
    return default if v is None else v not in ("0", "false", "False", "")


class DecodeEngine:
    MAX_BATCH = 8

    @staticmethod
    def eligible(model, B: int) -> bool:
        cfg = model.config
        if B > DecodeEngine.MAX_BATCH or cfg.n_embd % 16 != 0 or cfg.head_size not in (32, 64, 128):
            return False
        wte = model.transformer.wte.weight
        if wte.device.type != "cuda" or wte.dtype != torch.bfloat16:
            return False
        lins = [model.lm_head]
        for blk in model.transformer.h:
            lins += [blk.attn.c_attn, _unwrap(blk.attn.c_proj)[0], blk.mlp.c_fc1, blk.mlp.c_fc2, _unwrap(blk.mlp.c_proj)[0]]
            if blk.rms_1.scale.dtype != torch.bfloat16:
                return False
        for lin in lins:
            if not isinstance(lin, ColBlockQuantizedLinear) or lin.bias is not None:
                return False
            if lin.scales.dtype != torch.bfloat16 or lin.prepare(B) is None:
                return False
        for blk in model.transformer.h:
            if blk.mlp.c_fc1.tile_cols != blk.mlp.c_fc2.tile_cols:
                return False
        h0 = model.transformer.h[0]
        shapes = [(h0.attn.c_attn, 1), (_unwrap(h0.attn.c_proj)[0], 1), (h0.mlp.c_fc1, 2), (_unwrap(h0.mlp.c_proj)[0], 1),
                  (model.lm_head, 1)]
        return all(l.gemv_ok(B, nmat) for l, nmat in shapes)

    def __init__(self, model, B: int, max_seq_length: int):
        cfg = model.config
        self.model, self.B, self.S = model, B, max_seq_length
        dev = model.transformer.wte.weight.device
        h0 = model.transformer.h[0]
        self.tp_group = _unwrap(h0.attn.c_proj)[1] if getattr(model, "_tp", None) is not None else None
        self.tp = getattr(model, "_tp", None) is not None
        self.nh = h0.attn.n_head                      # heads on this rank
        E, V = cfg.n_embd, cfg.padded_vocab_size
        E_loc, H = self.nh * cfg.head_size, h0.mlp.c_fc1.out_features
        bf, f32, i32 = torch.bfloat16, torch.float32, torch.int32
        self.tok = torch.zeros(B, dtype=torch.int64, device=dev)
        self.pos = torch.zeros(3, dtype=i32, device=dev)
        self.x = torch.zeros((B, E), dtype=bf, device=dev)        # residual stream
        self.q = torch.zeros((B, E_loc), dtype=bf, device=dev)
        self.att = torch.zeros((B, E_loc), dtype=bf, device=dev)
        self.part = torch.zeros((B, E), dtype=bf, device=dev)     # row-parallel partial result (tensor parallel only)
        self.h = torch.zeros((B, H), dtype=bf, device=dev)
        self.logits = torch.zeros((B, V), dtype=bf, device=dev)
        self.rope = model._rope_f32
        self.use_pdl = _env_flag("LLB200_PDL", True)
        self.use_graph = _env_flag("LLB200_GRAPH", True)
        # measured (profiles/r01_step_probe.md): launching the attention kernel itself with the programmatic
        # attribute costs ~8 us per block (its 512 waiting CTAs crowd the SMs the GEMV pair shares)
        # LLB200_GEMV_PARITY=1: consecutive GEMVs alternate EVEN / ODD footprints (include/llb200.h) so that two CTAs of the
        # same GEMV can never share an SM; the attention kernel can then be a programmatic dependent too
        # (measured, profiles/r02_gemv_experiments.md section 6: every SM then holds exactly one CTA of each GEMV and the
        # 7B step drops 1360 -> 1339 us).  Default: on where every Linear of the chain takes the integer-MMA kernel (the
        # only one with the two footprints) on one GPU, and under tensor parallelism when the exchange kernel joins
        # the programmatic chain (below); with a plainly launched exchange kernel the GEMVs are separated anyway.
        lins = [model.lm_head]
        for blk in model.transformer.h:
            lins += [blk.attn.c_attn, _unwrap(blk.attn.c_proj)[0], blk.mlp.c_fc1, blk.mlp.c_fc2, _unwrap(blk.mlp.c_proj)[0]]
        all_u8 = all(lin.stream_for(B)[1] in (N.STREAM_U8, N.STREAM_W8) for lin in lins)   # the integer-MMA kernel (4- or 8-bit weights)
        # tensor parallel: the exchange kernel is itself a link of the programmatic chain (LLB200_TP_PDL, default on with the
        # peer exchange) - the footprints are then needed for the same placement reason
        self.tp_pdl = (self.tp and self.use_pdl and all_u8 and os.environ.get("LLB200_TP_EXCHANGE", "ll") in ("peer", "ll")
                       and _env_flag("LLB200_TP_PDL", True))
        self.parity = _env_flag("LLB200_GEMV_PARITY", (not self.tp) or self.tp_pdl) and self.use_pdl and all_u8
        self.tp_pdl = self.tp_pdl and self.parity
        self.attn_pdl = _env_flag("LLB200_PDL_ATTN", self.parity)
        self.n_launches = 0
        self.n_replays = 0

        nh, hs = self.nh, cfg.head_size
        self.splits = max(1, min(8, (2 * N._fn["ll_sm_count"]()) // max(1, B * nh), (max_seq_length + 63) // 64))   # <= 8: one cluster per head
        if os.environ.get("LLB200_ATTN_SPLITS"):
            self.splits = int(os.environ["LLB200_ATTN_SPLITS"])
        self.attn_scratch = torch.empty(N.attn_scratch_bytes(B, nh, hs, self.splits) // 4 + 16, dtype=f32, device=dev)
        self.attn_counters = torch.zeros(B * nh, dtype=i32, device=dev)
        # the attention kernel's per-head leader can emit attn.c_proj's input already as the GEMV's digit planes + records
        # (one head = one 128-element k-block): c_proj's prologue is then two bulk copies instead of the conversion every
        # CTA repeats.  Needs the K/V-staging attention kernel (bf16, batch 1, head size 128) and the integer-MMA GEMV.
        h0p = _unwrap(h0.attn.c_proj)[0]
        self.attn_digits = (_env_flag("LLB200_ATTN_DIGITS", True) and B == 1 and hs == 128
                            and h0p.stream_for(B)[1] in (N.STREAM_U8, N.STREAM_W8)
                            and N.call_ret("ll_decode_attn_digits_supported", N.LL_BF16, B, hs, max_seq_length, self.splits) == 1)
        if self.attn_digits:
            self.att_stride = E_loc + 16
            self.att_planes = torch.zeros(4 * self.att_stride, dtype=torch.uint8, device=dev)
            self.att_rec = torch.zeros(nh * 4, dtype=f32, device=dev)
        self.graph = None
        self.sample_graphs = {}                        # (temperature, top_k) -> graph of [draw, advance, step]
        self.out_tokens = torch.zeros(cfg.block_size + 2, dtype=torch.int64, device=dev)
        self.u = torch.zeros(cfg.block_size + 2, dtype=f32, device=dev)
        self.draw = torch.zeros(1, dtype=i32, device=dev)
        self.exchange = None
        if self.tp:
            self._setup_exchange(dev)
        self._build_args()

    def matches(self, B: int, max_seq_length: int) -> bool:
        """True if this engine serves (B, max_seq_length).  After ``LLaMA.reset_cache()`` the model's
        ``kv_caches`` list is empty: the engine's own cache tensors are re-attached (slots beyond the
        #valid count are never read, so no zeroing is needed) and the captured graph stays valid."""
        if B != self.B or max_seq_length != self.S:
            return False
        if not self.model.kv_caches:
            self.model.kv_caches.extend(self.kv)
        return self.model.kv_caches[0][0].data_ptr() == self._kv0_ptr

    def reset(self):
        pass

    # ------------------------------------------------------------------ argument blocks
    _PARITY = {"attn": N.GEMV_EVEN, "proj": N.GEMV_ODD, "fc": N.GEMV_EVEN, "down": N.GEMV_ODD, "head": N.GEMV_EVEN}

    def _gemv_args(self, lin, x, y, epi, plan, norm=None, lin2=None):
        a = N.GemvArgs()
        w, a.stream_kind = lin.stream_for(self.B)
        if epi == N.EPI_PUSH and a.stream_kind != N.STREAM_BF16:
            raise RuntimeError("the push exchange needs the bf16 weight stream: run with LLB200_GEMV_KIND=0")
        a.w = w.data_ptr()
        if lin2 is not None:
            a.w2 = lin2.stream_for(self.B)[0].data_ptr()
        a.N, a.K, a.group, a.M = lin.out_features, lin.in_features, lin.tile_cols, self.B
        a.x, a.y = x.data_ptr(), (y.data_ptr() if y is not None else 0)
        if norm is not None:
            a.norm_weight, a.eps = norm.scale.data_ptr(), norm.eps
        a.epilogue, a.flags, a.grid = epi, (N.GEMV_PDL if self.use_pdl else 0), 0
        if self.parity:
            a.flags |= self._PARITY[plan]
        return a

    def _build_args(self):
        m, cfg = self.model, self.model.config
        self.kv = list(m.kv_caches)
        self._kv0_ptr = m.kv_caches[0][0].data_ptr()
        self.layers = []
        for i, blk in enumerate(m.transformer.h):
            kc, vc = m.kv_caches[i]
            a_attn = self._gemv_args(blk.attn.c_attn, self.x, self.q, N.EPI_ROPE_KV, "attn", norm=blk.rms_1)
            a_attn.rope, a_attn.pos = self.rope.data_ptr(), self.pos.data_ptr()
            a_attn.k_cache, a_attn.v_cache = kc.data_ptr(), vc.data_ptr()
            a_attn.n_head, a_attn.head_size, a_attn.cache_len = self.nh, cfg.head_size, self.S
            # tensor parallel: the row-parallel GEMVs store their partial sums; all-reduce, then the residual add
            # (the two exchanges of a block use alternating partial-sum buffers, see csrc/tp_allreduce.cu)
            if not self.tp:
                outs, epi = (self.x, self.x), N.EPI_RESIDUAL
            elif self.exchange == "push":
                outs, epi = (None, None), N.EPI_PUSH
            else:
                outs, epi = self.parts, N.EPI_STORE
            a_proj = self._gemv_args(_unwrap(blk.attn.c_proj)[0], self.att, outs[0], epi, "proj")
            if self.attn_digits:
                a_proj.x_planes, a_proj.x_rec = self.att_planes.data_ptr(), self.att_rec.data_ptr()
            a_fc = self._gemv_args(blk.mlp.c_fc1, self.x, self.h, N.EPI_SWIGLU, "fc", norm=blk.rms_2, lin2=blk.mlp.c_fc2)
            a_down = self._gemv_args(_unwrap(blk.mlp.c_proj)[0], self.h, outs[1], epi, "down")
            if epi == N.EPI_PUSH:
                for j, a in enumerate((a_proj, a_down)):
                    a.push_world = self.world
                    for r in range(self.world):
                        a.push_dst[r], a.push_cnt[r] = self._push_dst[j][r], self._push_cnt[j][r]
                    self._arrivals[j] = self.world * N.gemv_plan(a.N, a.K, a.group, 1, self.B)[1]
            self.layers.append((a_attn, kc, vc, a_proj, a_fc, a_down))
        self.a_head = self._gemv_args(m.lm_head, self.x, self.logits, N.EPI_STORE, "head", norm=m.transformer.ln_f)

    # ------------------------------------------------------------------ one step
    def _run_step(self, embed: bool = True):
        cfg = self.model.config
        st = N.stream()
        wte = self.model.transformer.wte.weight
        if embed:      # (the sampled step's draw kernel has already copied the token's row into x)
            N.call("ll_embed_ssq", N.ptr(wte), N.ptr(self.tok), N.ptr(self.x), None, self.B, cfg.n_embd, st)
        flags = N.GEMV_PDL if (self.use_pdl and self.attn_pdl) else 0
        n = 1 if embed else 0
        for a_attn, kc, vc, a_proj, a_fc, a_down in self.layers:
            N.gemv(a_attn)
            if self.attn_digits:
                N.call("ll_decode_attn_digits", N.ptr(self.q), N.ptr(kc), N.ptr(vc), N.ptr(self.pos), N.ptr(self.att), self.B,
                       self.nh, cfg.head_size, self.S, self.splits, flags, N.ptr(self.att_planes), N.ptr(self.att_rec),
                       self.att_stride, st)
            else:
                N.call("ll_decode_attn", N.ptr(self.q), N.ptr(kc), N.ptr(vc), N.ptr(self.pos), N.ptr(self.att), N.LL_BF16,
                       self.B, self.nh, cfg.head_size, self.S, self.splits, N.ptr(self.attn_scratch),
                       N.ptr(self.attn_counters), flags, st)
            N.gemv(a_proj)
            if self.tp:
                self._reduce_add(0)
            N.gemv(a_fc)
            N.gemv(a_down)
            if self.tp:
                self._reduce_add(1)
            n += 5 + ((4 if self.exchange == "nccl" else 2) if self.tp else 0)
        N.gemv(self.a_head)
        self.n_launches = n + 1

    def _setup_exchange(self, dev):
        """Exchange after the row-parallel GEMVs.  "peer" (default): two symmetric-memory buffers + a flag array
        mapped into every rank, pulled by csrc/tp_allreduce.cu straight over NVLink - one small kernel per exchange.
        "push": the GEMV epilogue stores its partial sums into every rank's symmetric-memory receive buffer and
        signals arrival; a small kernel reduces from local memory (compute and the first half of the collective in
        one kernel; correct, but measured 4 % slower at 65B TP2: the system-scope release at the end of every GEMV
        CTA costs more than the pull it saves).  Both sit inside the step graph.  "nccl": dist.all_reduce + ll_add, run eagerly (a captured NCCL
        collective next to programmatic-launch kernels stalled on the 2-GPU box, profiles/r01_tp.md)."""
        import os
        B, E = self.x.shape
        rank, world, group = self.model._tp
        self.rank, self.world = rank, world
        mode = os.environ.get("LLB200_TP_EXCHANGE", "ll")
        if mode == "ll":
            # one NVLink traversal per exchange: every rank pushes {word, epoch} units into every peer's receive buffer and
            # reads the peers' units from its own memory (csrc/tp_allreduce.cu tp_ll_allreduce_add_kernel)
            import ctypes as C
            import torch.distributed._symmetric_memory as symm
            grp = group if group is not None else dist.group.WORLD
            nbytes = N.call_ret("ll_tp_ll_recv_bytes", world, B * E)
            recv = symm.empty((nbytes,), dtype=torch.uint8, device=dev)
            recv.zero_()
            hr = symm.rendezvous(recv, grp)
            torch.cuda.synchronize()
            dist.barrier(group=group)                     # every rank's buffer is zero (no epoch matches) before anyone sends
            self._peer_recv = (C.c_uint64 * world)(*[int(p) for p in hr.buffer_ptrs])
            self._symm = (recv, hr)
            self.parts = [torch.zeros((B, E), dtype=torch.bfloat16, device=dev) for _ in range(2)]
            self.epoch = torch.zeros(8, dtype=torch.int32, device=dev)   # one call counter per exchange CTA (LL_TP_MAX_CTAS)
            self.exchange = "ll"
            return
        if mode == "push":
            # the row-parallel GEMV's epilogue stores its partial sums into every rank's receive buffer and counts
            # its CTAs in; ll_tp_reduce_add then sums from local memory (csrc/gemv_int4.cu LL_EPI_PUSH, tp_allreduce.cu)
            import torch.distributed._symmetric_memory as symm
            grp = group if group is not None else dist.group.WORLD
            recv = symm.empty((2, world, B, E), dtype=torch.bfloat16, device=dev)
            cnt = symm.empty((64,), dtype=torch.int32, device=dev)
            recv.zero_()
            cnt.zero_()
            hb, hc = symm.rendezvous(recv, grp), symm.rendezvous(cnt, grp)
            torch.cuda.synchronize()
            dist.barrier(group=group)
            slot = B * E * 2
            self._push_dst = [[int(p) + (j * world + rank) * slot for p in hb.buffer_ptrs] for j in range(2)]
            self._push_cnt = [[int(p) + 4 * j for p in hc.buffer_ptrs] for j in range(2)]
            self._arrivals = [0, 0]
            self._symm = (recv, cnt, hb, hc)
            self.recv, self.cnt = recv, cnt
            self.epoch = torch.zeros(8, dtype=torch.int32, device=dev)   # one call counter per exchange CTA (LL_TP_MAX_CTAS)
            self.exchange = "push"
            return
        if mode == "peer":
            import torch.distributed._symmetric_memory as symm
            grp = group if group is not None else dist.group.WORLD
            buf = symm.empty((2, B, E), dtype=torch.bfloat16, device=dev)
            flg = symm.empty((64,), dtype=torch.int32, device=dev)
            buf.zero_()
            flg.zero_()
            hb, hf = symm.rendezvous(buf, grp), symm.rendezvous(flg, grp)
            torch.cuda.synchronize()
            dist.barrier(group=group)                     # every rank's flags are zero before anyone raises one
            import ctypes as C
            stride = B * E * 2
            self._peer_parts = [(C.c_uint64 * world)(*[int(p) + j * stride for p in hb.buffer_ptrs]) for j in range(2)]
            self._peer_flags = (C.c_uint64 * world)(*[int(p) for p in hf.buffer_ptrs])
            self._symm = (buf, flg, hb, hf)
            self.parts = [buf[0], buf[1]]
            self.epoch = torch.zeros(8, dtype=torch.int32, device=dev)   # one call counter per exchange CTA (LL_TP_MAX_CTAS)
            self.exchange = "peer"
        elif mode == "nccl":
            self.parts = [self.part, self.part]
            self.exchange = "nccl"
            self.use_graph = False
        else:
            raise ValueError(f"LLB200_TP_EXCHANGE must be 'll', 'peer', 'push' or 'nccl', not {mode!r}")

    def _reduce_add(self, j: int):
        """x += sum over ranks of the partial result in buffer ``j``: the path's only exchange step."""
        if self.exchange == "push":
            N.call("ll_tp_reduce_add", N.ptr(self.x), N.ptr(self.recv), N.ptr(self.cnt), N.ptr(self.epoch), self.world,
                   self.x.numel(), self._arrivals[j], N.stream())
        elif self.exchange == "ll":
            import ctypes as C
            N.call("ll_tp_ll_allreduce_add", N.ptr(self.x), N.ptr(self.parts[j]), C.cast(self._peer_recv, C.c_void_p),
                   N.ptr(self.epoch), self.rank, self.world, self.x.numel(), N.GEMV_PDL if self.tp_pdl else 0, N.stream())
        elif self.exchange == "peer":
            import ctypes as C
            N.call("ll_tp_allreduce_add", N.ptr(self.x), C.cast(self._peer_parts[j], C.c_void_p), C.cast(self._peer_flags, C.c_void_p),
                   N.ptr(self.epoch), self.rank, self.world, self.x.numel(), N.GEMV_PDL if self.tp_pdl else 0, N.stream())
        else:
            dist.all_reduce(self.part, op=dist.ReduceOp.SUM, group=self.tp_group)
            N.call("ll_add", N.ptr(self.x), N.ptr(self.part), N.ptr(self.x), N.LL_BF16, self.x.numel(), N.stream())

    def _capture(self):
        # warm up eagerly on a side stream (also sizes kernel attributes), then capture
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            self._run_step()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._run_step()
        self.graph = g

    def step(self):
        """Run one decode step on the tokens in ``self.tok`` at the position in ``self.pos``."""
        if not self.use_graph:
            self._run_step()
            return
        if self.graph is None:
            # the warm-up run inside _capture() writes a K/V row at the current slot and advances x;
            # harmless: the replay below rewrites the same slot from the same token and position
            self._capture()
        self.graph.replay()
        self.n_replays += 1

    # ------------------------------------------------------------------ sampled decoding (generate()'s loop on the device)
    def _draw(self, temperature: float, top_k):
        """token ~ top-k / temperature softmax of ``self.logits`` -> ``self.tok`` and ``out_tokens[pos + 1]``."""
        V = self.logits.shape[-1]
        N.call("ll_sample_topk", N.ptr(self.logits), V, float(temperature), int(top_k) if top_k is not None else 0,
               N.ptr(self.u), N.ptr(self.draw), N.ptr(self.pos), N.ptr(self.out_tokens), N.ptr(self.tok), N.stream())

    def _draw_advance_embed(self, temperature: float, top_k):
        """One kernel: the draw of ``_draw``, then the position update of ``ll_advance_pos`` and the embedding copy of
        ``ll_embed_ssq`` for the token just drawn (csrc/sampling.cu)."""
        V = self.logits.shape[-1]
        wte = self.model.transformer.wte.weight
        N.call("ll_sample_advance_embed", N.ptr(self.logits), V, float(temperature), int(top_k) if top_k is not None else 0,
               N.ptr(self.u), N.ptr(self.draw), N.ptr(self.pos), N.ptr(self.out_tokens), N.ptr(self.tok), self.S,
               self.rope.shape[0], N.ptr(wte), N.ptr(self.x), wte.shape[1], N.stream())

    def _run_sampled(self, temperature, top_k):
        self._draw_advance_embed(temperature, top_k)
        self._run_step(embed=False)
        self.n_launches += 1

    def begin_sampling(self, prompt: torch.Tensor, last_logits: torch.Tensor, n_draws: int) -> bool:
        """Device-side state for ``n_draws`` draws after a prompt of n tokens whose last logits are given:
        tokens in ``out_tokens[:n]``, position n - 1, fresh uniforms from torch's generator."""
        n = prompt.numel()
        if self.B != 1 or n + n_draws > self.out_tokens.numel() or n_draws > self.u.numel():
            return False
        self.out_tokens[:n] = prompt
        self.logits.copy_(last_logits.reshape(1, -1))
        self.u[:n_draws].uniform_()
        if self.tp:
            # every rank must draw the same tokens from its (identical) logits: rank 0's uniforms are the ones used,
            # however the ranks' generators were seeded or advanced
            grp = self.tp_group
            dist.broadcast(self.u[:n_draws], src=dist.get_global_rank(grp, 0) if grp is not None else 0, group=grp)
        self.draw.zero_()
        self.pos.copy_(torch.tensor([n - 1, (n - 1) % self.S, min(n, self.S)], dtype=torch.int32), non_blocking=True)
        return True

    def ensure_sample_graph(self, temperature: float, top_k):
        """Capture [draw, advance, step] for these sampling parameters.  MUST run before the prompt is prefilled:
        the warm-up runs write a K/V row at slot 0 and overwrite ``logits`` (the prefill then rewrites the row)."""
        key = (float(temperature), top_k)
        if not self.use_graph or key in self.sample_graphs:
            return
        if self.graph is None:
            self._capture()
        # run the draw kernel once outside capture (its first launch loads it)
        self.pos.zero_()
        self._draw_advance_embed(temperature, top_k)
        self.pos.zero_()
        self.draw.zero_()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._run_sampled(temperature, top_k)
        self.sample_graphs[key] = g

    def sample_and_step(self, temperature: float, top_k):
        """Draw the next token from the current logits, append it, run the decode step on it (one graph replay)."""
        if not self.use_graph:
            self._run_sampled(temperature, top_k)
            return
        self.sample_graphs[(float(temperature), top_k)].replay()   # KeyError: ensure_sample_graph() was not called
        self.n_replays += 1

    def sample_last(self, temperature: float, top_k):
        self._draw(temperature, top_k)

    def forward_token(self, idx: torch.Tensor, input_pos: torch.Tensor) -> torch.Tensor:
        """``LLaMA.forward(idx[B,1], max_seq_length, input_pos[1])`` (ref: model.py:84-128)."""
        if input_pos.device.type == "cpu":   # checked on the host when it costs no synchronisation (ref model.py:101 raises)
            assert 0 <= int(input_pos[0]) < self.rope.shape[0], f"input_pos {int(input_pos[0])} outside the RoPE table ({self.rope.shape[0]} rows)"
            input_pos = input_pos.to(self.tok.device)
        self.tok.copy_(idx.reshape(-1))
        N.call("ll_set_pos", N.ptr(self.pos), N.ptr(input_pos.to(torch.int64)), self.S, self.rope.shape[0], N.stream())
        self.step()
        return self.logits.clone().view(self.B, 1, -1)
