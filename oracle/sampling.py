"""Numpy restatement of the sampling tail of the reference's ``generate`` (/root/reference/generate.py:62-75).

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  The reference computes, on the last position's logits,

    logits = logits[0, -1] / temperature                        generate.py:62
    v, _ = torch.topk(logits, min(top_k, logits.size(-1)))      generate.py:65-67
    logits = torch.where(logits < v[[-1]], -float("Inf"), logits)
    probs = torch.nn.functional.softmax(logits, dim=-1)         generate.py:69
    idx_next = torch.multinomial(probs, num_samples=1)          generate.py:70

``probabilities`` follows those lines with bf16 rounding where a bf16 tensor is materialised (after the division and
on the softmax output; the softmax itself runs in fp32, as torch's does).  It is pinned against torch's own CPU ops in
tests/test_oracle_golden.py.  ``draw`` states the product's categorical draw: inverse CDF over the bf16
probabilities in index order from one uniform number - the same distribution as torch.multinomial, which however
consumes its generator differently (one exponential per entry), so no token-for-token comparison with the
reference's random stream is possible or claimed.
"""
from __future__ import annotations

import numpy as np

from .quant import F32, bf16_round


def probabilities(logits, temperature: float, top_k=None, dtype: str = "bf16") -> np.ndarray:
    rd = bf16_round if dtype == "bf16" else (lambda a: np.asarray(a, dtype=F32))
    l = rd(np.asarray(logits, dtype=F32) / F32(temperature))
    if top_k is not None and 0 < top_k < l.size:
        kth = np.sort(l)[::-1][top_k - 1]
        l = np.where(l < kth, F32(-np.inf), l)
    e = np.exp((l - l.max()).astype(F32)).astype(F32)
    return rd(e / e.sum(dtype=F32))


def draw(p: np.ndarray, u: float) -> int:
    """First index whose inclusive cumulative sum (fp32, index order) exceeds u * sum(p); the last kept index if none."""
    p = np.asarray(p, dtype=F32)
    cum = np.cumsum(p, dtype=F32)
    target = F32(u) * cum[-1]
    hit = np.nonzero(cum > target)[0]
    return int(hit[0]) if hit.size else int(np.nonzero(p > 0)[0][-1])


def draw_margin(p: np.ndarray, u: float) -> float:
    """Distance of u * sum(p) to the nearest CDF boundary, relative to sum(p): draws closer than fp32 summation
    noise to a boundary may legitimately land on either side."""
    cum = np.cumsum(np.asarray(p, dtype=np.float64))
    return float(np.abs(cum - u * cum[-1]).min() / cum[-1])
