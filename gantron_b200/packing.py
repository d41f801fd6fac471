"""Operand packing for the sm_100a kernels: K padding, bf16 conversion and the 3-plane fp32 split.

prec == 1  ("bf16 mode"): operands are bf16, K segments padded to a multiple of 32.
prec == 3  ("fp32 mode"): an fp32 value v is carried as three bf16 planes hi + lo + lo2 (24 mantissa
bits).  A product W.X is then evaluated by the same bf16 tensor-core kernels over a 6x longer K:

    weight sub-segments  (hi, lo, lo2, hi, lo, hi)
    activation sub-segs  (hi, hi, hi,  lo, lo, lo2)      -> hi*hi + lo*hi + lo2*hi + hi*lo + lo*lo + hi*lo2

(the dropped terms are <= 2^-24 relative).  The device twin is gtn_decoder.cuh:store_act.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import torch

BF16 = torch.bfloat16
W_PLANES = (0, 1, 2, 0, 1, 0)   # which of (hi, lo, lo2) each weight sub-segment holds
X_PLANES = (0, 0, 0, 1, 1, 2)   # ... and each activation sub-segment


def pad_to(n: int, m: int) -> int:
    return (n + m - 1) // m * m


def mult_of(prec: int) -> int:
    assert prec in (1, 3)
    return 1 if prec == 1 else 6


def split3(x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    x = x.float()
    hi = x.to(BF16)
    r1 = x - hi.float()
    lo = r1.to(BF16)
    lo2 = (r1 - lo.float()).to(BF16)
    return hi, lo, lo2


def expand(x: torch.Tensor, kpad: int, prec: int, is_weight: bool) -> torch.Tensor:
    """fp32 (..., K) -> bf16 (..., mult*kpad): zero-pad K to kpad and lay out the precision planes."""
    K = x.shape[-1]
    assert kpad >= K
    if prec == 1:
        if kpad == K:                       # nothing to pad: one conversion kernel (none if x is already bf16)
            return x.to(BF16).contiguous()
        out = torch.zeros(*x.shape[:-1], kpad, device=x.device, dtype=BF16)
        out[..., :K] = x.to(BF16)
        return out
    planes = split3(x)
    out = torch.zeros(*x.shape[:-1], 6 * kpad, device=x.device, dtype=BF16)
    order = W_PLANES if is_weight else X_PLANES
    for s, pl in enumerate(order):
        out[..., s * kpad:s * kpad + K] = planes[pl]
    return out


def pack_weight(segments: Sequence[Tuple[torch.Tensor, int]], prec: int, row_pad: int = 0) -> torch.Tensor:
    """Concatenate K segments [(W[:, cols], kpad), ...] into one packed bf16 matrix; optionally zero-pad rows."""
    parts: List[torch.Tensor] = [expand(w.detach(), kp, prec, True) for w, kp in segments]
    out = torch.cat(parts, dim=1)
    if row_pad and out.shape[0] < row_pad:
        out = torch.cat((out, torch.zeros(row_pad - out.shape[0], out.shape[1], device=out.device, dtype=BF16)), 0)
    return out.contiguous()
