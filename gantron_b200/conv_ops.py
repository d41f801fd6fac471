"""Shared host helpers of the implicit-GEMM conv paths (Postnet, Discriminator): weight packs, operand
layouts, per-module pack caches.

Operand layouts (see gtn_gemm.cu MODE_CONV / MODE_WGRAD):
  activations  channels-last bf16 [B][T][C_eff]        (C_eff = C in bf16 mode, 6 * pad64(C) plane blocks in fp32 mode)
  weights      bf16 [Cout][taps * cin_pad_eff]         K-major, tap-major then channel
fp32 mode carries every fp32 value as three bf16 planes (packing.py): along the CHANNEL axis for forward /
data-gradient convs, along the BATCH axis (the K axis of a weight gradient) for wgrad.
"""
from __future__ import annotations

from typing import Dict, Tuple

import torch

from . import packing
from .packing import BF16, W_PLANES, X_PLANES


def pad64(n: int) -> int:
    return (n + 63) // 64 * 64


def pack_conv_weight(w: torch.Tensor, prec: int, dgrad: bool = False) -> Tuple[torch.Tensor, int, int]:
    """(Cout,Cin,k) fp32 -> (bf16 [rows][k * cin_pad_eff], cin_eff, cin_pad_eff).

    dgrad=True packs the data-gradient operator W'[ci][tap][co] = W[co][ci][k-1-tap], so the same forward
    kernel computes dX from dY."""
    w = w.detach().float()
    if dgrad:
        w = w.flip(-1).permute(1, 0, 2)
    co, ci, k = w.shape
    cpad = pad64(ci)
    if prec == 1:
        wp = torch.zeros(co, k, cpad, device=w.device, dtype=BF16)
        wp[:, :, :ci] = w.permute(0, 2, 1).to(BF16)
        return wp.reshape(co, k * cpad).contiguous(), ci, cpad
    planes = packing.split3(w.permute(0, 2, 1))          # each (co,k,ci)
    wp = torch.zeros(co, k, 6, cpad, device=w.device, dtype=BF16)
    for s, pl in enumerate(W_PLANES):
        wp[:, :, s, :ci] = planes[pl]
    return wp.reshape(co, k * 6 * cpad).contiguous(), 6 * cpad, 6 * cpad


def act_operand(x: torch.Tensor, prec: int) -> torch.Tensor:
    """(B,T,C) fp32/bf16 channels-last -> conv operand (bf16 mode: the tensor itself as bf16; fp32 mode: planes)."""
    if prec == 1:
        if x.dtype == BF16 and x.stride(2) == 1 and x.stride(1) % 8 == 0 and x.data_ptr() % 16 == 0:
            return x
        C = x.shape[2]
        ld = (C + 7) // 8 * 8
        buf = torch.zeros(x.shape[0], x.shape[1], ld, device=x.device, dtype=BF16) if ld != C else \
            torch.empty(x.shape[0], x.shape[1], ld, device=x.device, dtype=BF16)
        buf[:, :, :C] = x
        return buf[:, :, :C]
    B, T, C = x.shape
    cpad = pad64(C)
    planes = packing.split3(x)
    out = torch.zeros(B, T, 6, cpad, device=x.device, dtype=BF16)
    for s, pl in enumerate(X_PLANES):
        out[:, :, s, :C] = planes[pl]
    return out.reshape(B, T, 6 * cpad)


def wgrad_operands(dy: torch.Tensor, x: torch.Tensor, prec: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """Operands of gtn_conv1d_wgrad_bf16 for dW = sum_{b,t} dy x: (dy', x') channels-last bf16, planes stacked on batch."""
    if prec == 1:
        return act_operand(dy, 1), act_operand(x, 1)
    dp, xp = packing.split3(dy), packing.split3(x)
    A = torch.cat([dp[i] for i in W_PLANES], 0)
    Bm = torch.cat([xp[i] for i in X_PLANES], 0)
    return act_operand(A, 1), act_operand(Bm, 1)


class PackCache:
    """Per-module cache of packed weights keyed by (name, kind), invalidated by the parameter's version counter."""

    def __init__(self):
        self._d: Dict[tuple, tuple] = {}

    def get(self, key, param: torch.Tensor, build):
        ver = (param._version, param.data_ptr())
        hit = self._d.get(key)
        if hit is not None and hit[0] == ver:
            return hit[1]
        val = build()
        self._d[key] = (ver, val)
        return val
