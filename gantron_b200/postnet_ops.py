"""Postnet.forward / backward (reference model.py:108-151) on the sm_100a kernels.

Per layer:  implicit-GEMM conv1d (tcgen05, bias + BatchNorm sum / sum-of-squares fused in the epilogue)
            -> gtn_bn_finalize (mean, rstd, running-stat update)
            -> gtn_bn_act_fwd  (normalise + tanh + Philox dropout, writes the next layer's bf16 operand)
Backward:   gtn_bn_act_bwd (two passes: channel reductions, then dy + bias gradient)
            -> tcgen05 weight-gradient GEMM (MN-major operands) + data-gradient conv (flipped/transposed pack).
Activations stay channels-last (B,T,C); the (B,n_mel,T) tensors of the API are transposed views.

Dropout sites: oracle/philox.py SITE_POSTNET0 + layer; element numbering (b*T + t)*C + c.
"""
from __future__ import annotations

from typing import List, Optional

import torch
from torch.autograd.function import once_differentiable

from . import ops
from .conv_ops import PackCache, act_operand, pack_conv_weight, wgrad_operands
from .packing import BF16

SITE_POSTNET0 = 16


def _cache(mod) -> PackCache:
    c = getattr(mod, "_gtn_pack_cache", None)
    if c is None:
        c = PackCache()
        object.__setattr__(mod, "_gtn_pack_cache", c)
    return c


def _layer_params(postnet):
    out = []
    for seq in postnet.convolutions:
        conv, bn = seq[0].conv, seq[1]
        out += [conv.weight, conv.bias, bn.weight, bn.bias]
    return out


def _stack_acts(postnet):
    """Activation per layer.  Postnet (model.py:146-151): tanh on all but the last; text-encoder convolutions
    (model.py:196-197, 216-217; `_gtn_act = 'relu'` on the owning module): relu on every layer."""
    n = len(postnet.convolutions)
    if getattr(postnet, "_gtn_act", "tanh") == "relu":
        return [ops.ACT_RELU] * n
    return [ops.ACT_TANH] * (n - 1) + [ops.ACT_NONE]


class _PostnetFn(torch.autograd.Function):
    """[conv1d -> BatchNorm1d -> activation -> dropout(0.5)] x n on the implicit-GEMM conv; shared by the Postnet and the
    text encoder's convolution bank (dropout sites `_gtn_site0` + layer, default SITE_POSTNET0)."""

    @staticmethod
    def forward(ctx, postnet, seed, prec, x, wgrad_input, *params):
        n_layers = len(postnet.convolutions)
        training = postnet.training
        acts = _stack_acts(postnet)
        site0 = getattr(postnet, "_gtn_site0", SITE_POSTNET0)
        B, C0, T = x.shape
        dev = x.device
        adt = BF16 if prec == 1 else torch.float32          # storage type of activations between layers
        cache = _cache(postnet)
        cur = x.detach().transpose(1, 2)                     # (B,T,C0) view
        cur = cur.to(adt).contiguous() if prec == 1 else cur.float().contiguous()
        saved = []
        for i in range(n_layers):
            conv, bn = postnet.convolutions[i][0].conv, postnet.convolutions[i][1]
            w, b, gamma, beta = params[4 * i:4 * i + 4]
            k, pad = conv.kernel_size[0], conv.padding[0]
            cout = w.shape[0]
            wp, cin_eff, cin_pad = cache.get(("fwd", i, prec), w, lambda w=w: pack_conv_weight(w, prec))
            xop = act_operand(cur, prec)
            last = i == n_layers - 1
            act = acts[i]
            y = torch.empty(B, T, cout, device=dev, dtype=adt)
            if training:
                stats = torch.zeros(2, cout, device=dev)
                ops.conv1d(wp, cin_eff, cin_pad, k, 1, pad, xop, y, None, bias=b.detach(), stats=stats)
                mean, rstd = ops.bn_finalize(stats, B * T, bn.eps, bn.momentum if bn.momentum is not None else 0.1,
                                             bn.running_mean, bn.running_var)
                bn.num_batches_tracked += 1
                drop = (0.5, seed, site0 + i)
            else:
                ops.conv1d(wp, cin_eff, cin_pad, k, 1, pad, xop, y, None, bias=b.detach())
                mean, rstd = bn.running_mean, torch.rsqrt(bn.running_var + bn.eps)
                drop = (0.0, 0, 0)
            out = torch.empty(B, T, cout, device=dev, dtype=torch.float32 if (last and site0 == SITE_POSTNET0) else adt)
            ops.bn_act_fwd(y, mean, rstd, gamma.detach(), beta.detach(), act, drop, out)
            saved.append((cur, y, mean, rstd, act, drop))
            cur = out
        if wgrad_input is not None:      # reference quirk: conv0's weight gradient sees the masked mel
            x0 = wgrad_input.detach().transpose(1, 2)
            x0 = x0.to(adt).contiguous() if prec == 1 else x0.float().contiguous()
            saved[0] = (x0,) + saved[0][1:]
        ctx.postnet, ctx.prec, ctx.saved, ctx.training = postnet, prec, saved, training
        ctx.shape = (B, C0, T)
        return cur.transpose(1, 2)

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        postnet, prec, saved = ctx.postnet, ctx.prec, ctx.saved
        B, C0, T = ctx.shape
        dev = g.device
        adt = BF16 if prec == 1 else torch.float32
        cache = _cache(postnet)
        n_layers = len(saved)
        dx = g.transpose(1, 2)
        dx = dx.to(adt).contiguous() if prec == 1 else dx.float().contiguous()
        grads: List[Optional[torch.Tensor]] = [None] * (4 * n_layers)
        params = _layer_params(postnet)
        for i in reversed(range(n_layers)):
            conv, bn = postnet.convolutions[i][0].conv, postnet.convolutions[i][1]
            w, b, gamma, beta = params[4 * i:4 * i + 4]
            x_in, y, mean, rstd, act, drop = saved[i]
            k, pad = conv.kernel_size[0], conv.padding[0]
            cout, cin = w.shape[0], w.shape[1]
            dy = torch.empty(B, T, cout, device=dev, dtype=adt)
            # eval mode: running statistics are constants (no batch-coupling terms)
            dgamma, dbeta, dbias = ops.bn_act_bwd(dx, y, mean, rstd, gamma.detach(), beta.detach(), act, drop, dy,
                                                  batch_stats=ctx.training)
            grads[4 * i + 1], grads[4 * i + 2], grads[4 * i + 3] = dbias, dgamma, dbeta
            dyo, xo = wgrad_operands(dy, x_in, prec)
            dw = torch.empty(k, cout, cin, device=dev)
            ops.conv1d_wgrad(dyo, xo, cin, k, 1, pad, dw)
            grads[4 * i] = dw.permute(1, 2, 0).contiguous()
            if i > 0 or ctx.needs_input_grad[3]:
                wT, ce, cp = cache.get(("dgrad", i, prec), w, lambda w=w: pack_conv_weight(w, prec, dgrad=True))
                dx = torch.empty(B, T, cin, device=dev, dtype=adt if i > 0 else torch.float32)
                ops.conv1d(wT, ce, cp, k, 1, (k - 1) - pad, act_operand(dy, prec), dx, None)
        ctx.saved = None
        gx = dx.transpose(1, 2) if ctx.needs_input_grad[3] else None
        return (None, None, None, gx, None) + tuple(grads)


def postnet_forward(postnet, x: torch.Tensor, seed: int, prec: int, wgrad_input: Optional[torch.Tensor] = None):
    return _PostnetFn.apply(postnet, seed, prec, x, wgrad_input, *_layer_params(postnet))
