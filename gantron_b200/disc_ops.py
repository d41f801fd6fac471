"""Discriminator.forward / adversarial_loss (reference model.py:500-540, layers.py:43-60) on the sm_100a kernels.

Forward: the reference's raw reshape (B,F',80) -> (B,1600,P) is kept as a zero-copy view; one transposing cast
makes the channels-last bf16 operand (B,P,1600); four implicit-GEMM convs (tcgen05) with bias -> Philox
dropout -> tanh fused in the epilogue (DiscConv1d order, layers.py:53-57); a warp-per-row head for the final
Conv1d(80 -> 1, k=1); one masked-mean kernel for the loss (no per-sample host syncs, cf. model.py:533-537).
Backward: gtn_drop_tanh_bwd + data-gradient convs (+ weight-gradient GEMMs when D's parameters require grad,
i.e. in the discriminator step; they are frozen during the generator step, train.py:366-367).

Dropout sites: oracle/philox.py SITE_DISC0 + layer; element numbering (b*P + p)*C + c.
"""
from __future__ import annotations

from typing import List, Optional

import torch
from torch.autograd.function import once_differentiable

from . import _lib, ops
from .conv_ops import PackCache, act_operand, pack_conv_weight, wgrad_operands
from .packing import BF16

SITE_DISC0 = 32


def _cache(mod) -> PackCache:
    c = getattr(mod, "_gtn_pack_cache", None)
    if c is None:
        c = PackCache()
        object.__setattr__(mod, "_gtn_pack_cache", c)
    return c


def _convs(disc):
    seq = disc.discriminator
    return [seq[i].module[0] for i in range(len(seq) - 1)], seq[len(seq) - 1]


def _geom(layer):
    """(kernel, dilation, padding) of a DiscConv1d conv, or of a DiscDense linear seen as a one-tap conv (layers.py:63-73:
    Linear -> Dropout -> Tanh is the same bias -> dropout -> tanh epilogue over rows instead of positions)."""
    if isinstance(layer, torch.nn.Linear):
        return 1, 1, 0
    return layer.kernel_size[0], layer.dilation[0], layer.padding[0]


def _w3(w: torch.Tensor) -> torch.Tensor:
    return w if w.dim() == 3 else w.unsqueeze(-1)


def _params(disc):
    convs, head = _convs(disc)
    out = []
    for c in convs:
        out += [c.weight, c.bias]
    return out + [head.weight, head.bias]


class _DiscFn(torch.autograd.Function):
    """x (B, window*n_mel, P) [the raw-reshaped input] -> (B, P).  Serves the conv Discriminator and, with one-tap layers
    and P = stacked windows, the LinearDiscriminator (model.py:543-583); dropout sites disc._gtn_site0 + layer."""

    @staticmethod
    def forward(ctx, disc, seed, prec, x, *params):
        convs, head = _convs(disc)
        B, C0, P = x.shape
        dev = x.device
        adt = BF16 if prec == 1 else torch.float32
        cache = _cache(disc)
        cur = x.detach().transpose(1, 2)
        cur = cur.to(adt).contiguous() if prec == 1 else cur.float().contiguous()
        saved = []
        for i, conv in enumerate(convs):
            w, b = params[2 * i], params[2 * i + 1]
            k, dil, pad = _geom(conv)
            wp, cin_eff, cin_pad = cache.get(("fwd", i, prec), w, lambda w=w: pack_conv_weight(_w3(w), prec))
            a = torch.empty(B, P, w.shape[0], device=dev, dtype=adt)
            p_drop = disc.discriminator[i].module[1].p if disc.discriminator[i].module[1].training else 0.0
            drop = (p_drop, seed, getattr(disc, "_gtn_site0", SITE_DISC0) + i)
            ops.conv1d(wp, cin_eff, cin_pad, k, dil, pad, act_operand(cur, prec), a, None, bias=b.detach(),
                       act=ops.ACT_TANH, dropout=drop, drop_first=True, block_n=64)
            saved.append((cur, a, drop))
            cur = a
        hw, hb = params[-2], params[-1]
        out = ops.disc_head_fwd(cur, hw.detach().reshape(-1).contiguous(), hb.detach())
        ctx.disc, ctx.prec, ctx.saved, ctx.dims = disc, prec, saved, (B, C0, P)
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        # (double backward -- gradient_penalty, train.py:32-69 -- never reaches this node: see discriminator_forward)
        disc, prec, saved = ctx.disc, ctx.prec, ctx.saved
        convs, head = _convs(disc)
        params = _params(disc)
        B, C0, P = ctx.dims
        dev = g.device
        adt = BF16 if prec == 1 else torch.float32
        cache = _cache(disc)
        n = len(convs)
        grads: List[Optional[torch.Tensor]] = [None] * len(params)
        need_w = any(ctx.needs_input_grad[4:])
        a_last = saved[-1][1]
        dcur = torch.empty(B, P, a_last.shape[2], device=dev, dtype=torch.float32)
        hw = params[-2]
        dw, db = ops.disc_head_bwd(g.contiguous().float(), a_last, hw.detach().reshape(-1).contiguous(), dcur, need_w)
        if need_w:
            grads[-2], grads[-1] = dw.reshape(hw.shape), db
        for i in reversed(range(n)):
            conv = convs[i]
            w = params[2 * i]
            x_in, a, drop = saved[i]
            k, dil, pad = _geom(conv)
            cout, cin = w.shape[0], w.shape[1]
            dpre = torch.empty(B, P, cout, device=dev, dtype=adt)
            dbias = ops.drop_tanh_bwd(dcur, a, drop, dpre, want_dbias=need_w)
            if need_w:
                dyo, xo = wgrad_operands(dpre, x_in, prec)
                dwt = torch.empty(k, cout, cin, device=dev)
                ops.conv1d_wgrad(dyo, xo, cin, k, dil, pad, dwt, block_n=128)
                grads[2 * i], grads[2 * i + 1] = dwt.permute(1, 2, 0).reshape(w.shape).contiguous(), dbias
            if i > 0 or ctx.needs_input_grad[3]:
                wT, ce, cp = cache.get(("dgrad", i, prec), w, lambda w=w: pack_conv_weight(_w3(w), prec, dgrad=True))
                dcur = torch.empty(B, P, cin, device=dev, dtype=torch.float32)
                ops.conv1d(wT, ce, cp, k, dil, dil * (k - 1) - pad, act_operand(dpre, prec), dcur, None, block_n=64)
        ctx.saved = None
        gx = dcur.transpose(1, 2) if ctx.needs_input_grad[3] else None
        return (None, None, None, gx) + tuple(grads)


def _keep_scale_mask(B: int, P: int, C: int, drop, dev) -> torch.Tensor:
    """(B,P,C) tensor holding keep / (1 - p) of the kernels' Philox stream for `drop` = (p, seed, site)."""
    ones = torch.ones(B, P, C, device=dev)
    out = torch.empty_like(ones)
    ops.bn_act_fwd(ones, None, None, None, None, ops.ACT_NONE, drop, out)
    return out


def _discriminator_composite(disc, x: torch.Tensor, seed: int) -> torch.Tensor:
    """The same network as ordinary differentiable torch ops (conv1d -> Philox dropout mask -> tanh, layers.py:53-57), for
    the one caller that differentiates TWICE through the Discriminator: gradient_penalty (train.py:32-69, create_graph=True
    at :55-57).  Same dropout stream as the kernel path; fp32 arithmetic."""
    import torch.nn.functional as F
    convs, head = _convs(disc)
    B, _, P = x.shape
    cur = x.float()
    for i, conv in enumerate(convs):
        dmod = disc.discriminator[i].module[1]
        y = F.conv1d(cur, conv.weight, conv.bias, padding=conv.padding[0], dilation=conv.dilation[0])
        if dmod.training and dmod.p > 0:
            y = y * _keep_scale_mask(B, P, conv.weight.shape[0], (dmod.p, seed, SITE_DISC0 + i), x.device).transpose(1, 2)
        cur = torch.tanh(y)
    return F.conv1d(cur, head.weight, head.bias).squeeze(1)


def discriminator_forward(disc, inputs: torch.Tensor, seed: int, prec: int) -> torch.Tensor:
    """inputs (B, frames, n_mel) -> (B, positions); tail handling and raw reshape as model.py:521-526."""
    win = disc.window
    r = inputs.shape[1] % win
    if r != 0:
        inputs = torch.cat((inputs[:, :-r], inputs[:, -win:]), dim=1)
    x = inputs.reshape(inputs.shape[0], disc.in_dim, -1)          # raw memory reinterpretation (copy if strided)
    params = _params(disc)
    if (getattr(disc, "_gp_lambda", 0) > 0 and torch.is_grad_enabled() and x.requires_grad
            and any(p.requires_grad for p in params)):
        # hparams.gradient_penalty_lambda > 0 and input AND parameters want gradients: in the reference's schedule that is
        # gradient_penalty only (the generator step freezes D, train.py:366-367; the discriminator step feeds detached mels,
        # train.py:313-321), and it goes on to differentiate the input gradient again.  The fused kernels are
        # once-differentiable, so this call takes the composite (fp32, TF32 off).
        with torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
            return _discriminator_composite(disc, x, seed)
    return _DiscFn.apply(disc, seed, prec, x, *params)


class _WindowMean(torch.autograd.Function):
    @staticmethod
    def forward(ctx, out, lengths, window):
        out = out.contiguous().float()
        B, P = out.shape
        lengths = lengths.to(device=out.device, dtype=torch.int64).contiguous()
        loss = torch.empty(1, device=out.device)
        _lib.check(_lib.load().gtn_window_mean_fwd(_lib.ptr(out), B, P, _lib.ptr(lengths), int(window), _lib.ptr(loss),
                                                   _lib.stream()), "gtn_window_mean_fwd")
        ctx.lengths, ctx.window, ctx.dims = lengths, int(window), (B, P)
        return loss.reshape(())

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        B, P = ctx.dims
        dout = torch.empty(B, P, device=g.device)
        gg = g.reshape(1).float().contiguous()
        _lib.check(_lib.load().gtn_window_mean_bwd(_lib.ptr(gg), B, P, _lib.ptr(ctx.lengths), ctx.window, _lib.ptr(dout),
                                                   _lib.stream()), "gtn_window_mean_bwd")
        return dout, None, None


def masked_window_mean(out: torch.Tensor, target_length: torch.Tensor, window: int) -> torch.Tensor:
    """mean_b mean(out[b, :ceil(len_b / window)])  (model.py:531-540) as one kernel, lengths stay on the device."""
    return _WindowMean.apply(out, target_length, window)


# ------------------------------------------------------------------------------------------------------------
# LinearDiscriminator (reference model.py:543-583)
# ------------------------------------------------------------------------------------------------------------
SITE_LIN0 = 40


def linear_discriminator_forward(disc, windows: torch.Tensor, seed: int, prec: int) -> torch.Tensor:
    """windows (R, window*n_mel) -> (R, 1): three DiscDense layers and the final Linear as one-tap layers of _DiscFn."""
    x = windows.t().unsqueeze(0)                                   # (1, C0, R); _DiscFn makes the (R, C0) bf16 operand
    return _DiscFn.apply(disc, seed, prec, x, *_params(disc)).reshape(-1, 1)


def linear_window_plan(lengths, window: int, max_overlap: int):
    """Window starts of every sample, drawn exactly like model.py:572-577 (module-level `random`, sample by sample, the last
    window aligned to the end) -> (sample index per window, start per window, 1 / (windows of the sample * B) per window)."""
    import random
    owner, starts, wts = [], [], []
    B = len(lengths)
    for i, n in enumerate(lengths):
        n = int(n)
        mine, start = [], 0
        while start + window < n:
            mine.append(start)
            start += window - random.randint(0, max_overlap)
        mine.append(n - window)
        owner += [i] * len(mine)
        starts += mine
        wts += [1.0 / (len(mine) * B)] * len(mine)
    return owner, starts, wts


def linear_adversarial_loss(disc, inputs: torch.Tensor, target_length: torch.Tensor, seed: int, prec: int) -> torch.Tensor:
    """mean_b mean_w D(window_w of sample b) (model.py:567-583).  The reference slices, stacks and runs the MLP once per
    sample with a host sync per window test; here the lengths come to the host ONCE, the window plan goes up as three small
    pinned tensors, one gather builds all windows of the batch and the MLP runs once over the stacked rows."""
    n_mel, W = disc.n_mel_channels, disc.windows_size
    x = inputs.transpose(1, 2) if inputs.size(1) == n_mel else inputs          # (B, T, n_mel), like model.py:570-571
    lens = target_length.tolist()
    assert min(lens) >= W, "LinearDiscriminator: every sample must be at least one window long"
    owner, starts, wts = linear_window_plan(lens, W, disc.max_window_overlap)
    dev = x.device
    up = lambda v, dt: torch.tensor(v, dtype=dt).pin_memory().to(dev, non_blocking=True)
    owner_t, starts_t, wts_t = up(owner, torch.int64), up(starts, torch.int64), up(wts, torch.float32)
    t_idx = starts_t[:, None] + torch.arange(W, device=dev)[None, :]
    windows = x[owner_t[:, None], t_idx].reshape(len(owner), W * n_mel)       # one gather (and one scatter-add backward)
    out = linear_discriminator_forward(disc, windows, seed, prec)
    return (out[:, 0] * wts_t).sum()
