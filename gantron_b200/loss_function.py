"""Drop-in `Tacotron2Loss` (reference loss_function.py:6-22) with the three reductions fused into one kernel.

    mel_loss  = MSE(mel_out, mel_target) + MSE(mel_out_postnet, mel_target)     (means over ALL elements, padding included)
    gate_loss = BCEWithLogits(gate_out, gate_target)

One forward launch produces the three sums; one backward launch writes dmel, dmel_postnet, dgate.
`attention_loss` (loss_function.py:25-52, the guided-attention term; SURVEY 8 "next" row f3) is one kernel over
the whole (B,T,T_in) alignment tensor instead of the reference's python loop over batch x characters of tiny GPU
launches, and its lengths stay on the device (no host sync).
"""
from __future__ import annotations

import ctypes as C

import torch
from torch import nn
from torch.autograd.function import once_differentiable

from . import _lib


def _strides9(mel, post, tgt):
    vals = []
    for t in (mel, post, tgt):
        vals += [t.stride(0), t.stride(1), t.stride(2)]
    return (C.c_longlong * 9)(*vals)


class _TacoLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mel, post, gate, mel_tgt, gate_tgt):
        for t in (mel, post, gate, mel_tgt, gate_tgt):
            if not t.is_cuda:
                raise RuntimeError("gantron_b200.Tacotron2Loss needs CUDA tensors (no CPU fallback)")
        mel, post, gate = mel.detach().float(), post.detach().float(), gate.detach().float()
        mel_tgt, gate_tgt = mel_tgt.float(), gate_tgt.float()
        if gate.stride(-1) != 1:
            gate = gate.contiguous()
        if gate_tgt.stride(-1) != 1:
            gate_tgt = gate_tgt.contiguous()
        B, Cm, T = mel.shape
        sums = torch.empty(3, device=mel.device)
        s9 = _strides9(mel, post, mel_tgt)
        _lib.check(_lib.load().gtn_taco_loss_fwd(_lib.ptr(mel), _lib.ptr(post), _lib.ptr(mel_tgt), s9, B, Cm, T,
                                                 _lib.ptr(gate), gate.stride(0), _lib.ptr(gate_tgt), gate_tgt.stride(0),
                                                 _lib.ptr(sums), _lib.stream()), "gtn_taco_loss_fwd")
        n = float(B * Cm * T)
        mel_loss = (sums[0] + sums[1]) / n
        gate_loss = sums[2] / float(B * T)
        ctx.saved = (mel, post, gate, mel_tgt, gate_tgt)
        return mel_loss, gate_loss

    @staticmethod
    @once_differentiable
    def backward(ctx, g_mel, g_gate):
        mel, post, gate, mel_tgt, gate_tgt = ctx.saved
        B, Cm, T = mel.shape
        dev = mel.device
        dmel = torch.empty(B, T, Cm, device=dev)
        dpost = torch.empty(B, T, Cm, device=dev)
        dgate = torch.empty(B, T, device=dev)
        s9 = _strides9(mel, post, mel_tgt)
        gm = g_mel.reshape(1).float().contiguous()
        gg = g_gate.reshape(1).float().contiguous()
        _lib.check(_lib.load().gtn_taco_loss_bwd(_lib.ptr(mel), _lib.ptr(post), _lib.ptr(mel_tgt), s9, B, Cm, T,
                                                 _lib.ptr(gate), gate.stride(0), _lib.ptr(gate_tgt), gate_tgt.stride(0),
                                                 _lib.ptr(gm), _lib.ptr(gg), _lib.ptr(dmel), _lib.ptr(dpost),
                                                 _lib.ptr(dgate), _lib.stream()), "gtn_taco_loss_bwd")
        ctx.saved = None
        return dmel.transpose(1, 2), dpost.transpose(1, 2), dgate, None, None


class _AttnGuideFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, align, enc_len, dec_len, width):
        if not align.is_cuda:
            raise RuntimeError("gantron_b200.attention_loss needs CUDA tensors (no CPU fallback)")
        a = align.detach().float().contiguous()
        B, T, Tin = a.shape
        el = enc_len.to(device=a.device, dtype=torch.int64).contiguous()
        dl = dec_len.to(device=a.device, dtype=torch.int64).contiguous()
        loss = torch.empty(1, device=a.device)
        _lib.check(_lib.load().gtn_attn_guide_loss_fwd(_lib.ptr(a), B, T, Tin, _lib.ptr(el), _lib.ptr(dl), float(width),
                                                       _lib.ptr(loss), _lib.stream()), "gtn_attn_guide_loss_fwd")
        ctx.saved, ctx.width = (a, el, dl), float(width)
        return loss.reshape(())

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        a, el, dl = ctx.saved
        B, T, Tin = a.shape
        d = torch.empty_like(a)
        gg = g.reshape(1).float().contiguous()
        _lib.check(_lib.load().gtn_attn_guide_loss_bwd(_lib.ptr(a), B, T, Tin, _lib.ptr(el), _lib.ptr(dl), ctx.width,
                                                       _lib.ptr(gg), _lib.ptr(d), _lib.stream()), "gtn_attn_guide_loss_bwd")
        ctx.saved = None
        return d, None, None, None


def attention_loss(attention_weights, encoded_length, decoded_length, width: float = 0.5):
    """mean_b BCE(align[b].T[:enc_len_b, :dec_len_b], gaussian guide)  (loss_function.py:25-32)."""
    return _AttnGuideFn.apply(attention_weights, encoded_length, decoded_length, width)


class Tacotron2Loss(nn.Module):
    """Same call convention as the reference (loss_function.py:10): forward(model_output, targets, input_lengths,
    output_lengths) -> (mel_loss, gate_loss, attn_loss)."""

    def forward(self, model_output, targets, input_lengths, output_lengths):
        mel_target, gate_target = targets[0], targets[1]
        mel_out, mel_out_postnet, gate_out, alignments = model_output
        mel_loss, gate_loss = _TacoLossFn.apply(mel_out, mel_out_postnet, gate_out, mel_target, gate_target)
        attn = attention_loss(alignments, input_lengths, output_lengths)
        return mel_loss, gate_loss, attn
