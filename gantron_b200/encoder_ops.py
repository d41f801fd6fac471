"""Text encoder (reference model.py:154-224) on the sm_100a kernels: the convolution bank runs on the implicit-GEMM conv +
fused BatchNorm / ReLU / Philox-dropout kernels it shares with the Postnet (postnet_ops._PostnetFn), the bidirectional LSTM
on the persistent register-resident kernel of gtn_bilstm.cu with its input projection and all weight gradients as batched
tcgen05 GEMMs.

Dropout sites: SITE_ENC0 + layer (oracle/philox.py), element numbering channels-last (b*T + t)*C + c.
"""
from __future__ import annotations

import ctypes as C

import torch
from torch.autograd.function import once_differentiable

from . import _lib
from .decoder_ops import _wgrad_tn, linear_nt
from .packing import BF16

SITE_ENC0 = 60


def bilstm_supported(lstm: torch.nn.LSTM, prec: int) -> bool:
    return (prec == 1 and lstm.hidden_size == 256 and lstm.bidirectional and lstm.num_layers == 1 and lstm.batch_first
            and lstm.bias and lstm.proj_size == 0)


def _lstm_params(lstm):
    return [lstm.weight_ih_l0, lstm.weight_hh_l0, lstm.bias_ih_l0, lstm.bias_hh_l0,
            lstm.weight_ih_l0_reverse, lstm.weight_hh_l0_reverse, lstm.bias_ih_l0_reverse, lstm.bias_hh_l0_reverse]


class _BiLstmFn(torch.autograd.Function):
    """out (B,T,2H) = BiLSTM(x (B,T,E)) on packed-sequence semantics (zero beyond lengths)."""

    @staticmethod
    def forward(ctx, x, lengths, save, wif, whf, bif, bhf, wir, whr, bir, bhr):
        B, T, E = x.shape
        H = whf.shape[1]
        dev = x.device
        w_ih = torch.cat((wif.detach().float(), wir.detach().float()), 0)                 # [8H][E]
        bias = torch.cat((bif.detach().float() + bhf.detach().float(), bir.detach().float() + bhr.detach().float()))
        x2 = x.detach().reshape(B * T, E)
        xproj = linear_nt(x2, w_ih, 1, bias=bias.contiguous())                            # [B*T][8H] fp32
        whh = torch.stack((whf.detach(), whr.detach())).to(BF16).contiguous()             # [2][4H][H]
        out = torch.empty(B, T, 2 * H, device=dev)
        p = _lib.BiLstmParams()
        p.B, p.T, p.H = B, T, H
        lens = lengths.to(device=dev, dtype=torch.int32).contiguous()
        gates = c_all = None
        if save:
            gates = torch.empty(2, T, B, H, 4, device=dev)
            c_all = torch.empty(2, T, B, H, device=dev)
            p.gates, p.c_all = gates.data_ptr(), c_all.data_ptr()
        p.xproj, p.whh, p.lengths, p.out = xproj.data_ptr(), whh.data_ptr(), lens.data_ptr(), out.data_ptr()
        _lib.check(_lib.load().gtn_bilstm_forward(C.byref(p), _lib.stream()), "gtn_bilstm_forward")
        ctx.saved = (x2, w_ih, whh, lens, out, gates, c_all)
        ctx.dims = (B, T, E, H)
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, dout):
        x2, w_ih, whh, lens, out, gates, c_all = ctx.saved
        B, T, E, H = ctx.dims
        dev = dout.device
        if gates is None:
            raise RuntimeError("gantron_b200 encoder LSTM: backward needs the forward pass to run with gradients enabled")
        dout = dout.float().contiguous()
        dgates = torch.empty(B * T, 8 * H, device=dev, dtype=BF16)
        p = _lib.BiLstmParams()
        p.B, p.T, p.H = B, T, H
        p.whh, p.lengths, p.gates, p.c_all = whh.data_ptr(), lens.data_ptr(), gates.data_ptr(), c_all.data_ptr()
        p.dout, p.dgates = dout.data_ptr(), dgates.data_ptr()
        _lib.check(_lib.load().gtn_bilstm_backward(C.byref(p), _lib.stream()), "gtn_bilstm_backward")
        # batched tcgen05 GEMMs over the B*T rows
        dx = linear_nt(dgates, w_ih.t().contiguous(), 1).reshape(B, T, E) if ctx.needs_input_grad[0] else None
        g_wih = torch.empty(8 * H, E, device=dev)
        _wgrad_tn(dgates, x2, g_wih, 1)
        g_b = dgates.sum(0, dtype=torch.float32)
        # recurrent weights: d W_hh = sum d gates(t) h(previous step of that direction)^T; h beyond a length is zero already
        hb = out.to(BF16)
        hp_f = torch.zeros(B, T, H, device=dev, dtype=BF16)
        hp_r = torch.zeros(B, T, H, device=dev, dtype=BF16)
        hp_f[:, 1:] = hb[:, :-1, :H]
        hp_r[:, :-1] = hb[:, 1:, H:]
        g_whf = torch.empty(4 * H, H, device=dev)
        g_whr = torch.empty(4 * H, H, device=dev)
        _wgrad_tn(dgates[:, :4 * H], hp_f.reshape(B * T, H), g_whf, 1)
        _wgrad_tn(dgates[:, 4 * H:], hp_r.reshape(B * T, H), g_whr, 1)
        ctx.saved = None
        return (dx, None, None, g_wih[:4 * H].contiguous(), g_whf, g_b[:4 * H].contiguous(), g_b[:4 * H].clone(),
                g_wih[4 * H:].contiguous(), g_whr, g_b[4 * H:].contiguous(), g_b[4 * H:].clone())


def bilstm_forward(lstm: torch.nn.LSTM, x: torch.Tensor, lengths: torch.Tensor) -> torch.Tensor:
    return _BiLstmFn.apply(x, lengths, torch.is_grad_enabled(), *_lstm_params(lstm))
