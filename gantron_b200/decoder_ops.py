"""Host drivers of the persistent decoder kernels: weight packing, workspace allocation, launches.

These mirror what `Decoder.forward` / `Decoder.inference` of the reference do around their python
loops (model.py:419-497) -- state initialisation (model.py:296-327), go frame (model.py:281-294),
prenet over all frames (model.py:438), output re-layout (model.py:349-376) -- but the loop itself is
one cooperative kernel launch.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Dict, Optional, Tuple

import torch

from . import _lib, ops, packing
from .packing import BF16, pad_to

SITE_PRENET1, SITE_PRENET2, SITE_ATT_H, SITE_DEC_H = 1, 2, 3, 4


@dataclass
class DecoderDims:
    n_mel: int
    prenet: int
    att_rnn: int
    dec_rnn: int
    att_dim: int
    loc_filters: int
    loc_kernel: int
    enc_dim: int        # time-varying columns of `memory` (encoder outputs)
    d_mem: int          # all columns (encoder + speaker + emotion + noise)

    @property
    def Pp(self): return pad_to(self.prenet, 32)
    @property
    def Hap(self): return pad_to(self.att_rnn, 32)
    @property
    def Hdp(self): return pad_to(self.dec_rnn, 32)
    @property
    def Dp(self): return pad_to(self.d_mem, 32)
    @property
    def NMp(self): return pad_to(self.n_mel, 32)


class PackedDecoder:
    """bf16 / plane-expanded copies of the decoder parameters in the layouts the kernels stream.

    Built from the module's fp32 nn.Parameters (state_dict names as in the reference);
    rebuilt whenever a parameter's version counter changes (i.e. after each optimizer step)."""

    def __init__(self, dec: torch.nn.Module, dims: DecoderDims, prec: int):
        self.dims, self.prec = dims, prec
        self._versions = None
        self.refresh(dec)

    @staticmethod
    def _params(dec):
        al = dec.attention_layer
        return dict(
            w_ih_a=dec.attention_rnn.weight_ih, w_hh_a=dec.attention_rnn.weight_hh,
            b_ih_a=dec.attention_rnn.bias_ih, b_hh_a=dec.attention_rnn.bias_hh,
            w_ih_d=dec.decoder_rnn.weight_ih, w_hh_d=dec.decoder_rnn.weight_hh,
            b_ih_d=dec.decoder_rnn.bias_ih, b_hh_d=dec.decoder_rnn.bias_hh,
            w_proj=dec.linear_projection.linear_layer.weight, b_proj=dec.linear_projection.linear_layer.bias,
            w_gate=dec.gate_layer.linear_layer.weight, b_gate=dec.gate_layer.linear_layer.bias,
            w_pre1=dec.prenet.layers[0].linear_layer.weight, w_pre2=dec.prenet.layers[1].linear_layer.weight,
            wq=al.query_layer.linear_layer.weight, wmem=al.memory_layer.linear_layer.weight,
            v=al.v.linear_layer.weight, wconv=al.location_layer.location_conv.conv.weight,
            wdense=al.location_layer.location_dense.linear_layer.weight)

    def stale(self, dec) -> bool:
        return self._versions != tuple(p._version for p in self._params(dec).values())

    @torch.no_grad()
    def refresh(self, dec) -> None:
        d, prec = self.dims, self.prec
        P = self._params(dec)
        self._versions = tuple(p._version for p in P.values())
        f = lambda t: t.detach().float()
        wa, wha = f(P["w_ih_a"]), f(P["w_hh_a"])
        self.w_att = packing.pack_weight([(wa[:, :d.prenet], d.Pp), (wa[:, d.prenet:], d.Dp), (wha, d.Hap)], prec)
        wd, whd = f(P["w_ih_d"]), f(P["w_hh_d"])
        self.w_dec = packing.pack_weight([(wd[:, :d.att_rnn], d.Hap), (wd[:, d.att_rnn:], d.Dp), (whd, d.Hdp)], prec)
        wpg = torch.cat((f(P["w_proj"]), f(P["w_gate"])), 0)           # rows 0..n_mel-1 mel, n_mel = gate
        self.w_proj = packing.pack_weight([(wpg[:, :d.dec_rnn], d.Hdp), (wpg[:, d.dec_rnn:], d.Dp)], prec,
                                          row_pad=pad_to(d.n_mel + 1, 16))
        self.b_proj = torch.cat((f(P["b_proj"]), f(P["b_gate"]))).contiguous()
        self.w_pre1 = packing.pack_weight([(f(P["w_pre1"]), d.NMp)], prec)
        self.w_pre2 = packing.pack_weight([(f(P["w_pre2"]), d.Pp)], prec)
        self.b_att = (f(P["b_ih_a"]) + f(P["b_hh_a"])).contiguous()
        self.b_dec = (f(P["b_ih_d"]) + f(P["b_hh_d"])).contiguous()
        self.wq = f(P["wq"]).contiguous()
        self.v = f(P["v"]).reshape(-1).contiguous()
        self.wconv = f(P["wconv"]).contiguous()
        self.wdense = f(P["wdense"]).contiguous()
        self.wmem = f(P["wmem"]).contiguous()


def linear_nt(x: torch.Tensor, w: torch.Tensor, prec: int, bias: Optional[torch.Tensor] = None,
              out_dtype=torch.float32) -> torch.Tensor:
    """y = x @ w.T (+ bias) on the tcgen05 GEMM.  x (M,K) fp32/bf16, w (N,K) fp32."""
    M, K = x.shape
    N = w.shape[0]
    kp = pad_to(K, 8) if prec == 1 else pad_to(K, 8)
    xa = packing.expand(x.float(), kp, prec, False)
    wa = packing.expand(w.float(), kp, prec, True)
    out = torch.empty(M, N, device=x.device, dtype=out_dtype)
    ops.gemm(xa, wa, out, bias_n=bias)
    return out


class FwdResult:
    pass


def _fill_common(p: _lib.DecoderFwdParams, d: DecoderDims, pk: PackedDecoder, B: int, Tin: int, n_steps: int,
                 E: int, prec: int) -> None:
    p.B, p.Tin, p.ldT, p.n_steps = B, Tin, pad_to(Tin, 4), n_steps
    p.E, p.Cc, p.D, p.Dp = E, d.d_mem - E, d.d_mem, d.Dp
    p.Ha, p.Hap, p.Hd, p.Hdp = d.att_rnn, d.Hap, d.dec_rnn, d.Hdp
    p.P, p.Pp, p.A, p.F, p.KS, p.NM, p.NMp = d.prenet, d.Pp, d.att_dim, d.loc_filters, d.loc_kernel, d.n_mel, d.NMp
    p.prec = prec
    for name in ("w_att", "w_dec", "w_proj", "w_pre1", "w_pre2", "b_att", "b_dec", "b_proj", "wq", "v", "wconv", "wdense"):
        setattr(p, name, getattr(pk, name).data_ptr())


def _alloc_state(p: _lib.DecoderFwdParams, d: DecoderDims, B: int, mult: int, dev) -> Dict[str, torch.Tensor]:
    z = lambda *s, dt=torch.float32: torch.zeros(*s, device=dev, dtype=dt)
    st = dict(
        xh_a=z(2, B, mult * d.Hap, dt=BF16), xh_d=z(2, B, mult * d.Hdp, dt=BF16), xctx=z(B, mult * d.Dp, dt=BF16),
        xmel=z(B, mult * d.NMp, dt=BF16), xp1=z(B, mult * d.Pp, dt=BF16), xp2=z(B, mult * d.Pp, dt=BF16),
        c_a=z(B, d.att_rnn), c_d=z(B, d.dec_rnn), w_prev=z(B, p.ldT), w_cum=z(B, p.ldT),
        qpart=z(d.att_rnn // 8, B, d.att_dim), barrier=z(4, dt=torch.int32), done=z(B, dt=torch.int32),
        lengths_out=z(B, dt=torch.int32), stop=z(4, dt=torch.int32), steps_done=z(4, dt=torch.int32))
    for k, v in st.items():
        setattr(p, k, v.data_ptr())
    return st


def split_memory(memory: torch.Tensor, enc_dim: int, const_tail: Optional[bool]) -> int:
    """How many leading columns of `memory` vary over encoder time.  The noise / speaker / emotion columns
    are `repeat`ed over T_in by the reference (model.py:275,469,658-663,707-712); if the caller passes a
    style that is NOT constant over T_in we treat every column as time-varying."""
    if memory.shape[2] == enc_dim:
        return enc_dim
    if const_tail is None:
        tail = memory[:, :, enc_dim:]
        const_tail = bool((tail == tail[:, :1]).all())
    return enc_dim if const_tail else memory.shape[2]


@torch.no_grad()
def decoder_inference(pk: PackedDecoder, memory: torch.Tensor, *, max_steps: int, gate_threshold: float,
                      seed: int, prec: int, const_tail: Optional[bool] = None, ignore_stop: bool = False,
                      mem_dtype: Optional[torch.dtype] = None):
    """Autoregressive decode (model.py:456-497) of `memory` (B,T_in,D) that already carries the style columns.
    Returns mel (B,n_mel,T'), gate (B,T',1), alignments (B,T',T_in), lengths (B,) int32, steps."""
    d = pk.dims
    B, Tin, D = memory.shape
    assert D == d.d_mem, f"memory has {D} columns, decoder expects {d.d_mem}"
    dev = memory.device
    mult = packing.mult_of(prec)
    E = split_memory(memory, d.enc_dim, const_tail)
    mdt = mem_dtype or (torch.float32 if prec == 3 else BF16)
    p = _lib.DecoderFwdParams()
    _fill_common(p, d, pk, B, Tin, max_steps, E, prec)
    p.train, p.use_dropout, p.mem_bf16, p.ignore_stop = 0, 0, int(mdt == BF16), int(ignore_stop)
    p.p_att = p.p_dec = 0.0
    p.gate_threshold = gate_threshold
    p.seed = seed
    st = _alloc_state(p, d, B, mult, dev)
    memory = memory.float()
    pm = linear_nt(memory.reshape(B * Tin, D), pk.wmem, prec).reshape(B, Tin, d.att_dim)
    mem_enc = memory[:, :, :E].to(mdt).contiguous()
    mem_const = memory[:, 0, E:].contiguous() if E < D else torch.zeros(B, 1, device=dev)
    pm = pm.to(mdt).contiguous()
    align = torch.zeros(B, max_steps, Tin, device=dev)
    mel = torch.zeros(B, d.n_mel, max_steps, device=dev)
    gate = torch.zeros(B, max_steps, device=dev)
    p.mem_enc, p.pm, p.mem_const = mem_enc.data_ptr(), pm.data_ptr(), mem_const.data_ptr()
    p.lengths = None
    p.align_out, p.mel_out, p.gate_out = align.data_ptr(), mel.data_ptr(), gate.data_ptr()
    _lib.check(_lib.load().gtn_decoder_forward(C.byref(p), _lib.stream()), "gtn_decoder_forward")
    steps = int(st["steps_done"][0].item())          # the single host sync of the whole decode
    lengths = st["lengths_out"].clone()
    lengths[st["done"] == 0] = steps
    return (mel[:, :, :steps], gate[:, :steps].unsqueeze(-1), align[:, :steps], lengths, steps)
