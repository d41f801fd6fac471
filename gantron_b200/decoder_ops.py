"""Host drivers of the persistent decoder kernels: weight packing, workspace allocation, launches.

These mirror what `Decoder.forward` / `Decoder.inference` of the reference do around their python
loops (model.py:419-497) -- state initialisation (model.py:296-327), go frame (model.py:281-294),
prenet over all frames (model.py:438), output re-layout (model.py:349-376) -- but the loop itself is
one cooperative kernel launch.
"""
from __future__ import annotations

import os

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
    def Pp(self): return pad_to(self.prenet, 64)
    @property
    def Hap(self): return pad_to(self.att_rnn, 64)
    @property
    def Hdp(self): return pad_to(self.dec_rnn, 64)
    @property
    def Dp(self): return pad_to(self.d_mem, 64)
    @property
    def NMp(self): return pad_to(self.n_mel, 64)


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

    @staticmethod
    def _key(dec):
        # version counters catch optimizer steps / in-place copies; data_ptr catches parameters that were re-pointed
        # (`p.data = ...`, module.to(), a deep copy of the module)
        return tuple((p._version, p.data_ptr()) for p in PackedDecoder._params(dec).values())

    def stale(self, dec) -> bool:
        return self._versions != self._key(dec)

    @torch.no_grad()
    def refresh(self, dec) -> None:
        d, prec = self.dims, self.prec
        P = self._params(dec)
        self._versions = self._key(dec)
        f = lambda t: t.detach().float()
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
        # location conv (F,2,KS) and dense (A,F) pre-multiplied: U[a][ch*KS + k] = sum_c dense[a][c] conv[c][ch][k]
        U = self.wdense @ self.wconv.reshape(d.loc_filters, 2 * d.loc_kernel)
        Up = torch.zeros(d.att_dim, 64, device=U.device)
        Up[:, :2 * d.loc_kernel] = U
        hi = Up.to(BF16)
        self.u_loc = torch.cat((hi, (Up - hi.float()).to(BF16)), 1).contiguous()      # [A][hi 64 | lo 64]
        self.wq_t = self.wq.t().contiguous()                                           # [Ha][A]


def linear_nt(x: torch.Tensor, w: torch.Tensor, prec: int, bias: Optional[torch.Tensor] = None,
              out_dtype=torch.float32) -> torch.Tensor:
    """y = x @ w.T (+ bias) on the tcgen05 GEMM.  x (M,K) fp32/bf16, w (N,K) fp32."""
    M, K = x.shape
    N = w.shape[0]
    kp = pad_to(K, 8) if prec == 1 else pad_to(K, 8)
    xa = packing.expand(x, kp, prec, False)          # (a bf16 x stays bf16 in bf16 mode)
    wa = packing.expand(w.float(), kp, prec, True)
    out = torch.empty(M, N, device=x.device, dtype=out_dtype)
    ops.gemm(xa, wa, out, bias_n=bias)
    return out


class PackV2:
    """Weight packs of the second-generation kernels for a given split of `memory` into E time-varying and
    D - E time-constant columns.  The constant columns never enter a per-frame GEMM: their products with the
    (per-sequence constant) context are folded into per-sequence biases (`seq_biases`)."""

    def __init__(self, dec, pk: "PackedDecoder", E: int):
        d, prec = pk.dims, pk.prec
        P = PackedDecoder._params(dec)
        f = lambda t: t.detach().float()
        self.E, self.Ep, self.versions = E, pad_to(E, 64), pk._versions
        Pn, Ha, Hd = d.prenet, d.att_rnn, d.dec_rnn
        wa, wha, wd, whd = f(P["w_ih_a"]), f(P["w_hh_a"]), f(P["w_ih_d"]), f(P["w_hh_d"])
        wpg = torch.cat((f(P["w_proj"]), f(P["w_gate"])), 0)
        self.w_att2 = _tile_rows(packing.pack_weight([(wa[:, :Pn], d.Pp), (wa[:, Pn:Pn + E], self.Ep), (wha, d.Hap)], prec), Ha)
        self.w_dec2 = _tile_rows(packing.pack_weight([(wd[:, :Ha], d.Hap), (wd[:, Ha:Ha + E], self.Ep), (whd, d.Hdp)], prec), Hd)
        self.w_proj2 = packing.pack_weight([(wpg[:, :Hd], d.Hdp), (wpg[:, Hd:Hd + E], self.Ep)], prec)
        self.wa_const = wa[:, Pn + E:].contiguous()
        self.wd_const = wd[:, Ha + E:].contiguous()
        self.wp_const = wpg[:, Hd + E:].contiguous()
        self.b_att, self.b_dec, self.b_proj = pk.b_att, pk.b_dec, pk.b_proj

    def backward_packs(self, dec, pk: "PackedDecoder"):
        """Transposed packs of the BPTT kernel (rows = LSTM input features without the prenet rows and without the
        time-constant context columns; K = gate rows g*H + u) and U^T for d[prev | cum] = dth . U."""
        if getattr(self, "_bwd", None) is None:
            d, prec = pk.dims, pk.prec
            P = PackedDecoder._params(dec)
            f = lambda t: t.detach().float()
            E, Pn, Ha, Hd = self.E, d.prenet, d.att_rnn, d.dec_rnn
            wa, wha, wd, whd = f(P["w_ih_a"]), f(P["w_hh_a"]), f(P["w_ih_d"]), f(P["w_hh_d"])
            wta = torch.cat((wa[:, Pn:Pn + E], wha), 1).t().contiguous()                 # [E + Ha][4Ha]
            wtd = torch.cat((wd[:, :Ha + E], whd), 1).t().contiguous()                   # [Ha + E + Hd][4Hd]
            U = pk.wdense @ pk.wconv.reshape(d.loc_filters, 2 * d.loc_kernel)            # [A][2KS]
            UT = torch.zeros(64, 128, device=U.device)
            UT[:2 * d.loc_kernel, :d.att_dim] = U.t()
            hi = UT.to(BF16)
            self._bwd = dict(
                wT_att2=packing.pack_weight([(wta, 4 * Ha)], prec), wT_dec2=packing.pack_weight([(wtd, 4 * Hd)], prec),
                ut_loc=torch.cat((hi, (UT - hi.float()).to(BF16)), 1).contiguous(),       # [64][hi 128 | lo 128]
                wa_pre_T=wa[:, :Pn].t().contiguous(),                                     # [P][4Ha]
                wq_au=(pk.wq.to(BF16) if prec == 1 else pk.wq.float()).contiguous(),      # [A][Ha] for the cell backward
                wa_const_T=self.wa_const.t().contiguous(), wd_const_T=self.wd_const.t().contiguous())
        return self._bwd

    def seq_biases(self, mem_const: torch.Tensor, prec: int, B: int, want_proj: bool):
        """gbias_a / gbias_d [B][H][4] and pbias [B][n_mel+1]: bias + W[:, const columns] . mem_const[b]."""
        def one(wc, bias):
            if wc.shape[1] == 0:
                return bias[None, :].expand(B, -1)
            return linear_nt(mem_const, wc, prec, bias=bias)
        ga = one(self.wa_const, self.b_att)
        gd = one(self.wd_const, self.b_dec)
        ga = ga.reshape(B, 4, -1).permute(0, 2, 1).contiguous()
        gd = gd.reshape(B, 4, -1).permute(0, 2, 1).contiguous()
        pb = one(self.wp_const, self.b_proj).contiguous() if want_proj else None
        return ga, gd, pb


def pack_v2(dec, pk: "PackedDecoder", E: int) -> PackV2:
    cache = getattr(pk, "_v2", None)
    if cache is None or cache.get("versions") != pk._versions:
        cache = {"versions": pk._versions}
        pk._v2 = cache
    if E not in cache:
        cache[E] = PackV2(dec, pk, E)
    return cache[E]


def _tile_rows(w: torch.Tensor, H: int) -> torch.Tensor:
    K = w.shape[1]
    return w.view(4, H // 32, 32, K).permute(1, 0, 2, 3).reshape(4 * H, K).contiguous()


PROF_BUFFER_BWD = None
KERNEL_EVENTS = None     # measurement hook (bench.py): dict name -> [(start_event, end_event)] around the persistent kernels


def _timed_call(name, fn):
    """Run fn(); when KERNEL_EVENTS is set, bracket it with CUDA events on the launching (current) stream."""
    if KERNEL_EVENTS is None:
        return fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    r = fn()
    e1.record()
    KERNEL_EVENTS.setdefault(name, []).append((e0, e1))
    return r
PROF_BUFFER = None     # dev tool: int64 [16*32] device tensor receiving clock64() stamps of CTA 0 (tools/time_infer.py)


def _fill_common(p: _lib.DecoderFwdParams, d: DecoderDims, pk: PackedDecoder, B: int, Tin: int, n_steps: int,
                 E: int, prec: int) -> None:
    p.B, p.Tin, p.ldT, p.n_steps = B, Tin, pad_to(Tin, 4), n_steps
    p.E, p.Cc, p.D, p.Dp = E, d.d_mem - E, d.d_mem, d.Dp
    p.Ha, p.Hap, p.Hd, p.Hdp = d.att_rnn, d.Hap, d.dec_rnn, d.Hdp
    p.P, p.Pp, p.A, p.F, p.KS, p.NM, p.NMp = d.prenet, d.Pp, d.att_dim, d.loc_filters, d.loc_kernel, d.n_mel, d.NMp
    p.prec = prec
    p.prof = PROF_BUFFER.data_ptr() if PROF_BUFFER is not None else None
    for name in ("w_pre1", "w_pre2", "b_att", "b_dec", "b_proj", "wq", "v", "wconv", "wdense", "u_loc", "wq_t"):
        setattr(p, name, getattr(pk, name).data_ptr())


def _fill_v2(p: _lib.DecoderFwdParams, v2: PackV2, mem_const: torch.Tensor, prec: int, B: int, want_proj: bool):
    """Second-generation packs + per-sequence biases; returns the tensors that must stay alive during the launch."""
    ga, gd, pb = v2.seq_biases(mem_const, prec, B, want_proj)
    p.w_att2, p.w_dec2, p.Ep = v2.w_att2.data_ptr(), v2.w_dec2.data_ptr(), v2.Ep
    p.w_proj = v2.w_proj2.data_ptr()
    p.gbias_a, p.gbias_d = ga.data_ptr(), gd.data_ptr()
    p.pbias = pb.data_ptr() if pb is not None else None
    return ga, gd, pb


def _alloc_state(p: _lib.DecoderFwdParams, d: DecoderDims, B: int, mult: int, dev, ctx_cols: Optional[int] = None
                 ) -> Dict[str, torch.Tensor]:
    z = lambda *s, dt=torch.float32: torch.zeros(*s, device=dev, dtype=dt)
    st = dict(
        xh_a=z(2, B, mult * d.Hap, dt=BF16), xh_d=z(2, B, mult * d.Hdp, dt=BF16),
        xctx=z(2, B, mult * (ctx_cols or d.Dp), dt=BF16),
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


def _infer1_packs(dec, pk: "PackedDecoder", E: int, ncta: int):
    """Utterance-independent operands of the batch-1 decode kernel (gtn_decoder_infer1.cu), cached per weight version."""
    cache = getattr(pk, "_i1", None)
    key = (pk._versions, E, ncta)
    if cache is not None and cache["key"] == key:
        return cache
    d = pk.dims
    P = PackedDecoder._params(dec)
    f = lambda t: t.detach().float()
    Pn, Ha, Hd = d.prenet, d.att_rnn, d.dec_rnn
    wa, wha, wd, whd = f(P["w_ih_a"]), f(P["w_hh_a"]), f(P["w_ih_d"]), f(P["w_hh_d"])
    wpg = torch.cat((f(P["w_proj"]), f(P["w_gate"])), 0)
    nu = (max(Ha, Hd) + ncta - 1) // ncta
    dev = wa.device
    # gate rows of CTA c in tensor-memory order: rho = gate * nu + local unit (32 lanes, unused ones point at a zero row)
    def row_index(H):
        idx = torch.full((ncta, 32), 4 * H, dtype=torch.long, device=dev)
        c = torch.arange(ncta, device=dev)[:, None, None]
        g = torch.arange(4, device=dev)[None, :, None]
        u = torch.arange(nu, device=dev)[None, None, :]
        unit = c * nu + u
        rows = torch.where(unit < H, g * H + unit, torch.full_like(unit + g, 4 * H))
        idx[:, :4 * nu] = rows.reshape(ncta, 4 * nu)
        return idx.reshape(-1)
    nm1 = pad_to(d.n_mel, 8)
    w1 = torch.zeros(Pn, nm1, device=dev)
    w1[:, :d.n_mel] = f(P["w_pre1"])
    U = torch.zeros(128, 64, device=dev)
    U[:d.att_dim, :2 * d.loc_kernel] = pk.wdense @ pk.wconv.reshape(d.loc_filters, 2 * d.loc_kernel)
    cache = dict(key=key, nu=nu, idx_a=row_index(Ha), idx_d=row_index(Hd),
                 wa_pre=wa[:, :Pn].to(BF16), wa_ctx=wa[:, Pn:Pn + E].contiguous(), wha=wha.to(BF16),
                 wd_h=wd[:, :Ha].to(BF16), wd_ctx=wd[:, Ha:Ha + E].contiguous(), whd=whd.to(BF16),
                 wp_h=wpg[:, :Hd].to(BF16).float().contiguous(), wp_ctx=wpg[:, Hd:Hd + E].contiguous(),
                 w1=w1.to(BF16).contiguous(), w2=f(P["w_pre2"]).to(BF16).contiguous(), U=U.contiguous(), nm1=nm1)
    pk._i1 = cache
    return cache


def infer1_supported(d: "DecoderDims", B: int, Tin: int, prec: int) -> bool:
    """Whether the batch-1 on-chip-weights kernel covers this call (else the clustered kernel runs it)."""
    if B != 1 or prec != 1 or os.environ.get("GANTRON_B200_INFER1", "1") == "0":
        return False
    ncta = _lib.load().gtn_decoder_infer1_ctas()
    if ncta <= 0:
        return False
    Tp = pad_to(Tin, 64)
    cols_a = pad_to(d.prenet + Tp + d.att_rnn, 256) // 8
    cols_d = pad_to(d.att_rnn + Tp + d.dec_rnn, 256) // 8
    nu = (max(d.att_rnn, d.dec_rnn) + ncta - 1) // ncta
    smem = d.prenet * d.prenet * 2 + 4 * (128 * 64 + 8 * (cols_a + cols_d) + 4 * Tp + 4096)
    return (nu <= 8 and cols_a + cols_d <= 512 and Tin <= min(2 * ncta, 512) and d.prenet % 64 == 0 and 2 * d.prenet <= 512
            and d.n_mel + 1 <= min(96, ncta) and d.att_dim <= 128 and d.loc_kernel % 2 == 1 and 2 * d.loc_kernel <= 64
            and d.att_rnn % 64 == 0 and d.dec_rnn % 64 == 0 and smem <= 226 * 1024)


def _decoder_inference_b1(dec, pk, memory, E, max_steps, gate_threshold, seed, ignore_stop):
    """Batch 1: every LSTM weight on chip (tensor memory) for the whole utterance; see gtn_decoder_infer1.cu."""
    d = pk.dims
    lib = _lib.load()
    ncta = lib.gtn_decoder_infer1_ctas()
    _, Tin, D = memory.shape
    dev = memory.device
    w = _infer1_packs(dec, pk, E, ncta)
    v2 = pack_v2(dec, pk, E)
    Pn, Ha, Hd, A, n_out = d.prenet, d.att_rnn, d.dec_rnn, d.att_dim, d.n_mel + 1
    Tp = pad_to(Tin, 64)
    Ka, Kd = pad_to(Pn + Tp + Ha, 256), pad_to(Ha + Tp + Hd, 256)
    mem = memory[0].float()
    mem_enc = mem[:, :E].contiguous()
    mem_const = mem[0:1, E:].contiguous() if E < D else torch.zeros(1, 1, device=dev)
    ga, gd, pb = v2.seq_biases(mem_const[:, :D - E], 1, 1, True)
    # everything linear in the attention context, pre-multiplied with the encoder memory (bf16 operands, fp32 accumulate)
    m_a = linear_nt(w["wa_ctx"], mem_enc, 1)                     # [4Ha][Tin]
    m_d = linear_nt(w["wd_ctx"], mem_enc, 1)                     # [4Hd][Tin]
    pc = linear_nt(w["wp_ctx"], mem_enc, 1).contiguous()         # [n_mel + 1][Tin]
    pm = linear_nt(mem, pk.wmem, 1).contiguous()                 # processed_memory [Tin][A] (model.py:326)
    z = lambda *s, dt=torch.float32: torch.zeros(*s, device=dev, dtype=dt)

    def image(pre, m, hh, K, idx):
        full = z(pre.shape[0] + 1, K, dt=BF16)                   # (+ one zero row for unused lanes)
        n0, n1 = pre.shape[1], pre.shape[1] + Tp
        full[:-1, :n0] = pre
        full[:-1, n0:n0 + Tin] = m.to(BF16)
        full[:-1, n1:n1 + hh.shape[1]] = hh
        return full[idx].view(ncta, 32, 4, K // 4).permute(0, 2, 1, 3)
    wimg = torch.cat((image(w["wa_pre"], m_a, w["wha"], Ka, w["idx_a"]), image(w["wd_h"], m_d, w["whd"], Kd, w["idx_d"])),
                     3).contiguous()
    q = _lib.DecoderInfer1Params()
    q.Tin, q.Tp, q.n_steps, q.Ha, q.Hd, q.P, q.A, q.KS, q.NM, q.NMp = Tin, Tp, max_steps, Ha, Hd, Pn, A, d.loc_kernel, d.n_mel, w["nm1"]
    q.nu, q.ncta, q.cols_a, q.cols_d = w["nu"], ncta, Ka // 8, Kd // 8
    q.ignore_stop, q.gate_threshold, q.seed = int(ignore_stop), gate_threshold, seed
    ws = dict(h_att_g=z(2, Ha), h_dec_g=z(2, Hd), e_g=z(2, Tp), qacc=z(2, 128, dt=torch.int64), macc=z(2, 96, dt=torch.int64),
              barrier=z(4, dt=torch.int32), mel_out=z(d.n_mel, max_steps), gate_out=z(max_steps), align_out=z(max_steps, Tin),
              steps_done=z(4, dt=torch.int32), done=z(4, dt=torch.int32), lengths_out=z(4, dt=torch.int32))
    for k, t in ws.items():
        setattr(q, k, t.data_ptr())
    keep = dict(wimg=wimg, w1=w["w1"], w2=w["w2"], u=w["U"], wq=pk.wq, v=pk.v, wp_h=w["wp_h"], pc=pc, pbias=pb.contiguous(),
                b_att=pk.b_att, gbias_a=ga.contiguous(), gbias_d=gd.contiguous(), pm=pm)
    for k, t in keep.items():
        setattr(q, k, t.data_ptr())
    q.prof = PROF_BUFFER.data_ptr() if PROF_BUFFER is not None else None
    _lib.check(lib.gtn_decoder_infer1(C.byref(q), _lib.stream()), "gtn_decoder_infer1")
    steps = int(ws["steps_done"][0].item())          # the single host sync of the whole decode
    lengths = torch.full((1,), steps, device=dev, dtype=torch.int32)
    return (ws["mel_out"][None, :, :steps], ws["gate_out"][None, :steps, None], ws["align_out"][None, :steps], lengths, steps)


PM_CHUNK_MAJOR = 32      # gtn_decoder_fwd_params.reserved0 flag: `pm` is bf16 [B][A/8][T_in][8]


@torch.no_grad()
def decoder_inference(dec, memory: torch.Tensor, *, max_steps: int, gate_threshold: float,
                      seed: int, prec: int, const_tail: Optional[bool] = None, ignore_stop: bool = False,
                      mem_dtype: Optional[torch.dtype] = None):
    """Autoregressive decode (model.py:456-497) of `memory` (B,T_in,D) that already carries the style columns, with the
    parameters of the Decoder module `dec`.
    Returns mel (B,n_mel,T'), gate (B,T',1), alignments (B,T',T_in), lengths (B,) int32, steps."""
    pk = dec.packed()
    if pk.prec != prec:
        raise ValueError(f"decoder_inference: prec={prec} but the module's packs were built for prec={pk.prec}")
    d = pk.dims
    B, Tin, D = memory.shape
    assert D == d.d_mem, f"memory has {D} columns, decoder expects {d.d_mem}"
    dev = memory.device
    mult = packing.mult_of(prec)
    E = split_memory(memory, d.enc_dim, const_tail)
    mdt = mem_dtype or (torch.float32 if prec == 3 else BF16)
    if mdt == BF16 and infer1_supported(d, B, Tin, prec):
        return _decoder_inference_b1(dec, pk, memory, E, max_steps, gate_threshold, seed, ignore_stop)
    p = _lib.DecoderFwdParams()
    _fill_common(p, d, pk, B, Tin, max_steps, E, prec)
    p.train, p.use_dropout, p.mem_bf16, p.ignore_stop = 0, 0, int(mdt == BF16), int(ignore_stop)
    p.p_att = p.p_dec = 0.0
    p.gate_threshold = gate_threshold
    p.seed = seed
    st = _alloc_state(p, d, B, mult, dev, pad_to(E, 64))
    memory = memory.float()
    pm = linear_nt(memory.reshape(B * Tin, D), pk.wmem, prec).reshape(B, Tin, d.att_dim)
    mem_enc = memory[:, :, :E].to(mdt).contiguous()
    mem_const = memory[:, 0, E:].contiguous() if E < D else torch.zeros(B, 1, device=dev)
    pm = pm.to(mdt).contiguous()
    if mdt == BF16 and d.att_dim % 8 == 0:
        pm = pm.view(B, Tin, d.att_dim // 8, 8).permute(0, 2, 1, 3).contiguous()      # chunk-major, see decoder_train_forward
        p.reserved0 = PM_CHUNK_MAJOR
    align = torch.zeros(B, max_steps, Tin, device=dev)
    mel = torch.zeros(B, d.n_mel, max_steps, device=dev)
    gate = torch.zeros(B, max_steps, device=dev)
    p.mem_enc, p.pm, p.mem_const = mem_enc.data_ptr(), pm.data_ptr(), mem_const.data_ptr()
    p.lengths = None
    p.align_out, p.mel_out, p.gate_out = align.data_ptr(), mel.data_ptr(), gate.data_ptr()
    keep = _fill_v2(p, pack_v2(dec, pk, E), mem_const[:, :D - E], prec, B, True)   # noqa: F841 (kept alive)
    _lib.check(_lib.load().gtn_decoder_forward2(C.byref(p), _lib.stream()), "gtn_decoder_forward2")
    steps = int(st["steps_done"][0].item())          # the single host sync of the whole decode
    lengths = st["lengths_out"].clone()
    lengths[st["done"] == 0] = steps
    return (mel[:, :, :steps], gate[:, :steps].unsqueeze(-1), align[:, :steps], lengths, steps)


# ==========================================================================================
# Training: teacher-forced forward + backward through time
# ==========================================================================================
def _bf(x: torch.Tensor, kpad: int, prec: int, is_weight: bool = False) -> torch.Tensor:
    """(rows, K) fp32/bf16 -> GEMM operand (rows, mult*kpad) bf16."""
    if prec == 1 and x.dtype == BF16 and x.shape[-1] == kpad and x.stride(-1) == 1:
        return x
    return packing.expand(x.float(), kpad, prec, is_weight)


def _wgrad_tn(dy: torch.Tensor, x: torch.Tensor, out: torch.Tensor, prec: int, accumulate: bool = False) -> None:
    """out[m][n] (+)= sum_r dy[r][m] x[r][n]: deferred weight gradient over r = (frame, batch) rows."""
    if prec == 1:
        # (.to is a no-op for operands the caller already converted once)
        ops.gemm_tn(_rows8(dy.to(BF16)), _rows8(x.to(BF16)), out, accumulate=accumulate,
                    block_n=256 if out.shape[1] >= 512 else 128)
        return
    # fp32 mode: stack the plane products along K (rows): (hi,hi),(lo,hi),(lo2,hi),(hi,lo),(lo,lo),(hi,lo2)
    dh = packing.split3(dy)
    xh = packing.split3(x)
    A = _rows8(torch.cat([dh[i] for i in packing.W_PLANES], 0))
    Bm = _rows8(torch.cat([xh[i] for i in packing.X_PLANES], 0))
    ops.gemm_tn(A, Bm, out, accumulate=accumulate)


def _rows8(t: torch.Tensor) -> torch.Tensor:
    """Make the row stride a multiple of 8 elements (TMA needs 16-byte strides)."""
    t = t if t.stride(-1) == 1 else t.contiguous()
    if t.stride(0) % 8 == 0 and t.data_ptr() % 16 == 0:
        return t
    n = t.shape[1]
    buf = torch.zeros(t.shape[0], pad_to(n, 8), device=t.device, dtype=t.dtype)
    buf[:, :n] = t
    return buf[:, :n]


class _PackedBackward:
    """Transposed packs used by the backward kernel; built lazily next to PackedDecoder."""

    def __init__(self, dec, pk: PackedDecoder):
        P = PackedDecoder._params(dec)
        f = lambda t: t.detach().float()
        wpg = torch.cat((f(P["w_proj"]), f(P["w_gate"])), 0)                       # [n_mel+1][Hd+D]
        self.wpgT = wpg.t().contiguous()                                           # [Hd+D][n_mel+1]
        self.wmemT = f(P["wmem"]).t().contiguous()                                 # [D][A]
        self.w2T = f(P["w_pre2"]).t().contiguous()                                 # [P_in][P_out] -> rows = inputs
        self.versions = pk._versions


def _packed_bwd(dec, pk: PackedDecoder) -> _PackedBackward:
    pb = getattr(pk, "_bwd", None)
    if pb is None or pb.versions != pk._versions:
        pb = _PackedBackward(dec, pk)
        pk._bwd = pb
    return pb


class _PrenetFn(torch.autograd.Function):
    """Stand-alone Prenet (model.py:102-105): two tcgen05 GEMMs with fused relu + always-on Philox dropout, and the
    matching backward (y = relu(x W^T) * keep * 2  =>  d(pre-activation) = dy * 2 * [y > 0])."""

    @staticmethod
    def forward(ctx, x2, w1, w2, seed, prec):
        h1 = _prenet_layer(x2, w1, seed, SITE_PRENET1, prec)
        h2 = _prenet_layer(h1, w2, seed, SITE_PRENET2, prec)
        ctx.save_for_backward(x2, w1, w2, h1, h2)
        ctx.prec = prec
        return h2

    @staticmethod
    def backward(ctx, g):
        x2, w1, w2, h1, h2 = ctx.saved_tensors
        prec = ctx.prec
        d2 = g.float() * 2.0 * (h2 > 0)
        g_w2 = torch.empty_like(w2, dtype=torch.float32)
        _wgrad_tn(d2, h1, g_w2, prec)
        d1 = linear_nt(d2, w2.detach().float().t().contiguous(), prec) * 2.0 * (h1 > 0)
        g_w1 = torch.empty_like(w1, dtype=torch.float32)
        _wgrad_tn(d1, x2.float(), g_w1, prec)
        g_x = linear_nt(d1, w1.detach().float().t().contiguous(), prec) if ctx.needs_input_grad[0] else None
        return g_x, g_w1, g_w2, None, None


def prenet_forward(prenet_mod, x: torch.Tensor, seed: int, prec: int = 1) -> torch.Tensor:
    """Stand-alone Prenet.forward (model.py:102-105); differentiable w.r.t. the input and both weights."""
    shp = x.shape
    x2 = x.reshape(-1, shp[-1]).float()
    w1 = prenet_mod.layers[0].linear_layer.weight
    w2 = prenet_mod.layers[1].linear_layer.weight
    return _PrenetFn.apply(x2, w1, w2, seed, prec).reshape(*shp[:-1], w2.shape[0])


def _prenet_layer(x: torch.Tensor, w: torch.Tensor, seed: int, site: int, prec: int) -> torch.Tensor:
    kp = pad_to(x.shape[1], 8)
    out = torch.empty(x.shape[0], w.shape[0], device=x.device, dtype=torch.float32)
    ops.gemm(packing.expand(x.float(), kp, prec, False), packing.expand(w.detach().float(), kp, prec, True), out,
             act=ops.ACT_RELU, dropout=(0.5, seed, site, 0))
    return out


def _decoder_params_list(dec):
    P = PackedDecoder._params(dec)
    names = ["w_ih_a", "w_hh_a", "b_ih_a", "b_hh_a", "w_ih_d", "w_hh_d", "b_ih_d", "b_hh_d", "w_proj", "b_proj",
             "w_gate", "b_gate", "w_pre1", "w_pre2", "wq", "wmem", "v", "wconv", "wdense"]
    return names, [P[n] for n in names]


class _DecoderTrainFn(torch.autograd.Function):
    """Teacher-forced decoder pass as ONE autograd node (the reference builds ~22k nodes)."""

    @staticmethod
    def forward(ctx, dec, seed, memory, mels, memory_lengths, *params):
        pk = dec.packed()
        d, prec = pk.dims, pk.prec
        mult = packing.mult_of(prec)
        B, Tin, D = memory.shape
        T = mels.shape[2]
        dev = memory.device
        training = dec.training
        E = d.enc_dim if D > d.enc_dim else D
        mdt = torch.float32 if prec == 3 else BF16
        memory_f = memory.detach().float().contiguous()

        # ---- time-independent GEMMs: prenet over all frames (model.py:435-438), memory layer (model.py:326)
        frames = torch.cat((torch.zeros(1, B, d.n_mel, device=dev), mels.detach().float().permute(2, 0, 1)), 0)
        frames2 = frames.reshape((T + 1) * B, d.n_mel)
        h1 = _prenet_layer(frames2, params[12], seed, SITE_PRENET1, prec)          # fp32 [(T+1)B][P]
        pre = _prenet_layer(h1, params[13], seed, SITE_PRENET2, prec)              # fp32 [(T+1)B][P]
        pre_x = packing.expand(pre, d.Pp, prec, False)                             # kernel operand
        pm = linear_nt(memory_f.reshape(B * Tin, D), pk.wmem, prec).reshape(B, Tin, d.att_dim)

        p = _lib.DecoderFwdParams()
        _fill_common(p, d, pk, B, Tin, T, E, prec)
        p.train, p.use_dropout, p.mem_bf16, p.ignore_stop = 1, int(training), int(mdt == BF16), 0
        p.p_att, p.p_dec = dec.p_attention_dropout, dec.p_decoder_dropout
        p.gate_threshold = dec.gate_threshold
        p.seed = seed
        st = _alloc_state(p, d, B, mult, dev, pad_to(E, 64))
        mem_enc = memory_f[:, :, :E].to(mdt).contiguous()
        mem_const = memory_f[:, 0, E:].contiguous() if E < D else torch.zeros(B, 1, device=dev)
        pm_k = pm.to(mdt).contiguous()
        if mdt == BF16 and d.att_dim % 8 == 0:
            # chunk-major processed_memory [B][A / 8][T_in][8] (flag PM_CHUNK_MAJOR): an attention thread owns one encoder
            # position, so its 16-byte loads are contiguous across a warp in this order (32 different rows otherwise)
            pm_k = pm_k.view(B, Tin, d.att_dim // 8, 8).permute(0, 2, 1, 3).contiguous()
            p.reserved0 = PM_CHUNK_MAJOR
        lengths = memory_lengths.to(device=dev, dtype=torch.int32).contiguous()
        z = lambda *s: torch.zeros(*s, device=dev)
        e = lambda *s: torch.empty(*s, device=dev)
        sv = dict(
            # (rows 1 .. T of the three state buffers are written by the kernel for every (frame, sequence, column): only row 0,
            # the zero initial state, is cleared -- three 210 MB fills per step otherwise)
            gates_a=e(T, B, 4 * d.att_rnn), c_a_all=e(T, B, d.att_rnn), h_a_buf=e(T + 1, B, d.att_rnn),
            gates_d=e(T, B, 4 * d.dec_rnn), c_d_all=e(T, B, d.dec_rnn), h_d_buf=e(T + 1, B, d.dec_rnn),
            ctx_buf=e(T + 1, B, D), q_all=e(T, B, d.att_dim), th_all=e(T, B, Tin, d.att_dim),
            f_all=e(1), cum_all=e(T, B, Tin), align=z(B, T, Tin))
        for k in ("h_a_buf", "h_d_buf", "ctx_buf"):
            sv[k][0].zero_()
        p.mem_enc, p.pm, p.mem_const, p.lengths = mem_enc.data_ptr(), pm_k.data_ptr(), mem_const.data_ptr(), lengths.data_ptr()
        p.pre_all = pre_x.data_ptr()
        p.align_out = sv["align"].data_ptr()
        p.gates_a, p.c_a_all, p.h_a_all = sv["gates_a"].data_ptr(), sv["c_a_all"].data_ptr(), sv["h_a_buf"][1:].data_ptr()
        p.gates_d, p.c_d_all, p.h_d_all = sv["gates_d"].data_ptr(), sv["c_d_all"].data_ptr(), sv["h_d_buf"][1:].data_ptr()
        p.ctx_all = sv["ctx_buf"][1:].data_ptr()
        p.q_all, p.th_all, p.f_all, p.cum_all = (sv["q_all"].data_ptr(), sv["th_all"].data_ptr(), sv["f_all"].data_ptr(),
                                                 sv["cum_all"].data_ptr())
        keep = _fill_v2(p, pack_v2(dec, pk, E), mem_const[:, :D - E], prec, B, False)   # noqa: F841 (kept alive)
        _lib.check(_timed_call("decoder_fwd3_kernel", lambda: _lib.load().gtn_decoder_forward2(C.byref(p), _lib.stream())),
                   "gtn_decoder_forward2")

        # ---- deferred mel/gate projection over all frames (model.py:411-416): batched over b so the output is
        # written directly as [B][T][n_mel+1] (channels-last); the API tensors are views of it
        n_out = d.n_mel + 1
        # bf16 mode: the hidden-state and context saves are converted ONCE here, for the projection now and for the deferred
        # weight-gradient GEMMs of the backward pass
        hd_b = sv["h_d_buf"].to(BF16) if prec == 1 else None
        cx_b = sv["ctx_buf"].to(BF16) if prec == 1 else None
        hd = (hd_b if prec == 1 else sv["h_d_buf"])[1:]            # [T][B][Hd]
        cx = (cx_b if prec == 1 else sv["ctx_buf"])[1:]            # [T][B][D]
        out_cl = torch.empty(B, T, n_out, device=dev)
        wpg = torch.cat((params[8].detach().float(), params[10].detach().float()), 0)
        bpg = pk.b_proj
        _proj_gemm(hd, wpg[:, :d.dec_rnn], out_cl, prec, bias=bpg, accumulate=False)
        _proj_gemm(cx, wpg[:, d.dec_rnn:], out_cl, prec, bias=None, accumulate=True)
        mel = out_cl[:, :, :d.n_mel].transpose(1, 2)      # (B,n_mel,T) view, like the reference's transpose
        gate = out_cl[:, :, d.n_mel]                      # (B,T)
        align = sv["align"]

        ctx.dec, ctx.seed, ctx.training = dec, seed, training
        ctx.dims = (B, Tin, D, T, E, prec)
        ctx.saved = dict(sv=sv, frames2=frames2, h1=h1, pre=pre, memory_f=memory_f, mem_enc=mem_enc, mem_const=mem_const,
                         lengths=lengths, out_cl=out_cl, hd_b=hd_b, cx_b=cx_b)
        return mel, gate, align

    @staticmethod
    def backward(ctx, d_mel, d_gate, d_align):
        dec = ctx.dec
        pk = dec.packed()
        pb = _packed_bwd(dec, pk)
        d = pk.dims
        B, Tin, D, T, E, prec = ctx.dims
        mult = packing.mult_of(prec)
        S = ctx.saved
        sv = S["sv"]
        dev = sv["align"].device
        n_out = d.n_mel + 1
        z = lambda *s, dt=torch.float32: torch.zeros(*s, device=dev, dtype=dt)

        # d(out)[t][b][n_mel+1]
        dout = z(T, B, n_out)
        if d_mel is not None:
            dout[:, :, :d.n_mel] = d_mel.permute(2, 0, 1)
        if d_gate is not None:
            dout[:, :, d.n_mel] = d_gate.t()
        dout2 = dout.reshape(T * B, n_out)
        # hoisted projection backward: d[h_dec'|ctx'] for every frame
        dproj = linear_nt(dout2, pb.wpgT, prec)                                   # [TB][Hd+D]

        wk = _bptt_v2(dec, pk, d, ctx, S, sv, dproj, d_align, prec)

        # ---- deferred weight gradients: K = T*B GEMMs on MN-major operands --------------------------------
        TB = T * B
        dga = wk["dga_all"].reshape(TB, -1)
        dgd = wk["dgd_all"].reshape(TB, -1)
        pre_T = S["pre"][:TB]                                   # prenet outputs of frames 0..T-1
        bf = (lambda x: x.to(BF16)) if prec == 1 else (lambda x: x)    # bf16 mode: every GEMM operand is converted ONCE
        dga_o, dgd_o, dout_o = bf(dga), bf(dgd), bf(dout2)
        ha_o = bf(sv["h_a_buf"])
        hd_o = S["hd_b"] if S.get("hd_b") is not None else bf(sv["h_d_buf"])
        cx_o = S["cx_b"] if S.get("cx_b") is not None else bf(sv["ctx_buf"])
        h_a_prev, h_a_cur = ha_o[:T].reshape(TB, -1), ha_o[1:].reshape(TB, -1)
        h_d_prev, h_d_cur = hd_o[:T].reshape(TB, -1), hd_o[1:].reshape(TB, -1)
        ctx_prev, ctx_cur = cx_o[:T].reshape(TB, -1), cx_o[1:].reshape(TB, -1)
        Ha, Hd, P = d.att_rnn, d.dec_rnn, d.prenet
        g_w_ih_a = torch.empty(4 * Ha, P + D, device=dev)
        _wgrad_tn(dga_o, pre_T, g_w_ih_a[:, :P], prec)
        _wgrad_tn(dga_o, ctx_prev, g_w_ih_a[:, P:], prec)
        g_w_hh_a = torch.empty(4 * Ha, Ha, device=dev)
        _wgrad_tn(dga_o, h_a_prev, g_w_hh_a, prec)
        g_w_ih_d = torch.empty(4 * Hd, Ha + D, device=dev)
        _wgrad_tn(dgd_o, h_a_cur, g_w_ih_d[:, :Ha], prec)
        _wgrad_tn(dgd_o, ctx_cur, g_w_ih_d[:, Ha:], prec)
        g_w_hh_d = torch.empty(4 * Hd, Hd, device=dev)
        _wgrad_tn(dgd_o, h_d_prev, g_w_hh_d, prec)
        if prec == 1:
            # bias gradients = column sums of the (bf16-rounded, as in the GEMMs) gate gradients, fp32 accumulate
            g_b_a = dga_o.sum(0, dtype=torch.float32)
            g_b_d = dgd_o.sum(0, dtype=torch.float32)
            g_bpg = dout_o.sum(0, dtype=torch.float32)
        else:
            ones = torch.ones(TB, 8, device=dev)
            g_b_a = torch.empty(4 * Ha, 8, device=dev); _wgrad_tn(dga, ones, g_b_a, prec); g_b_a = g_b_a[:, 0].contiguous()
            g_b_d = torch.empty(4 * Hd, 8, device=dev); _wgrad_tn(dgd, ones, g_b_d, prec); g_b_d = g_b_d[:, 0].contiguous()
            g_bpg = torch.empty(n_out, 8, device=dev); _wgrad_tn(dout2, ones, g_bpg, prec); g_bpg = g_bpg[:, 0]
        g_wpg = torch.empty(n_out, Hd + D, device=dev)
        _wgrad_tn(dout_o, h_d_cur, g_wpg[:, :Hd], prec)
        _wgrad_tn(dout_o, ctx_cur, g_wpg[:, Hd:], prec)
        g_wq = torch.empty(d.att_dim, Ha, device=dev)
        _wgrad_tn(wk["dq_all"].reshape(TB, -1), h_a_cur, g_wq, prec)
        # prenet backward (relu + always-on dropout: y = relu(x) * keep * 2  =>  dx = dy * 2 * [y > 0])
        # The factors 2 (exact in floating point) are applied to the small weight-gradient matrices instead of the [TB][P]
        # activations: one masked copy per layer instead of a compare, two multiplies and two conversions; same bits.
        tb = torch.ops.aten.threshold_backward
        d2 = bf(tb(wk["dpre_all"].reshape(TB, P), pre_T, 0.0))                    # = true d2 / 2
        g_w_pre2 = torch.empty(P, P, device=dev)
        _wgrad_tn(d2, S["h1"][:TB], g_w_pre2, prec)
        g_w_pre2.mul_(2.0)
        d1 = bf(tb(linear_nt(d2, pb.w2T, prec), S["h1"][:TB], 0.0))               # = true d1 / 4
        g_w_pre1 = torch.empty(P, d.n_mel, device=dev)
        _wgrad_tn(d1, S["frames2"][:TB], g_w_pre1, prec)
        g_w_pre1.mul_(4.0)
        # memory layer and memory gradients
        dpm2 = wk["dpm"].reshape(B * Tin, d.att_dim)
        mem2 = S["memory_f"].reshape(B * Tin, D)
        g_wmem = torch.empty(d.att_dim, D, device=dev)
        _wgrad_tn(dpm2, mem2, g_wmem, prec)
        g_memory = None
        if ctx.needs_input_grad[2]:
            g_memory = linear_nt(dpm2, pb.wmemT, prec).reshape(B, Tin, D)        # through processed_memory
            _ctx_memory_grad(sv["align"], wk["dctx_all"], g_memory, prec)        # through the context bmm
        g_v = wk["dv_part"].sum(0).reshape(1, -1)
        # chain rule through the pre-multiplied location filter U[a][ch*KS+k] = sum_c dense[a][c] conv[c][ch][k]
        dU = wk["dU_part"].sum(0)[:d.att_dim, :2 * d.loc_kernel]
        g_wd = dU @ pk.wconv.reshape(d.loc_filters, 2 * d.loc_kernel).t()
        g_wc = (pk.wdense.t() @ dU).reshape(d.loc_filters, 2, d.loc_kernel)
        grads = [g_w_ih_a, g_w_hh_a, g_b_a, g_b_a.clone(), g_w_ih_d, g_w_hh_d, g_b_d, g_b_d.clone(),
                 g_wpg[:d.n_mel].contiguous(), g_bpg[:d.n_mel].contiguous(), g_wpg[d.n_mel:].contiguous(),
                 g_bpg[d.n_mel:].contiguous(), g_w_pre1, g_w_pre2, g_wq, g_wmem, g_v, g_wc, g_wd]
        ctx.saved = None
        return (None, None, g_memory, None, None) + tuple(grads)


def _bptt_v2(dec, pk, d, ctx, S, sv, dproj, d_align, prec):
    """Launch gtn_decoder_backward2 and derive what it leaves to K = T*B GEMMs: prenet-output gradients and the
    gradients of the time-constant context columns.  Returns the same dict layout as the v1 path."""
    B, Tin, D, T, E, _ = ctx.dims
    mult = packing.mult_of(prec)
    dev = sv["align"].device
    v2p = pack_v2(dec, pk, E)
    bp = v2p.backward_packs(dec, pk)
    Ha, Hd, A, Ep = d.att_rnn, d.dec_rnn, d.att_dim, v2p.Ep
    z = lambda *s, dt=torch.float32: torch.zeros(*s, device=dev, dtype=dt)
    lib = _lib.load()
    grid = lib.gtn_decoder_bwd2_grid(B, Ha, Hd, E)
    if grid <= 0:
        raise RuntimeError("gtn_decoder_bwd2_grid failed: " + _lib.last_error())
    q = _lib.DecoderBwd2Params()
    q.B, q.Tin, q.ldT, q.T = B, Tin, pad_to(Tin, 4), T
    q.E, q.Ep, q.D, q.Ha, q.Hd, q.A, q.KS = E, Ep, D, Ha, Hd, A, d.loc_kernel
    q.prec, q.use_dropout, q.mem_bf16, q.grid = prec, int(ctx.training), int(S["mem_enc"].dtype == BF16), grid
    q.p_att, q.p_dec, q.seed = dec.p_attention_dropout, dec.p_decoder_dropout, ctx.seed
    wk = dict(
        xdg_a=z(B, mult * 4 * Ha, dt=BF16), xdg_d=z(2, B, mult * 4 * Hd, dt=BF16),
        carry_ha=z(B, Ha), dha_dec=z(B, Ha), dc_a=z(B, Ha), dc_d=z(B, Hd), dhd_rec=z(B, Hd), dctx_att=z(B, Ep), dctx_dec=z(B, Ep),
        dwprev=z(B, q.ldT), dcum=z(B, q.ldT), barrier=z(4, dt=torch.int32),
        # pre-activation gate gradients: consumed only as GEMM operands, so bf16 mode stores them as bf16
        dga_all=torch.empty(T, B, 4 * Ha, device=dev, dtype=BF16 if prec == 1 else torch.float32),
        dgd_all=torch.empty(T, B, 4 * Hd, device=dev, dtype=BF16 if prec == 1 else torch.float32),
        dq_all=torch.empty(T, B, A, device=dev), dctx_all=torch.empty(B, T, E, device=dev), dpm=z(B, Tin, A),
        dv_part=z(grid, A), dU_part=z(grid, 128, 64))
    for k, v in wk.items():
        setattr(q, k, v.data_ptr())
    q.wT_att2, q.wT_dec2, q.ut_loc = bp["wT_att2"].data_ptr(), bp["wT_dec2"].data_ptr(), bp["ut_loc"].data_ptr()
    q.wq_au, q.v = bp["wq_au"].data_ptr(), pk.v.data_ptr()
    q.mem_enc, q.lengths = S["mem_enc"].data_ptr(), S["lengths"].data_ptr()
    q.gates_a, q.c_a_all, q.gates_d, q.c_d_all = (sv["gates_a"].data_ptr(), sv["c_a_all"].data_ptr(),
                                                  sv["gates_d"].data_ptr(), sv["c_d_all"].data_ptr())
    q.align, q.th_all, q.cum_all = sv["align"].data_ptr(), sv["th_all"].data_ptr(), sv["cum_all"].data_ptr()
    q.dproj_all = dproj.data_ptr()
    q.prof = PROF_BUFFER_BWD.data_ptr() if PROF_BUFFER_BWD is not None else None
    q.flags = (int(os.environ.get("GTN_BWD_FLAGS", "1")) & 1) | (2 if prec == 1 else 0)   # bit 0: L2 prefetch, bit 1: bf16 dga/dgd
    d_align_c = None
    if d_align is not None:
        d_align_c = d_align.float().contiguous()
        q.dalign = d_align_c.data_ptr()
    _lib.check(_timed_call("decoder_bwd2_kernel", lambda: lib.gtn_decoder_backward2(C.byref(q), _lib.stream())),
               "gtn_decoder_backward2")
    TB = T * B
    dga, dgd = wk["dga_all"].reshape(TB, 4 * Ha), wk["dgd_all"].reshape(TB, 4 * Hd)
    # d prenet outputs: dGates_att . W_ih_att[:, :P]
    wk["dpre_all"] = linear_nt(dga, bp["wa_pre_T"], prec).reshape(T, B, d.prenet)
    # total context gradient incl. the time-constant columns (needed for d(memory) of the speaker / emotion columns)
    Cc = D - E
    dctx_full = torch.empty(B, T, D, device=dev)
    dctx_full[:, :, :E] = wk["dctx_all"]
    if Cc > 0:
        dc = dproj.reshape(T, B, Hd + D)[:, :, Hd + E:] + linear_nt(dgd, bp["wd_const_T"], prec).reshape(T, B, Cc)
        if T > 1:   # attention LSTM of frame t+1 consumed the context of frame t
            dc[:T - 1] += linear_nt(dga[B:], bp["wa_const_T"], prec).reshape(T - 1, B, Cc)
        dctx_full[:, :, E:] = dc.permute(1, 0, 2)
    wk["dctx_all"] = dctx_full
    return wk


def _proj_gemm(x_tb: torch.Tensor, w: torch.Tensor, out_cl: torch.Tensor, prec: int, bias, accumulate: bool) -> None:
    """out_cl[b][t][:] (+)= x_tb[t][b][:] @ w.T (+ bias): batched over b so that rows land in (b, t) order."""
    T, B, K = x_tb.shape
    kp = pad_to(K, 8)
    xa = packing.expand(x_tb, kp, prec, False)                 # [T][B][mult*kp]
    wa = packing.expand(w.float(), kp, prec, True)             # [N][mult*kp]
    A = xa.permute(1, 0, 2)                                    # [B][T][K'] view: batch stride = K', row stride = B*K'
    wz = wa.unsqueeze(0).expand(B, -1, -1)                     # batch stride 0
    ops.gemm(A, wz, out_cl, bias_n=bias, accumulate=accumulate)


def _ctx_memory_grad(align: torch.Tensor, dctx_all: torch.Tensor, g_memory: torch.Tensor, prec: int) -> None:
    """g_memory[b] += align[b]^T (T_in x T) . dctx[b] (T x D)   (backward of the context bmm, model.py:88)."""
    B, T, Tin = align.shape
    D = dctx_all.shape[2]
    tp, dp = pad_to(Tin, 8), pad_to(D, 8)

    def planes(x, kp, which):
        if prec == 1:
            buf = torch.zeros(B, T, kp, device=x.device, dtype=BF16)
            buf[:, :, :x.shape[2]] = x.to(BF16)
            return buf[:, :, :x.shape[2]]
        pl = packing.split3(x)
        buf = torch.zeros(B, 6 * T, kp, device=x.device, dtype=BF16)
        for s, i in enumerate(which):
            buf[:, s * T:(s + 1) * T, :x.shape[2]] = pl[i]
        return buf[:, :, :x.shape[2]]
    ops.gemm_tn(planes(align, tp, packing.W_PLANES), planes(dctx_all, dp, packing.X_PLANES), g_memory, accumulate=True)


def decoder_train_forward(dec, memory: torch.Tensor, decoder_inputs: torch.Tensor, memory_lengths: torch.Tensor, seed: int):
    """Decoder.forward (model.py:419-454) on a `memory` that already has the noise columns appended."""
    _, params = _decoder_params_list(dec)
    return _DecoderTrainFn.apply(dec, seed, memory, decoder_inputs, memory_lengths, *params)
