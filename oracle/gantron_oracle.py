"""CPU oracle for the GANtron decoder / Postnet / Discriminator hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``gantron_b200/`` imports this file; only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs do, and there only as the checker / CPU baseline.

It is a *functional restatement* (plain torch CPU ops on a flat ``state_dict``-style
weight dict, explicit dropout keep-masks, explicit state tuples) of the algorithm the
reference implements with stateful ``nn.Module``s.  Each function cites the reference
lines it follows.  All arithmetic is torch's (the reference's own arithmetic also lives in
PyTorch: torch==1.5.1 in the author's environment, /root/reference/unins:136, not pinned
by requirements.txt; 2.11.0+cu128 here).

PARITY PINNING: the reference has no tests, golden vectors or known-answer fixtures of
its own.  This oracle is pinned instead against *outputs of the
reference itself run in the build container* -- ``oracle/make_golden.py`` imports
/root/reference/model.py (with import stubs + CPU shims), injects the same Philox masks,
and writes ``tests/golden/*.npz``; ``tests/test_oracle_vs_reference.py`` re-checks live
whenever /root/reference is present, and ``tests/test_oracle_golden.py`` checks the
committed fixtures anywhere.

Gradients are obtained with torch autograd over these functions (they are differentiable
torch ops); the only hand-written backward is `_Conv0Quirk`, which reproduces a reference
side effect (see its docstring).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor


# ----------------------------------------------------------------------------------------
# dims
# ----------------------------------------------------------------------------------------
@dataclass
class Dims:
    """Shape bookkeeping derived the way /root/reference/model.py:228-271,501-512,587-599 does."""
    n_mel: int = 80
    prenet: int = 256
    att_rnn: int = 1024
    dec_rnn: int = 1024
    att_dim: int = 128
    loc_filters: int = 32
    loc_kernel: int = 31
    enc_dim: int = 512
    noise: int = 512            # decoder noise size (0 when use_noise=False or encoder_inputs)
    spk: int = 0                # 64 when vesus_path is set
    emo: int = 0                # 5 when use_labels and vesus and not encoder_inputs
    postnet_dim: int = 512
    postnet_k: int = 5
    postnet_n: int = 5
    disc_window: int = 20
    disc_dim: int = 512
    p_att: float = 0.1
    p_dec: float = 0.1
    gate_threshold: float = 0.5
    max_decoder_steps: int = 1000

    @property
    def d_mem(self) -> int:       # model.py:242-246
        return self.enc_dim + self.noise + self.spk + self.emo

    @classmethod
    def from_hparams(cls, hp) -> "Dims":
        vesus = bool(getattr(hp, "vesus_path", None))
        use_labels = bool(hp.use_labels) and vesus          # model.py:591
        noise = hp.noise_size if getattr(hp, "use_noise", True) else 0   # model.py:589-590
        if hp.encoder_inputs:
            noise = 0                                      # model.py:240
        return cls(
            n_mel=hp.n_mel_channels, prenet=hp.prenet_dim, att_rnn=hp.attention_rnn_dim,
            dec_rnn=hp.decoder_rnn_dim, att_dim=hp.attention_dim,
            loc_filters=hp.attention_location_n_filters, loc_kernel=hp.attention_location_kernel_size,
            enc_dim=hp.encoder_embedding_dim, noise=noise,
            spk=hp.speakers_embedding if vesus else 0,
            emo=5 if (vesus and use_labels and not hp.encoder_inputs) else 0,
            postnet_dim=hp.postnet_embedding_dim, postnet_k=hp.postnet_kernel_size,
            postnet_n=hp.postnet_n_convolutions, disc_window=hp.discriminator_window,
            disc_dim=hp.discriminator_dim, p_att=hp.p_attention_dropout, p_dec=hp.p_decoder_dropout,
            gate_threshold=hp.gate_threshold, max_decoder_steps=hp.max_decoder_steps)


def _drop(x: Tensor, keep: Optional[Tensor], p: float) -> Tensor:
    """F.dropout(x, p, training=True) with an explicit {0,1} keep mask (None = no dropout)."""
    if keep is None or p == 0.0:
        return x
    return x * keep * (1.0 / (1.0 - p))


# ----------------------------------------------------------------------------------------
# Prenet  (model.py:94-105)
# ----------------------------------------------------------------------------------------
def prenet(x: Tensor, w1: Tensor, w2: Tensor, keep1: Tensor, keep2: Tensor) -> Tensor:
    """Two bias-free linears, each relu -> dropout(0.5) that is ALWAYS on (model.py:104)."""
    h = _drop(torch.relu(x @ w1.t()), keep1, 0.5)
    return _drop(torch.relu(h @ w2.t()), keep2, 0.5)


# ----------------------------------------------------------------------------------------
# LSTMCell  (torch.nn.LSTMCell as used at model.py:252-254,261-263,391-392,406-407)
# ----------------------------------------------------------------------------------------
def lstm_cell(x: Tensor, h: Tensor, c: Tensor, w_ih: Tensor, w_hh: Tensor,
              b_ih: Tensor, b_hh: Tensor) -> Tuple[Tensor, Tensor]:
    """gates = x W_ih^T + b_ih + h W_hh^T + b_hh, chunked (i, f, g, o)."""
    gates = x @ w_ih.t() + b_ih + h @ w_hh.t() + b_hh
    i, f, g, o = gates.chunk(4, dim=1)
    c2 = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(g)
    h2 = torch.sigmoid(o) * torch.tanh(c2)
    return h2, c2


# ----------------------------------------------------------------------------------------
# Location-sensitive attention  (model.py:15-91)
# ----------------------------------------------------------------------------------------
def location_features(w_prev: Tensor, w_cum: Tensor, conv_w: Tensor, dense_w: Tensor) -> Tensor:
    """(B,T_in) x2 -> (B,T_in,att_dim): conv1d(2->F,k,pad=(k-1)/2, no bias) then linear(F->A).
    model.py:27-31 with the cat of model.py:396-398."""
    cat = torch.stack((w_prev, w_cum), dim=1)                     # (B,2,T_in)
    pad = (conv_w.shape[-1] - 1) // 2
    f = F.conv1d(cat, conv_w, padding=pad)                        # (B,F,T_in)
    return f.transpose(1, 2) @ dense_w.t()                        # (B,T_in,A)


def attention_step(h_att: Tensor, memory: Tensor, processed_memory: Tensor, w_prev: Tensor,
                   w_cum: Tensor, pad_mask: Optional[Tensor], query_w: Tensor, v_w: Tensor,
                   conv_w: Tensor, dense_w: Tensor, score_mask_value: float = -float("inf")
                   ) -> Tuple[Tensor, Tensor]:
    """energies = v . tanh(Wq h + loc + pm) (model.py:62-68); padded positions get
    score_mask_value without autograd seeing it (model.py:84-85); softmax over T_in and the
    context bmm (model.py:87-89).  Returns (context (B,D_mem), weights (B,T_in))."""
    q = (h_att @ query_w.t()).unsqueeze(1)                        # (B,1,A)
    loc = location_features(w_prev, w_cum, conv_w, dense_w)
    e = (torch.tanh(q + loc + processed_memory) @ v_w.t()).squeeze(-1)   # (B,T_in)
    if pad_mask is not None:
        # `.data.masked_fill_`: value replaced, gradient path untouched
        e = _fill_no_grad(e, pad_mask, score_mask_value)
    w = torch.softmax(e, dim=1)
    ctx = torch.bmm(w.unsqueeze(1), memory).squeeze(1)
    return ctx, w


def _fill_no_grad(x: Tensor, mask: Tensor, value: float) -> Tensor:
    """Value-level masked fill that leaves d(out)/d(x) = 1 everywhere (what `.data.masked_fill_` does)."""
    filled = x.detach().masked_fill(mask, value)
    return filled + (x - x.detach())


# ----------------------------------------------------------------------------------------
# Decoder  (model.py:227-497)
# ----------------------------------------------------------------------------------------
@dataclass
class DecoderState:
    """What the reference keeps as module attributes (model.py:296-327)."""
    h_att: Tensor
    c_att: Tensor
    h_dec: Tensor
    c_dec: Tensor
    w: Tensor
    w_cum: Tensor
    ctx: Tensor


def init_state(memory: Tensor, d: Dims) -> DecoderState:
    B, T_in, D = memory.shape
    z = lambda n: memory.new_zeros(B, n)
    return DecoderState(z(d.att_rnn), z(d.att_rnn), z(d.dec_rnn), z(d.dec_rnn), z(T_in), z(T_in), z(D))


def _dec_w(sd: Dict[str, Tensor], name: str, prefix: str = "decoder.") -> Tensor:
    return sd[prefix + name]


def decoder_step(sd: Dict[str, Tensor], st: DecoderState, x_prenet: Tensor, memory: Tensor,
                 processed_memory: Tensor, pad_mask: Optional[Tensor], d: Dims,
                 keep_att: Optional[Tensor], keep_dec: Optional[Tensor],
                 score_mask_value: float = -float("inf"), prefix: str = "decoder."
                 ) -> Tuple[Tensor, Tensor, DecoderState]:
    """One frame (model.py:378-417).  keep_att / keep_dec are the recurrent-state dropout masks
    (None in eval mode: model.py:393-394,408-409 use self.training)."""
    g = lambda n: sd[prefix + n]
    h_att, c_att = lstm_cell(torch.cat((x_prenet, st.ctx), -1), st.h_att, st.c_att,
                             g("attention_rnn.weight_ih"), g("attention_rnn.weight_hh"),
                             g("attention_rnn.bias_ih"), g("attention_rnn.bias_hh"))
    h_att = _drop(h_att, keep_att, d.p_att)            # the DROPPED h is the recurrent state
    ctx, w = attention_step(
        h_att, memory, processed_memory, st.w, st.w_cum, pad_mask,
        g("attention_layer.query_layer.linear_layer.weight"),
        g("attention_layer.v.linear_layer.weight"),
        g("attention_layer.location_layer.location_conv.conv.weight"),
        g("attention_layer.location_layer.location_dense.linear_layer.weight"),
        score_mask_value)
    w_cum = st.w_cum + w                               # model.py:403
    h_dec, c_dec = lstm_cell(torch.cat((h_att, ctx), -1), st.h_dec, st.c_dec,
                             g("decoder_rnn.weight_ih"), g("decoder_rnn.weight_hh"),
                             g("decoder_rnn.bias_ih"), g("decoder_rnn.bias_hh"))
    h_dec = _drop(h_dec, keep_dec, d.p_dec)
    hc = torch.cat((h_dec, ctx), 1)
    mel = hc @ g("linear_projection.linear_layer.weight").t() + g("linear_projection.linear_layer.bias")
    gate = hc @ g("gate_layer.linear_layer.weight").t() + g("gate_layer.linear_layer.bias")
    return mel, gate, DecoderState(h_att, c_att, h_dec, c_dec, w, w_cum, ctx)


def pad_mask_from_lengths(lengths: Tensor, max_len: Optional[int] = None) -> Tensor:
    """True at PADDED positions: ~get_mask_from_lengths (utils.py:20-24, model.py:441,634)."""
    n = int(lengths.max()) if max_len is None else max_len
    return torch.arange(n, device=lengths.device)[None, :] >= lengths[:, None]


def decoder_teacher_forced(sd: Dict[str, Tensor], memory: Tensor, mels: Tensor, memory_lengths: Tensor,
                           d: Dims, keep_pre1: Tensor, keep_pre2: Tensor,
                           keep_att: Optional[Tensor], keep_dec: Optional[Tensor],
                           score_mask_value: float = -float("inf"), prefix: str = "decoder.",
                           return_states: bool = False):
    """Decoder.forward (model.py:419-454) on a `memory` that already has the noise appended.

    memory (B,T_in,D_mem); mels (B,n_mel,T); keep_pre* (T+1,B,prenet); keep_att/keep_dec (T,B,rnn)
    or None (eval).  Returns mel (B,n_mel,T), gate (B,T), alignments (B,T,T_in)."""
    g = lambda n: sd[prefix + n]
    B, _, T = mels.shape
    frames = torch.cat((mels.new_zeros(1, B, d.n_mel), mels.permute(2, 0, 1)), 0)      # go frame + (T,B,80)
    pre = prenet(frames, g("prenet.layers.0.linear_layer.weight"), g("prenet.layers.1.linear_layer.weight"),
                 keep_pre1, keep_pre2)                                                   # (T+1,B,256)
    pm = memory @ g("attention_layer.memory_layer.linear_layer.weight").t()              # model.py:326
    pad = pad_mask_from_lengths(memory_lengths, memory.shape[1])
    st = init_state(memory, d)
    mel_l, gate_l, w_l, states = [], [], [], []
    for t in range(T):                                  # the T+1-th prenet row is never consumed (model.py:444)
        mel, gate, st = decoder_step(sd, st, pre[t], memory, pm, pad, d,
                                     None if keep_att is None else keep_att[t],
                                     None if keep_dec is None else keep_dec[t],
                                     score_mask_value, prefix)
        mel_l.append(mel); gate_l.append(gate.squeeze(1)); w_l.append(st.w)
        if return_states:
            states.append(st)
    out = (torch.stack(mel_l, 2), torch.stack(gate_l, 1), torch.stack(w_l, 1))
    return out + (states,) if return_states else out


def decoder_inference(sd: Dict[str, Tensor], memory: Tensor, d: Dims, keep_fn,
                      max_steps: Optional[int] = None, prefix: str = "decoder.",
                      force_steps: Optional[int] = None):
    """Decoder.inference (model.py:456-497) on a `memory` that already carries the style columns.

    keep_fn(step) -> (keep1, keep2), each (B,prenet): the two always-on prenet masks of that step.
    Batch semantics (the reference only runs at B=1, model.py:486): a row is *done* at the first
    step whose sigmoid(gate) > threshold; the loop ends after the step at which every row is done, or
    at max_steps; done rows keep decoding (their frames past `lengths[b]` are don't-care).  At B=1
    this is exactly the reference loop.  force_steps overrides the stop test (benchmarks).
    Returns mel (B,n_mel,T'), gate (B,T',1), alignments (B,T',T_in), lengths (B,), states."""
    g = lambda n: sd[prefix + n]
    max_steps = d.max_decoder_steps if max_steps is None else max_steps
    B = memory.shape[0]
    pm = memory @ g("attention_layer.memory_layer.linear_layer.weight").t()
    st = init_state(memory, d)
    x = memory.new_zeros(B, d.n_mel)
    done = torch.zeros(B, dtype=torch.bool)
    lengths = torch.full((B,), 0, dtype=torch.long)
    mel_l, gate_l, w_l = [], [], []
    while True:
        k1, k2 = keep_fn(len(mel_l))
        xp = prenet(x, g("prenet.layers.0.linear_layer.weight"), g("prenet.layers.1.linear_layer.weight"), k1, k2)
        mel, gate, st = decoder_step(sd, st, xp, memory, pm, None, d, None, None, prefix=prefix)
        mel_l.append(mel); gate_l.append(gate); w_l.append(st.w)
        fired = torch.sigmoid(gate.detach()).squeeze(1) > d.gate_threshold
        newly = fired & ~done
        lengths[newly] = len(mel_l)
        done |= fired
        if force_steps is not None:
            if len(mel_l) == force_steps:
                break
        elif bool(done.all()):
            break
        elif len(mel_l) == max_steps:
            break
        x = mel
    lengths[~done] = len(mel_l)
    return torch.stack(mel_l, 2), torch.stack(gate_l, 1), torch.stack(w_l, 1), lengths, st


# ----------------------------------------------------------------------------------------
# Postnet  (model.py:108-151)
# ----------------------------------------------------------------------------------------
class _Conv0Quirk(torch.autograd.Function):
    """conv1d whose weight gradient is computed from a DIFFERENT input than the forward used.

    Reference side effect: `parse_output` (model.py:638) masks
    `outputs[0].data` in place AFTER the postnet ran; that storage is the tensor autograd saved as
    the first postnet conv's input, so d(loss)/d(conv0.weight) is evaluated on the *masked* mel
    while the forward pass (and d/d(input)) used the unmasked one."""

    @staticmethod
    def forward(ctx, x, x_for_wgrad, w, b, pad):
        ctx.save_for_backward(x_for_wgrad, w)
        ctx.pad = pad
        return F.conv1d(x, w, b, padding=pad)

    @staticmethod
    def backward(ctx, gy):
        xw, w = ctx.saved_tensors
        gx = torch.nn.grad.conv1d_input(xw.shape, w, gy, padding=ctx.pad)
        gw = torch.nn.grad.conv1d_weight(xw, w.shape, gy, padding=ctx.pad)
        return gx, None, gw, gy.sum((0, 2)), None


def batch_norm_train(x: Tensor, gamma: Tensor, beta: Tensor, eps: float = 1e-5):
    """BatchNorm1d in training mode over (B,C,T): biased batch variance for normalisation; returns
    also (mean, unbiased var) for the running-stat update (momentum 0.1)."""
    mean = x.mean((0, 2))
    var = x.var((0, 2), unbiased=False)
    y = (x - mean[None, :, None]) * torch.rsqrt(var[None, :, None] + eps) * gamma[None, :, None] + beta[None, :, None]
    n = x.shape[0] * x.shape[2]
    return y, mean.detach(), (var * n / max(n - 1, 1)).detach()


def postnet(sd: Dict[str, Tensor], x: Tensor, d: Dims, keeps: Optional[Sequence[Tensor]], training: bool,
            x_masked_for_wgrad: Optional[Tensor] = None, prefix: str = "postnet.",
            stats_out: Optional[dict] = None) -> Tensor:
    """5 x (conv k5 pad2 -> BatchNorm1d), tanh on all but the last, dropout(0.5) when training
    (model.py:146-151).  keeps: postnet_n masks (B,C_out,T) or None."""
    pad = (d.postnet_k - 1) // 2
    for i in range(d.postnet_n):
        w, b = sd[f"{prefix}convolutions.{i}.0.conv.weight"], sd[f"{prefix}convolutions.{i}.0.conv.bias"]
        if i == 0 and x_masked_for_wgrad is not None:
            y = _Conv0Quirk.apply(x, x_masked_for_wgrad, w, b, pad)
        else:
            y = F.conv1d(x, w, b, padding=pad)
        gam, bet = sd[f"{prefix}convolutions.{i}.1.weight"], sd[f"{prefix}convolutions.{i}.1.bias"]
        if training:
            y, m, v = batch_norm_train(y, gam, bet)
            if stats_out is not None:
                stats_out[i] = (m, v)
        else:
            rm, rv = sd[f"{prefix}convolutions.{i}.1.running_mean"], sd[f"{prefix}convolutions.{i}.1.running_var"]
            y = (y - rm[None, :, None]) * torch.rsqrt(rv[None, :, None] + 1e-5) * gam[None, :, None] + bet[None, :, None]
        if i < d.postnet_n - 1:
            y = torch.tanh(y)
        x = _drop(y, keeps[i] if (training and keeps is not None) else None, 0.5)
    return x


# ----------------------------------------------------------------------------------------
# Text encoder  (model.py:154-224) -- SURVEY 8f1
# ----------------------------------------------------------------------------------------
def encoder_forward(sd: Dict[str, Tensor], emb: Tensor, lengths: Optional[Tensor], keeps: Optional[Sequence[Tensor]],
                    training: bool, prefix: str = "encoder.", n_conv: int = 3) -> Tensor:
    """emb (B,C,T) embedded text -> (B,T,2H).  Three conv k5 + BatchNorm1d + relu + dropout(0.5) (model.py:196-197), then the
    bidirectional LSTM on PACKED sequences (model.py:199-213; `lengths` None = Encoder.inference, model.py:216-224).
    keeps: one (B,C,T) mask per conv layer or None."""
    x = emb
    for i in range(n_conv):
        w, b = sd[f"{prefix}convolutions.{i}.0.conv.weight"], sd[f"{prefix}convolutions.{i}.0.conv.bias"]
        y = F.conv1d(x, w, b, padding=(w.shape[-1] - 1) // 2)
        gam, bet = sd[f"{prefix}convolutions.{i}.1.weight"], sd[f"{prefix}convolutions.{i}.1.bias"]
        if training:
            y, _, _ = batch_norm_train(y, gam, bet)
        else:
            rm, rv = sd[f"{prefix}convolutions.{i}.1.running_mean"], sd[f"{prefix}convolutions.{i}.1.running_var"]
            y = (y - rm[None, :, None]) * torch.rsqrt(rv[None, :, None] + 1e-5) * gam[None, :, None] + bet[None, :, None]
        x = _drop(torch.relu(y), keeps[i] if (training and keeps is not None) else None, 0.5)
    x = x.transpose(1, 2)
    H = sd[f"{prefix}lstm.weight_hh_l0"].shape[1]
    lstm = torch.nn.LSTM(x.shape[2], H, 1, batch_first=True, bidirectional=True)
    params = {k: sd[f"{prefix}lstm.{k}"] for k in ("weight_ih_l0", "weight_hh_l0", "bias_ih_l0", "bias_hh_l0",
                                                   "weight_ih_l0_reverse", "weight_hh_l0_reverse", "bias_ih_l0_reverse",
                                                   "bias_hh_l0_reverse")}
    if lengths is None:
        return torch.func.functional_call(lstm, params, (x,))[0]
    packed = torch.nn.utils.rnn.pack_padded_sequence(x, lengths.cpu(), batch_first=True, enforce_sorted=False)
    out, _ = torch.func.functional_call(lstm, params, (packed,))
    return torch.nn.utils.rnn.pad_packed_sequence(out, batch_first=True, total_length=x.shape[1])[0]


# ----------------------------------------------------------------------------------------
# Output masking  (model.py:632-642)
# ----------------------------------------------------------------------------------------
def parse_output(mel: Tensor, mel_post: Tensor, gate: Tensor, output_lengths: Optional[Tensor]):
    """Value-level masking through `.data` (no autograd cut): mel/mel_post -> 0, gate -> 1e3."""
    if output_lengths is None:
        return mel, mel_post, gate
    pad = pad_mask_from_lengths(output_lengths, mel.shape[-1])            # (B,T)
    pad3 = pad[:, None, :].expand_as(mel)
    return _fill_no_grad(mel, pad3, 0.0), _fill_no_grad(mel_post, pad3, 0.0), _fill_no_grad(gate, pad, 1e3)


def generator_from_memory(sd: Dict[str, Tensor], memory: Tensor, mels: Tensor, text_lengths: Tensor,
                          output_lengths: Optional[Tensor], d: Dims, masks: dict, training: bool = True,
                          mask_padding: bool = True, score_mask_value: float = -float("inf")):
    """Tacotron2.forward from the encoder outputs on (model.py:665-673): decoder -> postnet ->
    residual -> parse_output.  `memory` already has speaker/emotion/noise columns appended.
    masks: pre1, pre2, att, dec, post (list)."""
    mel, gate, align = decoder_teacher_forced(
        sd, memory, mels, text_lengths, d, masks["pre1"], masks["pre2"],
        masks.get("att") if training else None, masks.get("dec") if training else None, score_mask_value)
    lens = output_lengths if mask_padding else None
    mel_masked_val = parse_output(mel, mel, gate, lens)[0].detach() if lens is not None else None
    post = postnet(sd, mel, d, masks.get("post"), training, x_masked_for_wgrad=mel_masked_val)
    mel_post = mel + post
    mel_o, mel_post_o, gate_o = parse_output(mel, mel_post, gate, lens)
    return mel_o, mel_post_o, gate_o, align


# ----------------------------------------------------------------------------------------
# Discriminator  (model.py:500-540, layers.py:43-60)
# ----------------------------------------------------------------------------------------
def disc_layer_dims(d: Dims) -> List[Tuple[int, int, int, int]]:
    """(c_in, c_out, kernel, dilation) per conv (model.py:503-512)."""
    in_dim = d.disc_window * d.n_mel
    first = min((in_dim // d.disc_dim + 1) * d.disc_dim, 1024)
    return [(in_dim, first, 5, 1), (first, d.disc_dim, 5, 2), (d.disc_dim, d.disc_dim, 5, 2),
            (d.disc_dim, d.n_mel, 5, 2), (d.n_mel, 1, 1, 1)]


def discriminator_forward(sd: Dict[str, Tensor], inputs: Tensor, d: Dims, keeps: Optional[Sequence[Tensor]],
                          prefix: str = "") -> Tensor:
    """inputs (B,F,n_mel) -> (B,P).  Tail handling (model.py:523-524) then the RAW reshape
    (model.py:526): (B,F',80) memory reinterpreted as (B,1600,F'/20) -- channel c, position p is flat
    element c*P+p of the sample, not a window transpose.  Each DiscConv1d is conv -> dropout(0.5,
    always training: train.py:274) -> tanh (layers.py:53-57); last is a plain k=1 conv."""
    win = d.disc_window
    r = inputs.shape[1] % win
    if r != 0:
        inputs = torch.cat((inputs[:, :-r], inputs[:, -win:]), 1)
    x = inputs.reshape(inputs.shape[0], win * d.n_mel, -1)
    dims = disc_layer_dims(d)
    for i, (_, _, k, dil) in enumerate(dims[:-1]):
        w, b = sd[f"{prefix}discriminator.{i}.module.0.weight"], sd[f"{prefix}discriminator.{i}.module.0.bias"]
        x = F.conv1d(x, w, b, padding=dil * (k - 1) // 2, dilation=dil)
        x = torch.tanh(_drop(x, None if keeps is None else keeps[i], 0.5))
    w, b = sd[f"{prefix}discriminator.4.weight"], sd[f"{prefix}discriminator.4.bias"]
    return F.conv1d(x, w, b).squeeze(1)


def adversarial_loss(sd: Dict[str, Tensor], mel: Tensor, lengths: Tensor, d: Dims,
                     keeps: Optional[Sequence[Tensor]], prefix: str = "") -> Tensor:
    """mean_b mean(out[b, :ceil(len_b / window)])  (model.py:529-540)."""
    out = discriminator_forward(sd, mel.transpose(1, 2), d, keeps, prefix)
    total = 0.0
    for b in range(lengths.shape[0]):
        n = int(math.ceil(int(lengths[b]) / d.disc_window))
        total = total + out[b, :n].mean()
    return total / lengths.shape[0]


# ----------------------------------------------------------------------------------------
# LinearDiscriminator  (model.py:543-583, layers.py:63-73)
# ----------------------------------------------------------------------------------------
def linear_window_plan(lengths: Sequence[int], window: int, max_overlap: int = 6) -> List[List[int]]:
    """Window start frames per sample, drawn like model.py:572-577: module-level `random.randint`, sample by sample; the
    last window is aligned to the end of the sample."""
    import random
    plan = []
    for n in lengths:
        n, start, mine = int(n), 0, []
        while start + window < n:
            mine.append(start)
            start += window - random.randint(0, max_overlap)
        mine.append(n - window)
        plan.append(mine)
    return plan


def linear_discriminator_forward(sd: Dict[str, Tensor], windows: Tensor, keeps: Optional[Sequence[Tensor]],
                                 prefix: str = "") -> Tensor:
    """windows (R, window*n_mel) -> (R, 1): three DiscDense = Linear -> Dropout(0.5) -> Tanh, then Linear (model.py:546-551)."""
    x = windows
    for i in range(3):
        w, b = sd[f"{prefix}discriminator.{i}.module.0.weight"], sd[f"{prefix}discriminator.{i}.module.0.bias"]
        x = torch.tanh(_drop(F.linear(x, w, b), None if keeps is None else keeps[i], 0.5))
    return F.linear(x, sd[f"{prefix}discriminator.3.weight"], sd[f"{prefix}discriminator.3.bias"])


def linear_adversarial_loss(sd: Dict[str, Tensor], mel: Tensor, lengths: Tensor, d: Dims, keeps: Optional[Sequence[Tensor]],
                            prefix: str = "") -> Tensor:
    """mean_b mean_w D(window_w of sample b) (model.py:567-583).  `keeps[l]` is the dropout mask of layer l over ALL windows of
    the batch stacked sample after sample (the numbering of the kernels, which run the MLP once over the stacked rows); the
    reference runs sample by sample, which is the same arithmetic on row slices."""
    x = mel.transpose(1, 2) if mel.size(1) == d.n_mel else mel
    plan = linear_window_plan([int(v) for v in lengths], d.disc_window)
    rows = [x[i, s:s + d.disc_window].flatten() for i, starts in enumerate(plan) for s in starts]
    out = linear_discriminator_forward(sd, torch.stack(rows), keeps, prefix)[:, 0]
    total, r0 = 0.0, 0
    for starts in plan:
        total = total + out[r0:r0 + len(starts)].mean()
        r0 += len(starts)
    return total / lengths.shape[0]


# ----------------------------------------------------------------------------------------
# Losses  (loss_function.py:6-52)
# ----------------------------------------------------------------------------------------
def tacotron2_losses(mel: Tensor, mel_post: Tensor, gate: Tensor, mel_target: Tensor, gate_target: Tensor):
    """mel_loss = MSE(mel)+MSE(mel_post) over ALL elements incl. padding; gate_loss = BCE-with-logits
    (loss_function.py:18-20)."""
    mel_loss = F.mse_loss(mel, mel_target) + F.mse_loss(mel_post, mel_target)
    gate_loss = F.binary_cross_entropy_with_logits(gate.reshape(-1, 1), gate_target.reshape(-1, 1))
    return mel_loss, gate_loss


def attention_guide(n_chars: int, n_frames: int, width: float = 0.5) -> Tensor:
    """Guided-attention target (loss_function.py:42-52): row n is exp(-(x-c_n)^2/(width*3)) with
    c_n = n*(n_frames-1)//(n_chars-1) (integer floor), x = 0..n_frames-1."""
    x = torch.linspace(0, n_frames - 1, n_frames)
    centers = torch.tensor([n * (n_frames - 1) // (n_chars - 1) for n in range(n_chars)], dtype=torch.float32)
    return torch.exp(-(x[None, :] - centers[:, None]) ** 2 / (width * 3.0))


def attention_loss(alignments: Tensor, enc_len: Tensor, dec_len: Tensor) -> Tensor:
    """mean_b BCE(align[b].T[:enc_len_b,:dec_len_b], guide) (loss_function.py:25-32)."""
    total = 0.0
    for b in range(alignments.shape[0]):
        ne, nd = int(enc_len[b]), int(dec_len[b])
        total = total + F.binary_cross_entropy(alignments[b].t()[:ne, :nd], attention_guide(ne, nd).to(alignments.dtype))
    return total / alignments.shape[0]


# ----------------------------------------------------------------------------------------
# Synthetic LJSpeech-shaped batches -- shared by tests and bench
# ----------------------------------------------------------------------------------------
def synthetic_batch(B: int, T_in: int, T_out: int, n_mel: int = 80, seed: int = 1234, ragged: bool = False,
                    n_symbols: int = 148):
    """7-tuple in TextMelCollate order (data_utils.py:94-131):
    text_padded, input_lengths, mel_padded, gate_padded, speaker_ids, emotions, output_lengths."""
    gen = torch.Generator().manual_seed(seed)
    if ragged:
        in_len = torch.tensor([max(T_in - 7 * i, max(T_in * 2 // 3, 2)) for i in range(B)])
        out_len = torch.tensor([max(T_out - 31 * i, max(T_out // 2, 1)) for i in range(B)])
        if T_out < 64:
            out_len = torch.tensor([max(T_out - i, 1) for i in range(B)])
        if T_in < 64:
            in_len = torch.tensor([max(T_in - i, 2) for i in range(B)])
    else:
        in_len = torch.full((B,), T_in)
        out_len = torch.full((B,), T_out)
    text = torch.randint(1, n_symbols, (B, T_in), generator=gen)
    mel = (torch.randn(B, n_mel, T_out, generator=gen) * 2 - 5).clamp(-11.5, 2.0)
    gate = torch.zeros(B, T_out)
    for b in range(B):
        text[b, in_len[b]:] = 0
        mel[b, :, out_len[b]:] = 0
        gate[b, out_len[b] - 1:] = 1
    spk = torch.randint(0, 123, (B,), generator=gen).float()
    emo = torch.softmax(torch.randn(B, 5, generator=gen) * 3, 1)
    return text, in_len.long(), mel, gate, spk, emo, out_len.long()


def seeded_state_dict(shapes: Dict[str, Tuple[int, ...]], seed: int = 7, scale: float = 1.0) -> Dict[str, Tensor]:
    """Deterministic weights for fixtures, independent of torch's module-init order: keys in sorted
    order, uniform(-a, a) with Xavier-like a = sqrt(6/(fan_in+fan_out)); biases small; BN gamma near 1,
    running_var near 1."""
    gen = torch.Generator().manual_seed(seed)
    out = {}
    for k in sorted(shapes):
        shp = tuple(shapes[k])
        if k.endswith("num_batches_tracked"):
            out[k] = torch.zeros(shp, dtype=torch.long)
            continue
        u = torch.rand(shp, generator=gen) * 2 - 1
        if len(shp) >= 2:
            rf = int(torch.tensor(shp[2:]).prod()) if len(shp) > 2 else 1
            a = math.sqrt(6.0 / (shp[1] * rf + shp[0] * rf))
            out[k] = u * a * scale
        elif k.endswith("running_var") or (k.endswith(".1.weight") and "convolutions" in k):
            out[k] = 1.0 + 0.1 * u
        else:
            out[k] = 0.05 * u
    return out
