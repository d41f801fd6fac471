"""Drop-in replacement for the reference's `model.py` surface, executed by sm_100a kernels.

Same importable names, constructor arguments, attribute paths, state_dict keys / shapes and return
conventions as /root/reference/model.py, so `train.py`, `study_model.py` and
`inference_samples.py` run against it unchanged (`from model import Tacotron2, Discriminator,
LinearDiscriminator`).  What differs is HOW the hot path executes:

  Decoder.forward / Decoder.inference  -> one persistent cooperative kernel per pass (+ tcgen05 GEMMs
                                          for everything time-independent), decoder_ops.py
  Postnet.forward                      -> implicit-GEMM conv1d on tcgen05 + fused BN/tanh/dropout, postnet_ops.py
  Discriminator.forward/adversarial_loss -> implicit-GEMM convs + fused dropout/tanh + masked mean, disc_ops.py

Out of scope by the north star and left on stock PyTorch: Encoder (convs + BiLSTM), LinearDiscriminator,
embeddings.  There is no CPU or eager fallback for the hot path: CUDA tensors + libgantron_b200.so or an error.
"""
from __future__ import annotations

import os
import random
from math import sqrt

import torch
from torch import nn
from torch.nn import functional as F

from .layers import ConvNorm, DiscConv1d, DiscDense, LinearNorm


def _precision_of(hparams) -> int:
    """1 = bf16 operands (default), 3 = fp32-accurate 3-plane mode."""
    name = os.environ.get("GANTRON_B200_PRECISION") or getattr(hparams, "b200_precision", "bf16")
    if name not in ("bf16", "fp32"):
        raise ValueError(f"b200_precision must be 'bf16' or 'fp32', got {name!r}")
    return 1 if name == "bf16" else 3


def _new_seed(mod=None) -> int:
    """Philox key for one pass, drawn from torch's CPU generator (so torch.manual_seed reproduces runs).
    Tests pin it by setting `module._forced_seed`."""
    forced = getattr(mod, "_forced_seed", None) if mod is not None else None
    if forced is not None:
        return int(forced)
    return int(torch.randint(0, 2 ** 62, (1,), dtype=torch.int64).item())


def _require_cuda(t: torch.Tensor, what: str) -> None:
    if not t.is_cuda:
        raise RuntimeError(f"gantron_b200.{what}: the sm_100a kernels need CUDA tensors; there is no CPU fallback")


def get_mask_from_lengths(lengths: torch.Tensor, max_len=None) -> torch.Tensor:
    """True at valid positions (utils.py:20-24), device-agnostic.  Callers that already know the padded length (the
    time axis of the tensor they mask) pass it as max_len: the reference's max(lengths).item() is a host<->device
    synchronisation in the middle of the training step, and the result is the same mask."""
    if max_len is None:
        max_len = int(torch.max(lengths).item())
    return torch.arange(max_len, device=lengths.device)[None, :] < lengths[:, None]


class LocationLayer(nn.Module):
    """conv1d(2 -> n_filters, k, no bias) + linear(n_filters -> attention_dim, no bias) (model.py:15-31)."""

    def __init__(self, attention_n_filters, attention_kernel_size, attention_dim):
        super().__init__()
        self.location_conv = ConvNorm(2, attention_n_filters, kernel_size=attention_kernel_size,
                                      padding=int((attention_kernel_size - 1) / 2), bias=False, stride=1, dilation=1)
        self.location_dense = LinearNorm(attention_n_filters, attention_dim, bias=False, w_init_gain='tanh')

    def forward(self, attention_weights_cat):
        # stand-alone convenience (the decoder kernels fuse this layer; see gtn_decoder_fwd2.cu phase B)
        return self.location_dense(self.location_conv(attention_weights_cat).transpose(1, 2))


class Attention(nn.Module):
    """Location-sensitive attention parameters (model.py:34-91)."""

    def __init__(self, attention_rnn_dim, embedding_dim, attention_dim, attention_location_n_filters,
                 attention_location_kernel_size):
        super().__init__()
        self.query_layer = LinearNorm(attention_rnn_dim, attention_dim, bias=False, w_init_gain='tanh')
        self.memory_layer = LinearNorm(embedding_dim, attention_dim, bias=False, w_init_gain='tanh')
        self.v = LinearNorm(attention_dim, 1, bias=False)
        self.location_layer = LocationLayer(attention_location_n_filters, attention_location_kernel_size, attention_dim)
        self.score_mask_value = -float("inf")   # written by train.py:119; padded positions always get weight 0

    def get_alignment_energies(self, query, processed_memory, attention_weights_cat):
        q = self.query_layer(query.unsqueeze(1))
        loc = self.location_layer(attention_weights_cat)
        return self.v(torch.tanh(q + loc + processed_memory)).squeeze(-1)

    def forward(self, attention_hidden_state, memory, processed_memory, attention_weights_cat, mask):
        # stand-alone convenience API; Decoder never calls it (the step is fused in the persistent kernel)
        e = self.get_alignment_energies(attention_hidden_state, processed_memory, attention_weights_cat)
        if mask is not None:
            e.data.masked_fill_(mask, self.score_mask_value)
        w = F.softmax(e, dim=1)
        return torch.bmm(w.unsqueeze(1), memory).squeeze(1), w


class Prenet(nn.Module):
    """Two bias-free linears, relu + dropout(0.5) that is on even in eval() (model.py:94-105)."""

    def __init__(self, in_dim, sizes):
        super().__init__()
        self.layers = nn.ModuleList([LinearNorm(i, o, bias=False) for i, o in zip([in_dim] + sizes[:-1], sizes)])
        self._prec = _precision_of(None)       # the owning Decoder overwrites this with its hparams' precision

    def forward(self, x):
        from . import decoder_ops
        _require_cuda(x, "Prenet")
        return decoder_ops.prenet_forward(self, x, seed=_new_seed(self), prec=self._prec)


class Postnet(nn.Module):
    """Five conv1d(k=5) + BatchNorm1d, tanh on all but the last, dropout 0.5 in training (model.py:108-151)."""

    def __init__(self, hparams):
        super().__init__()
        n, e, k = hparams.n_mel_channels, hparams.postnet_embedding_dim, hparams.postnet_kernel_size
        pad = int((k - 1) / 2)
        chans = [n] + [e] * (hparams.postnet_n_convolutions - 1) + [n]
        self.convolutions = nn.ModuleList()
        for i in range(hparams.postnet_n_convolutions):
            last = i == hparams.postnet_n_convolutions - 1
            self.convolutions.append(nn.Sequential(
                ConvNorm(chans[i], chans[i + 1], kernel_size=k, stride=1, padding=pad, dilation=1,
                         w_init_gain='linear' if last else 'tanh'),
                nn.BatchNorm1d(chans[i + 1])))
        self._prec = _precision_of(hparams)

    def forward(self, x, wgrad_input=None):
        from . import postnet_ops
        _require_cuda(x, "Postnet")
        return postnet_ops.postnet_forward(self, x, seed=_new_seed(self), prec=self._prec, wgrad_input=wgrad_input)


class Encoder(nn.Module):
    """Text encoder: 3 x (conv k5 + BN + relu + dropout) + BiLSTM (model.py:154-224), SURVEY 8f1.
    On CUDA tensors the convolution bank runs on the implicit-GEMM conv + fused BN / ReLU / dropout kernels and (bf16 mode,
    256 units per direction) the BiLSTM on the persistent kernel of gtn_bilstm.cu (encoder_ops.py); packed-sequence
    semantics without the reference's `input_lengths.cpu().numpy()` host sync (model.py:202).  CPU tensors, fp32 mode or
    other LSTM sizes take the stock PyTorch ops of the reference (the encoder is not part of the north star's hot path)."""

    def __init__(self, hparams):
        super().__init__()
        self.noise_size = hparams.noise_size if hparams.encoder_inputs else 0
        self.fp16 = hparams.fp16_run
        dim, k = hparams.encoder_embedding_dim, hparams.encoder_kernel_size
        convs = []
        for i in range(hparams.encoder_n_convolutions):
            in_dim = dim
            if i == 0 and hparams.encoder_inputs:
                in_dim += (hparams.n_labels if hparams.use_labels else 0) + (hparams.noise_size if hparams.use_noise else 0)
            convs.append(nn.Sequential(
                ConvNorm(in_dim, dim, kernel_size=k, stride=1, padding=int((k - 1) / 2), dilation=1, w_init_gain='relu'),
                nn.BatchNorm1d(dim)))
        self.convolutions = nn.ModuleList(convs)
        self.lstm = nn.LSTM(dim, int(dim / 2), 1, batch_first=True, bidirectional=True)
        self._prec = _precision_of(hparams)
        self._gtn_act, self._gtn_site0 = "relu", 60        # conv stack: relu, dropout sites 60.. (postnet_ops._PostnetFn)

    def append_noise(self, x, noise=None):
        if self.noise_size > 0:
            if noise is None:   # CPU draw then move, like the reference (reproducible under torch.manual_seed)
                noise = torch.rand(x.size(0), self.noise_size, 1).repeat_interleave(x.size(2), dim=2).to(x.device)
            x = torch.cat([x, noise.to(x.dtype)], dim=1)
        return x

    def _use_kernels(self, x):
        return x.is_cuda and os.environ.get("GANTRON_B200_ENCODER", "1") != "0"

    def _convs(self, x):
        if self._use_kernels(x):
            from . import postnet_ops
            return postnet_ops.postnet_forward(self, x, seed=_new_seed(self), prec=self._prec).transpose(1, 2)
        for conv in self.convolutions:
            x = F.dropout(F.relu(conv(x)), 0.5, self.training)
        return x.transpose(1, 2)

    def _lstm(self, x, lengths):
        """x (B,T,E) -> (B,T,2H), zero beyond `lengths` (None: every position is valid)."""
        from . import encoder_ops
        if self._use_kernels(x) and encoder_ops.bilstm_supported(self.lstm, self._prec):
            if lengths is None:
                lengths = torch.full((x.size(0),), x.size(1), device=x.device, dtype=torch.int32)
            return encoder_ops.bilstm_forward(self.lstm, x, lengths)
        x = x.float()
        self.lstm.flatten_parameters()
        if lengths is None:
            return self.lstm(x)[0]
        packed = nn.utils.rnn.pack_padded_sequence(x, lengths.cpu().numpy(), batch_first=True)
        out, _ = self.lstm(packed)
        return nn.utils.rnn.pad_packed_sequence(out, batch_first=True)[0]

    def forward(self, x, input_lengths):
        return self._lstm(self._convs(self.append_noise(x)), input_lengths)

    def inference(self, x, style=None):
        return self._lstm(self._convs(self.append_noise(x, style)), None)


class Decoder(nn.Module):
    """Tacotron2 decoder (model.py:227-497).  Parameters as in the reference; both loops run as a single
    persistent clustered sm_100a kernel (gtn_decoder_fwd2.cu / gtn_decoder_bwd2.cu)."""

    def __init__(self, hparams):
        super().__init__()
        hp = hparams
        self.n_mel_channels = hp.n_mel_channels
        self.n_frames_per_step = hp.n_frames_per_step
        if hp.n_frames_per_step != 1:
            raise NotImplementedError("n_frames_per_step=1 only (as the reference: hparams.py:71)")
        self.encoder_embedding_dim = hp.encoder_embedding_dim
        self.attention_rnn_dim = hp.attention_rnn_dim
        self.decoder_rnn_dim = hp.decoder_rnn_dim
        self.prenet_dim = hp.prenet_dim
        self.max_decoder_steps = hp.max_decoder_steps
        self.gate_threshold = hp.gate_threshold
        self.p_attention_dropout = hp.p_attention_dropout
        self.p_decoder_dropout = hp.p_decoder_dropout
        self.noise_size = 0 if hp.encoder_inputs else hp.noise_size
        self.fp16 = hp.fp16_run
        d_mem = hp.encoder_embedding_dim + self.noise_size          # model.py:242-246
        if hp.vesus_path:
            d_mem += hp.speakers_embedding
            if hp.use_labels and not hp.encoder_inputs:
                d_mem += 5
        self.decoder_in_dim = d_mem
        self.prenet = Prenet(hp.n_mel_channels * hp.n_frames_per_step, [hp.prenet_dim, hp.prenet_dim])
        self.attention_rnn = nn.LSTMCell(hp.prenet_dim + d_mem, hp.attention_rnn_dim)
        self.attention_layer = Attention(hp.attention_rnn_dim, d_mem, hp.attention_dim,
                                         hp.attention_location_n_filters, hp.attention_location_kernel_size)
        self.decoder_rnn = nn.LSTMCell(hp.attention_rnn_dim + d_mem, hp.decoder_rnn_dim, 1)
        self.linear_projection = LinearNorm(hp.decoder_rnn_dim + d_mem, hp.n_mel_channels * hp.n_frames_per_step)
        self.gate_layer = LinearNorm(hp.decoder_rnn_dim + d_mem, 1, bias=True, w_init_gain='sigmoid')
        self._prec = _precision_of(hp)
        self.prenet._prec = self._prec
        self._attention_dim = hp.attention_dim
        self._loc = (hp.attention_location_n_filters, hp.attention_location_kernel_size)
        self._packed = None
        self.last_lengths = None      # per-row frame counts of the last inference() (batch extension)

    # ---- kernel-side views of the parameters -------------------------------------------------
    def _dims(self):
        from .decoder_ops import DecoderDims
        return DecoderDims(n_mel=self.n_mel_channels, prenet=self.prenet_dim, att_rnn=self.attention_rnn_dim,
                           dec_rnn=self.decoder_rnn_dim, att_dim=self._attention_dim, loc_filters=self._loc[0],
                           loc_kernel=self._loc[1], enc_dim=self.decoder_in_dim - self._const_cols(),
                           d_mem=self.decoder_in_dim)

    def _const_cols(self):
        # columns appended after the encoder outputs (speaker, emotions, noise): constant over encoder time
        return self.decoder_in_dim - self.encoder_embedding_dim

    def packed(self):
        from .decoder_ops import PackedDecoder
        if self._packed is None or self._packed.prec != self._prec:
            self._packed = PackedDecoder(self, self._dims(), self._prec)
        elif self._packed.stale(self):
            self._packed.refresh(self)
        return self._packed

    def append_noise(self, x):
        """cat a per-sequence U(0,1) noise vector to every encoder position (model.py:273-279); drawn on the
        CPU generator like the reference so seeds reproduce."""
        if self.noise_size > 0:
            noise = torch.rand(x.size(0), 1, self.noise_size)
            if x.is_cuda:
                noise = noise.pin_memory().to(x.device, non_blocking=True)   # no host stall behind the encoder
            x = torch.cat([x, noise.to(x.dtype).expand(-1, x.size(1), -1)], dim=-1)
        return x

    def forward(self, memory, decoder_inputs, memory_lengths):
        """Teacher-forced pass (model.py:419-454) -> mel (B,n_mel,T), gate (B,T), alignments (B,T,T_in)."""
        from . import decoder_ops
        _require_cuda(memory, "Decoder.forward")
        memory = self.append_noise(memory)
        return decoder_ops.decoder_train_forward(self, memory, decoder_inputs, memory_lengths, seed=_new_seed(self))

    def inference(self, memory, style):
        """Autoregressive pass (model.py:456-497) -> mel (B,n_mel,T'), gate (B,T',1), alignments (B,T',T_in).
        Batch > 1 is an extension (the reference raises at model.py:486): rows stop independently, the loop
        ends when all rows have fired or at max_decoder_steps; per-row lengths are left in `self.last_lengths`."""
        from . import decoder_ops
        _require_cuda(memory, "Decoder.inference")
        const_tail = None
        if style is None:
            style = torch.rand(memory.size(0), 1, self.noise_size).repeat_interleave(memory.size(1), dim=1).to(memory.device)
        if self.noise_size > 0 or style.numel() > 0:
            memory = torch.cat([memory, style.to(memory.dtype)], dim=-1)
        mel, gate, align, lengths, steps = decoder_ops.decoder_inference(
            self, memory, max_steps=self.max_decoder_steps, gate_threshold=self.gate_threshold,
            seed=_new_seed(self), prec=self._prec, const_tail=const_tail)
        self.last_lengths = lengths
        if steps == self.max_decoder_steps and bool((lengths >= steps).any()) and not bool(
                (torch.sigmoid(gate[:, -1, 0]) > self.gate_threshold).all()):
            print("Warning! Reached max decoder steps")
        return mel, gate, align


class Discriminator(nn.Module):
    """Conv discriminator over 20-frame windows (model.py:500-540)."""

    def __init__(self, hparams):
        super().__init__()
        self.window = hparams.discriminator_window
        self.in_dim = self.window * hparams.n_mel_channels
        dd = hparams.discriminator_dim
        first_hidden = min((int(self.in_dim / dd) + 1) * dd, 1024)
        self.discriminator = nn.Sequential(
            DiscConv1d(self.in_dim, first_hidden, 5),
            DiscConv1d(first_hidden, dd, 5, dilation=2),
            DiscConv1d(dd, dd, 5, dilation=2),
            DiscConv1d(dd, hparams.n_mel_channels, 5, dilation=2),
            nn.Conv1d(hparams.n_mel_channels, 1, 1, bias=True))
        self._prec = _precision_of(hparams)
        self._gp_lambda = getattr(hparams, "gradient_penalty_lambda", 0)     # > 0: gradient_penalty calls get a twice-
                                                                             # differentiable composite (disc_ops)

    def forward(self, inputs):
        """inputs (B, frames, n_mel) -> (B, positions)."""
        from . import disc_ops
        _require_cuda(inputs, "Discriminator.forward")
        return disc_ops.discriminator_forward(self, inputs, seed=_new_seed(self), prec=self._prec)

    def adversarial_loss(self, inputs, target_length):
        """mean over the batch of the mean of the first ceil(len/window) outputs (model.py:529-540);
        one masked-mean kernel instead of B host syncs."""
        from . import disc_ops
        _require_cuda(inputs, "Discriminator.adversarial_loss")
        out = self.forward(inputs.transpose(1, 2))
        return disc_ops.masked_window_mean(out, target_length, self.window)


class LinearDiscriminator(nn.Module):
    """MLP discriminator over random overlapping windows (model.py:543-583; `discriminator_type='linear'`).
    Same kernels as the conv Discriminator: a DiscDense (Linear -> Dropout -> Tanh) is a one-tap layer of disc_ops._DiscFn."""

    def __init__(self, hparams):
        super().__init__()
        dd = hparams.discriminator_dim
        self.discriminator = nn.Sequential(DiscDense(hparams.discriminator_window * hparams.n_mel_channels, dd),
                                           DiscDense(dd, dd), DiscDense(dd, dd), nn.Linear(dd, 1))
        self.n_mel_channels = hparams.n_mel_channels
        self.windows_size = hparams.discriminator_window
        self.max_window_overlap = 6
        self._prec = _precision_of(hparams)
        self._gtn_site0 = 40                                               # disc_ops.SITE_LIN0

    def forward(self, inputs):
        """inputs (rows, window * n_mel) -> (rows, 1)."""
        from . import disc_ops
        _require_cuda(inputs, "LinearDiscriminator.forward")
        return disc_ops.linear_discriminator_forward(self, inputs, seed=_new_seed(self), prec=self._prec)

    def adversarial_loss(self, inputs, target_length):
        from . import disc_ops
        _require_cuda(inputs, "LinearDiscriminator.adversarial_loss")
        return disc_ops.linear_adversarial_loss(self, inputs, target_length, seed=_new_seed(self), prec=self._prec)


class Tacotron2(nn.Module):
    """Generator wrapper (model.py:586-723): embeddings -> Encoder (stock) -> [speaker | emotion] columns ->
    Decoder -> Postnet + residual -> output masking."""

    def __init__(self, hparams):
        super().__init__()
        if not hparams.use_noise:                         # the reference mutates hparams here (model.py:589-591)
            hparams.noise_size = 0
        hparams.use_labels = hparams.use_labels if hparams.vesus_path else False
        self.mask_padding = hparams.mask_padding
        self.fp16_run = hparams.fp16_run
        self.n_mel_channels = hparams.n_mel_channels
        self.n_frames_per_step = hparams.n_frames_per_step
        self.encoder_inputs = hparams.encoder_inputs
        self.embedding = nn.Embedding(hparams.n_symbols, hparams.symbols_embedding_dim)
        val = sqrt(3.0) * sqrt(2.0 / (hparams.n_symbols + hparams.symbols_embedding_dim))
        self.embedding.weight.data.uniform_(-val, val)
        self.vesus = False
        self.use_labels = hparams.use_labels
        if hparams.vesus_path:
            self.vesus = True
            self.speaker_embedding = nn.Embedding(123, hparams.speakers_embedding)
            self.speaker_embedding.weight.data.uniform_(-val, val)
        self.encoder = Encoder(hparams)
        self.decoder = Decoder(hparams)
        self.postnet = Postnet(hparams)

    def parse_batch(self, batch):
        """Collate 7-tuple -> ((text, in_len, mel, max_len, speaker, emotions, out_len), (mel, gate)) on the GPU
        (model.py:617-630)."""
        text, in_len, mel, gate, spk, emo, out_len = batch
        dev = torch.device("cuda") if torch.cuda.is_available() else torch.device("cpu")
        g = lambda t: t.contiguous().to(dev, non_blocking=True)
        text, in_len, spk, out_len = g(text).long(), g(in_len).long(), g(spk).long(), g(out_len).long()
        mel, gate, emo = g(mel).float(), g(gate).float(), g(emo).float()
        max_len = torch.max(in_len.data).item()
        return (text, in_len, mel, max_len, spk, emo, out_len), (mel, gate)

    def parse_output(self, outputs, output_lengths=None):
        """Padding frames -> mel 0 / gate 1e3, written through .data like the reference (model.py:632-642):
        values change, the autograd graph does not see it."""
        if self.mask_padding and output_lengths is not None:
            pad = ~get_mask_from_lengths(output_lengths, outputs[0].shape[-1])
            outputs[0].data.masked_fill_(pad[:, None, :], 0.0)
            outputs[1].data.masked_fill_(pad[:, None, :], 0.0)
            outputs[2].data.masked_fill_(pad, 1e3)
        return outputs

    def _append_conditioning(self, enc, speaker_ids, emotions):
        if self.vesus:
            spk = self.speaker_embedding(speaker_ids)[:, None].expand(-1, enc.size(1), -1)
            if self.use_labels and not self.encoder_inputs:
                emo = emotions[:, None].expand(-1, enc.size(1), -1)
                return torch.cat((enc, spk, emo), dim=2)
            return torch.cat((enc, spk), dim=2)
        return enc

    def forward(self, inputs):
        text, text_lengths, mels, max_len, speaker_ids, emotions, output_lengths = inputs
        text_lengths, output_lengths = text_lengths.data, output_lengths.data
        emb = self.embedding(text).transpose(1, 2)
        if self.encoder_inputs and self.use_labels:
            emb = torch.cat((emb, emotions[:, :, None].expand(-1, -1, emb.size(-1))), dim=1)
        enc = self._append_conditioning(self.encoder(emb, text_lengths), speaker_ids, emotions)
        mel, gate, align = self.decoder(enc, mels, memory_lengths=text_lengths)
        # the reference's conv0 weight gradient sees the MASKED mel
        wgrad_in = None
        if self.mask_padding:
            pad = ~get_mask_from_lengths(output_lengths, mel.shape[-1])
            wgrad_in = mel.detach().masked_fill(pad[:, None, :], 0.0)
        mel_post = mel + self.postnet(mel, wgrad_input=wgrad_in)
        return self.parse_output([mel, mel_post, gate, align], output_lengths)

    def inference(self, text_inputs, style=None, emotions=None, speaker=None):
        """-> [mel, mel_postnet, gate (B,T,1), alignments] (model.py:675-723)."""
        dev = self.embedding.weight.device
        if isinstance(text_inputs, str):
            from text import text_to_sequence     # the reference's text front-end (out of scope, used as-is)
            import numpy as np
            seq = np.array(text_to_sequence(text_inputs, ['english_cleaners']))[None, :]
            text_inputs = torch.from_numpy(seq).to(dev).long()
        if self.use_labels and emotions is None:
            emotions = torch.rand(1, 5).to(dev)
        emb = self.embedding(text_inputs).transpose(1, 2)
        if self.encoder_inputs and emotions is not None:
            emb = torch.cat((emb, emotions[:, :, None].expand(-1, -1, emb.size(-1))), dim=1)
        enc = self.encoder.inference(emb, style)
        if self.encoder_inputs:
            style = None
        if speaker is not None and hasattr(self, 'speaker_embedding'):
            enc = torch.cat((enc, self.speaker_embedding(speaker)[:, None].expand(-1, enc.size(1), -1)), dim=2)
        if emotions is not None and not self.encoder_inputs:
            enc = torch.cat((enc, emotions[:, None].expand(-1, enc.size(1), -1)), dim=2)
        mel, gate, align = self.decoder.inference(enc, style)
        mel_post = mel + self.postnet(mel)
        return self.parse_output([mel, mel_post, gate, align])
