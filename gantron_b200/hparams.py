"""Hyper-parameter bag with the reference's field names and defaults (/root/reference/hparams.py:7-149).

The reference's own `HParams` object works unchanged with gantron_b200.model (it is duck-typed); this
class exists so that tests, smoke() and bench.py can build the same configuration on a machine where the
reference tree is absent.  Extra, optional B200 fields (read with getattr, so a reference HParams without
them is fine):
    b200_precision : 'bf16' (default) or 'fp32' (3-plane split, fp32-accurate parity mode)
"""
from __future__ import annotations

import ast

N_SYMBOLS = 148   # len(text.symbols) in the reference (probed)

_DEFAULTS = dict(
    version=0.6, epochs=100, iterations=None, iters_per_checkpoint=5000, seed=1234, dynamic_loss_scaling=True,
    fp16_run=False, distributed_run=False, dist_backend="nccl", dist_url="tcp://localhost:54321",
    cudnn_enabled=True, cudnn_benchmark=False,
    ignore_layers=['decoder.attention_rnn.weight_ih', 'decoder.attention_layer.memory_layer.linear_layer.weight',
                   'decoder.decoder_rnn.weight_ih', 'decoder.linear_projection.linear_layer.weight',
                   'decoder.gate_layer.linear_layer.weight'],
    attn_steps=5000, reduce_lr_steps_every=5e4, vesus_path=None, speakers_embedding=64, use_labels=True,
    use_noise=False, use_intended_labels=True,
    load_mel_from_disk=False, text_cleaners=['english_cleaners'], n_labels=5,
    max_wav_value=32768.0, sampling_rate=22050, filter_length=1024, hop_length=256, win_length=1024, n_ftt=1024,
    n_mel_channels=80, mel_fmin=0.0, mel_fmax=8000.0,
    n_symbols=N_SYMBOLS, symbols_embedding_dim=512,
    encoder_kernel_size=5, encoder_n_convolutions=3, encoder_embedding_dim=512,
    n_frames_per_step=1, decoder_rnn_dim=1024, prenet_dim=256, max_decoder_steps=500, gate_threshold=0.5,
    p_attention_dropout=0.1, p_decoder_dropout=0.1,
    attention_rnn_dim=1024, attention_dim=128, attention_location_n_filters=32, attention_location_kernel_size=31,
    postnet_embedding_dim=512, postnet_kernel_size=5, postnet_n_convolutions=5,
    discriminator_window=20, discriminator_dim=512, g_freq=2, d_freq=1, clipping_value=0.001,
    gradient_penalty_lambda=0, noise_size=512, disc_warmp_up=500, discriminator_type='conv', encoder_inputs=False,
    use_saved_learning_rate=False, g_learning_rate=0.001, d_learning_rate=0.0007, weight_decay=1e-6,
    grad_clip_thresh=1.0, batch_size=32, mask_padding=True,
    b200_precision='bf16',
)


class HParams:
    def __init__(self, hparams_string=None, **overrides):
        for k, v in _DEFAULTS.items():
            setattr(self, k, list(v) if isinstance(v, list) else v)
        for k, v in overrides.items():
            setattr(self, k, v)
        if hparams_string:
            self.add_params_string(hparams_string)

    def add_params_string(self, hparams_string):
        """'k=v,k=v' overrides; values go through ast.literal_eval unless they look like paths."""
        for item in hparams_string.split(','):
            key, value = item.split('=')
            if '/' not in value:
                try:
                    value = ast.literal_eval(value)
                except (ValueError, SyntaxError):
                    pass
            setattr(self, key, value)

    def add_param(self, param, value):
        setattr(self, param, value)
