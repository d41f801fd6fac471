"""Seeded synthetic LJSpeech-shaped batches in the reference's collate order (data_utils.py:94-131), for bench.py,
smoke() and the tests: no datasets or checkpoints are reachable offline."""
from __future__ import annotations

import torch


def synthetic_batch(B: int, T_in: int, T_out: int, n_mel: int = 80, seed: int = 1234, ragged: bool = False,
                    n_symbols: int = 148):
    """-> (text_padded, input_lengths, mel_padded, gate_padded, speaker_ids, emotions, output_lengths) on the CPU.

    text ids uniform in [1, n_symbols), zero padded; input lengths sorted descending; mel ~ N(-5, 2) clipped to the
    log-mel range [-11.5, 2], zero beyond each length; gate target 1 from the last valid frame on."""
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
