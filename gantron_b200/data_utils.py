"""The batch producer in front of the path (reference data_utils.py:82-131): `TextMelCollate` with the
reference's semantics writing straight into pinned host buffers, and `DevicePrefetcher`, which uploads batch k + 1 on a
side stream while step k runs.  Together with `Tacotron2.parse_batch` (non-blocking copies from pinned memory) the
training step sees its inputs without a host stall; `bench.py`'s `e2e` leg times exactly this hand-over.
"""
from __future__ import annotations

from typing import Iterable, Iterator, Optional, Sequence, Tuple

import torch


class TextMelCollate:
    """Zero-pads a list of (text ids, mel (n_mel, frames)[, speaker, emotions]) items into one batch, sorted by decreasing
    text length (the encoder's `pack_padded_sequence` contract), frames rounded up to `n_frames_per_step`
    (data_utils.py:82-131).  Returns (text_padded, input_lengths, mel_padded, gate_padded, speaker_ids, emotions,
    output_lengths) with the reference's dtypes.  Differences: the outputs live in pinned memory when CUDA is available
    (`pin_memory=None`) so that `parse_batch` can copy them asynchronously, and `speaker_ids` / `emotions` are zero when the
    items carry none (the reference leaves them uninitialised, data_utils.py:102-103)."""

    def __init__(self, n_frames_per_step: int, pin_memory: Optional[bool] = None):
        self.n_frames_per_step = int(n_frames_per_step)
        self.pin_memory = torch.cuda.is_available() if pin_memory is None else bool(pin_memory)

    def _zeros(self, shape, dtype):
        t = torch.zeros(shape, dtype=dtype)
        return t.pin_memory() if self.pin_memory else t

    def __call__(self, batch: Sequence[Sequence[torch.Tensor]]):
        n = len(batch)
        text_len = torch.tensor([item[0].numel() for item in batch], dtype=torch.long)
        input_lengths, order = torch.sort(text_len, dim=0, descending=True)
        frames = [int(item[1].size(1)) for item in batch]
        t_max = max(frames)
        step = self.n_frames_per_step
        t_max += (-t_max) % step
        n_mel = int(batch[0][1].size(0))
        text = self._zeros((n, int(input_lengths[0])), torch.long)
        mel = self._zeros((n, n_mel, t_max), torch.float32)
        gate = self._zeros((n, t_max), torch.float32)
        speaker = self._zeros((n,), torch.float32)
        emotions = self._zeros((n, 5), torch.float32)
        out_len = self._zeros((n,), torch.long)
        labelled = len(batch[0]) == 4
        for row, src in enumerate(order.tolist()):
            item = batch[src]
            text[row, :item[0].numel()] = item[0]
            f = frames[src]
            mel[row, :, :f] = item[1]
            gate[row, f - 1:] = 1.0                      # the stop target: last real frame and all padding
            out_len[row] = f
            if labelled:
                speaker[row] = float(item[-2])
                emotions[row] = torch.as_tensor(item[-1], dtype=torch.float32)
        if self.pin_memory:
            input_lengths = input_lengths.pin_memory()
        return text, input_lengths, mel, gate, speaker, emotions, out_len


class DevicePrefetcher:
    """Iterates over collated batches and yields `model.parse_batch(batch)` results whose host-to-device copies were issued
    one batch ahead on a side stream.  `for x, y in DevicePrefetcher(loader, generator): ...` replaces
    `for batch in loader: x, y = generator.parse_batch(batch)` (train.py:291-294)."""

    def __init__(self, batches: Iterable, model, device: Optional[torch.device] = None):
        self.batches, self.model = batches, model
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else device
        self.stream = torch.cuda.Stream(device=self.device)

    def _upload(self, batch):
        with torch.cuda.stream(self.stream):
            parsed = self.model.parse_batch(batch)
        return parsed, batch                              # (the pinned source stays alive until the copy is consumed)

    def __iter__(self) -> Iterator[Tuple]:
        it = iter(self.batches)
        try:
            nxt = self._upload(next(it))
        except StopIteration:
            return
        while nxt is not None:
            cur, _src = nxt
            torch.cuda.current_stream(self.device).wait_stream(self.stream)
            for group in cur:
                for t in group:
                    if torch.is_tensor(t):
                        t.record_stream(torch.cuda.current_stream(self.device))
            try:
                nxt = self._upload(next(it))
            except StopIteration:
                nxt = None
            yield cur
