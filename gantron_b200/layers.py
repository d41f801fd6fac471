"""Parameter containers whose attribute names define the reference's state_dict keys.

Same public names / constructor arguments as /root/reference/layers.py:9-73 (LinearNorm, ConvNorm,
DiscConv1d, DiscDense) so that checkpoints load with strict=True (train.py:139,147).  The hot path never
calls these modules' forward(): Decoder / Postnet / Discriminator read the .weight / .bias Parameters and
run the sm_100a kernels; forward() here exists for callers that use a layer on its own (e.g. the stock
Encoder, which is out of scope and stays on PyTorch).
"""
from __future__ import annotations

import torch
from torch import nn


def _xavier(weight: torch.Tensor, gain_name: str) -> None:
    nn.init.xavier_uniform_(weight, gain=nn.init.calculate_gain(gain_name))


class LinearNorm(nn.Module):
    """nn.Linear under `.linear_layer` with Xavier-uniform init (layers.py:9-19)."""

    def __init__(self, in_dim, out_dim, bias=True, w_init_gain='linear'):
        super().__init__()
        self.linear_layer = nn.Linear(in_dim, out_dim, bias=bias)
        _xavier(self.linear_layer.weight, w_init_gain)

    def forward(self, x):
        return self.linear_layer(x)


class ConvNorm(nn.Module):
    """nn.Conv1d under `.conv` with Xavier-uniform init; 'same' padding by default (layers.py:22-40)."""

    def __init__(self, in_channels, out_channels, kernel_size=1, stride=1, padding=None, dilation=1, bias=True,
                 w_init_gain='linear'):
        super().__init__()
        if padding is None:
            assert kernel_size % 2 == 1
            padding = int(dilation * (kernel_size - 1) / 2)
        self.conv = nn.Conv1d(in_channels, out_channels, kernel_size=kernel_size, stride=stride, padding=padding,
                              dilation=dilation, bias=bias)
        _xavier(self.conv.weight, w_init_gain)

    def forward(self, signal):
        return self.conv(signal)


class DiscConv1d(nn.Module):
    """Conv1d -> Dropout -> Tanh under `.module` (layers.py:43-60); note the order."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=None, dilation=1, bias=True,
                 dropout=0.5):
        super().__init__()
        if padding is None:
            assert kernel_size % 2 == 1
            padding = int(dilation * (kernel_size - 1) / 2)
        conv = nn.Conv1d(in_channels, out_channels, kernel_size=kernel_size, stride=stride, padding=padding,
                         dilation=dilation, bias=bias)
        _xavier(conv.weight, 'tanh')
        self.module = nn.Sequential(conv, nn.Dropout(dropout), nn.Tanh())

    def forward(self, inputs):
        return self.module(inputs)


class DiscDense(nn.Module):
    """Linear -> Dropout -> Tanh under `.module` (layers.py:63-73); used by LinearDiscriminator only."""

    def __init__(self, in_dim, out_dim, bias=True, dropout=0.5):
        super().__init__()
        self.module = nn.Sequential(nn.Linear(in_dim, out_dim, bias=bias), nn.Dropout(dropout), nn.Tanh())

    def forward(self, inputs):
        return self.module(inputs)
