"""Parameter containers mirroring the reference's `models/common_layers.py` surface.

Same class names, constructor arguments and attribute (=> state_dict) names as the
reference (common_layers.py:12-223) so checkpoints load unchanged; the `forward` methods
here are stock-PyTorch (used by the un-accelerated training `forward(batch)` path, SURVEY.md
§3.4), while inference goes through the sm_100a programs in `forwardtacotron_b200.engine`.
"""
import math
from typing import List, Optional

import torch
import torch.nn as nn
import torch.nn.functional as F


class LengthRegulator(nn.Module):
    """Training-path (stock torch) expansion; inference uses ftb_lr_scan / ftb_lr_gather."""

    def forward(self, x: torch.Tensor, dur: torch.Tensor) -> torch.Tensor:
        dur.clamp_(min=0.)                                       # in place, like the reference (:18)
        reps = (dur + 0.5).long()
        rows = [torch.repeat_interleave(x[b], reps[b], dim=0) for b in range(x.size(0))]
        return nn.utils.rnn.pad_sequence(rows, batch_first=True, padding_value=0.)


class HighwayNetwork(nn.Module):

    def __init__(self, size: int) -> None:
        super().__init__()
        self.W1 = nn.Linear(size, size)
        self.W2 = nn.Linear(size, size)
        nn.init.zeros_(self.W1.bias)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        gate = torch.sigmoid(self.W2(x))
        return gate * F.relu(self.W1(x)) + (1. - gate) * x


class BatchNormConv(nn.Module):
    """conv (no bias, padding k//2) -> optional ReLU -> BatchNorm1d."""

    def __init__(self, in_channels: int, out_channels: int, kernel: int, relu=True) -> None:
        super().__init__()
        self.conv = nn.Conv1d(in_channels, out_channels, kernel, stride=1, padding=kernel // 2, bias=False)
        self.bnorm = nn.BatchNorm1d(out_channels)
        self.relu = relu

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        y = self.conv(x)
        if self.relu is True:
            y = F.relu(y)
        return self.bnorm(y)


class CBHG(nn.Module):

    def __init__(self, K: int, in_channels: int, channels: int, proj_channels: list,
                 num_highways: int, dropout: float = 0.5) -> None:
        super().__init__()
        self.dropout = dropout
        self.bank_kernels = list(range(1, K + 1))
        self.conv1d_bank = nn.ModuleList([BatchNormConv(in_channels, channels, k) for k in self.bank_kernels])
        self.maxpool = nn.MaxPool1d(kernel_size=2, stride=1, padding=1)
        self.conv_project1 = BatchNormConv(K * channels, proj_channels[0], 3)
        self.conv_project2 = BatchNormConv(proj_channels[0], proj_channels[1], 3, relu=False)
        self.pre_highway = nn.Linear(proj_channels[-1], channels, bias=False)
        self.highways = nn.ModuleList([HighwayNetwork(channels) for _ in range(num_highways)])
        self.rnn = nn.GRU(channels, channels, batch_first=True, bidirectional=True)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        n = x.size(-1)
        y = torch.cat([conv(x)[..., :n] for conv in self.conv1d_bank], dim=1)
        y = self.maxpool(y)[..., :n]
        y = F.dropout(y, p=self.dropout, training=self.training)
        y = F.dropout(self.conv_project1(y), p=self.dropout, training=self.training)
        y = self.conv_project2(y) + x
        y = self.pre_highway(y.transpose(1, 2))
        for hw in self.highways:
            y = hw(y)
        return self.rnn(y)[0]


class PositionalEncoding(nn.Module):

    def __init__(self, d_model: int, dropout=0.1, max_len=5000) -> None:
        super().__init__()
        self.dropout = nn.Dropout(p=dropout)
        self.scale = nn.Parameter(torch.ones(1))
        pos = torch.arange(max_len, dtype=torch.float).unsqueeze(1)
        freq = torch.exp(torch.arange(0, d_model, 2).float() * (-math.log(10000.0) / d_model))
        table = torch.zeros(max_len, d_model)
        table[:, 0::2] = torch.sin(pos * freq)
        table[:, 1::2] = torch.cos(pos * freq)
        self.register_buffer('pe', table.unsqueeze(1))          # [max_len, 1, d_model]

    def forward(self, x: torch.Tensor) -> torch.Tensor:        # x [T, N, E]
        return self.dropout(x + self.scale * self.pe[:x.size(0)])


class FFTBlock(nn.Module):

    def __init__(self, d_model: int, nhead: int, conv1_kernel: int, conv2_kernel: int, d_fft: int,
                 dropout: float = 0.1):
        super().__init__()
        self.self_attn = nn.MultiheadAttention(d_model, nhead, dropout=dropout)
        self.dropout = nn.Dropout(dropout)
        self.conv1 = nn.Conv1d(d_model, d_fft, conv1_kernel, stride=1, padding=conv1_kernel // 2)
        self.conv2 = nn.Conv1d(d_fft, d_model, conv2_kernel, stride=1, padding=conv2_kernel // 2)
        self.norm1 = nn.LayerNorm(d_model)
        self.norm2 = nn.LayerNorm(d_model)
        self.dropout1 = nn.Dropout(dropout)
        self.dropout2 = nn.Dropout(dropout)
        self.activation = nn.ReLU()

    def forward(self, src: torch.Tensor, src_pad_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        att = self.self_attn(src, src, src, attn_mask=None, key_padding_mask=src_pad_mask)[0]
        src = self.norm1(src + self.dropout1(att))
        y = src.permute(1, 2, 0)                                  # [N, E, T]
        y = self.conv2(self.activation(self.conv1(y)))
        return self.norm2(src + self.dropout2(y.permute(2, 0, 1)))


class ForwardTransformer(nn.Module):

    def __init__(self, d_model: int, d_fft: int, layers: int, heads: int, conv1_kernel: int,
                 conv2_kernel: int, dropout: float = 0.1) -> None:
        super().__init__()
        self.d_model = d_model
        self.heads = heads
        self.pos_encoder = PositionalEncoding(d_model, dropout)
        self.layers = nn.ModuleList([FFTBlock(d_model=d_model, nhead=heads, d_fft=d_fft,
                                              conv1_kernel=conv1_kernel, conv2_kernel=conv2_kernel,
                                              dropout=dropout) for _ in range(layers)])
        self.norm = nn.LayerNorm(d_model)

    def forward(self, x: torch.Tensor, src_pad_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        y = self.pos_encoder(x.transpose(0, 1))
        for layer in self.layers:
            y = layer(y, src_pad_mask=src_pad_mask)
        return self.norm(y).transpose(0, 1)


def make_token_len_mask(x: torch.Tensor) -> torch.Tensor:
    return (x == 0).transpose(0, 1)


def make_mel_len_mask(x: torch.Tensor, mel_lens: torch.Tensor) -> torch.Tensor:
    pos = torch.arange(x.size(1), device=x.device)[None, :]
    return pos >= mel_lens.to(x.device)[:, None]
