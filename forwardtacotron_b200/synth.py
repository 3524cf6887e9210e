"""Deterministic synthetic weights and inputs, keyed by state_dict entry name.

The reference publishes no checkpoint and pins no numeric golden for the model
path, so every parity fixture is generated from weights that can
be rebuilt bit-identically on any box from (key, shape, seed) alone: the golden
generator (oracle/make_golden.py, run where /root/reference is importable)
loads them into the reference model, and the GPU tests load the very same
tensors into this package's models.  Values are drawn with torch's CPU
generator, which is platform independent.

Unlike default init, BatchNorm running stats / affine and LayerNorm affine are
randomised (a wrong BN epilogue is invisible at the near-identity default,
SURVEY.md §7.3-5) and the duration head can be shifted so that durations spread
over ~[1.5, 4.5] instead of triggering the all-2 fallback
(forward_tacotron.py:176-177).
"""
import zlib
from typing import Dict, Mapping, Tuple

import torch


def _gen(key: str, seed: int) -> torch.Generator:
    g = torch.Generator(device='cpu')
    g.manual_seed((zlib.crc32(key.encode()) ^ (seed * 0x9E3779B1)) & 0x7FFFFFFF)
    return g


def _uniform(shape, lo, hi, g):
    return torch.rand(shape, generator=g, dtype=torch.float32) * (hi - lo) + lo


def synth_tensor(key: str, shape: Tuple[int, ...], dtype: torch.dtype, seed: int) -> torch.Tensor:
    g = _gen(key, seed)
    leaf = key.rsplit('.', 1)[-1]
    if dtype in (torch.int64, torch.int32):
        return torch.zeros(shape, dtype=dtype)
    if leaf == 'running_var':
        return _uniform(shape, 0.5, 1.5, g)
    if leaf == 'running_mean':
        return torch.randn(shape, generator=g) * 0.1
    is_norm = ('bnorm' in key) or ('.norm' in key)
    if is_norm and leaf == 'weight':
        return _uniform(shape, 0.6, 1.4, g)
    if is_norm and leaf == 'bias':
        return torch.randn(shape, generator=g) * 0.1
    if leaf == 'scale':                       # PositionalEncoding.scale
        return _uniform(shape, 0.7, 1.3, g)
    if 'embedding' in key:
        return torch.randn(shape, generator=g)
    if len(shape) >= 2:
        fan_in = 1
        for s in shape[1:]:
            fan_in *= int(s)
        a = (3.0 / fan_in) ** 0.5             # unit-gain uniform
        return _uniform(shape, -a, a, g)
    return _uniform(shape, -0.1, 0.1, g)


def synth_state_dict(template: Mapping[str, torch.Tensor], seed: int = 0,
                     spread_durations: bool = True) -> Dict[str, torch.Tensor]:
    """Return a state_dict with the template's keys/shapes/dtypes and synthetic values.

    `pos_encoder.pe` buffers (the fixed sinusoid table, common_layers.py:132-140)
    are copied from the template unchanged.
    """
    out = {}
    for key in sorted(template.keys()):
        ref = template[key]
        if key.endswith('pos_encoder.pe'):
            out[key] = ref.detach().clone()
            continue
        out[key] = synth_tensor(key, tuple(ref.shape), ref.dtype, seed).to(ref.dtype)
    if spread_durations:
        for key in out:
            if key.startswith('dur_pred.lin.'):
                if key.endswith('weight'):
                    out[key] = out[key] * 4.0
                else:
                    out[key] = torch.full_like(out[key], 3.0)
    return out


def synth_tokens(batch: int, length: int, seed: int = 0, num_chars: int = 135,
                 with_pad: bool = False) -> torch.Tensor:
    """Token ids in [1, num_chars) (id 0 is the pad symbol); `with_pad` zero-fills
    ragged tails to exercise FastPitch's key-padding mask (fast_pitch.py:200)."""
    g = _gen(f'tokens/{batch}x{length}', seed)
    x = torch.randint(1, num_chars, (batch, length), generator=g, dtype=torch.int64)
    if with_pad:
        for b in range(batch):
            keep = length - (b * 3) % max(1, length // 2)
            x[b, keep:] = 0
    return x


def synth_speaker_emb(batch: int, dims: int = 256, seed: int = 0) -> torch.Tensor:
    g = _gen(f'semb/{batch}x{dims}', seed)
    e = torch.randn((batch, dims), generator=g)
    return e / e.norm(dim=1, keepdim=True)


def synth_batch(batch: int, length: int, seed: int = 0, n_mels: int = 80, speaker_dims: int = 0) -> dict:
    """A collated training-style batch (utils/dataset.py:239-263 layout) for `forward(batch)` in eval mode: ragged
    token lengths (zero padded), integer durations (0 on the padding), mel_len = sum of durations, per-token
    pitch / energy, and a mel target whose only used property is its length."""
    g = _gen(f'batch/{batch}x{length}', seed)
    x = torch.randint(1, 135, (batch, length), generator=g, dtype=torch.int64)
    dur = torch.randint(0, 5, (batch, length), generator=g).to(torch.float32)
    for b in range(batch):
        keep = length - (b * 5) % max(1, length // 2)
        x[b, keep:] = 0
        dur[b, keep:] = 0
        dur[b, 0] = max(float(dur[b, 0]), 1.0)
    mel_len = dur.sum(dim=1).long()
    out = {'x': x, 'dur': dur, 'mel_len': mel_len,
           'pitch': torch.randn((batch, length), generator=g), 'energy': torch.randn((batch, length), generator=g),
           'mel': torch.zeros((batch, n_mels, int(mel_len.max()) + 3))}
    if speaker_dims:
        out['speaker_emb'] = synth_speaker_emb(batch, speaker_dims, seed)
        # utils/dataset.py:141-142: 1 = unvoiced (zero pitch), 2 = voiced, 0 = padding
        cond = torch.randint(1, 3, (batch, length), generator=g, dtype=torch.int64)
        cond[x == 0] = 0
        out['pitch_cond'] = cond
    return out
