"""Batched caller of `generate()`: the reference CLI synthesises one sentence per call
(`gen_forward.py:88-116`); this module runs lists of tokenised sentences through the batch path instead:

    sort by length -> buckets of at most `max_batch` sentences whose lengths differ by at most `max_len_ratio`
    -> zero-pad (token id 0 = padding, utils/text/symbols.py:21) -> model.generate(...) -> per-sentence un-padding
    -> `[1, n_mels, M_i]` fp32 arrays in the caller's order (the `.npy` layout of gen_forward.py:112-113).

Parity note: the reference applies no padding mask in `generate()` -- the BiGRU/BiLSTM run over the
padded tail and the length regulator pads to the longest sentence of the batch -- so a sentence synthesised inside a
padded batch is NOT bit-identical to the same sentence synthesised alone, in the reference as much as here.  The
oracle for a bucket is the reference applied to that padded bucket; bucketing by length bounds the effect, and
`max_len_ratio=1.0` restricts buckets to sentences of identical token length.
"""
from typing import Callable, Dict, List, Optional, Sequence

import numpy as np
import torch


def bucket_by_length(lengths: Sequence[int], max_batch: int, max_len_ratio: float = 1.25) -> List[List[int]]:
    """Indices grouped into buckets: ascending length, at most `max_batch` per bucket, longest / shortest length
    inside a bucket <= `max_len_ratio`."""
    if max_batch < 1 or max_len_ratio < 1.0:
        raise ValueError('max_batch >= 1 and max_len_ratio >= 1.0 required')
    order = sorted(range(len(lengths)), key=lambda i: (lengths[i], i))
    buckets: List[List[int]] = []
    for i in order:
        if lengths[i] < 1:
            raise ValueError(f'sentence {i} is empty')
        cur = buckets[-1] if buckets else None
        if cur is not None and len(cur) < max_batch and lengths[i] <= max_len_ratio * lengths[cur[0]]:
            cur.append(i)
        else:
            buckets.append([i])
    return buckets


def pad_tokens(seqs: Sequence[Sequence[int]]) -> torch.Tensor:
    """[B, T_max] int64, zero padded (0 is the padding symbol)."""
    T = max(len(s) for s in seqs)
    x = torch.zeros((len(seqs), T), dtype=torch.int64)
    for b, s in enumerate(seqs):
        x[b, :len(s)] = torch.as_tensor(list(s), dtype=torch.int64)
    return x


def own_frames(dur: torch.Tensor, token_lens: Sequence[int]) -> List[int]:
    """Mel frames that belong to each sentence's REAL tokens: sum of trunc(dur + 0.5) over t < len_b
    (common_layers.py:21).  The padded tokens' frames follow them in the regulated sequence and are dropped."""
    r = (dur.detach().clamp_min(0) + 0.5).long().cpu()
    return [int(r[b, :n].sum()) for b, n in enumerate(token_lens)]


def synthesize_batched(model, token_seqs: Sequence[Sequence[int]], max_batch: int = 64,
                       speaker_emb: Optional[torch.Tensor] = None, alpha: float = 1.0,
                       pitch_function: Callable[[torch.Tensor], torch.Tensor] = lambda x: x,
                       energy_function: Callable[[torch.Tensor], torch.Tensor] = lambda x: x,
                       device: Optional[torch.device] = None, max_len_ratio: float = 1.25,
                       key: str = 'mel_post') -> List[np.ndarray]:
    """Synthesise every sentence of `token_seqs` through `model.generate` in length buckets.

    `speaker_emb`: [S] (one speaker for all sentences) or [N, S] (per sentence) for the multi-speaker models.
    Returns one `[1, n_mels, M_i]` fp32 array per sentence, in input order."""
    n = len(token_seqs)
    lengths = [len(s) for s in token_seqs]
    out: List[Optional[np.ndarray]] = [None] * n
    for bucket in bucket_by_length(lengths, max_batch, max_len_ratio):
        x = pad_tokens([token_seqs[i] for i in bucket])
        if device is not None:
            x = x.to(device)
        args = [x]
        if speaker_emb is not None:
            se = speaker_emb if speaker_emb.dim() == 2 else speaker_emb[None, :].expand(n, -1)
            se = se[torch.as_tensor(bucket)].contiguous()
            args.append(se.to(device) if device is not None else se)
        res: Dict[str, torch.Tensor] = model.generate(*args, alpha=alpha, pitch_function=pitch_function,
                                                      energy_function=energy_function)
        mel = res[key]
        frames = own_frames(res['dur'], [lengths[i] for i in bucket])
        mel_host = mel.detach().to('cpu', torch.float32)
        for b, i in enumerate(bucket):
            out[i] = np.ascontiguousarray(mel_host[b:b + 1, :, :frames[b]].numpy())
    return out  # type: ignore[return-value]


def save_npy(path: str, mel: np.ndarray) -> None:
    """The reference CLI's output format: `np.save(..., allow_pickle=False)` of `[1, n_mels, M]` (gen_forward.py:112-113)."""
    np.save(path, mel, allow_pickle=False)
