"""Sentence-sharded multi-GPU synthesis: one process per GPU, replicated weights, every rank
synthesises its own contiguous slice of the sentence batch, and the only communication is the gather of the
variable-length mel outputs (NCCL over NVLink on the GPUs, gloo in the CPU tests).

The reference has no distributed code at all; its results depend on batch composition
(`pad_sequence` to the batch max and the batch-global fallback, common_layers.py:23, forward_tacotron.py:176),
so a G-way sharded run equals the reference applied to each shard separately, not to the un-sharded batch.
"""
from typing import Dict, Optional, Tuple

import torch
import torch.distributed as dist


def shard_slice(n: int, rank: int, world: int) -> slice:
    """Contiguous, balanced slice of `n` sentences for `rank` (the first n % world ranks get one more)."""
    base, extra = divmod(n, world)
    start = rank * base + min(rank, extra)
    return slice(start, start + base + (1 if rank < extra else 0))


def mel_lengths(dur: torch.Tensor) -> torch.Tensor:
    """Valid mel frames per sentence from the (clamped) durations `generate()` returns: sum of
    trunc(dur + 0.5) (common_layers.py:21)."""
    return (dur.clamp_min(0) + 0.5).long().sum(dim=1)


_HOST_GROUPS = {}


def _host_group(group):
    """A gloo twin of the default (NCCL) process group for host-side metadata, created on first use by a collective
    call of all ranks; None when `group` is a sub-group or the twin cannot be created (then the metadata travels
    through NCCL and costs a device sync)."""
    if group is not None and group is not dist.group.WORLD:
        return None
    if 'world' not in _HOST_GROUPS:
        try:
            _HOST_GROUPS['world'] = dist.new_group(backend='gloo')
        except Exception:
            _HOST_GROUPS['world'] = None
    return _HOST_GROUPS['world']


def gather_variable_mels(mel: torch.Tensor, lens: torch.Tensor, group=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """All-gather `[b_local, C, M_local]` mels whose batch and frame extents differ per rank.

    The extents (b_local, M_local) of every shard are host-known shapes: they are exchanged over a host-side (gloo)
    group, so the GPU stream is never drained -- the NCCL all_gather of the zero-padded `[b_max, C, M_max]` blocks (plus
    the lengths) is enqueued behind the still-running frame program of `generate()`.  Returns (`[B_total, C, M_max]`,
    `[B_total]` int64) with the sentences in rank order, i.e. the order of the un-sharded batch under `shard_slice`.
    Frames beyond a shard's own M_local are zero (the shard never computed them).
    """
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return mel, lens.to(torch.int64)
    world = dist.get_world_size(group)
    dev = mel.device
    b_local, C, m_local = mel.shape
    host = _host_group(group) if mel.is_cuda else group
    if host is not None or not mel.is_cuda:
        meta = torch.tensor([b_local, m_local], dtype=torch.int64)
        metas = torch.empty((world, 2), dtype=torch.int64)
        _all_gather(metas, meta, host)
    else:
        meta = torch.tensor([b_local, m_local], dtype=torch.int64, device=dev)
        metas = torch.empty((world, 2), dtype=torch.int64, device=dev)
        _all_gather(metas, meta, group)
        metas = metas.cpu()                              # sub-group without a host twin: one device sync
    counts = [int(v) for v in metas[:, 0]]
    b_max, m_max = max(counts), int(metas[:, 1].max())
    uniform = all(c == b_max for c in counts) and bool((metas[:, 1] == m_max).all())
    if uniform and mel.is_contiguous():                  # equal shards (the weak-scaling case): no padding copies
        block, lens_block = mel, lens.to(torch.int64)
    else:
        block = torch.zeros((b_max, C, m_max), dtype=mel.dtype, device=dev)
        block[:b_local, :, :m_local] = mel
        lens_block = torch.zeros((b_max,), dtype=torch.int64, device=dev)
        lens_block[:b_local] = lens.to(torch.int64)
    blocks = torch.empty((world, b_max, C, m_max), dtype=mel.dtype, device=dev)
    lens_all = torch.empty((world, b_max), dtype=torch.int64, device=dev)
    _all_gather(blocks, block, group)
    _all_gather(lens_all, lens_block, group)
    if uniform:
        return blocks.view(world * b_max, C, m_max), lens_all.view(-1)
    out = torch.cat([blocks[r, :counts[r]] for r in range(world)], dim=0)
    out_lens = torch.cat([lens_all[r, :counts[r]] for r in range(world)], dim=0)
    return out, out_lens


def _all_gather(dst: torch.Tensor, src: torch.Tensor, group) -> None:
    """dst[r] = rank r's src.  NCCL: one all_gather_into_tensor; gloo (CPU tests): the list form."""
    if src.is_cuda:
        dist.all_gather_into_tensor(dst, src, group=group)
        return
    parts = [torch.empty_like(src) for _ in range(dst.shape[0])]
    dist.all_gather(parts, src, group=group)
    for r, p in enumerate(parts):
        dst[r].copy_(p)


def generate_sharded(model, x: torch.Tensor, speaker_emb: Optional[torch.Tensor] = None, group=None,
                     **generate_kwargs) -> Dict[str, torch.Tensor]:
    """Run `model.generate` on this rank's slice of the global batch `x` [B, T] and gather the results.

    Every rank passes the same global `x` (and `speaker_emb` [B, S] for the multi-speaker models); returns
    `{'mel_post': [B, n_mels, M], 'mel': ..., 'mel_lens': [B]}` on every rank.  With no process group this is a
    plain `generate` call on the whole batch.
    """
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    sl = shard_slice(x.shape[0], rank, world)
    args = (x[sl],) if speaker_emb is None else (x[sl], speaker_emb[sl])
    if sl.stop > sl.start:
        out = model.generate(*args, **generate_kwargs)
        lens = mel_lengths(out['dur'])
        mel, mel_post = out['mel'], out['mel_post']
    else:                                  # more ranks than sentences: contribute an empty block
        n_mels = getattr(model, 'n_mels', 80)
        mel = mel_post = torch.zeros((0, n_mels, 0), device=x.device)
        lens = torch.zeros((0,), dtype=torch.int64, device=x.device)
    post_all, lens_all = gather_variable_mels(mel_post, lens, group)
    mel_all, _ = gather_variable_mels(mel, lens, group) if mel is not mel_post else (post_all, lens_all)
    return {'mel': mel_all, 'mel_post': post_all, 'mel_lens': lens_all}
