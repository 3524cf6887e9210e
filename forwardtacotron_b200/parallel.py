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


def _meta_exchange(b_local: int, m_local: int, lens: torch.Tensor, b_cap: int, group) -> torch.Tensor:
    """[world, 2 + b_cap] int32 on the data device: (b_local, M_local, lens...) of every rank, exchanged by ONE small
    collective of the data-path backend itself (NCCL on the GPUs) -- enqueued on the stream behind the still-running
    frame program, no host-side (gloo / TCP) round trip.  Only the ranks that must size a receive buffer read it."""
    world = dist.get_world_size(group)
    dev = lens.device
    meta = torch.zeros((2 + b_cap,), dtype=torch.int32, device=dev)
    meta[0], meta[1] = b_local, m_local
    meta[2:2 + b_local] = lens.to(torch.int32)
    metas = torch.empty((world, 2 + b_cap), dtype=torch.int32, device=dev)
    _all_gather(metas, meta, group)
    return metas


def gather_variable_mels(mel: torch.Tensor, lens: torch.Tensor, group=None, dst: Optional[int] = None,
                         b_cap: Optional[int] = None) -> Tuple[Optional[torch.Tensor], Optional[torch.Tensor]]:
    """Gather `[b_local, C, M_local]` mels whose batch and frame extents differ per rank.

    dst = r   : gather to rank r only (the consumer): every other rank SENDS its own contiguous block, exactly its own
                size, point to point (NVLink through NCCL send/recv) and returns (None, None) without ever draining
                its stream; rank r reads the tiny metadata tensor once (its only host sync, after which every peer is
                done generating anyway), sizes `[B_total, C, M_max]` and receives each block straight into its slice
                when the frame extents agree (no staging copy).
    dst = None: every rank receives everything (all_gather of zero-padded blocks), for callers that need it everywhere.
    `b_cap` is an upper bound of b_local over the ranks (default: exchanged first -- costs one more tiny collective and
    a sync; `generate_sharded` knows it from the shard split).  Returns (`[B_total, C, M_max]`, `[B_total]` int64) with
    the sentences in rank order; frames beyond a shard's own M_local are zero.
    """
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return mel, lens.to(torch.int64)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dev = mel.device
    b_local, C, m_local = mel.shape
    if b_cap is None:
        t = torch.tensor([b_local], dtype=torch.int32, device=dev)
        _all_reduce_max(t, group)
        b_cap = int(t.item())
    metas = _meta_exchange(b_local, m_local, lens, b_cap, group)
    if dst is not None and rank != dst:
        if b_local > 0 and m_local > 0:
            _p2p([dist.P2POp(dist.isend, mel.contiguous(), _global_rank(dst, group), group)])
        return None, None
    metas = metas.cpu()                                   # the consumer's one host sync
    counts = [int(v) for v in metas[:, 0]]
    ms = [int(v) for v in metas[:, 1]]
    m_max = max(ms)
    lens_all = torch.cat([metas[r, 2:2 + counts[r]] for r in range(world)]).to(torch.int64).to(dev)
    if dst is None:
        b_max = max(counts)
        uniform = all(c == b_max for c in counts) and all(m == m_max for m in ms)
        if uniform and mel.is_contiguous():               # equal shards (the weak-scaling case): no padding copies
            block = mel
        else:
            block = torch.zeros((b_max, C, m_max), dtype=mel.dtype, device=dev)
            block[:b_local, :, :m_local] = mel
        blocks = torch.empty((world, b_max, C, m_max), dtype=mel.dtype, device=dev)
        _all_gather(blocks, block, group)
        if uniform:
            return blocks.view(world * b_max, C, m_max), lens_all
        return torch.cat([blocks[r, :counts[r]] for r in range(world)], dim=0), lens_all
    total = sum(counts)
    exact = all(m == m_max or c == 0 for m, c in zip(ms, counts))
    out = (torch.empty if exact else torch.zeros)((total, C, m_max), dtype=mel.dtype, device=dev)
    ops, staged, b0 = [], [], 0
    for r in range(world):
        sl = slice(b0, b0 + counts[r])
        b0 += counts[r]
        if counts[r] == 0 or ms[r] == 0:
            continue
        if r == rank:
            out[sl, :, :m_local] = mel
        elif ms[r] == m_max:
            ops.append(dist.P2POp(dist.irecv, out[sl], _global_rank(r, group), group))       # contiguous slice: in place
        else:
            buf = torch.empty((counts[r], C, ms[r]), dtype=mel.dtype, device=dev)
            staged.append((sl, ms[r], buf))
            ops.append(dist.P2POp(dist.irecv, buf, _global_rank(r, group), group))
    _p2p(ops)
    for sl, m, buf in staged:
        out[sl, :, :m] = buf
    return out, lens_all


def _global_rank(r: int, group) -> int:
    return r if group is None or group is dist.group.WORLD else dist.get_global_rank(group, r)


def _p2p(ops) -> None:
    """One batch of point-to-point transfers.  NCCL: stream-ordered on the current stream (work.wait() only makes the
    stream wait, the host does not block); gloo (CPU tests): wait() blocks until the bytes are there."""
    if not ops:
        return
    for w in dist.batch_isend_irecv(ops):
        w.wait()


def _all_gather(dst: torch.Tensor, src: torch.Tensor, group) -> None:
    """dst[r] = rank r's src.  NCCL: one all_gather_into_tensor; gloo (CPU tests): the list form."""
    if src.is_cuda:
        dist.all_gather_into_tensor(dst, src, group=group)
        return
    parts = [torch.empty_like(src) for _ in range(dst.shape[0])]
    dist.all_gather(parts, src, group=group)
    for r, p in enumerate(parts):
        dst[r].copy_(p)


def _all_reduce_max(t: torch.Tensor, group) -> None:
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)


def generate_sharded(model, x: torch.Tensor, speaker_emb: Optional[torch.Tensor] = None, group=None,
                     dst: Optional[int] = None, gather_mel: bool = True, **generate_kwargs) -> Dict[str, torch.Tensor]:
    """Run `model.generate` on this rank's slice of the global batch `x` [B, T] and gather the results.

    Every rank passes the same global `x` (and `speaker_emb` [B, S] for the multi-speaker models); returns
    `{'mel_post': [B, n_mels, M], 'mel': ..., 'mel_lens': [B]}` on every rank (dst = None) or on rank `dst` only (the
    other ranks then get None values and never synchronise with the host); `gather_mel=False` gathers `mel_post`
    alone (what `gen_forward.py:109` consumes).  With no process group this is a plain `generate` call on the whole
    batch.
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
    b_cap = -(-x.shape[0] // world)                      # ceil: the largest shard of shard_slice
    post_all, lens_all = gather_variable_mels(mel_post, lens, group, dst, b_cap)
    if not gather_mel:
        mel_all = None
    elif mel is mel_post or mel.data_ptr() == mel_post.data_ptr():
        mel_all = post_all
    else:
        mel_all, _ = gather_variable_mels(mel, lens, group, dst, b_cap)
    return {'mel': mel_all, 'mel_post': post_all, 'mel_lens': lens_all}
