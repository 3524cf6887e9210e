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


def bind_host_to_device(device: torch.device) -> Optional[str]:
    """Pin this process to the CPUs of the NUMA node its GPU hangs off (sysfs `local_cpulist` of the PCI device), so the
    pinned staging buffers of the H2D / D2H copies are first-touched on that node.  One process per GPU under torchrun is
    not NUMA-bound by default; with 8 ranks copying their mels to the host at once the far-node traffic shows.  Returns
    the cpu list it bound to, or None when sysfs does not say (then nothing changes)."""
    import os
    try:
        p = torch.cuda.get_device_properties(device)
        path = f'/sys/bus/pci/devices/{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0/local_cpulist'
        with open(path) as f:
            spec = f.read().strip()
        cpus = set()
        for part in spec.split(','):
            lo, _, hi = part.partition('-')
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return spec
    except Exception:
        return None


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
    """A gloo twin of the default (NCCL) process group for host-known shapes, created on first use by a collective call
    of all ranks; None when `group` is a sub-group or the twin cannot be created (the shapes then travel through NCCL
    and the consumer pays one device sync)."""
    if group is not None and group is not dist.group.WORLD:
        return None
    if 'world' not in _HOST_GROUPS:
        try:
            _HOST_GROUPS['world'] = dist.new_group(backend='gloo')
        except Exception:
            _HOST_GROUPS['world'] = None
    return _HOST_GROUPS['world']


class LengthExchange:
    """Device-side exchange of the regulated lengths M_local between the ranks of a sharded generate().

    `with LengthExchange(model_device, group) as ex: out = model.generate(...)`: while active, the accelerated models
    call `ex._hook(meta, lens)` right after the LengthRegulator scan has been enqueued (models/_accel.scan_hook), where
    this class enqueues ONE 16-byte NCCL all_gather of `meta` and an asynchronous copy of the result into pinned host
    memory.  The collective runs on the GPU at a point where this rank's host is waiting for the scan result anyway;
    `ex.lengths()` afterwards only waits for that copy's event -- no host-side (gloo / TCP) round trip and no stream
    drain between generate() and the gather of the mels."""

    def __init__(self, device: torch.device, group=None):
        self.device, self.group = device, group
        self.world = dist.get_world_size(group)
        self._dev = torch.zeros((self.world, 4), dtype=torch.int32, device=device)
        self._host = torch.zeros((self.world, 4), dtype=torch.int32).pin_memory()
        self._event = torch.cuda.Event()
        self._scanned = torch.cuda.Event()
        self._side = torch.cuda.Stream(device)
        self._fired = False

    def _hook(self, meta: torch.Tensor, lens: torch.Tensor) -> None:
        if self._fired:                       # one exchange per generate() (the first scan)
            return
        self._fired = True
        # On a SIDE stream that waits for the scan only: the collective is a rendezvous of all ranks, and the compute
        # stream (frame program next) must not wait for the slowest rank's token program.
        main = torch.cuda.current_stream(self.device)
        self._scanned.record(main)
        with torch.cuda.stream(self._side):
            self._side.wait_event(self._scanned)
            dist.all_gather_into_tensor(self._dev, meta, group=self.group)
            self._host.copy_(self._dev, non_blocking=True)
            self._event.record(self._side)
        meta.record_stream(self._side)

    def __enter__(self):
        from .models import _accel
        self._fired = False
        _accel.scan_hook = self._hook
        return self

    def __exit__(self, *exc):
        from .models import _accel
        _accel.scan_hook = None
        if not self._fired and exc[0] is None:      # a rank with an empty shard still takes part in the collective
            self._hook(torch.zeros((4,), dtype=torch.int32, device=self.device), None)

    def lengths(self):
        """M_local of every rank (host integers)."""
        self._event.synchronize()
        return [int(v) for v in self._host[:, 0]]


def _shapes(b_local: int, m_local: int, dev, on_cuda: bool, group):
    """(b_local, M_local) of every rank as host integers.  Both are HOST-known on every rank the moment generate()
    returns (tensor shapes; generate() itself read M from the device while the frame program was being enqueued), so
    they are exchanged host to host (gloo) while the GPUs are still busy: no device sync, nothing on the GPU timeline.
    The per-sentence lengths live on the device and travel with the payload instead."""
    world = dist.get_world_size(group)
    host = _host_group(group) if on_cuda else group
    if host is not None:
        meta = torch.tensor([b_local, m_local], dtype=torch.int64)
        metas = torch.empty((world, 2), dtype=torch.int64)
        _all_gather(metas, meta, host)
    else:                                                # sub-group without a host twin: one device sync
        meta = torch.tensor([b_local, m_local], dtype=torch.int64, device=dev)
        metas = torch.empty((world, 2), dtype=torch.int64, device=dev)
        _all_gather(metas, meta, group)
        metas = metas.cpu()
    return [int(v) for v in metas[:, 0]], [int(v) for v in metas[:, 1]]


def gather_variable_mels(mel: torch.Tensor, lens: torch.Tensor, group=None, dst: Optional[int] = None,
                         b_cap: Optional[int] = None, shapes=None) -> Tuple[Optional[torch.Tensor], Optional[torch.Tensor]]:
    """Gather `[b_local, C, M_local]` mels whose batch and frame extents differ per rank.

    dst = r   : gather to rank r only (the consumer): every other rank SENDS its own contiguous block and its lengths,
                exactly their own size, point to point (NVLink through NCCL send/recv) and returns (None, None); rank r
                sizes `[B_total, C, M_max]` from the host-exchanged shapes and receives each block straight into its
                slice when the frame extents agree (no staging copy).  Root ingest = the payload once, not world x.
    dst = None: every rank receives everything (all_gather of zero-padded blocks), for callers that need it everywhere.
    No rank drains its stream: the receives / collectives are enqueued behind the still-running frame program.
    Returns (`[B_total, C, M_max]`, `[B_total]` int64) with the sentences in rank order; frames beyond a shard's own
    M_local are zero.  `shapes` = ([b_local of every rank], [M_local of every rank]) when the caller already has them
    (generate_sharded: shard split + LengthExchange) -- then nothing is exchanged on the host.  (`b_cap`: unused.)
    """
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return mel, lens.to(torch.int64)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dev = mel.device
    b_local, C, m_local = mel.shape
    counts, ms = shapes if shapes is not None else _shapes(b_local, m_local, dev, mel.is_cuda, group)
    m_max, total = max(ms), sum(counts)
    lens = lens.to(torch.int64)
    if dst is not None and rank != dst:
        if b_local > 0:
            ops = [dist.P2POp(dist.isend, lens.contiguous(), _global_rank(dst, group), group)]
            if m_local > 0:
                ops.append(dist.P2POp(dist.isend, mel.contiguous(), _global_rank(dst, group), group))
            _p2p(ops)
        return None, None
    if dst is None:
        b_max = max(counts)
        uniform = all(c == b_max for c in counts) and all(m == m_max for m in ms)
        if uniform and mel.is_contiguous():               # equal shards (the weak-scaling case): no padding copies
            block, lens_block = mel, lens
        else:
            block = torch.zeros((b_max, C, m_max), dtype=mel.dtype, device=dev)
            block[:b_local, :, :m_local] = mel
            lens_block = torch.zeros((b_max,), dtype=torch.int64, device=dev)
            lens_block[:b_local] = lens
        blocks = torch.empty((world, b_max, C, m_max), dtype=mel.dtype, device=dev)
        lens_blocks = torch.empty((world, b_max), dtype=torch.int64, device=dev)
        _all_gather(blocks, block, group)
        _all_gather(lens_blocks, lens_block, group)
        if uniform:
            return blocks.view(world * b_max, C, m_max), lens_blocks.view(-1)
        return (torch.cat([blocks[r, :counts[r]] for r in range(world)], dim=0),
                torch.cat([lens_blocks[r, :counts[r]] for r in range(world)], dim=0))
    exact = all(m == m_max or c == 0 for m, c in zip(ms, counts))
    out = (torch.empty if exact else torch.zeros)((total, C, m_max), dtype=mel.dtype, device=dev)
    lens_all = torch.empty((total,), dtype=torch.int64, device=dev)
    ops, staged, b0 = [], [], 0
    for r in range(world):
        sl = slice(b0, b0 + counts[r])
        b0 += counts[r]
        if counts[r] == 0:
            continue
        if r == rank:
            lens_all[sl] = lens
            if m_local > 0:
                out[sl, :, :m_local] = mel
            continue
        ops.append(dist.P2POp(dist.irecv, lens_all[sl], _global_rank(r, group), group))
        if ms[r] == 0:
            continue
        if ms[r] == m_max:
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
    # accelerated models on the GPUs: the regulated lengths are exchanged on the device from inside generate()
    ex = LengthExchange(x.device, group) if (world > 1 and x.is_cuda and hasattr(model, 'program_for')) else None

    def run():
        if sl.stop > sl.start:
            o = model.generate(*args, **generate_kwargs)
            return mel_lengths(o['dur']), o['mel'], o['mel_post']
        n_mels = getattr(model, 'n_mels', 80)          # more ranks than sentences: contribute an empty block
        empty = torch.zeros((0, n_mels, 0), device=x.device)
        return torch.zeros((0,), dtype=torch.int64, device=x.device), empty, empty
    if ex is not None:
        with ex:
            lens, mel, mel_post = run()
        counts = [(lambda s: s.stop - s.start)(shard_slice(x.shape[0], r, world)) for r in range(world)]
        shapes = (counts, ex.lengths())
    else:
        lens, mel, mel_post = run()
        shapes = None
    b_cap = -(-x.shape[0] // world)                      # ceil: the largest shard of shard_slice
    post_all, lens_all = gather_variable_mels(mel_post, lens, group, dst, b_cap, shapes)
    if not gather_mel:
        mel_all = None
    elif mel is mel_post or mel.data_ptr() == mel_post.data_ptr():
        mel_all = post_all
    else:
        mel_all, _ = gather_variable_mels(mel, lens, group, dst, b_cap, shapes)
    return {'mel': mel_all, 'mel_post': post_all, 'mel_lens': lens_all}
