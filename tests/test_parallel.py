"""World-size-2 gloo tests (CPU) of the sentence-sharding plumbing: shard slices, the variable-length mel
gather and `generate_sharded` over a stand-in model (the real models need CUDA; their sharded run is covered by
bench.py --gpus N and tests/test_gpu_models.py on the GPU box)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from forwardtacotron_b200 import parallel as par


def test_shard_slice_partitions_exactly():
    for n in (0, 1, 5, 64, 127, 128):
        for world in (1, 2, 3, 8):
            sl = [par.shard_slice(n, r, world) for r in range(world)]
            assert sl[0].start == 0 and sl[-1].stop == n
            assert all(a.stop == b.start for a, b in zip(sl, sl[1:]))
            sizes = [s.stop - s.start for s in sl]
            assert max(sizes) - min(sizes) <= 1


def test_mel_lengths_round_half_up():
    dur = torch.tensor([[0.49, 0.5, 2.5, -1.0], [2.0, 2.0, 2.0, 2.0]])
    assert par.mel_lengths(dur).tolist() == [0 + 1 + 3 + 0, 8]


class _StubModel:
    """Deterministic stand-in with the generate() contract: M depends on the shard (ragged across ranks)."""
    n_mels = 4

    def generate(self, x, alpha=1.0):
        B, T = x.shape
        dur = (x % 3).float() / alpha                           # durations 0..2
        lens = par.mel_lengths(dur)
        M = int(lens.max())
        mel = torch.zeros(B, self.n_mels, M)
        for b in range(B):
            mel[b, :, :lens[b]] = x[b].sum().float() + torch.arange(int(lens[b])).float()
        return {'mel': mel, 'mel_post': mel + 0.5, 'dur': dur}


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, ret):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(0)
        x = torch.randint(1, 135, (5, 11), generator=g)          # 5 sentences over 2 ranks: 3 + 2
        out = par.generate_sharded(_StubModel(), x, alpha=1.0)
        # reference: the stub applied to each shard separately, padded to the global max
        model = _StubModel()
        parts = [model.generate(x[par.shard_slice(5, r, world)]) for r in range(world)]
        M = max(p['mel'].shape[2] for p in parts)
        want = torch.cat([torch.nn.functional.pad(p['mel_post'], (0, M - p['mel_post'].shape[2])) for p in parts])
        lens = torch.cat([par.mel_lengths(p['dur']) for p in parts])
        ok = torch.equal(out['mel_post'], want) and torch.equal(out['mel_lens'], lens) and \
            out['mel'].shape == want.shape
        # ragged direct gather, including a rank with an empty block
        mel = torch.full((rank * 2, 3, 4 + rank), float(rank + 1))
        ln = torch.full((rank * 2,), 4 + rank, dtype=torch.int64)
        got, gl = par.gather_variable_mels(mel, ln)
        ok = ok and got.shape == (2, 3, 5) and bool((got == 2.0).all()) and gl.tolist() == [5, 5]
        # equal shards (the weak-scaling case): gathered without padding copies, rank order kept
        mel = torch.arange(2 * 3 * 4, dtype=torch.float32).reshape(2, 3, 4) + 100.0 * rank
        got, gl = par.gather_variable_mels(mel, torch.tensor([4, 3 + rank]))
        want_u = torch.cat([torch.arange(24, dtype=torch.float32).reshape(2, 3, 4) + 100.0 * r for r in range(world)])
        ok = ok and torch.equal(got, want_u) and gl.tolist() == [4, 3, 4, 4]
        # gather to ONE consumer (dst = 0): the same result there, nothing (and no host sync) anywhere else
        one = par.generate_sharded(_StubModel(), x, alpha=1.0, dst=0)
        if rank == 0:
            ok = ok and torch.equal(one['mel_post'], want) and torch.equal(one['mel_lens'], lens) and \
                torch.equal(one['mel'], out['mel'])
        else:
            ok = ok and one['mel_post'] is None and one['mel_lens'] is None and one['mel'] is None
        post_only = par.generate_sharded(_StubModel(), x, alpha=1.0, dst=1, gather_mel=False)
        ok = ok and post_only['mel'] is None and ((post_only['mel_post'] is None) == (rank != 1))
        if rank == 1:
            ok = ok and torch.equal(post_only['mel_post'], want)
        # equal frame extents: the consumer receives every block straight into its slice of the result
        mel = torch.arange(2 * 3 * 4, dtype=torch.float32).reshape(2, 3, 4) + 100.0 * rank
        got, gl = par.gather_variable_mels(mel, torch.tensor([4, 3 + rank]), dst=0)
        if rank == 0:
            ok = ok and torch.equal(got, want_u) and gl.tolist() == [4, 3, 4, 4]
        else:
            ok = ok and got is None and gl is None
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_generate_sharded_world2_gloo():
    world = 2
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_worker, args=(world, _free_port(), ret), nprocs=world, join=True)
        assert dict(ret) == {0: True, 1: True}
