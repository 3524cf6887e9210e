"""ftb_lr_regulate_static: the no-allocation LengthRegulator entry point (include/ftb200.h) against the existing
ftb_lr_scan + ftb_lr_gather pair, through the C ABI, bit-exact (the tolerance of every LengthRegulator test here), on
two calls in a row -- the second smaller than the first, so stale contents of the shared static scratch would show.

Collected last on purpose (the file name sorts after every other test file).
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from forwardtacotron_b200 import runtime as rt            # noqa: E402

DEV = torch.device('cuda', 0)
MAX_TOKENS, MAX_BATCH = 65536, 1024                       # FTB_LR_STATIC_MAX_TOKENS / _BATCH


def _existing(h, x, dur, fallback):
    B, T, C = x.shape
    d = dur.clone()
    cum = torch.empty((B, T), dtype=torch.int32, device=DEV)
    lens = torch.empty((B,), dtype=torch.int32, device=DEV)
    meta = torch.empty((4,), dtype=torch.int32, device=DEV)
    h.lr_scan(d, B, T, fallback, cum, lens, meta)
    M = int(meta.cpu()[0])
    out = torch.full((B, M, C), 7.0, device=DEV)
    idx = torch.full((B, M), -7, dtype=torch.int32, device=DEV)
    a = rt.LrGatherArgs()
    a.x, a.x_stride, a.x_ld, a.c1, a.c2 = x.data_ptr(), T, C, C, 0
    a.cum, a.lens, a.batch, a.T, a.M = cum.data_ptr(), lens.data_ptr(), B, T, M
    a.out, a.out_stride, a.out_ld, a.idx, a.round_tf32 = out.data_ptr(), M, C, idx.data_ptr(), 0
    h.lr_gather(a)
    torch.cuda.synchronize()
    return d, out, idx, lens, meta, M


def _static(h, x, dur, fallback, M):
    B, T, C = x.shape
    d = dur.clone()
    out = torch.full((B, M, C), 9.0, device=DEV)
    idx = torch.full((B, M), -9, dtype=torch.int32, device=DEV)
    lens = torch.full((B,), -9, dtype=torch.int32, device=DEV)
    meta = torch.full((4,), -9, dtype=torch.int32, device=DEV)
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats(DEV)['allocation.all.allocated']
    h.lr_regulate_static(x.view(B * T, C), T, C, d, B, T, fallback, M, out.view(B * M, C), M,
                         idx=idx, lens_out=lens, meta_out=meta)
    assert torch.cuda.memory_stats(DEV)['allocation.all.allocated'] == before     # torch's allocator untouched
    torch.cuda.synchronize()
    return d, out, idx, lens, meta


def _case(seed, B, T, C, lo, hi):
    gen = torch.Generator().manual_seed(seed)
    x = torch.randn(B, T, C, generator=gen).to(DEV)
    dur = (torch.rand(B, T, generator=gen) * (hi - lo) + lo).to(DEV).contiguous()
    return x, dur


def _bits(t):
    return t.cpu().numpy().view(np.uint32) if t.dtype == torch.float32 else t.cpu().numpy()


def test_static_scratch_matches_existing_entry_points_twice():
    h = rt.Handle.get(DEV)
    cases = [(_case(11, 16, 120, 512, -0.5, 4.5), True),        # ragged batch, negatives clamped, zero durations
             (_case(12, 5, 33, 256, -0.3, 0.6), True),          # smaller 2nd call; truncated sum <= 0 -> fallback fill
             (_case(13, 7, 50, 64, 0.0, 3.0), False)]
    for (x, dur), fallback in cases:
        d0, out0, idx0, lens0, meta0, M = _existing(h, x, dur, fallback)
        d1, out1, idx1, lens1, meta1 = _static(h, x, dur, fallback, M)
        assert np.array_equal(_bits(d1), _bits(d0))
        assert np.array_equal(_bits(out1), _bits(out0))
        assert np.array_equal(_bits(idx1), _bits(idx0))
        assert np.array_equal(_bits(lens1), _bits(lens0))
        assert np.array_equal(_bits(meta1)[:3], _bits(meta0)[:3])
    assert int(meta1.cpu()[1]) == 0 and M > 0


def test_static_scratch_rejects_sizes_beyond_capacity():
    h = rt.Handle.get(DEV)
    x = torch.zeros((8, 64), device=DEV)
    d = torch.zeros((8,), device=DEV)
    out = torch.zeros((8, 64), device=DEV)
    for batch, T in [(MAX_BATCH + 1, 1), (2, MAX_TOKENS // 2 + 1)]:
        with pytest.raises(RuntimeError, match='capacity of the no-allocation entry point'):
            h.lr_regulate_static(x, T, 64, d, batch, T, False, 1, out, 1)
