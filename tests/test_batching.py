"""Host logic of the batched caller, on CPU with a stub model; the GPU test compares a real
bucket with the oracle applied to the same padded bucket."""
import numpy as np
import pytest
import torch

from forwardtacotron_b200 import batching as bt
from forwardtacotron_b200 import parallel as par


def test_bucket_by_length_limits_size_and_ratio():
    lens = [5, 50, 7, 6, 52, 9, 5, 49, 100]
    buckets = bt.bucket_by_length(lens, max_batch=3, max_len_ratio=1.25)
    assert sorted(i for b in buckets for i in b) == list(range(len(lens)))
    for b in buckets:
        ls = [lens[i] for i in b]
        assert len(b) <= 3 and max(ls) <= 1.25 * min(ls) and ls == sorted(ls)
    same = bt.bucket_by_length(lens, max_batch=8, max_len_ratio=1.0)
    assert all(len({lens[i] for i in b}) == 1 for b in same)
    with pytest.raises(ValueError):
        bt.bucket_by_length([3, 0], 4)


class _Stub:
    """generate() contract; mel value encodes (sentence checksum, frame index); padded tokens get duration 1."""
    n_mels = 3

    def __init__(self):
        self.calls = []

    def generate(self, x, speaker_emb=None, alpha=1.0, pitch_function=lambda v: v, energy_function=lambda v: v):
        self.calls.append(tuple(x.shape))
        dur = torch.where(x > 0, (x % 3).float() + 0.6, torch.ones(()))      # real tokens: 0.6 / 1.6 / 2.6
        lens = par.mel_lengths(dur)
        mel = torch.zeros(x.shape[0], self.n_mels, int(lens.max()))
        for b in range(x.shape[0]):
            base = float(x[b].sum()) + (0.0 if speaker_emb is None else float(speaker_emb[b, 0]))
            mel[b] = base + torch.arange(mel.shape[2]).float()
        return {'mel': mel, 'mel_post': mel + 0.25, 'dur': dur}


def test_synthesize_batched_order_unpadding_and_speakers(tmp_path):
    g = torch.Generator().manual_seed(0)
    seqs = [torch.randint(1, 100, (n,), generator=g).tolist() for n in (12, 5, 13, 40, 6, 11)]
    model = _Stub()
    semb = torch.arange(6, dtype=torch.float32)[:, None].repeat(1, 4) * 1000.0
    mels = bt.synthesize_batched(model, seqs, max_batch=3, speaker_emb=semb, max_len_ratio=1.3)
    assert len(model.calls) == 3 and all(c[0] <= 3 for c in model.calls)          # buckets {5,6} {11,12,13} {40}
    for i, (s, m) in enumerate(zip(seqs, mels)):
        own = sum(int(t % 3 + 0.6 + 0.5) for t in s)                               # frames of the real tokens only
        assert m.shape == (1, 3, own) and m.dtype == np.float32
        assert m[0, 0, 0] == pytest.approx(sum(s) + 1000.0 * i + 0.25)              # right sentence, right speaker
    p = tmp_path / 'a.npy'
    bt.save_npy(str(p), mels[0])
    assert np.array_equal(np.load(str(p), allow_pickle=False), mels[0])
    one = bt.synthesize_batched(_Stub(), seqs, max_batch=8, speaker_emb=torch.zeros(4), key='mel')
    assert one[1][0, 0, 0] == pytest.approx(sum(seqs[1]))
