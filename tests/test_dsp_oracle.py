"""CPU checks of the mel -> waveform restatement (oracle/dsp_oracle.py, parity unpinned: librosa is absent) and of the
host-side mel filterbank of forwardtacotron_b200.vocoder: internal consistency only -- perfect reconstruction of the
STFT pair, Griffin-Lim monotonically approaching the target magnitudes, filterbank shape / ordering / normalisation."""
import numpy as np

from forwardtacotron_b200 import vocoder as voc
from oracle import dsp_oracle as D


def _signal(n, seed=0):
    rng = np.random.RandomState(seed)
    t = np.arange(n) / 22050.0
    return 0.4 * np.sin(2 * np.pi * 220 * t) + 0.2 * np.sin(2 * np.pi * 1333 * t + 1.0) + 0.05 * rng.randn(n)


def test_stft_pair_reconstructs():
    y = _signal(256 * 20)
    S = D.stft(y)
    assert S.shape == (513, 21)
    assert np.abs(D.istft(S) - y).max() < 1e-10


def test_griffinlim_restatement_converges():
    y = _signal(256 * 24, 1)
    S = np.abs(D.stft(y))
    ang = np.exp(2j * np.pi * np.random.RandomState(0).rand(*S.shape))
    sc = [D.spectral_convergence(D.griffinlim(S, ang, n_iter=n), S) for n in (0, 8, 32)]
    assert sc[2] < sc[1] < sc[0] and sc[2] < 0.35 * sc[0]


def test_mel_filterbank_slaney():
    b = voc.mel_filterbank(22050, 1024, 80, 0, 8000)
    assert b.shape == (80, 513) and b.dtype == np.float32 and b.min() >= 0
    # Slaney normalisation: every triangle has (about) the same area in Hz: sum * bin width ~ 1
    area = b.sum(axis=1) * (22050 / 1024)
    assert np.all(np.abs(area - 1.0) < 0.25)
    assert np.all(np.diff(b.argmax(axis=1)) >= 0) and b[:, 372 + 2:].max() == 0.0
    x = np.abs(np.random.RandomState(2).randn(513, 3))
    target = b.astype(np.float64) @ x
    sol = D.nnls_lbfgs(b.astype(np.float64), target)
    assert sol.min() >= 0.0 and np.linalg.norm(b @ sol - target) < 3e-2 * np.linalg.norm(target)
