"""mel -> waveform on the GPU (csrc/ftb_gl.cu through the C ABI) against the numpy restatement of
librosa's algorithms in oracle/dsp_oracle.py.  Parity with the reference itself is UNPINNED (librosa is absent; see the
oracle's header); these tests pin the kernels to the restatement and check the properties the algorithm must have."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from forwardtacotron_b200 import vocoder as voc           # noqa: E402
from oracle import dsp_oracle as D                        # noqa: E402

DEV = torch.device('cuda', 0)


def _signal(n, seed):
    rng = np.random.RandomState(seed)
    t = np.arange(n) / 22050.0
    y = 0.4 * np.sin(2 * np.pi * 220 * t) + 0.2 * np.sin(2 * np.pi * 1333 * t + 1.0) + 0.05 * rng.randn(n)
    return y.astype(np.float32)


def test_stft_istft_match_numpy_fft():
    gl = voc.GriffinLim(device=DEV)
    frames = 37
    ys = np.stack([_signal(256 * (frames - 1), s) for s in (1, 2, 3)])
    y = torch.from_numpy(ys).to(DEV)
    spec = gl.stft(y, frames)
    torch.cuda.synchronize()
    got = torch.view_as_complex(spec).cpu().numpy()                         # [B, frames, bins]
    for b in range(3):
        ref = D.stft(ys[b]).T
        assert got[b].shape == ref.shape
        assert np.abs(got[b] - ref).max() < 2e-4 * np.abs(ref).max()
    back = gl.istft(spec).cpu().numpy()
    assert back.shape == ys.shape
    assert np.abs(back - ys).max() < 1e-4                                   # a Hann window at hop = n/4 reconstructs exactly
    for b in range(3):
        assert np.abs(back[b] - D.istft(D.stft(ys[b]))).max() < 1e-4


def test_griffinlim_matches_restatement_and_converges():
    gl = voc.GriffinLim(device=DEV)
    frames, bins = 41, 513
    ys = np.stack([_signal(256 * (frames - 1), s) for s in (4, 5)])
    S = np.stack([np.abs(D.stft(y)) for y in ys])                           # [B, bins, frames]
    rng = np.random.RandomState(0)
    phi = rng.rand(2, bins, frames) * 2 * np.pi
    ang = np.exp(1j * phi)
    mag = torch.from_numpy(S.transpose(0, 2, 1).astype(np.float32)).contiguous().to(DEV)
    a_t = torch.from_numpy(np.stack([ang.real, ang.imag], axis=-1).transpose(0, 2, 1, 3).astype(np.float32)).contiguous().to(DEV)
    # two iterations: the same trajectory as the float64 restatement
    got2 = gl.griffinlim_from_stft(mag, n_iter=2, angles=a_t).cpu().numpy()
    for b in range(2):
        ref2 = D.griffinlim(S[b], ang[b], n_iter=2)
        assert np.abs(got2[b] - ref2).max() < 2e-3 * max(1.0, np.abs(ref2).max())
    # 32 iterations (the reference's n_iter): as consistent with the magnitudes as the restatement's result
    got = gl.griffinlim_from_stft(mag, n_iter=32, angles=a_t).cpu().numpy()
    again = gl.griffinlim_from_stft(mag, n_iter=32, angles=a_t).cpu().numpy()
    assert np.array_equal(got, again)                                       # deterministic
    for b in range(2):
        ref = D.griffinlim(S[b], ang[b], n_iter=32)
        sc_got, sc_ref = D.spectral_convergence(got[b], S[b]), D.spectral_convergence(ref, S[b])
        sc_0 = D.spectral_convergence(D.istft(S[b] * ang[b]), S[b])
        assert sc_got < 0.5 * sc_0 and sc_got < 1.15 * sc_ref + 1e-3, (sc_got, sc_ref, sc_0)


def test_mel_inversion_and_end_to_end_shapes():
    """mel filterbank properties + the clipped-pseudo-inverse / projected-gradient inversion against L-BFGS-B."""
    gl = voc.GriffinLim(device=DEV)
    basis = voc.mel_filterbank(22050, 1024, 80, 0, 8000)
    assert basis.shape == (80, 513) and basis.min() >= 0 and np.all(basis.sum(axis=1) > 0)
    assert basis[:, int(8000 / (22050 / 2) * 512) + 2:].max() == 0.0        # nothing above fmax
    peaks = basis.argmax(axis=1)
    assert np.all(np.diff(peaks) >= 0)                                      # triangles ordered in frequency
    frames = 12
    y = _signal(256 * (frames - 1), 9)
    S = np.abs(D.stft(y))                                                   # [bins, frames]
    mel = np.log(np.clip(basis.astype(np.float64) @ S, 1e-5, None)).astype(np.float32)
    got = gl.mel_to_stft(torch.from_numpy(mel)[None].to(DEV))[0].cpu().numpy().T          # [bins, frames]
    ref = D.nnls_lbfgs(basis.astype(np.float64), np.exp(mel.astype(np.float64)))
    res = lambda x: np.linalg.norm(basis @ x - np.exp(mel))                 # noqa: E731
    assert got.min() >= 0.0
    assert res(got) < 1.5 * res(ref) + 1e-3 * np.linalg.norm(np.exp(mel))   # as good a fit as the L-BFGS-B solve
    wav = gl.griffinlim(torch.from_numpy(mel).to(DEV), n_iter=8)
    assert wav.shape == (256 * (frames - 1),) and bool(torch.isfinite(wav).all())
