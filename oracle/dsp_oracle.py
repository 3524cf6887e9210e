"""CPU restatement (numpy / scipy) of the mel -> waveform step of the reference: DSP.griffinlim (utils/dsp.py:80-94).

TEST INFRASTRUCTURE ONLY (imported by tests/; nothing under forwardtacotron_b200/ imports it).

PARITY UNPINNED.  The arithmetic lives in librosa (`librosa.feature.inverse.mel_to_stft`, `librosa.core.griffinlim`), a
third-party dependency that is neither vendored under /root/reference nor pinned by its requirements nor installed in
this image, and the reference's tests hold no golden waveform.  The functions below restate librosa's published
algorithms (0.10 defaults: momentum 0.99, Hann window, centred zero-padded frames, NNLS by L-BFGS-B from the clipped
pseudo-inverse); they have not been checked against librosa itself.  librosa also starts from random phases of an unseeded
generator, so even the reference cannot reproduce its own samples -- callers here pass the starting phases explicitly.
"""
import numpy as np
import scipy.optimize


def hann(n_fft):
    n = np.arange(n_fft, dtype=np.float64)
    return 0.5 - 0.5 * np.cos(2.0 * np.pi * n / n_fft)            # scipy.signal.get_window('hann', n, fftbins=True)


def stft(y, n_fft=1024, hop=256):
    """librosa.stft(center=True, pad_mode='constant'): [bins, frames] complex."""
    y = np.pad(np.asarray(y, dtype=np.float64), n_fft // 2)
    frames = 1 + (len(y) - n_fft) // hop
    w = hann(n_fft)
    return np.stack([np.fft.rfft(w * y[f * hop:f * hop + n_fft]) for f in range(frames)], axis=1)


def istft(S, n_fft=1024, hop=256):
    """librosa.istft(center=True, length=None): overlap-add of window * irfft, divided by the window sum-square."""
    frames = S.shape[1]
    w = hann(n_fft)
    total = n_fft + hop * (frames - 1)
    y, wss = np.zeros(total), np.zeros(total)
    for f in range(frames):
        y[f * hop:f * hop + n_fft] += w * np.fft.irfft(S[:, f], n=n_fft)
        wss[f * hop:f * hop + n_fft] += w * w
    nz = wss > np.finfo(np.float32).tiny
    y[nz] /= wss[nz]
    return y[n_fft // 2:total - n_fft // 2]


def griffinlim(S, angles, n_iter=32, n_fft=1024, hop=256, momentum=0.99):
    """librosa.core.griffinlim with explicit starting phases `angles` [bins, frames] (unit complex)."""
    S = np.asarray(S, dtype=np.float64)
    angles = np.asarray(angles, dtype=np.complex128)
    rebuilt = np.zeros_like(angles)
    eps = np.finfo(np.float32).tiny
    for _ in range(n_iter):
        tprev = rebuilt
        inverse = istft(S * angles, n_fft, hop)
        rebuilt = stft(inverse, n_fft, hop)
        angles = rebuilt - (momentum / (1 + momentum)) * tprev
        angles = angles / (np.abs(angles) + eps)
    return istft(S * angles, n_fft, hop)


def spectral_convergence(y, S, n_fft=1024, hop=256):
    """|| |STFT(y)| - S ||_F / ||S||_F: how consistent a waveform is with the magnitudes it was built from."""
    return float(np.linalg.norm(np.abs(stft(y, n_fft, hop)) - S) / np.linalg.norm(S))


def nnls_lbfgs(A, B):
    """librosa.util.nnls for a 2-D right-hand side: minimise 0.5 ||A x - B||^2 / B.size over x >= 0 with L-BFGS-B,
    started from the clipped pseudo-inverse solution."""
    x0 = np.clip(np.linalg.pinv(A) @ B, 0, None)
    shape = x0.shape

    def obj(x):
        x = x.reshape(shape)
        diff = A @ x - B
        return 0.5 * np.sum(diff ** 2) / B.size, (A.T @ diff).ravel() / B.size
    x, _, _ = scipy.optimize.fmin_l_bfgs_b(obj, x0.ravel(), bounds=[(0, None)] * x0.size, m=A.shape[1])
    return x.reshape(shape)
