"""mel -> waveform on the GPU: the Griffin-Lim vocoder of the reference's DSP class (utils/dsp.py:80-94), batched.

    DSP.griffinlim(mel, n_iter=32):  S = librosa.feature.inverse.mel_to_stft(exp(mel), power=1, sr, n_fft, fmin, fmax)
                                     wav = librosa.core.griffinlim(S, n_iter, hop_length, win_length)

librosa is a third-party dependency that is neither vendored in the reference nor installed in this image; both calls are
restated from its published algorithms (oracle/dsp_oracle.py holds the numpy twin used by the tests).  PARITY UNPINNED:
librosa starts Griffin-Lim from random phases drawn from an unseeded generator and solves the mel inversion with SciPy's
L-BFGS-B, so the reference's own output is not reproducible sample by sample; what is tested is agreement with the numpy
restatement on the same magnitudes and starting phases, and the spectral convergence of the result.

The 32-iteration loop (inverse STFT, STFT, phase update) runs on the kernels of csrc/ftb_gl.cu through the C ABI; the
one-off mel inversion (an 80 x 513 non-negative least-squares problem per frame) is a projected-gradient solve started
from the clipped pseudo-inverse, written with torch matmuls (a plain library GEMM, not a hot loop).
"""
import math
from typing import Optional

import numpy as np
import torch

from . import runtime as rt


def _hz_to_mel(f):
    """Slaney mel scale (librosa.hz_to_mel, htk=False): linear below 1 kHz, logarithmic above."""
    f = np.asarray(f, dtype=np.float64)
    f_sp = 200.0 / 3
    mels = f / f_sp
    min_log_hz, min_log_mel, logstep = 1000.0, 1000.0 / f_sp, math.log(6.4) / 27.0
    return np.where(f >= min_log_hz, min_log_mel + np.log(np.maximum(f, 1e-10) / min_log_hz) / logstep, mels)


def _mel_to_hz(m):
    m = np.asarray(m, dtype=np.float64)
    f_sp = 200.0 / 3
    min_log_hz, min_log_mel, logstep = 1000.0, 1000.0 / f_sp, math.log(6.4) / 27.0
    return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), f_sp * m)


def mel_filterbank(sr: int, n_fft: int, n_mels: int, fmin: float, fmax: float) -> np.ndarray:
    """librosa.filters.mel(sr, n_fft, n_mels, fmin, fmax, htk=False, norm='slaney') -> [n_mels, n_fft//2 + 1] float32."""
    fftfreqs = np.linspace(0.0, sr / 2.0, n_fft // 2 + 1)
    mel_f = _mel_to_hz(np.linspace(_hz_to_mel(fmin), _hz_to_mel(fmax), n_mels + 2))
    fdiff = np.diff(mel_f)
    ramps = mel_f[:, None] - fftfreqs[None, :]
    lower = -ramps[:-2] / fdiff[:-1, None]
    upper = ramps[2:] / fdiff[1:, None]
    weights = np.maximum(0.0, np.minimum(lower, upper))
    weights *= (2.0 / (mel_f[2:n_mels + 2] - mel_f[:n_mels]))[:, None]           # Slaney area normalisation
    return weights.astype(np.float32)


class GriffinLim:
    """Batched GPU twin of DSP.griffinlim.  Constructor arguments = the `dsp` block of the reference's yaml."""

    def __init__(self, sample_rate: int = 22050, n_fft: int = 1024, num_mels: int = 80, hop_length: int = 256,
                 win_length: int = 1024, fmin: float = 0.0, fmax: float = 8000.0, device=None, momentum: float = 0.99,
                 nnls_iters: int = 64, **_ignored):
        if win_length != n_fft:
            raise ValueError('win_length != n_fft is not supported (the reference configs use 1024 / 1024)')
        self.device = torch.device(device if device is not None else 'cuda')
        self.n_fft, self.hop, self.n_mels, self.momentum, self.nnls_iters = n_fft, hop_length, num_mels, momentum, nnls_iters
        self.handle = rt.Handle.get(self.device)
        basis = mel_filterbank(sample_rate, n_fft, num_mels, fmin, fmax)
        self.basis = torch.from_numpy(basis).to(self.device)                                    # [n_mels, bins]
        self.pinv = torch.from_numpy(np.linalg.pinv(basis.astype(np.float64)).astype(np.float32)).to(self.device)
        self.lipschitz = float(np.linalg.norm(basis.astype(np.float64), 2) ** 2)
        n = np.arange(n_fft, dtype=np.float64)
        self.window = torch.from_numpy((0.5 - 0.5 * np.cos(2.0 * np.pi * n / n_fft)).astype(np.float32)).to(self.device)
        k = np.arange(n_fft // 2, dtype=np.float64)
        tw = np.stack([np.cos(2.0 * np.pi * k / n_fft), -np.sin(2.0 * np.pi * k / n_fft)], axis=1)
        self.twiddles = torch.from_numpy(tw.astype(np.float32)).contiguous().to(self.device)    # exp(-2 pi i k / n)

    # -------------------------------------------------------------------------------------------
    def mel_to_stft(self, mel: torch.Tensor) -> torch.Tensor:
        """exp(mel) [B, n_mels, M] -> magnitudes [B, M, bins] >= 0 minimising ||basis @ S - exp(mel)||^2 (librosa's
        mel_to_stft with power = 1): clipped pseudo-inverse, then `nnls_iters` accelerated projected-gradient steps."""
        target = torch.exp(mel.to(self.device, torch.float32)).transpose(1, 2).contiguous()      # [B, M, n_mels]
        x = torch.clamp(target @ self.pinv.t(), min=0.0)                                         # [B, M, bins]
        y, t = x, 1.0
        step = 1.0 / self.lipschitz
        for _ in range(self.nnls_iters):
            grad = (y @ self.basis.t() - target) @ self.basis
            x_new = torch.clamp(y - step * grad, min=0.0)
            t_new = 0.5 * (1.0 + math.sqrt(1.0 + 4.0 * t * t))
            y = x_new + ((t - 1.0) / t_new) * (x_new - x)
            x, t = x_new, t_new
        return x.contiguous()

    def stft(self, y: torch.Tensor, frames: int) -> torch.Tensor:
        spec = torch.empty((y.shape[0], frames, self.n_fft // 2 + 1, 2), dtype=torch.float32, device=self.device)
        self.handle.stft(y, y.shape[1], self.window, self.twiddles, self.n_fft, self.hop, frames, spec)
        return spec

    def istft(self, spec: torch.Tensor, frame_buf: Optional[torch.Tensor] = None) -> torch.Tensor:
        B, frames = spec.shape[0], spec.shape[1]
        if frame_buf is None:
            frame_buf = torch.empty((B, frames, self.n_fft), dtype=torch.float32, device=self.device)
        y = torch.empty((B, max(1, self.hop * (frames - 1))), dtype=torch.float32, device=self.device)
        self.handle.istft(spec, self.window, self.twiddles, self.n_fft, self.hop, frames, frame_buf, y)
        return y

    def griffinlim_from_stft(self, mag: torch.Tensor, n_iter: int = 32, angles: Optional[torch.Tensor] = None,
                             seed: int = 0) -> torch.Tensor:
        """mag [B, M, bins] -> wav [B, hop * (M - 1)].  `angles` [B, M, bins, 2]: unit phasors to start from (default:
        uniformly random phases from `seed`, what librosa draws from its unseeded generator)."""
        B, M, bins = mag.shape
        mag = mag.contiguous()
        if angles is None:
            g = torch.Generator(device=self.device).manual_seed(seed)
            phi = torch.rand((B, M, bins), generator=g, device=self.device) * (2.0 * math.pi)
            angles = torch.stack([torch.cos(phi), torch.sin(phi)], dim=-1)
        spec = (mag.unsqueeze(-1) * angles.to(self.device, torch.float32)).contiguous()
        tprev = torch.zeros_like(spec)
        frame_buf = torch.empty((B, M, self.n_fft), dtype=torch.float32, device=self.device)
        for _ in range(n_iter):
            y = self.istft(spec, frame_buf)
            rebuilt = self.stft(y, M)
            self.handle.gl_update(mag, rebuilt, tprev, spec, self.momentum)
        return self.istft(spec, frame_buf)

    def griffinlim(self, mel: torch.Tensor, n_iter: int = 32, seed: int = 0) -> torch.Tensor:
        """DSP.griffinlim for a batch: mel [B, n_mels, M] (or [n_mels, M]) log-mel -> wav [B, hop * (M - 1)]."""
        single = mel.dim() == 2
        if single:
            mel = mel.unsqueeze(0)
        wav = self.griffinlim_from_stft(self.mel_to_stft(mel), n_iter=n_iter, seed=seed)
        return wav[0] if single else wav
