"""TEST INFRASTRUCTURE ONLY - CPU restatement of the reference's wave -> log-mel front-end.

Follows audio/stft.py:51-81 (STFT.transform: reflect pad, strided conv with the windowed Fourier basis, magnitude),
audio/stft.py:160-178 (TacotronSTFT.mel_spectrogram: mel_basis @ magnitude, log(clamp(., 1e-5)), energy = L2 norm
over bins) and audio/audio_processing.py:85-91, but through an independent formulation (framing + numpy rfft in
float64) so that it cross-checks the convolution form the product uses.

The mel filterbank is `librosa.filters.mel(sr, n_fft, n_mels, fmin, fmax)` (stft.py:146-148).  librosa 0.7.2 is
pinned by the reference (environment.yaml) and is NOT installed in this image, so `slaney_mel` below restates its
published algorithm (Slaney auditory-toolbox mel scale, htk=False, norm="slaney" area normalisation, float32):
PARITY UNPINNED for the filterbank values.  Everything else is pinned against the reference's own STFT /
TacotronSTFT classes run in the build container with this filterbank injected (oracle/make_golden.py,
tests/golden/mel_frontend.npz).
"""
import numpy as np


def _hz_to_mel(f):
    f = np.asanyarray(f, dtype=np.float64)
    f_sp = 200.0 / 3
    mels = f / f_sp
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(f >= min_log_hz, min_log_mel + np.log(np.maximum(f, 1e-30) / min_log_hz) / logstep, mels)


def _mel_to_hz(m):
    m = np.asanyarray(m, dtype=np.float64)
    f_sp = 200.0 / 3
    freqs = f_sp * m
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), freqs)


def slaney_mel(sr, n_fft, n_mels=128, fmin=0.0, fmax=None):
    """librosa.filters.mel (htk=False, norm='slaney'): [n_mels, 1 + n_fft//2] float32."""
    if fmax is None:
        fmax = sr / 2.0
    n_bins = 1 + n_fft // 2
    fftfreqs = np.linspace(0.0, sr / 2.0, n_bins)
    mel_f = _mel_to_hz(np.linspace(_hz_to_mel(fmin), _hz_to_mel(fmax), n_mels + 2))
    fdiff = np.diff(mel_f)
    ramps = mel_f[:, None] - fftfreqs[None, :]
    w = np.zeros((n_mels, n_bins))
    for i in range(n_mels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        w[i] = np.maximum(0.0, np.minimum(lower, upper))
    enorm = 2.0 / (mel_f[2:n_mels + 2] - mel_f[:n_mels])
    return (w * enorm[:, None]).astype(np.float32)


def hann_periodic(n):
    """scipy.signal.get_window('hann', n, fftbins=True)"""
    return 0.5 - 0.5 * np.cos(2.0 * np.pi * np.arange(n) / n)


def mel_spectrogram(y, n_fft=1024, hop=256, win=1024, n_mels=80, sr=22050, fmin=0.0, fmax=8000.0, mel_basis=None):
    """y [B, T] in [-1, 1] -> (log-mel [B, n_mels, F] f32, energy [B, F] f32), F = T // hop + 1."""
    y = np.asarray(y, dtype=np.float64)
    assert win == n_fft, "pad_center of a shorter window is not restated"
    half = n_fft // 2
    ypad = np.pad(y, ((0, 0), (half, half)), mode="reflect")
    F = (ypad.shape[1] - n_fft) // hop + 1
    idx = np.arange(n_fft)[None, :] + hop * np.arange(F)[:, None]
    frames = ypad[:, idx] * hann_periodic(win).astype(np.float32).astype(np.float64)  # [B, F, n_fft]
    mag = np.abs(np.fft.rfft(frames, axis=-1))  # [B, F, bins]
    if mel_basis is None:
        mel_basis = slaney_mel(sr, n_fft, n_mels, fmin, fmax)
    mel = np.einsum("mk,bfk->bmf", mel_basis.astype(np.float64), mag)
    return np.log(np.maximum(mel, 1e-5)).astype(np.float32), np.sqrt((mag ** 2).sum(-1)).astype(np.float32)
