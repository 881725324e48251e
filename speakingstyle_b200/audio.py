"""Reference-mel front-end on the GPU: drop-in for the reference's ``audio/stft.py`` classes.

``TacotronSTFT(filter_length, hop_length, win_length, n_mel_channels, sampling_rate, mel_fmin, mel_fmax)`` with
``mel_spectrogram(y[B,T]) -> (log-mel [B, n_mel, F], energy [B, F])`` and the same buffer names
(``stft_fn.forward_basis [2*(n_fft/2+1), 1, n_fft]``, ``mel_basis [n_mel, n_fft/2+1]``), so a reference
``state_dict`` loads.  The wave never leaves the device: reflect pad -> Fourier-basis convolution -> magnitude ->
mel filterbank -> log, with both contractions on the 3xTF32 tensor-core convolution kernel (fp32-accurate, as the
reference's fp32 conv1d / matmul are).  A 1024-sample frame at hop 256 is four consecutive 256-sample rows, so
the strided convolution of stft.py:67-72 is run as a k = n_fft/hop "valid" convolution over hop-sized rows.

The mel filterbank is built at construction like the reference's (librosa.filters.mel, Slaney scale and area
normalisation, restated here because librosa is not a dependency); pass your own ``mel_basis`` via
``load_state_dict`` to reproduce a specific librosa version bit for bit.
"""
import math

import numpy as np
import torch
import torch.nn as nn

from . import ops


def _hz_to_mel(f):
    f = np.asanyarray(f, dtype=np.float64)
    f_sp, min_log_hz = 200.0 / 3, 1000.0
    logstep = math.log(6.4) / 27.0
    return np.where(f >= min_log_hz, min_log_hz / f_sp + np.log(np.maximum(f, 1e-30) / min_log_hz) / logstep, f / f_sp)


def _mel_to_hz(m):
    m = np.asanyarray(m, dtype=np.float64)
    f_sp, min_log_hz = 200.0 / 3, 1000.0
    min_log_mel, logstep = min_log_hz / f_sp, math.log(6.4) / 27.0
    return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), f_sp * m)


def mel_filterbank(sr, n_fft, n_mels, fmin, fmax):
    """Triangular Slaney-scale filters with area normalisation: [n_mels, n_fft//2 + 1] float32 (stft.py:146-148)."""
    fmax = sr / 2.0 if fmax is None else fmax
    fftfreqs = np.linspace(0.0, sr / 2.0, 1 + n_fft // 2)
    mel_f = _mel_to_hz(np.linspace(_hz_to_mel(fmin), _hz_to_mel(fmax), n_mels + 2))
    fdiff = np.diff(mel_f)
    ramps = mel_f[:, None] - fftfreqs[None, :]
    lower = -ramps[:-2] / fdiff[:-1, None]
    upper = ramps[2:] / fdiff[1:, None]
    w = np.maximum(0.0, np.minimum(lower, upper))
    return (w * (2.0 / (mel_f[2:] - mel_f[:-2]))[:, None]).astype(np.float32)


class STFT(nn.Module):
    """audio/stft.py:14-81, forward transform only (the inverse belongs to Griffin-Lim, off the path)."""

    def __init__(self, filter_length, hop_length, win_length, window="hann"):
        super().__init__()
        if window != "hann" or win_length != filter_length or filter_length % hop_length or hop_length % 4:
            raise NotImplementedError("STFT: hann window with win_length == filter_length, hop | filter_length, hop % 4 == 0")
        self.filter_length, self.hop_length, self.win_length = filter_length, hop_length, win_length
        n = np.arange(filter_length)
        cutoff = filter_length // 2 + 1
        ang = 2.0 * np.pi * np.outer(np.arange(cutoff), n) / filter_length
        basis = np.vstack([np.cos(ang), -np.sin(ang)])  # rows of fft(eye): real part, then imaginary part
        inverse = np.linalg.pinv(filter_length / hop_length * basis).T  # audio/stft.py:33-35 (inverse transform: unused here)
        basis = torch.from_numpy(basis[:, None, :]).float()
        basis = basis * torch.from_numpy(0.5 - 0.5 * np.cos(2.0 * np.pi * n / win_length)).float()  # periodic hann
        self.register_buffer("forward_basis", basis)
        # the reference also registers the windowed pseudo-inverse basis of its inverse transform (audio/stft.py:33-49);
        # it is not used on this path, but the buffer is kept so that state_dicts are interchangeable (strict=True)
        self.register_buffer("inverse_basis", torch.from_numpy(inverse[:, None, :]).float()
                             * torch.from_numpy(0.5 - 0.5 * np.cos(2.0 * np.pi * n / win_length)).float())
        self._packed = None

    def _weights(self):
        fb = self.forward_basis
        key = (fb.data_ptr(), fb._version)
        if self._packed is None or self._packed[0] != key:
            rows, n_fft, hop = fb.shape[0], self.filter_length, self.hop_length
            rows4 = (rows + 3) // 4 * 4  # the kernel wants Cout % 4 == 0: pad with zero rows
            w = torch.zeros(rows4, n_fft, dtype=torch.float32, device=fb.device)
            w[:rows] = fb[:, 0, :]
            # conv weight [Cout, Cin = hop, k = n_fft / hop]: tap j multiplies samples j*hop .. (j+1)*hop - 1 of the frame
            self._packed = (key, ops.pack_conv(w.view(rows4, n_fft // hop, hop).permute(0, 2, 1).contiguous()), rows4)
        return self._packed[1], self._packed[2]

    @torch.no_grad()
    def transform(self, input_data, want_phase=False):
        """input_data [B, T] on the device -> magnitude [B, n_fft/2+1, F] (a transposed view of the channels-last
        result; the phase the reference also returns is only needed by Griffin-Lim and is not produced)."""
        mag, _ = self._magnitude(input_data, want_energy=False)
        if want_phase:
            raise NotImplementedError("STFT.transform: phase is not produced (Griffin-Lim is off the path)")
        return mag[..., : self.filter_length // 2 + 1].transpose(1, 2)

    def _magnitude(self, y, want_energy=True):
        if not y.is_cuda:
            raise RuntimeError("speakingstyle_b200 STFT needs CUDA tensors; there is no CPU fallback")
        y = y.to(torch.float32).contiguous()
        B, T = y.shape
        n_fft, hop = self.filter_length, self.hop_length
        F = T // hop + 1
        rows = F + n_fft // hop - 1
        ypad = ops.reflect_pad(y, n_fft // 2, rows * hop)
        w, cout = self._weights()
        spec = ops.conv1d(ypad.view(B, rows, hop), w, None, pad=0, s_out=F)  # [B, F, cout]: re | im | zero pad
        nb = n_fft // 2 + 1
        return ops.stft_magnitude(spec, nb, want_energy=want_energy)


class TacotronSTFT(nn.Module):
    """audio/stft.py:130-178."""

    def __init__(self, filter_length, hop_length, win_length, n_mel_channels, sampling_rate, mel_fmin, mel_fmax):
        super().__init__()
        self.n_mel_channels, self.sampling_rate = n_mel_channels, sampling_rate
        self.stft_fn = STFT(filter_length, hop_length, win_length)
        self.register_buffer("mel_basis", torch.from_numpy(
            mel_filterbank(sampling_rate, filter_length, n_mel_channels, mel_fmin, mel_fmax)))
        self._packed = None

    def _weights(self, width):
        mb = self.mel_basis
        key = (mb.data_ptr(), mb._version, width)
        if self._packed is None or self._packed[0] != key:
            w = torch.zeros(mb.shape[0], width, dtype=torch.float32, device=mb.device)
            w[:, : mb.shape[1]] = mb
            self._packed = (key, ops.pack_conv(w))
        return self._packed[1]

    def spectral_normalize(self, magnitudes):
        return torch.log(torch.clamp(magnitudes, min=1e-5))

    def spectral_de_normalize(self, magnitudes):
        return torch.exp(magnitudes)

    @torch.no_grad()
    def mel_spectrogram(self, y, channels_last=False):
        """y [B, T] float in [-1, 1] on the device -> (log-mel [B, n_mel, F], energy [B, F]); channels_last=True
        returns the mel as [B, F, n_mel], the layout FastSpeech2.forward takes its reference mels in."""
        with torch.cuda.device(y.device):  # launches go to the wave's own device, whatever the current one is
            if ops.out_of_range(y.contiguous(), -1.0, 1.0):  # stft.py:168-169
                raise AssertionError("mel_spectrogram: wave must be in [-1, 1]")
            mag, energy = self.stft_fn._magnitude(y)
            mel = ops.conv1d(mag, self._weights(mag.shape[-1]), None)  # [B, F, n_mel]
            ops.log_clamp_(mel, 1e-5)
            return (mel if channels_last else mel.transpose(1, 2)), energy


def get_mel_from_wav(audio, _stft):
    """audio/tools.py:8-15 for a single 1-D wave (numpy or tensor): returns numpy (mel [n_mel, F], energy [F])."""
    dev = _stft.mel_basis.device
    a = torch.clip(torch.as_tensor(np.asarray(audio), dtype=torch.float32), -1, 1).unsqueeze(0).to(dev)
    mel, energy = _stft.mel_spectrogram(a)
    return mel[0].contiguous().cpu().numpy().astype(np.float32), energy[0].cpu().numpy().astype(np.float32)
