"""Reference-mel front-end: wave -> reflect pad -> STFT -> magnitude -> mel -> log.

CPU: the oracle restatement (oracle/mel_oracle.py) against fixtures written by the reference's own TacotronSTFT
(tests/golden/mel_frontend.npz, oracle/make_golden.py), and the product's constant buffers against the reference's.
GPU: the CUDA path through the C ABI against the same fixtures.  The mel FILTERBANK values are a restatement of
librosa's algorithm on both sides (librosa is not installed): parity unpinned for those 80 x 513 numbers only."""
import numpy as np
import pytest
import torch

from oracle import mel_oracle as M

DEV = "cuda:0"


def test_oracle_matches_reference_fixture(golden):
    g = golden("mel_frontend")
    for ykey, mkey, ekey in (("y", "mel", "energy"), ("y_odd", "mel_odd", "energy_odd")):
        mel, en = M.mel_spectrogram(g[ykey])
        assert mel.shape == g[mkey].shape and en.shape == g[ekey].shape
        assert np.abs(mel - g[mkey]).max() <= 2e-5
        assert (np.abs(en - g[ekey]) / np.maximum(g[ekey], 1e-3)).max() <= 2e-5
    assert np.array_equal(M.slaney_mel(22050, 1024, 80, 0.0, 8000.0), g["mel_basis"])


def test_reference_state_dict_loads_strictly_and_buffers_agree():
    """Live reference (build container only): a reference TacotronSTFT state_dict loads with strict=True and this
    package's own buffers equal the reference's."""
    from oracle import ref_harness as rh

    if not rh.available():
        pytest.skip("reference checkout not present")
    from speakingstyle_b200.audio import TacotronSTFT

    ref = rh.load_audio().TacotronSTFT(1024, 256, 1024, 80, 22050, 0.0, 8000.0)
    ours = TacotronSTFT(1024, 256, 1024, 80, 22050, 0.0, 8000.0)
    own = {k: v.clone() for k, v in ours.state_dict().items()}
    ours.load_state_dict(ref.state_dict(), strict=True)
    for k, v in ref.state_dict().items():
        assert (own[k] - v).abs().max().item() <= 1e-6, k


def test_filterbank_has_the_published_shape():
    """Properties of librosa's slaney filterbank that do not depend on its version: triangular, area-normalised,
    band edges on the Slaney mel scale (linear below 1 kHz, log above), nothing above fmax."""
    w = M.slaney_mel(22050, 1024, 80, 0.0, 8000.0).astype(np.float64)
    freqs = np.linspace(0, 22050 / 2, 513)
    assert w.shape == (80, 513) and (w >= 0).all() and (w[:, freqs > 8000.0] == 0).all()
    peaks = freqs[w.argmax(1)]
    assert (np.diff(peaks) > 0).all()
    area = (w * (freqs[1] - freqs[0])).sum(1)
    assert np.abs(area - 1.0).max() < 0.08  # unit area per filter up to bin quantisation
    lin = peaks[peaks < 900]
    assert np.abs(np.diff(lin) - np.diff(lin).mean()).max() < 2 * (freqs[1] - freqs[0])  # evenly spaced below 1 kHz


def test_product_buffers_match_the_reference(golden):
    from speakingstyle_b200.audio import TacotronSTFT

    g = golden("mel_frontend")
    t = TacotronSTFT(1024, 256, 1024, 80, 22050, 0.0, 8000.0)
    # the reference's buffer names (audio/stft.py:48-49, 149): state_dicts are interchangeable, strict=True
    assert sorted(t.state_dict().keys()) == ["mel_basis", "stft_fn.forward_basis", "stft_fn.inverse_basis"]
    assert tuple(t.stft_fn.inverse_basis.shape) == (1026, 1, 1024)
    assert np.array_equal(t.mel_basis.numpy(), g["mel_basis"])
    fb = t.stft_fn.forward_basis
    assert tuple(fb.shape) == (1026, 1, 1024)
    assert np.abs(fb[::19, 0, ::7].numpy() - g["forward_basis_s"]).max() <= 1e-6
    assert abs(fb.double().abs().sum().item() - float(g["forward_basis_abs"])) / float(g["forward_basis_abs"]) <= 1e-6


@pytest.mark.gpu
def test_gpu_mel_spectrogram_matches_reference_fixture(golden):
    from speakingstyle_b200.audio import TacotronSTFT, get_mel_from_wav

    g = golden("mel_frontend")
    t = TacotronSTFT(1024, 256, 1024, 80, 22050, 0.0, 8000.0).to(DEV)
    for ykey, mkey, ekey in (("y", "mel", "energy"), ("y_odd", "mel_odd", "energy_odd")):
        mel, en = t.mel_spectrogram(torch.from_numpy(g[ykey]).to(DEV))
        assert tuple(mel.shape) == g[mkey].shape and tuple(en.shape) == g[ekey].shape
        assert np.abs(mel.cpu().numpy() - g[mkey]).max() <= 1e-4
        assert (np.abs(en.cpu().numpy() - g[ekey]) / np.maximum(g[ekey], 1e-3)).max() <= 1e-5
    # channels-last result is what FastSpeech2.forward takes as reference mels
    cl, _ = t.mel_spectrogram(torch.from_numpy(g["y"]).to(DEV), channels_last=True)
    assert torch.equal(cl.transpose(1, 2), t.mel_spectrogram(torch.from_numpy(g["y"]).to(DEV))[0])
    # single-wave helper of audio/tools.py
    m1, e1 = get_mel_from_wav(g["y"][1], t)
    assert m1.shape == (80, g["mel"].shape[2]) and np.abs(m1 - g["mel"][1]).max() <= 1e-4 and e1.shape == (g["mel"].shape[2],)
    # magnitude alone (STFT.transform)
    mag = t.stft_fn.transform(torch.from_numpy(g["y"]).to(DEV))
    assert tuple(mag.shape) == (3, 513, g["mel"].shape[2])
    with pytest.raises(AssertionError):
        t.mel_spectrogram(torch.full((1, 3000), 1.5, device=DEV))


@pytest.mark.gpu
def test_gpu_mel_front_end_feeds_the_reference_encoder():
    """wave -> mel -> FastSpeech2 (the loop README.md:7 of the reference describes): a long batch at full size runs
    and matches the oracle's mel within the log-mel tolerance."""
    from speakingstyle_b200.audio import TacotronSTFT

    t = TacotronSTFT(1024, 256, 1024, 80, 22050, 0.0, 8000.0).to(DEV)
    gen = torch.Generator().manual_seed(11)
    y = (torch.rand(8, 22050 * 3, generator=gen) * 2 - 1) * 0.7
    mel, en = t.mel_spectrogram(y.to(DEV), channels_last=True)
    want, want_en = M.mel_spectrogram(y.numpy())
    assert tuple(mel.shape) == (8, 22050 * 3 // 256 + 1, 80)
    assert np.abs(mel.cpu().numpy() - want.transpose(0, 2, 1)).max() <= 1e-4
    assert (np.abs(en.cpu().numpy() - want_en) / want_en).max() <= 1e-5
