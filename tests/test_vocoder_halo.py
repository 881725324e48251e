"""The ragged vocoder schedule's halos (hifigan.Generator._halo_by_rate) on the CPU: against an exact integer
receptive-field calculation over the generator's layers, and against the oracle's HiFi-GAN - frames past
`n + halo_frames() + 3` (the 3 = conv_pre's own k = 7 reach into the mel) must not reach any of an utterance's n*hop samples
(utils/model.py:109-113 is what makes them unobservable)."""
import torch

from oracle import fs2_oracle as O
from speakingstyle_b200 import configs as C
from speakingstyle_b200 import synth as S
from speakingstyle_b200.hifigan import Generator


def exact_reach(h, n_frames):
    """{rows per frame: largest row index of a tensor at that rate that can reach a sample < n_frames*hop}, walking the
    layers backwards in exact integer arithmetic (hifigan/models.py:149-165)."""
    rates = [1]
    for u in h["upsample_rates"]:
        rates.append(rates[-1] * u)
    t = n_frames * rates[-1] - 1 + 3  # conv_post, k = 7: stage output rows
    need = {}
    for i in reversed(range(len(h["upsample_rates"]))):
        # three ResBlocks in parallel, three units each: y' = y + c2(c1(y)), c1 dilated
        rb = max(sum((k - 1) // 2 * (d + 1) for d in ds)
                 for k, ds in zip(h["resblock_kernel_sizes"], h["resblock_dilation_sizes"]))
        t += rb
        need[rates[i + 1]] = t  # every layer of this stage (and the ups output) needs at most this row
        u, ku = h["upsample_rates"][i], h["upsample_kernel_sizes"][i]
        t = (t + (ku - u) // 2) // u  # ConvTranspose1d: output t reads inputs s with s*u - pad <= t
    need[1] = t
    return need


def test_halos_cover_the_exact_receptive_field():
    h = C.hifigan_config()
    gen = Generator(h)
    halo = gen._halo_by_rate()
    assert gen.halo_frames() == max(halo.values())
    for n in (0, 1, 2, 7, 100, 101, 589, 1000):
        need = exact_reach(h, n) if n > 0 else {}
        for rate, last_row in need.items():
            computed = (n + halo[rate]) * rate  # rows [0, computed) of that tensor are computed for this utterance
            assert last_row < computed, (n, rate, last_row, computed)
            # and the bound is not wasteful: within three frames of the exact need
            assert computed - 1 - last_row <= 3 * rate, (n, rate)


def test_frames_past_the_halo_do_not_reach_the_kept_samples():
    h = C.hifigan_config()
    sd = S.fold_weight_norm(S.hifigan_state_dict(h))
    gen = Generator(h)
    n, T = 5, 40
    cut = n + gen.halo_frames() + 3
    g = torch.Generator().manual_seed(3)
    mel_a = torch.randn(1, 80, T, generator=g)
    mel_b = mel_a.clone()
    mel_b[:, :, cut:] = torch.randn(1, 80, T - cut, generator=g) * 3.0
    with torch.no_grad():
        wa, wb = O.hifigan_forward(sd, h, mel_a), O.hifigan_forward(sd, h, mel_b)
    hop = 256
    assert torch.allclose(wa[..., : n * hop], wb[..., : n * hop], rtol=0, atol=1e-7)
    assert not torch.allclose(wa[..., (cut + 4) * hop:], wb[..., (cut + 4) * hop:], atol=1e-3)  # the change is real
