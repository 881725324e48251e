"""TEST INFRASTRUCTURE ONLY - writes tests/golden/*.npz from the REAL reference (/root/reference).

Run in the build container (the reference cannot travel to the GPU box):

    python -m oracle.make_golden            # tiny cases + cfg1
    python -m oracle.make_golden --large    # also cfg2 / cfg3 (minutes of CPU)

Weights and batches come from speakingstyle_b200/synth.py (seeded, reproducible anywhere) and are
loaded into the reference modules with strict=True.  Small cases store every output and a few
intermediates captured with forward hooks; the BASELINE.json-sized cases store the integer
outputs in full and a strided sample of the float tensors (files stay well under 1 MB each).
"""
import argparse
import os
import sys
import tempfile
import time

import numpy as np
import torch

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, _ROOT)

from oracle import ref_harness as rh  # noqa: E402
from speakingstyle_b200 import configs as C  # noqa: E402
from speakingstyle_b200 import synth as S  # noqa: E402

GOLDEN = os.path.join(_ROOT, "tests", "golden")
FS2_SEED, VOC_SEED = 1234, 4321
STRIDE = {"cfg1": 16, "cfg2": 32, "cfg3": 32}
WAVE_STRIDE = 64


def run_reference(name, controls=(1.0, 1.0, 1.0), batch_seed=1):
    ds, batch, voc = S.make_workload(name, seed=batch_seed)
    d = tempfile.mkdtemp()
    C.write_preprocessed_dir(d, ds)
    mc, pc = C.model_config(ds), C.preprocess_config(ds, d)
    sd = S.fastspeech2_state_dict(mc, ds, seed=FS2_SEED)
    model = rh.build_fastspeech2(pc, mc, sd)
    cap = {}
    hooks = [
        model.reference_encoder.register_forward_hook(lambda m, i, o: cap.update(gammas=o[0], betas=o[1])),
        model.encoder.register_forward_hook(lambda m, i, o: cap.update(encoder=o)),
        model.variance_adaptor.register_forward_hook(lambda m, i, o: cap.update(lr_out=o[0])),
        model.decoder.register_forward_hook(lambda m, i, o: cap.update(decoder=o[0])),
    ]
    t0 = time.time()
    with torch.no_grad(), rh.quiet():
        out = model(batch["speakers"], batch["texts"], batch["src_lens"], batch["max_src_len"], batch["mels"],
                    batch["mel_lens"], batch["max_mel_len"], p_control=controls[0], e_control=controls[1],
                    d_control=controls[2])
    dt = time.time() - t0
    for hk in hooks:
        hk.remove()
    wave = None
    if voc:
        h = C.hifigan_config()
        gen = rh.build_generator(h, S.hifigan_state_dict(h, seed=VOC_SEED))
        with torch.no_grad():
            wave = gen(out[1].transpose(1, 2))
    return out, cap, wave, dt


def save_full(name, tag, controls=(1.0, 1.0, 1.0)):
    out, cap, wave, dt = run_reference(name, controls)
    rec = {
        "mel": out[0], "postnet_mel": out[1], "pitch": out[2], "energy": out[3], "log_d": out[4],
        "d_rounded": out[5], "src_masks": out[6], "mel_masks": out[7], "src_lens": out[8], "mel_lens": out[9],
        "gammas": cap["gammas"], "betas": cap["betas"], "encoder": cap["encoder"], "lr_out": cap["lr_out"],
        "decoder": cap["decoder"], "controls": torch.tensor(controls, dtype=torch.float64),
    }
    if wave is not None:
        rec["wave"] = wave
        rec["wave_int16"] = torch.from_numpy(
            (wave.squeeze(1).numpy() * 32768.0).astype("int16"))  # utils/model.py:105-108
    path = os.path.join(GOLDEN, f"{tag}.npz")
    np.savez_compressed(path, **{k: v.numpy() for k, v in rec.items()})
    print(f"{tag}: reference {dt:.2f}s -> {path} ({os.path.getsize(path) / 1024:.0f} KiB) mel_lens={out[9].tolist()}")


def save_compact(name):
    out, cap, wave, dt = run_reference(name)
    st = STRIDE[name]
    rec = {
        "pitch": out[2], "energy": out[3], "log_d": out[4], "d_rounded": out[5], "src_lens": out[8],
        "mel_lens": out[9], "gammas": cap["gammas"], "betas": cap["betas"],
        "mel_s": out[0][:, ::st], "postnet_mel_s": out[1][:, ::st], "stride": torch.tensor(st),
        "postnet_mel_sum": out[1].double().sum(dim=(1, 2)), "postnet_mel_abs": out[1].double().abs().sum(dim=(1, 2)),
    }
    if wave is not None:
        rec["wave_s"] = wave[:, 0, ::WAVE_STRIDE]
        rec["wave_stride"] = torch.tensor(WAVE_STRIDE)
        rec["wave_abs"] = wave.double().abs().sum(dim=(1, 2))
    path = os.path.join(GOLDEN, f"{name}.npz")
    np.savez_compressed(path, **{k: v.numpy() for k, v in rec.items()})
    print(f"{name}: reference {dt:.2f}s -> {path} ({os.path.getsize(path) / 1024:.0f} KiB) "
          f"T={out[0].shape[1]} frames={int(out[9].sum())}")


def save_teacher_forced():
    """Teacher-forced / fine-control call (p, e, d targets given) on the real reference."""
    ds = "BC2013"
    batch, tg = S.make_teacher_forced_batch()
    d = tempfile.mkdtemp()
    C.write_preprocessed_dir(d, ds)
    mc, pc = C.model_config(ds), C.preprocess_config(ds, d)
    model = rh.build_fastspeech2(pc, mc, S.fastspeech2_state_dict(mc, ds, seed=FS2_SEED))
    with torch.no_grad(), rh.quiet():
        out = model(batch["speakers"], batch["texts"], batch["src_lens"], batch["max_src_len"], batch["mels"],
                    batch["mel_lens"], batch["max_mel_len"], p_targets=tg["p_targets"], e_targets=tg["e_targets"],
                    d_targets=tg["d_targets"])
    names = ["mel", "postnet_mel", "pitch", "energy", "log_d", "d_rounded", "src_masks", "mel_masks", "src_lens",
             "mel_lens"]
    path = os.path.join(GOLDEN, "tiny_tf.npz")
    np.savez_compressed(path, **{n: o.numpy() for n, o in zip(names, out)})
    print(f"tiny_tf -> {path} ({os.path.getsize(path) / 1024:.0f} KiB) mel_lens={out[9].tolist()} dtype={out[9].dtype}")


def save_hifigan_small():
    h = C.hifigan_config()
    gen = rh.build_generator(h, S.hifigan_state_dict(h, seed=VOC_SEED))
    g = torch.Generator().manual_seed(77)
    mel = torch.randn(2, 80, 24, generator=g) * 1.5
    with torch.no_grad():
        wave = gen(mel)
    path = os.path.join(GOLDEN, "hifigan_small.npz")
    np.savez_compressed(path, mel=mel.numpy(), wave=wave.numpy(),
                        wave_int16=(wave.squeeze(1).numpy() * 32768.0).astype("int16"))
    print(f"hifigan_small -> {path} ({os.path.getsize(path) / 1024:.0f} KiB)")


def save_length_regulator():
    """LengthRegulator / bucketize / get_mask_from_lengths alone, on awkward inputs."""
    ns = rh.load()
    g = torch.Generator().manual_seed(9)
    x = torch.randn(4, 7, 16, generator=g)
    d = torch.tensor([[0, 3, 1, 0, 2, 5, 1], [2.7, -1.0, 0.0, 3.2, 1.9, -0.0, 4.0], [0] * 7,
                      [1, 1, 1, 1, 1, 1, 9]], dtype=torch.float32)
    lr = ns.modules.LengthRegulator()
    out, mel_len = lr(x, d, None)
    out_fix, mel_len_fix = lr(x, d, 24)
    dt = torch.tensor([[1, 0, 2, 3, 0, 0, 1], [0, 0, 0, 0, 0, 0, 4], [2] * 7, [1, 2, 3, 0, 2, 1, 0]])
    out_t, mel_len_t = lr(x, dt, None)
    edges = torch.linspace(-2.0, 3.0, 255)
    v = torch.cat([torch.randn(500, generator=g) * 2.0, edges[::17], torch.tensor([-5.0, 9.0, float("nan"), -2.0, 3.0])])
    idx = torch.bucketize(v, edges)
    lens = torch.tensor([3, 0, 7, 5])
    np.savez_compressed(
        os.path.join(GOLDEN, "length_regulator.npz"), x=x.numpy(), d=d.numpy(), out=out.numpy(),
        mel_len=mel_len.numpy(), out_fix=out_fix.numpy(), mel_len_fix=mel_len_fix.numpy(), d_int=dt.numpy(),
        out_int=out_t.numpy(), mel_len_int=mel_len_t.numpy(), edges=edges.numpy(), v=v.numpy(), idx=idx.numpy(),
        lens=lens.numpy(), mask=ns.get_mask_from_lengths(lens).numpy(),
        mask9=ns.get_mask_from_lengths(lens, 9).numpy(),
        pe_256=ns.get_sinusoid_encoding_table(40, 256).numpy(), pe_1024=ns.get_sinusoid_encoding_table(12, 1024).numpy())
    print("length_regulator.npz written")


def save_mel_frontend():
    """wave -> log-mel through the reference's own TacotronSTFT (audio/stft.py) - librosa's filterbank replaced by
    the restatement in oracle/mel_oracle.py (librosa is not installed: filterbank values unpinned, the rest pinned)."""
    st = rh.load_audio()
    t = st.TacotronSTFT(1024, 256, 1024, 80, 22050, 0.0, 8000.0)
    g = torch.Generator().manual_seed(505)
    n = torch.arange(9000, dtype=torch.float32)
    tone = 0.4 * torch.sin(2 * np.pi * 440.0 * n / 22050) + 0.2 * torch.sin(2 * np.pi * 3000.0 * n / 22050)
    y = torch.stack([tone + 0.05 * torch.randn(9000, generator=g), (torch.rand(9000, generator=g) * 2 - 1) * 0.9,
                     torch.cat([torch.zeros(4500), 0.5 * torch.randn(4500, generator=g).clamp(-1.9, 1.9)])]).clamp(-1, 1)
    y_odd = y[:2, :4099].contiguous()
    with torch.no_grad():
        mel, en = t.mel_spectrogram(y)
        mel_o, en_o = t.mel_spectrogram(y_odd)
    fb = t.stft_fn.forward_basis
    path = os.path.join(GOLDEN, "mel_frontend.npz")
    np.savez_compressed(path, y=y.numpy(), mel=mel.numpy(), energy=en.numpy(), y_odd=y_odd.numpy(), mel_odd=mel_o.numpy(),
                        energy_odd=en_o.numpy(), mel_basis=t.mel_basis.numpy(), forward_basis_s=fb[::19, 0, ::7].numpy(),
                        forward_basis_abs=np.float64(fb.double().abs().sum().item()))
    print(f"mel_frontend -> {path} ({os.path.getsize(path) / 1024:.0f} KiB) mel {tuple(mel.shape)}")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--large", action="store_true")
    ap.add_argument("--only", default=None)
    args = ap.parse_args()
    torch.set_num_threads(os.cpu_count() or 1)
    os.makedirs(GOLDEN, exist_ok=True)
    jobs = {
        "lr": save_length_regulator,
        "hifigan_small": save_hifigan_small,
        "tiny": lambda: save_full("tiny", "tiny"),
        "tiny_ms": lambda: save_full("tiny_ms", "tiny_ms"),
        "tiny_ctl": lambda: save_full("tiny", "tiny_ctl", controls=(1.2, 0.8, 1.3)),
        "tiny_tf": save_teacher_forced,
        "mel_frontend": save_mel_frontend,
        "cfg1": lambda: save_compact("cfg1"),
    }
    if args.large:
        jobs["cfg2"] = lambda: save_compact("cfg2")
        jobs["cfg3"] = lambda: save_compact("cfg3")
    for k, fn in jobs.items():
        if args.only is None or k in args.only.split(","):
            fn()


if __name__ == "__main__":
    main()
