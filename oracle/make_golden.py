"""TEST INFRASTRUCTURE ONLY - writes tests/golden/*.npz from the REAL reference (/root/reference).

Run in the build container (the reference cannot travel to the GPU box):

    python -m oracle.make_golden            # tiny cases + cfg1
    python -m oracle.make_golden --large    # also cfg2 / cfg3 / cfg4s / cfg5s (minutes of CPU)

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


def _reference_model(ds):
    d = tempfile.mkdtemp()
    C.write_preprocessed_dir(d, ds)
    mc, pc = C.model_config(ds), C.preprocess_config(ds, d)
    return rh.build_fastspeech2(pc, mc, S.fastspeech2_state_dict(mc, ds, seed=FS2_SEED))


CFG5S_UTTS, CFG5S_MEL_STRIDE, CFG5S_WAVE_STRIDE, CFG5S_WAVE_SHARDS = 256, 64, 256, (0, 5)
CFG4S_MELS, CFG4S_RANKS, CFG4S_FULL_EVERY = 1024, 8, 4


def save_cfg5_shards():
    """BASELINE configs[4] at reduced size: the first 256 utterances of the cfg5 corpus (L 40-110, Tr 200-1000) planned
    by shard.plan_shards (batch <= 32, bucket width 16) - 9 shards - each run through the real reference as one batch.
    Per shard: integers in full, pitch / energy predictions in full (bucket ids), a strided sample of postnet_mel and
    per-utterance checksums; for two shards also the HiFi-GAN wave (strided sample + per-utterance |.| sums over the
    samples vocoder_infer keeps, utils/model.py:109-113)."""
    corpus = S.Corpus("cfg5", seed=1, n=CFG5S_UTTS)
    plan = corpus.plan(1)
    model = _reference_model(corpus.dataset)
    h = C.hifigan_config()
    gen = rh.build_generator(h, S.hifigan_state_dict(h, seed=VOC_SEED))
    rec = {"n_shards": np.int64(len(plan.shards)), "mel_stride": np.int64(CFG5S_MEL_STRIDE),
           "wave_stride": np.int64(CFG5S_WAVE_STRIDE), "wave_shards": np.array(CFG5S_WAVE_SHARDS)}
    t0 = time.time()
    for sh in plan.shards:
        b = corpus.shard_batch(sh)
        with torch.no_grad(), rh.quiet():
            out = model(b["speakers"], b["texts"], b["src_lens"], b["max_src_len"], b["mels"], b["mel_lens"], b["max_mel_len"])
        i = sh.index
        rec[f"utts_{i}"] = np.array(sh.utterances)
        rec[f"d_rounded_{i}"] = out[5].numpy()
        rec[f"mel_lens_{i}"] = out[9].numpy()
        rec[f"pitch_{i}"] = out[2].numpy()
        rec[f"energy_{i}"] = out[3].numpy()
        rec[f"postnet_s_{i}"] = out[1][:, ::CFG5S_MEL_STRIDE].numpy()
        rec[f"postnet_abs_{i}"] = out[1].double().abs().sum(dim=(1, 2)).numpy()
        if i in CFG5S_WAVE_SHARDS:
            with torch.no_grad():
                wave = gen(out[1].transpose(1, 2))[:, 0]
            n = (out[9] * 256).tolist()
            rec[f"wave_s_{i}"] = wave[:, ::CFG5S_WAVE_STRIDE].numpy()
            rec[f"wave_abs_{i}"] = np.array([wave[r, : n[r]].double().abs().sum().item() for r in range(len(n))])
        print(f"cfg5s shard {i}: B={len(sh.utterances)} L={sh.max_src_len} Tr={sh.max_ref_len} T={out[1].shape[1]} "
              f"frames={int(out[9].sum())} ({time.time() - t0:.0f}s)", flush=True)
    path = os.path.join(GOLDEN, "cfg5s.npz")
    np.savez_compressed(path, **rec)
    print(f"cfg5s -> {path} ({os.path.getsize(path) / 1024:.0f} KiB)")


def save_cfg4_sweep():
    """BASELINE configs[3] at reduced size: ReferenceEncoder alone over the first 1024 mels of the cfg4 corpus
    (Tr 100-1000), length-bucketed (width 64, batch <= 64) and planned for 8 ranks; every shard through the real
    reference's ReferenceEncoder.  gammas / betas in full for every 4th utterance, their sums for all."""
    corpus = S.Corpus("cfg4", seed=1, n=CFG4S_MELS)
    plan = corpus.plan(CFG4S_RANKS)
    enc = _reference_model(corpus.dataset).reference_encoder
    ns = rh.load()
    ga = torch.zeros(CFG4S_MELS, 256)
    be = torch.zeros(CFG4S_MELS, 256)
    t0 = time.time()
    for sh in plan.shards:
        b = corpus.shard_batch(sh)
        with torch.no_grad(), rh.quiet():
            g_, b_ = enc(b["mels"], b["max_mel_len"], ns.get_mask_from_lengths(b["mel_lens"], b["max_mel_len"]))
        ids = torch.tensor(sh.utterances)
        ga[ids], be[ids] = g_[:, 0], b_[:, 0]
    path = os.path.join(GOLDEN, "cfg4s.npz")
    np.savez_compressed(path, gammas=ga[::CFG4S_FULL_EVERY].numpy(), betas=be[::CFG4S_FULL_EVERY].numpy(),
                        gamma_sum=ga.double().sum(1).numpy(), beta_sum=be.double().sum(1).numpy(),
                        gamma_abs=ga.double().abs().sum(1).numpy(), beta_abs=be.double().abs().sum(1).numpy(),
                        every=np.int64(CFG4S_FULL_EVERY), n_shards=np.int64(len(plan.shards)))
    print(f"cfg4s: {len(plan.shards)} shards, {time.time() - t0:.0f}s -> {path} ({os.path.getsize(path) / 1024:.0f} KiB)")


def save_long_reference():
    """ReferenceEncoder with reference mels LONGER than the 1001-row position table (model/modules.py:373-380: the
    table is rebuilt on the fly for max_len > max_seq_len) - Tr = 1100 and 1003 in one batch."""
    enc = _reference_model("BC2013").reference_encoder
    ns = rh.load()
    g = torch.Generator().manual_seed(606)
    lens = torch.tensor([1100, 1003, 640])
    mels = torch.randn(3, 1100, 80, generator=g) * (torch.arange(1100)[None, :, None] < lens[:, None, None])
    with torch.no_grad(), rh.quiet():
        ga, be = enc(mels, 1100, ns.get_mask_from_lengths(lens, 1100))
    path = os.path.join(GOLDEN, "refenc_long.npz")
    np.savez_compressed(path, mel_lens=lens.numpy(), gammas=ga.numpy(), betas=be.numpy())
    print(f"refenc_long -> {path} ({os.path.getsize(path) / 1024:.0f} KiB)")


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
        "tiny_fl": lambda: save_full("tiny_fl", "tiny_fl"),
        "mel_frontend": save_mel_frontend,
        "cfg1": lambda: save_compact("cfg1"),
    }
    jobs["refenc_long"] = save_long_reference
    if args.large:
        jobs["cfg2"] = lambda: save_compact("cfg2")
        jobs["cfg3"] = lambda: save_compact("cfg3")
        jobs["cfg4s"] = save_cfg4_sweep
        jobs["cfg5s"] = save_cfg5_shards
    for k, fn in jobs.items():
        if args.only is None or k in args.only.split(","):
            fn()


if __name__ == "__main__":
    main()
