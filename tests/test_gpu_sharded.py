"""BASELINE configs[3] / [4] shapes through the shard planner against fixtures produced by the REAL reference
(oracle/make_golden.py --large: cfg5s, cfg4s) and the ReferenceEncoder beyond its 1001-row position table.

A shard is a reference batch (results depend on how a batch is padded), so the fixtures were
produced shard by shard on the very plan `synth.Corpus(...).plan()` returns here; the GPU path is driven through
`shard.run_rank`, one rank after the other on the one device.  Integers bit-exact; floats <= 1e-3."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from speakingstyle_b200 import shard as SH  # noqa: E402
from speakingstyle_b200 import synth as S  # noqa: E402
from test_gpu_model import DEV, TOL, _report, get_models, get_vocoder, maxabs  # noqa: E402


def test_cfg5_shards_fastspeech2_hifigan_vs_reference(golden):
    """9 cfg5-sized shards (batch <= 32, L 40-110, Tr 200-1000): FastSpeech2 + ReferenceEncoder on every shard,
    HiFi-GAN through synthesize_batch (ragged schedule, trimmed int16 PCM) on the shards the fixture holds waves for."""
    from speakingstyle_b200 import tools

    g = golden("cfg5s")
    corpus = S.Corpus("cfg5", seed=1, n=256)
    model, mc, sd = get_models(corpus.dataset)
    gen, h, vsd = get_vocoder()
    n_ranks = 4
    plan = corpus.plan(n_ranks)
    assert len(plan.shards) == int(g["n_shards"]) >= 8
    ms, ws = int(g["mel_stride"]), int(g["wave_stride"])
    wave_shards = set(int(v) for v in g["wave_shards"])
    edges_p, edges_e = sd["variance_adaptor.pitch_bins"], sd["variance_adaptor.energy_bins"]
    worst_mel = worst_wave = 0.0
    flips = n_ids = 0
    seen = set()
    for rank in range(n_ranks):
        for sh in plan.shards_of(rank):
            i = sh.index
            assert list(g[f"utts_{i}"]) == sh.utterances  # the plan is part of the contract
            batch = corpus.shard_batch(sh)
            if i in wave_shards:
                out, waves = tools.synthesize_batch(model, gen, batch)
            else:
                out = model(batch["speakers"], batch["texts"], batch["src_lens"], batch["max_src_len"], batch["mels"],
                            batch["mel_lens"], batch["max_mel_len"])
            assert np.array_equal(out[5].cpu().numpy(), g[f"d_rounded_{i}"]), f"durations of shard {i}"
            assert np.array_equal(out[9].cpu().numpy(), g[f"mel_lens_{i}"]), f"mel_lens of shard {i}"
            # Bucket ids (recomputed from the returned predictions with the same edges) must equal the reference's; a
            # flip is only accepted where the reference's own value is within 2e-5 of a bin edge (fp32 summation-order
            # noise, SURVEY.md 7.2 - adjudicated against an fp64 oracle run in test_gpu_model.py).  A pitch flip changes
            # the energy predictor's input, so those utterances are left out of the float comparison (and counted).
            bad_rows = []
            for which, got_v, edges in (("pitch", out[2], edges_p), ("energy", out[3], edges_e)):
                ref_v = torch.from_numpy(g[f"{which}_{i}"])
                fl = torch.bucketize(got_v.cpu(), edges) != torch.bucketize(ref_v, edges)
                if bad_rows:
                    fl = fl & ~bad_rows[0][:, None]  # energy ids downstream of a pitch flip are not comparable
                if bool(fl.any()):
                    margin = (ref_v[fl][:, None] - edges[None, :]).abs().min(dim=1).values
                    assert margin.max().item() < 2e-5, f"shard {i}: {which} bucket flip away from an edge: {margin.tolist()}"
                    flips += int(fl.sum())
                n_ids += fl.numel()
                bad_rows.append(fl.any(1))
            ok = ~(bad_rows[0] | bad_rows[1])
            assert maxabs(out[2], g[f"pitch_{i}"]) <= 2e-4
            worst_mel = max(worst_mel, maxabs(out[1][:, ::ms][ok], g[f"postnet_s_{i}"][ok.numpy()]))
            ref_abs = torch.from_numpy(g[f"postnet_abs_{i}"])
            got_abs = out[1].double().abs().sum(dim=(1, 2)).cpu()
            assert (((got_abs - ref_abs).abs() / ref_abs)[ok]).max().item() <= 1e-5
            if i in wave_shards:
                mel_lens = out[9].cpu().tolist()
                assert [len(w) for w in waves] == [m * 256 for m in mel_lens]
                for r, w in enumerate(waves):
                    if not bool(ok[r]) or len(w) == 0:
                        continue
                    ref_s = g[f"wave_s_{i}"][r][: (len(w) + ws - 1) // ws]
                    got_s = w[::ws].astype(np.float64) / 32768.0
                    worst_wave = max(worst_wave, float(np.abs(got_s - ref_s).max()))
                    ref_sum = float(g[f"wave_abs_{i}"][r])
                    got_sum = float(np.abs(w.astype(np.float64)).sum() / 32768.0)
                    assert abs(got_sum - ref_sum) <= 1e-3 * ref_sum + len(w) / 32768.0  # int16 truncation: < 1 LSB per sample
            seen.add(i)
    assert seen == set(range(len(plan.shards)))
    _report(f"cfg5s: {len(plan.shards)} shards over {n_ranks} planned ranks; postnet max-abs {worst_mel:.3e}; "
            f"wave max-abs {worst_wave:.3e} (int16 grid 3.1e-5); near-edge bucket flips (margin < 2e-5): {flips} of {n_ids} ids")
    assert worst_mel <= TOL and worst_wave <= TOL
    assert flips <= max(1, n_ids // 1000)


def test_cfg4_sweep_over_eight_planned_ranks_vs_reference(golden):
    """ReferenceEncoder over 1024 mels (Tr 100-1000, bucket width 64, batch <= 64), planned for 8 ranks."""
    g = golden("cfg4s")
    corpus = S.Corpus("cfg4", seed=1, n=1024)
    model, mc, sd = get_models(corpus.dataset)
    plan = corpus.plan(8)
    assert len(plan.shards) == int(g["n_shards"])
    ga = torch.zeros(1024, 256, dtype=torch.float64)
    be = torch.zeros(1024, 256, dtype=torch.float64)

    def gpu(batch):
        a, b = model.reference_encoder(batch["mels"].to(DEV), batch["max_mel_len"], None,
                                       lens=batch["mel_lens"].to(DEV, torch.int32))
        return {"g": a[:, 0].double().cpu(), "b": b[:, 0].double().cpu()}

    texts = {i: torch.ones(1, dtype=torch.int64) for i in range(1024)}
    done = 0
    for rank in range(8):
        mels = {i: corpus.utterance(i)[1] for s_ in plan.shards_of(rank) for i in s_.utterances}
        res = SH.run_rank(plan, rank, gpu, corpus.speakers, texts, mels)
        for uid, r in res.items():
            ga[uid], be[uid] = r["g"], r["b"]
        done += len(res)
    assert done == 1024
    ev = int(g["every"])
    err = max(maxabs(ga[::ev], g["gammas"]), maxabs(be[::ev], g["betas"]))
    err_sum = max((ga.sum(1) - torch.from_numpy(g["gamma_sum"])).abs().max().item(),
                  (be.sum(1) - torch.from_numpy(g["beta_sum"])).abs().max().item())
    _report(f"cfg4s: 1024 mels, {len(plan.shards)} shards, 8 planned ranks (imbalance {plan.imbalance():.3f}): "
            f"gamma/beta max-abs {err:.2e}, per-utterance sum error {err_sum:.2e}")
    assert err <= 2e-4 and err_sum <= 2e-3


def test_reference_encoder_beyond_the_position_table(golden):
    """Tr = 1100 > 1001 rows of position_enc: the 1024-wide table is rebuilt on the fly (model/modules.py:373-380)."""
    g = golden("refenc_long")
    model, mc, sd = get_models("BC2013")
    lens = torch.from_numpy(g["mel_lens"])
    gen = torch.Generator().manual_seed(606)
    mels = torch.randn(3, 1100, 80, generator=gen) * (torch.arange(1100)[None, :, None] < lens[:, None, None])
    ga, be = model.reference_encoder(mels.to(DEV), 1100, None, lens=lens.to(DEV, torch.int32))
    err = max(maxabs(ga, g["gammas"]), maxabs(be, g["betas"]))
    _report(f"refenc_long (Tr 1100 / 1003 / 640): gamma/beta max-abs {err:.2e}")
    assert ga.shape == g["gammas"].shape and err <= 2e-4
