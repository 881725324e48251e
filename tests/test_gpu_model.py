"""End-to-end parity of the drop-in modules (which run only C-ABI kernels) against
(a) the committed golden fixtures produced by the real reference and (b) the oracle run live on the
same seeded weights and batches.  Tolerances are north_star's: integers bit-exact (durations,
mel_lens, masks, bucket ids, mel2phone); fp32 mel / wave max-abs <= 1e-3."""
import tempfile

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import fs2_oracle as O  # noqa: E402
from speakingstyle_b200 import configs as C  # noqa: E402
from speakingstyle_b200 import synth as S  # noqa: E402

DEV = "cuda:0"
TOL = 1e-3


_models = {}


def get_models(ds):
    import speakingstyle_b200 as ss

    if ds not in _models:
        d = C.write_preprocessed_dir(tempfile.mkdtemp(), ds)
        mc, pc = C.model_config(ds), C.preprocess_config(ds, d)
        sd = S.fastspeech2_state_dict(mc, ds)
        _models[ds] = (ss.get_model(pc, mc, DEV, sd), mc, sd)
    return _models[ds]


_voc = {}


def get_vocoder():
    import speakingstyle_b200 as ss

    if not _voc:
        h = C.hifigan_config()
        vsd = S.hifigan_state_dict(h)
        _voc["v"] = (ss.get_vocoder(h, DEV, vsd), h, S.fold_weight_norm(vsd))
    return _voc["v"]


def run(model, batch, controls=(1.0, 1.0, 1.0), extra=False):
    return model(batch["speakers"], batch["texts"], batch["src_lens"], batch["max_src_len"], batch["mels"],
                 batch["mel_lens"], batch["max_mel_len"], p_control=controls[0], e_control=controls[1],
                 d_control=controls[2], return_extra=extra)


def maxabs(a, b):
    return (a.detach().cpu().double() - torch.as_tensor(b).double()).abs().max().item()


@pytest.mark.parametrize("name,tag,controls", [("tiny", "tiny", (1.0, 1.0, 1.0)), ("tiny_ms", "tiny_ms", (1.0, 1.0, 1.0)),
                                                ("tiny", "tiny_ctl", (1.2, 0.8, 1.3))])
def test_fastspeech2_small_vs_golden(golden, name, tag, controls):
    g = golden(tag)
    ds, batch, voc = S.make_workload(name)
    model, mc, sd = get_models(ds)
    out, extra = run(model, batch, controls, extra=True)
    mel, post, pitch, energy, log_d, d_rounded, src_masks, mel_masks, src_lens, mel_lens = out
    # dtypes / shapes of the 10-tuple
    assert mel.dtype == torch.float32 and d_rounded.dtype == torch.float32
    assert src_masks.dtype == torch.bool and mel_masks.dtype == torch.bool and mel_lens.dtype == torch.int32
    # integers: bit-exact
    assert np.array_equal(d_rounded.cpu().numpy(), g["d_rounded"])
    assert np.array_equal(src_masks.cpu().numpy(), g["src_masks"])
    assert np.array_equal(src_lens.cpu().numpy(), g["src_lens"])
    if controls[2] == 1.0:
        assert np.array_equal(mel_lens.cpu().numpy(), g["mel_lens"])
        assert np.array_equal(mel_masks.cpu().numpy(), g["mel_masks"])
    else:
        # non-integer durations: mel_len = int(float32 sum) depends on the summation order, which differs
        # between ATen's vectorised CPU sum and an index-order sum when the sum sits on an integer
        assert np.abs(mel_lens.cpu().numpy().astype(np.int64) - g["mel_lens"]).max() <= 1
    assert mel.shape == g["mel"].shape
    # floats
    for key, val in (("gammas", extra["gammas"]), ("betas", extra["betas"]), ("pitch", pitch), ("energy", energy),
                     ("log_d", log_d), ("mel", mel), ("postnet_mel", post)):
        assert maxabs(val, g[key]) <= TOL, key
    # bucket ids and frame->phoneme map against the oracle
    with torch.no_grad():
        r = O.fastspeech2_forward(sd, mc, batch, *controls, intermediates=True)
    assert torch.equal(extra["pitch_idx"].cpu().long(), r["pitch_idx"])
    assert torch.equal(extra["energy_idx"].cpu().long(), r["energy_idx"])
    assert torch.equal(extra["mel2phone"].cpu().long(), r["mel2phone"])
    if voc:
        gen, h, vsd = get_vocoder()
        wave, pcm = gen(post, channels_last=True, pcm=True)
        assert wave.shape == g["wave"].shape
        assert maxabs(wave, g["wave"]) <= TOL
        # mean-removed relative L2 (a DC-dominated wave would hide errors in max-abs)
        gw = torch.from_numpy(g["wave"]).double()
        ac = gw - gw.mean()
        rel = ((wave.cpu().double() - gw).norm() / ac.norm()).item()
        assert rel <= 1e-3, rel
        assert np.abs(pcm.cpu().numpy().astype(np.int32) - g["wave_int16"].astype(np.int32)).max() <= 33


def test_frame_level_pitch_energy_vs_golden(golden):
    """config/LJSpeech_paper (preprocess.yaml:26,29 frame_level, model.yaml:19 log pitch bins, 4 decoder layers): the
    pitch / energy predictors run on the EXPANDED sequence after the LengthRegulator (model/modules.py:146-155) and the
    decoder's position table is added after their embeddings.  Fixture from the real reference."""
    g = golden("tiny_fl")
    ds, batch, voc = S.make_workload("tiny_fl")
    model, mc, sd = get_models(ds)
    assert model.variance_adaptor.pitch_feature_level == "frame_level" and len(model.decoder.layer_stack) == 4
    out, extra = run(model, batch, extra=True)
    assert np.array_equal(out[5].cpu().numpy(), g["d_rounded"])
    assert np.array_equal(out[9].cpu().numpy(), g["mel_lens"])
    assert np.array_equal(out[7].cpu().numpy(), g["mel_masks"])
    assert out[2].shape == g["pitch"].shape == out[0].shape[:2]  # [B, T]: one value per frame
    assert maxabs(out[4], g["log_d"]) <= 1e-4
    # un-normalised features are O(100): 2e-3 absolute is ~1e-5 relative
    assert maxabs(out[2], g["pitch"]) <= 2e-3 and maxabs(out[3], g["energy"]) <= 2e-3
    with torch.no_grad():
        r = O.fastspeech2_forward(sd, mc, batch, intermediates=True, levels=("frame_level", "frame_level"))
    p_same = extra["pitch_idx"].cpu().long() == r["pitch_idx"]
    e_same = extra["energy_idx"].cpu().long() == r["energy_idx"]
    assert p_same.float().mean().item() >= 0.995 and e_same.float().mean().item() >= 0.995  # [B, T] ids; log bins are narrow at the low end
    ok = (p_same.all(1) & e_same.all(1))
    assert int(ok.sum()) >= 2
    assert maxabs(out[0][ok], g["mel"][ok.numpy()]) <= TOL and maxabs(out[1][ok], g["postnet_mel"][ok.numpy()]) <= TOL


def test_per_phoneme_control_vectors_match_the_oracle():
    """p_control / e_control / d_control as [B, L] tensors (notebooks/control.ipynb cells 17-23): the reference's
    `prediction * control` and `round(...) * d_control` broadcast a tensor just as they do a number
    (model/modules.py:88,99,137-140)."""
    ds, batch, voc = S.make_workload("tiny")
    model, mc, sd = get_models(ds)
    B, L = batch["texts"].shape
    g = torch.Generator().manual_seed(99)
    pc = 0.5 + torch.rand(B, L, generator=g)
    ec = 0.5 + torch.rand(B, L, generator=g)
    dc = torch.randint(1, 4, (B, L), generator=g).float()  # integer factors: mel_len = int(sum) is unambiguous
    out, extra = run(model, batch, (pc, ec, dc), extra=True)
    with torch.no_grad():
        r = O.fastspeech2_forward(sd, mc, batch, pc, ec, dc, intermediates=True)
    ref = r["out"]
    assert torch.equal(out[5].cpu(), ref[5]) and torch.equal(out[9].cpu(), ref[9])
    assert torch.equal(extra["pitch_idx"].cpu().long(), r["pitch_idx"])
    assert torch.equal(extra["energy_idx"].cpu().long(), r["energy_idx"])
    assert maxabs(out[2], ref[2]) <= 1e-4 and maxabs(out[3], ref[3]) <= 1e-4
    assert maxabs(out[1], ref[1]) <= TOL
    # a [L] vector broadcasts over the batch like the reference's multiply
    out2 = run(model, batch, (pc[0], 1.0, 1.0))
    with torch.no_grad():
        ref2 = O.fastspeech2_forward(sd, mc, batch, pc[0], 1.0, 1.0)
    assert torch.equal(out2[5].cpu(), ref2[5]) and maxabs(out2[2], ref2[2]) <= 1e-4


def test_max_wav_value_reaches_the_pcm_epilogue(golden):
    """utils/model.py:105-108 scales by preprocess_config audio.max_wav_value before astype(int16)."""
    from speakingstyle_b200 import tools
    g = golden("hifigan_small")
    gen, h, vsd = get_vocoder()
    mel = torch.from_numpy(g["mel"]).to(DEV)
    pc = C.preprocess_config("BC2013")
    pc["preprocessing"]["audio"]["max_wav_value"] = 16384.0
    wavs = tools.vocoder_infer(mel, gen, None, pc)
    want = (g["wave"][:, 0] * 16384.0).astype(np.int16)
    assert all(np.abs(w.astype(np.int32) - want[i].astype(np.int32)).max() <= 2 for i, w in enumerate(wavs))


def test_model_on_another_device_than_the_current_one():
    """The library launches on the current device: forward() makes the model's own device current for the call."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import speakingstyle_b200 as ss
    ds, batch, voc = S.make_workload("tiny")
    d = C.write_preprocessed_dir(tempfile.mkdtemp(), ds)
    mc, pc = C.model_config(ds), C.preprocess_config(ds, d)
    sd = S.fastspeech2_state_dict(mc, ds)
    m1 = ss.get_model(pc, mc, "cuda:1", sd)
    assert torch.cuda.current_device() == 0
    out1 = run(m1, batch)
    out0 = run(get_models(ds)[0], batch)
    assert out1[1].device.index == 1 and torch.cuda.current_device() == 0
    assert torch.equal(out1[1].cpu(), out0[1].cpu()) and torch.equal(out1[5].cpu(), out0[5].cpu())


def _report(line):
    import os

    d = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    if os.path.isdir(d):
        with open(os.path.join(d, "parity_report.txt"), "a") as f:
            f.write(line + "\n")


def _bucket_flips(name, which, got_idx, golden_values, edges, exclude_rows=None, fp64=None):
    """Bucket ids must equal the reference's.  The one exception: a value so close to a bin edge that two fp32
    summation orders land on different sides.  Such a flip is only accepted if (a) the reference's own
    fp32 value lies within MARGIN of the edge AND (b) an fp64 run of the oracle on the same batch adjudicates it: the
    id computed here equals the fp64 id, or the fp64 value itself is within MARGIN of the edge (nobody is "right").
    ``fp64`` is a callable returning the fp64 oracle's intermediates (run only when there is a flip).
    Returns the per-row 'has a flip' mask."""
    MARGIN = 2e-5  # 0.04 % of a bin width (0.056); observed flips sit at 1e-6 .. 7e-6
    gv = torch.from_numpy(golden_values)
    want = torch.bucketize(gv, edges)
    got = got_idx.cpu().long()
    flips = got != want
    if exclude_rows is not None:
        flips = flips & ~exclude_rows[:, None]
    n = int(flips.sum())
    if n:
        margin = (gv[flips][:, None] - edges[None, :]).abs().min(dim=1).values
        _report(f"{name} {which}: {n} of {flips.numel()} bucket ids differ; fp32-reference margins {margin.tolist()}")
        assert n <= max(1, flips.numel() // 1000), f"{n} {which} bucket flips"
        assert margin.max().item() < MARGIN, f"{which} bucket flip away from an edge: margins {margin.tolist()}"
        assert fp64 is not None, "a bucket flip needs the fp64 adjudication"
        r64 = fp64()
        v64 = r64[which][flips]
        e64 = edges.double()
        id64 = torch.bucketize(v64, e64)
        m64 = (v64[:, None] - e64[None, :]).abs().min(dim=1).values
        agree = id64 == got[flips]
        _report(f"{name} {which}: fp64 adjudication - this implementation agrees with fp64 on {int(agree.sum())} of {n}; "
                f"fp64 margins {m64.tolist()}")
        assert bool((agree | (m64 < MARGIN)).all()), f"{which} flip not explained by an edge: fp64 margins {m64.tolist()}"
    return flips.any(dim=1)


@pytest.mark.parametrize("name", ["cfg1", "cfg2", "cfg3"])
def test_fastspeech2_baseline_configs_vs_golden(golden, name):
    """BASELINE.json configs 1-3 at full size against the strided golden samples."""
    g = golden(name)
    ds, batch, voc = S.make_workload(name)
    model, mc, sd = get_models(ds)
    out, extra = run(model, batch, extra=True)
    st = int(g["stride"])
    d_equal = np.array_equal(out[5].cpu().numpy(), g["d_rounded"])
    if not d_equal:  # report the rounding margin of every mismatching token before failing
        bad = np.argwhere(out[5].cpu().numpy() != g["d_rounded"])
        v = np.exp(g["log_d"][tuple(bad.T)]) - 1
        pytest.fail(f"{len(bad)} duration mismatches; distance to .5 boundary: {np.abs((v % 1) - 0.5)}")
    assert np.array_equal(out[9].cpu().numpy(), g["mel_lens"])
    assert np.array_equal(out[8].cpu().numpy(), g["src_lens"])
    assert maxabs(extra["gammas"], g["gammas"]) <= 1e-4 and maxabs(extra["betas"], g["betas"]) <= 1e-4
    assert maxabs(out[4], g["log_d"]) <= 1e-4
    assert maxabs(out[2], g["pitch"]) <= 2e-4
    # a pitch bucket flip changes the energy predictor's input for that utterance, so rows with one
    # are left out of the energy / mel / wave comparison (and reported)
    cache = {}

    def fp64():  # the oracle in float64 on the same batch, only if a flip has to be adjudicated
        if not cache:
            with torch.no_grad():
                cache.update(O.fastspeech2_forward(sd, mc, batch, dtype=torch.float64, intermediates=True))
        return cache

    p_rows = _bucket_flips(name, "pitch", extra["pitch_idx"], g["pitch"], sd["variance_adaptor.pitch_bins"], fp64=fp64)
    e_rows = _bucket_flips(name, "energy", extra["energy_idx"], g["energy"], sd["variance_adaptor.energy_bins"],
                           exclude_rows=p_rows, fp64=fp64)
    ok = ~(p_rows | e_rows)
    assert int(ok.sum()) >= int(0.9 * ok.numel())
    _report(f"{name}: rows compared {int(ok.sum())}/{ok.numel()}; mel max-abs "
            f"{maxabs(out[0][:, ::st][ok], g['mel_s'][ok.numpy()]):.3e}; postnet "
            f"{maxabs(out[1][:, ::st][ok], g['postnet_mel_s'][ok.numpy()]):.3e}")
    assert maxabs(out[3][~p_rows], g["energy"][(~p_rows).numpy()]) <= 2e-4
    assert maxabs(out[0][:, ::st][ok], g["mel_s"][ok.numpy()]) <= TOL
    assert maxabs(out[1][:, ::st][ok], g["postnet_mel_s"][ok.numpy()]) <= TOL
    got_sum = out[1].double().sum(dim=(1, 2)).cpu()
    assert (got_sum - torch.from_numpy(g["postnet_mel_sum"]))[ok].abs().max().item() <= 0.5
    if voc:
        gen, h, vsd = get_vocoder()
        wave = gen(out[1], channels_last=True)
        ws = int(g["wave_stride"])
        _report(f"{name}: wave max-abs {maxabs(wave[:, 0, ::ws][ok], g['wave_s'][ok.numpy()]):.3e}")
        assert maxabs(wave[:, 0, ::ws][ok], g["wave_s"][ok.numpy()]) <= TOL
        wabs = torch.from_numpy(g["wave_abs"])
        rel = ((wave.double().abs().sum(dim=(1, 2)).cpu() - wabs).abs() / wabs)[ok].max().item()
        assert rel <= 1e-4


def test_hifigan_small_vs_golden_and_stages(golden):
    g = golden("hifigan_small")
    gen, h, vsd = get_vocoder()
    mel = torch.from_numpy(g["mel"]).to(DEV)
    wave, pcm = gen(mel, pcm=True)
    assert wave.shape == g["wave"].shape
    assert maxabs(wave, g["wave"]) <= 1e-4
    assert np.abs(pcm.cpu().numpy().astype(np.int32) - g["wave_int16"].astype(np.int32)).max() <= 4
    # channels_last entry gives the same result
    wave2 = gen(mel.transpose(1, 2).contiguous(), channels_last=True)
    assert torch.equal(wave, wave2)
    # odd length (not a multiple of any tile) against the oracle
    mel3 = torch.randn(3, 80, 37, generator=torch.Generator().manual_seed(8))
    with torch.no_grad():
        ref = O.hifigan_forward(vsd, h, mel3)
    assert maxabs(gen(mel3.to(DEV)), ref) <= 1e-4


def test_vocoder_infer_and_synthesize_batch_return_trimmed_int16(golden):
    """utils/model.py:97-115 / utils/tools.py:222-230 through the ragged PCM epilogue: same samples as
    the padded int16 path, trimmed per utterance, delivered in pinned host memory."""
    from speakingstyle_b200 import tools
    g = golden("hifigan_small")
    gen, h, vsd = get_vocoder()
    mel = torch.from_numpy(g["mel"]).to(DEV)
    _, pcm = gen(mel, pcm=True)
    pcm = pcm.cpu().numpy()
    lengths = [pcm.shape[1] - 300, 256 * 3 + 1]
    wavs = tools.vocoder_infer(mel, gen, lengths=lengths)
    assert [w.dtype for w in wavs] == [np.int16] * 2 and [len(w) for w in wavs] == lengths
    for i, w in enumerate(wavs):
        assert np.array_equal(w, pcm[i, :lengths[i]])
        assert np.abs(w.astype(np.int32) - g["wave_int16"][i, :lengths[i]].astype(np.int32)).max() <= 4
    full = tools.vocoder_infer(mel, gen)  # lengths=None: whole padded waves, as the reference returns them
    assert all(np.array_equal(full[i], pcm[i]) for i in range(2))
    # end to end: FastSpeech2 -> Generator, waves trimmed to mel_len * hop
    ds, batch, _ = S.make_workload("tiny")
    model = get_models(ds)[0]
    out, waves = tools.synthesize_batch(model, gen, batch)
    mel_lens = out[9].cpu().tolist()
    assert [len(w) for w in waves] == [m * 256 for m in mel_lens]
    _, pcm2 = gen(out[1], channels_last=True, pcm=True)
    for i, w in enumerate(waves):
        assert np.array_equal(w, pcm2[i, :mel_lens[i] * 256].cpu().numpy())


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_ragged_vocoder_schedule_returns_the_same_samples(precision):
    """vocoder_infer / synthesize_batch trim wave b to lengths[b] (utils/model.py:109-113); the generator then runs every
    layer only on the frames that can reach a returned sample (Generator.forward frame_lens).  The returned samples
    must be bit-identical to trimming the full padded computation - for lengths near 0, mid-batch and the full length."""
    from speakingstyle_b200 import tools
    gen, h, vsd = get_vocoder()
    T = 150
    mel = torch.randn(12, 80, T, generator=torch.Generator().manual_seed(31)).to(DEV)
    # full length, empty, a few frames, mid-batch, just inside the largest halo, and odd sample counts around tile /
    # frame boundaries of the deep stages (their halos are only 2-3 frames, _halo_by_rate)
    lengths = [T * 256, 0, 256 * 3 + 1, 256 * 61, 256 * (T - gen.halo_frames() + 1) - 5, 256 * 100 + 255, 256 * 127,
               256 * 128 + 1, 1, 256 * 33 - 1, 256 * (T - 2), 256 * (T - 3) + 17]
    gen.precision = precision
    try:
        _, pcm = gen(mel, pcm=True)
        wavs = tools.vocoder_infer(mel, gen, lengths=lengths)
    finally:
        gen.precision = "fp32"
    pcm = pcm.cpu().numpy()
    assert [len(w) for w in wavs] == lengths
    for i, w in enumerate(wavs):
        assert np.array_equal(w, pcm[i, :lengths[i]]), i
    # the halo really is what bounds the schedule: stage-0 rows of utterance 2 stop well before the padded length
    rows = gen._ragged_rows([4], T, mel.device)
    halo = gen._halo_by_rate()
    assert int(rows[8][0]) == (4 + halo[8]) * 8 < T * 8 and int(rows[256][0]) == (4 + halo[256]) * 256
    assert halo[256] <= halo[128] <= halo[64] <= halo[8] <= halo[1] == gen.halo_frames()


@pytest.mark.parametrize("name", ["tiny", "tiny_ms"])
def test_batch_invariant_mode_is_independent_of_batch_composition(name):
    """SURVEY 8 (f4), opt-in: with the padding leaks closed, an utterance synthesised alone and inside a batch gives
    the same integers and the same mel, bit for bit - and the mode is off by default (reference behaviour, which
    test_outputs_depend_on_batch_composition_like_the_reference pins)."""
    import speakingstyle_b200 as ss

    ds, batch, voc = S.make_workload(name)
    d = C.write_preprocessed_dir(tempfile.mkdtemp(), ds)
    mc, pc = C.model_config(ds), C.preprocess_config(ds, d)
    model = ss.get_model(pc, mc, DEV, S.fastspeech2_state_dict(mc, ds))
    assert model.batch_invariant is False
    model.set_batch_invariant(True)
    out = run(model, batch)
    B = batch["texts"].shape[0]
    for i in range(B):
        L, Tr = int(batch["src_lens"][i]), int(batch["mel_lens"][i])
        one = {"speakers": batch["speakers"][i:i + 1], "texts": batch["texts"][i:i + 1, :L].contiguous(),
               "src_lens": batch["src_lens"][i:i + 1], "max_src_len": L,
               "mels": batch["mels"][i:i + 1, :Tr].contiguous(), "mel_lens": batch["mel_lens"][i:i + 1],
               "max_mel_len": Tr}
        o1 = run(model, one)
        n = int(o1[9][0])
        assert n == int(out[9][i])                                            # mel_len
        assert torch.equal(o1[5][0, :L], out[5][i, :L])                       # d_rounded
        for k in (2, 3, 4):                                                   # pitch, energy, log-duration
            assert torch.equal(o1[k][0, :L], out[k][i, :L]), k
        for k in (0, 1):                                                      # mel, postnet mel
            assert torch.equal(o1[k][0, :n], out[k][i, :n]), k
            assert float(out[k][i, n:].abs().sum()) == 0.0                    # pads are zeros in this mode
    # and it does change something: the default mode differs on the shorter utterances
    model.set_batch_invariant(False)
    ref = run(model, batch)
    assert not torch.equal(ref[1], out[1])


def test_reference_sweep_cfg4_sharded_matches_oracle():
    """BASELINE configs[3] at reduced size: ReferenceEncoder alone over length-bucketed mels (bucket width 64,
    batch <= 64), planned for 2 ranks, every shard checked against the oracle on the same shard (results depend
    on the shard's padding, so the plan is part of the contract)."""
    from speakingstyle_b200 import shard as SH

    model, mc, sd = get_models("BC2013")
    g = torch.Generator().manual_seed(404)
    n = 90
    lens = torch.randint(100, 1001, (n,), generator=g).tolist()
    mels = [torch.randn(t, 80, generator=g) for t in lens]
    plan = SH.plan_shards([0] * n, lens, max_batch=64, bucket_width=64, n_ranks=2, vocoder=False, text=False)
    assert sorted(u for s_ in plan.shards for u in s_.utterances) == list(range(n))
    texts = [torch.ones(1, dtype=torch.int64)] * n

    def gpu(batch):
        ga, be, pooled = model.reference_encoder(batch["mels"].to(DEV), batch["max_mel_len"], None,
                                                 lens=batch["mel_lens"].to(DEV, torch.int32), return_pooled=True)
        return {"gammas": ga[:, 0].cpu(), "betas": be[:, 0].cpu(), "pooled": pooled.cpu()}

    def cpu(batch):
        with torch.no_grad():
            ga, be, pooled = O.reference_encoder(sd, mc, batch["mels"], O.mask_from_lengths(batch["mel_lens"],
                                                                                           batch["max_mel_len"]),
                                                 return_pooled=True)
        return {"gammas": ga[:, 0], "betas": be[:, 0], "pooled": pooled}

    worst = 0.0
    for rank in range(2):
        got = SH.run_rank(plan, rank, gpu, [0] * n, texts, mels)
        want = SH.run_rank(plan, rank, cpu, [0] * n, texts, mels)
        assert sorted(got) == sorted(want) and len(got) > 0
        for uid in got:
            for k in ("gammas", "betas", "pooled"):
                worst = max(worst, maxabs(got[uid][k], want[uid][k]))
    _report(f"cfg4-style sweep ({n} mels, {len(plan.shards)} shards, 2 ranks): max-abs {worst:.2e}")
    assert worst <= 2e-4


def test_modules_standalone_api():
    """The sub-modules keep the reference's own signatures."""
    import speakingstyle_b200 as ss

    ds, batch, _ = S.make_workload("tiny")
    model, mc, sd = get_models(ds)
    with torch.no_grad():
        r = O.fastspeech2_forward(sd, mc, batch, intermediates=True)
    mels = batch["mels"].to(DEV)
    ref_mask = ss.get_mask_from_lengths(batch["mel_lens"].to(DEV), batch["max_mel_len"])
    gammas, betas = model.reference_encoder(mels, batch["max_mel_len"], ref_mask)
    assert gammas.shape == (3, 1, 256) and maxabs(gammas, r["gammas"]) <= 1e-4 and maxabs(betas, r["betas"]) <= 1e-4
    src_mask = ss.get_mask_from_lengths(batch["src_lens"].to(DEV), batch["max_src_len"])
    enc = model.encoder(batch["texts"].to(DEV), src_mask, gammas, betas)
    assert maxabs(enc, r["encoder"]) <= 1e-4
    va = model.variance_adaptor(enc, src_mask, None, None, None, None, None, 1.0, 1.0, 1.0, gammas, betas)
    assert len(va) == 7 and maxabs(va[0], r["x"]) <= 1e-4 and torch.equal(va[4].cpu(), r["d_rounded"])
    assert maxabs(enc, r["encoder"]) <= 1e-4  # the adaptor did not modify its input
    dec, dec_mask = model.decoder(va[0], va[6], gammas, betas)
    assert maxabs(dec, r["decoder"]) <= 2e-4
    # LengthRegulator(x, duration, max_len) -> (out, mel_len) with integer targets (teacher forcing)
    lr = ss.LengthRegulator()
    x = torch.randn(2, 5, 8, generator=torch.Generator().manual_seed(2))
    dur = torch.tensor([[1, 0, 3, 2, 0], [0, 0, 4, 0, 1]])
    out, mel_len = lr(x.to(DEV), dur.to(DEV), None)
    want, want_len, _ = O.length_regulate(x, dur.float())
    assert torch.equal(out.cpu(), want) and torch.equal(mel_len.cpu(), want_len) and mel_len.dtype == torch.int64
    out, _ = lr(x.to(DEV), dur.to(DEV), 9)
    assert out.shape == (2, 9, 8)
    # PostNet on its own
    post = model.postnet(torch.from_numpy(np.ascontiguousarray(r["out"][0].numpy())).to(DEV))
    assert maxabs(post, r["out"][1] - r["out"][0]) <= 1e-4


def test_outputs_depend_on_batch_composition_like_the_reference():
    """Padding leaks are part of the contract: the same utterance alone and inside
    a padded batch gives different results in the reference; the drop-in must follow the oracle in both."""
    ds, batch, _ = S.make_workload("tiny")
    model, mc, sd = get_models(ds)
    j = int(torch.argmin(batch["src_lens"]))
    L, Tr = int(batch["src_lens"][j]), int(batch["mel_lens"][j])
    solo = {"speakers": batch["speakers"][j:j + 1], "texts": batch["texts"][j:j + 1, :L].contiguous(),
            "src_lens": batch["src_lens"][j:j + 1], "max_src_len": L,
            "mels": batch["mels"][j:j + 1, :Tr].contiguous(), "mel_lens": batch["mel_lens"][j:j + 1], "max_mel_len": Tr}
    out = run(model, solo)
    with torch.no_grad():
        ref = O.fastspeech2_forward(sd, mc, solo)
    assert torch.equal(out[5].cpu(), ref[5]) and maxabs(out[1], ref[1]) <= TOL


def _custom_batch(src_lens, mel_lens, seed=3):
    g = torch.Generator().manual_seed(seed)
    B, L, Tr = len(src_lens), max(src_lens), max(mel_lens)
    sl, ml = torch.tensor(src_lens), torch.tensor(mel_lens)
    texts = torch.randint(1, C.N_SRC_VOCAB, (B, L), generator=g) * (torch.arange(L)[None, :] < sl[:, None])
    mels = torch.randn(B, Tr, 80, generator=g) * (torch.arange(Tr)[None, :, None] < ml[:, None, None])
    return {"speakers": torch.zeros(B, dtype=torch.int64), "texts": texts, "src_lens": sl, "max_src_len": L,
            "mels": mels, "mel_lens": ml, "max_mel_len": Tr}


@pytest.mark.parametrize("src_lens,mel_lens", [
    ([1, 120, 37], [1, 400, 50]),      # one-token utterance with a one-frame reference next to long ones
    ([170, 90], [300, 1000]),          # > 1000 output frames (on-the-fly position table, Models.py:154-162) and the
                                       # longest reference the encoder's table covers
    ([7], [1]),                        # batch of one
])
def test_edge_shapes_against_the_oracle(src_lens, mel_lens):
    """Ragged and extreme shapes: every integer output exact, floats within the fp32 tolerance, against the oracle
    (itself pinned to the reference by tests/test_oracle.py)."""
    model, mc, sd = get_models("BC2013")
    batch = _custom_batch(src_lens, mel_lens)
    out = run(model, batch)
    with torch.no_grad():
        ref = O.fastspeech2_forward(sd, mc, batch)
    assert out[0].shape == ref[0].shape
    assert torch.equal(out[5].cpu(), ref[5])                                   # d_rounded
    assert torch.equal(out[9].cpu().to(torch.int64), ref[9].to(torch.int64))   # mel_lens
    assert torch.equal(out[6].cpu(), ref[6]) and torch.equal(out[7].cpu(), ref[7])
    for k in (0, 1, 2, 3, 4):
        assert maxabs(out[k], ref[k]) <= TOL, k
    _report(f"edge shapes src={src_lens} ref={mel_lens}: T={out[0].shape[1]}, mel max-abs {maxabs(out[1], ref[1]):.2e}")


# ---------------------------------------------------------------------------------------------
# bf16 precision mode (tcgen05 convolutions): integers identical, floats within relative L2 1e-2
# ---------------------------------------------------------------------------------------------
def rel_l2(a, b):
    a, b = a.detach().cpu().double(), torch.as_tensor(b).double()
    return ((a - b).norm() / b.norm()).item()


@pytest.mark.parametrize("name", ["tiny", "tiny_ms", "cfg1", "cfg3"])
def test_bf16_mode_matches_within_relative_l2(golden, name):
    g = golden(name)
    ds, batch, voc = S.make_workload(name)
    model, mc, sd = get_models(ds)
    model.precision = "bf16"
    try:
        out = run(model, batch)
    finally:
        model.precision = "fp32"
    # the integer-deciding chain is fp32 in both modes: identical integers
    assert np.array_equal(out[5].cpu().numpy(), g["d_rounded"])
    assert np.array_equal(out[9].cpu().numpy(), g["mel_lens"])
    if "mel" in g:
        mel_ref, post_ref, sl = g["mel"], g["postnet_mel"], slice(None)
    else:
        st = int(g["stride"])
        mel_ref, post_ref, sl = g["mel_s"], g["postnet_mel_s"], slice(None, None, st)
    r_mel, r_post = rel_l2(out[0][:, sl], mel_ref), rel_l2(out[1][:, sl], post_ref)
    _report(f"{name} bf16 mode: mel rel-L2 {r_mel:.3e}, postnet rel-L2 {r_post:.3e}")
    assert r_mel <= 1e-2 and r_post <= 1e-2
    if voc:
        gen, h, vsd = get_vocoder()
        gen.precision = "bf16"
        try:
            wave = gen(out[1], channels_last=True)
        finally:
            gen.precision = "fp32"
        if "wave" in g:
            ref = torch.from_numpy(g["wave"])
            got = wave.cpu()
        else:
            ws = int(g["wave_stride"])
            ref, got = torch.from_numpy(g["wave_s"]), wave[:, 0, ::ws].cpu()
        r = rel_l2(got, ref)
        ac = ref.double() - ref.double().mean()
        r_ac = ((got.double() - ref.double()).norm() / ac.norm()).item()  # against the mean-removed signal
        _report(f"{name} bf16 mode: wave rel-L2 {r:.3e} (mean-removed {r_ac:.3e})")
        assert r <= 1e-2 and r_ac <= 1.5e-2


def test_hifigan_bf16_alone(golden):
    g = golden("hifigan_small")
    gen, h, vsd = get_vocoder()
    gen.precision = "bf16"
    try:
        wave, pcm = gen(torch.from_numpy(g["mel"]).to(DEV), pcm=True)
    finally:
        gen.precision = "fp32"
    r = rel_l2(wave, g["wave"])
    _report(f"hifigan_small bf16 mode: wave rel-L2 {r:.3e}")
    assert wave.shape == g["wave"].shape and r <= 1e-2


def test_sharded_run_matches_oracle_per_shard():
    """A shard is a reference batch: the GPU path run shard by shard (one rank's view of the plan)
    equals the oracle run on the very same shards."""
    from speakingstyle_b200 import shard as SH

    model, mc, sd = get_models("BC2013")
    g = torch.Generator().manual_seed(21)
    n = 7
    texts = [torch.randint(1, C.N_SRC_VOCAB, (int(torch.randint(5, 14, (1,), generator=g)),), generator=g) for _ in range(n)]
    mels = [torch.randn(int(torch.randint(10, 50, (1,), generator=g)), 80, generator=g) for _ in range(n)]
    plan = SH.plan_shards([t.numel() for t in texts], [m.shape[0] for m in mels], max_batch=3, bucket_width=4, n_ranks=2,
                          vocoder=False)

    def gpu(batch):
        out = run(model, batch)
        return {"postnet_mel": out[1].cpu(), "d_rounded": out[5].cpu(), "mel_len": out[9].cpu()}

    def cpu(batch):
        with torch.no_grad():
            out = O.fastspeech2_forward(sd, mc, batch)
        return {"postnet_mel": out[1], "d_rounded": out[5], "mel_len": out[9]}

    for rank in range(2):
        a = SH.run_rank(plan, rank, gpu, [0] * n, texts, mels)
        b = SH.run_rank(plan, rank, cpu, [0] * n, texts, mels)
        assert sorted(a) == sorted(b) and len(a) > 0
        for uid in a:
            assert torch.equal(a[uid]["d_rounded"], b[uid]["d_rounded"])
            assert int(a[uid]["mel_len"]) == int(b[uid]["mel_len"])
            assert maxabs(a[uid]["postnet_mel"], b[uid]["postnet_mel"]) <= TOL


def test_teacher_forced_path_vs_golden(golden):
    """p_targets / e_targets / d_targets: embeddings from the targets, predictions returned
    unscaled, LengthRegulator expands by the given integer durations to max_mel_len, decoder mask = the
    reference mel's mask, mel_lens int64 from the regulator."""
    g = golden("tiny_tf")
    model, mc, sd = get_models("BC2013")
    batch, tg = S.make_teacher_forced_batch()
    out = model(batch["speakers"], batch["texts"], batch["src_lens"], batch["max_src_len"], batch["mels"],
                batch["mel_lens"], batch["max_mel_len"], p_targets=tg["p_targets"], e_targets=tg["e_targets"],
                d_targets=tg["d_targets"])
    names = ["mel", "postnet_mel", "pitch", "energy", "log_d", "d_rounded", "src_masks", "mel_masks", "src_lens", "mel_lens"]
    for n, o in zip(names, out):
        if o.dtype.is_floating_point:
            assert o.shape == g[n].shape and maxabs(o, g[n]) <= TOL, n
        else:
            assert np.array_equal(o.cpu().numpy(), g[n]), n
    assert out[9].dtype == torch.int64 and out[5].dtype == torch.int64


def test_fp32_mode_on_3xtf32_kernels_only(golden, monkeypatch):
    """SS_FP32_CONVS=tf32x3: the fp32 mode without the split-bf16 kernels - HiFi-GAN, decoder and PostNet on
    the 3xTF32 kernels, as before they existed.  Same fixtures, the tighter bars that path reaches."""
    import speakingstyle_b200 as ss
    from speakingstyle_b200 import ops

    monkeypatch.setenv("SS_FP32_CONVS", "tf32x3")
    assert ops.fp32_convs() == "tf32x3"
    g = golden("hifigan_small")
    h = C.hifigan_config()
    gen = ss.get_vocoder(h, DEV, S.hifigan_state_dict(h))
    assert gen.fp32_convs == "tf32x3"
    timer, ops.conv_timer = ops.conv_timer, ops.KernelTimer()
    try:
        wave = gen(torch.from_numpy(g["mel"]).to(DEV))
        kinds = {m[0] for _, _, _, _, m in ops.conv_timer.rows()}
    finally:
        ops.conv_timer = timer
    assert kinds == {"tf32x3"}, kinds
    assert maxabs(wave, g["wave"]) <= 2e-5
    g = golden("tiny")
    ds, batch, _ = S.make_workload("tiny")
    model, mc, sd = get_models(ds)
    out = run(model, batch)  # FastSpeech2 reads the environment variable per call
    assert maxabs(out[0], g["mel"]) <= 2e-5 and maxabs(out[1], g["postnet_mel"]) <= 2e-5
    monkeypatch.delenv("SS_FP32_CONVS")
    wave_default = get_vocoder()[0](torch.from_numpy(golden("hifigan_small")["mel"]).to(DEV))
    assert maxabs(wave_default, wave.cpu()) <= 1e-4  # the two fp32 forms agree far inside the 1e-3 bar
