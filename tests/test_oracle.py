"""Pins the oracle: (1) against tests/golden (outputs of the real reference, made by
oracle/make_golden.py), (2) against the live reference where /root/reference exists."""
import ctypes
import tempfile

import numpy as np
import pytest
import torch

from oracle import fs2_oracle as O
from oracle import intops, ref_harness
from speakingstyle_b200 import configs as C
from speakingstyle_b200 import synth as S

F32P = ctypes.POINTER(ctypes.c_float)
I64P = ctypes.POINTER(ctypes.c_int64)


def _p(a, t):
    return a.ctypes.data_as(t)


def _oracle_run(name, controls=(1.0, 1.0, 1.0), dtype=torch.float32):
    ds, batch, voc = S.make_workload(name)
    mc = C.model_config(ds)
    sd = S.fastspeech2_state_dict(mc, ds)
    pp = C.preprocess_config(ds)["preprocessing"]
    with torch.no_grad():
        r = O.fastspeech2_forward(sd, mc, batch, *controls, dtype=dtype, intermediates=True,
                                  levels=(pp["pitch"]["feature"], pp["energy"]["feature"]))
    return r, voc


@pytest.mark.parametrize("name,tag,controls", [("tiny", "tiny", (1.0, 1.0, 1.0)),
                                                ("tiny_ms", "tiny_ms", (1.0, 1.0, 1.0)),
                                                ("tiny", "tiny_ctl", (1.2, 0.8, 1.3)),
                                                # config/LJSpeech_paper: frame-level pitch / energy, log pitch bins
                                                ("tiny_fl", "tiny_fl", (1.0, 1.0, 1.0))])
def test_fastspeech2_oracle_matches_golden(golden, name, tag, controls):
    g = golden(tag)
    r, voc = _oracle_run(name, controls)
    out = r["out"]
    # integers: exact
    assert np.array_equal(out[5].numpy(), g["d_rounded"])
    assert np.array_equal(out[9].numpy(), g["mel_lens"]) and out[9].dtype == torch.int32
    assert np.array_equal(out[6].numpy(), g["src_masks"])
    assert np.array_equal(out[7].numpy(), g["mel_masks"])
    assert np.array_equal(out[8].numpy(), g["src_lens"])
    # floats: same ATen kernels -> tight tolerance (not bit-exact across thread counts)
    for key, val in (("mel", out[0]), ("postnet_mel", out[1]), ("pitch", out[2]), ("energy", out[3]),
                     ("log_d", out[4]), ("gammas", r["gammas"]), ("betas", r["betas"]),
                     ("encoder", r["encoder"]), ("lr_out", r["x"]), ("decoder", r["decoder"])):
        # un-normalised frame-level features are O(100): the same relative accuracy is a larger absolute one
        atol = 2e-3 if (name == "tiny_fl" and key in ("pitch", "energy")) else 2e-5
        np.testing.assert_allclose(val.numpy(), g[key], atol=atol, rtol=0, err_msg=key)
    if voc and "wave" in g:
        h = C.hifigan_config()
        vsd = S.fold_weight_norm(S.hifigan_state_dict(h))
        with torch.no_grad():
            wave = O.hifigan_forward(vsd, h, out[1].transpose(1, 2))
        np.testing.assert_allclose(wave.numpy(), g["wave"], atol=2e-5, rtol=0)


def test_fastspeech2_oracle_float64_agrees_on_integers(golden):
    """fp32 vs fp64 oracle agree on every duration: the rounding margins are wide enough for an
    fp32-accurate kernel to be bit-exact."""
    g = golden("tiny")
    r, _ = _oracle_run("tiny", dtype=torch.float64)
    assert np.array_equal(r["out"][5].numpy(), g["d_rounded"])
    assert np.array_equal(r["out"][9].numpy(), g["mel_lens"])


def test_cfg1_compact_golden(golden):
    g = golden("cfg1")
    r, _ = _oracle_run("cfg1")
    out = r["out"]
    st = int(g["stride"])
    assert np.array_equal(out[5].numpy(), g["d_rounded"])
    assert np.array_equal(out[9].numpy(), g["mel_lens"])
    np.testing.assert_allclose(out[1][:, ::st].numpy(), g["postnet_mel_s"], atol=5e-5, rtol=0)
    np.testing.assert_allclose(r["gammas"].numpy(), g["gammas"], atol=2e-5, rtol=0)


def test_hifigan_oracle_matches_golden(golden):
    g = golden("hifigan_small")
    h = C.hifigan_config()
    sd = S.fold_weight_norm(S.hifigan_state_dict(h))
    with torch.no_grad():
        wave = O.hifigan_forward(sd, h, torch.from_numpy(g["mel"]))
    np.testing.assert_allclose(wave.numpy(), g["wave"], atol=2e-5, rtol=0)
    w16 = np.stack(O.wave_to_int16(wave))
    assert np.abs(w16.astype(np.int32) - g["wave_int16"].astype(np.int32)).max() <= 1
    # C restatement of the int16 epilogue on the golden float wave: exact
    flat = np.ascontiguousarray(g["wave"].reshape(-1))
    out = np.empty(flat.size, dtype=np.int16)
    intops().ss_oracle_wave_int16(_p(flat, F32P), flat.size, 32768.0, out.ctypes.data_as(ctypes.POINTER(ctypes.c_int16)))
    assert np.array_equal(out, g["wave_int16"].reshape(-1))


def test_integer_steps_match_golden(golden):
    g = golden("length_regulator")
    lib = intops()
    x, d = g["x"], g["d"]
    B, L, Cc = x.shape
    for dur, want_out, want_len, max_len in ((d, g["out"], g["mel_len"], None),
                                             (d, g["out_fix"], g["mel_len_fix"], 24),
                                             (g["d_int"].astype(np.float32), g["out_int"], g["mel_len_int"], None)):
        # torch restatement
        out, mel_len, m2p = O.length_regulate(torch.from_numpy(x), torch.from_numpy(dur), max_len)
        assert np.array_equal(out.numpy(), want_out)
        assert np.array_equal(mel_len.numpy(), want_len)
        # C restatement
        dur = np.ascontiguousarray(dur, dtype=np.float32)
        ml = np.zeros(B, dtype=np.int64)
        T = lib.ss_oracle_length_regulate(_p(x, F32P), _p(dur, F32P), B, L, Cc, -1 if max_len is None else max_len,
                                          None, _p(ml, I64P), None)
        assert T == want_out.shape[1]
        o = np.empty((B, T, Cc), dtype=np.float32)
        m2 = np.empty((B, T), dtype=np.int64)
        lib.ss_oracle_length_regulate(_p(x, F32P), _p(dur, F32P), B, L, Cc, -1 if max_len is None else max_len,
                                      _p(o, F32P), _p(ml, I64P), _p(m2, I64P))
        assert np.array_equal(o, want_out) and np.array_equal(ml, want_len)
        assert np.array_equal(m2, m2p.numpy())
    # bucketize incl. exact edges, out-of-range and NaN
    v, edges = g["v"], g["edges"]
    idx = np.empty(v.size, dtype=np.int64)
    lib.ss_oracle_bucketize(_p(v, F32P), v.size, _p(edges, F32P), edges.size, _p(idx, I64P))
    assert np.array_equal(idx, g["idx"])
    assert np.array_equal(O.bucketize(torch.from_numpy(v), torch.from_numpy(edges)).numpy(), g["idx"])
    # masks and position table
    assert np.array_equal(O.mask_from_lengths(torch.from_numpy(g["lens"])).numpy(), g["mask"])
    assert np.array_equal(O.mask_from_lengths(torch.from_numpy(g["lens"]), 9).numpy(), g["mask9"])
    assert np.array_equal(O.sinusoid_table(40, 256).numpy(), g["pe_256"])
    assert np.array_equal(S.sinusoid_table(12, 1024).numpy(), g["pe_1024"])


def test_round_durations_c_vs_torch(golden):
    g = golden("tiny")
    logd = np.ascontiguousarray(g["log_d"].reshape(-1))
    out = np.empty_like(logd)
    intops().ss_oracle_round_durations(_p(logd, F32P), 1.0, logd.size, _p(out, F32P))
    assert np.array_equal(out, g["d_rounded"].reshape(-1))
    ml = np.empty(g["d_rounded"].shape[0], dtype=np.int32)
    dr = np.ascontiguousarray(g["d_rounded"])
    intops().ss_oracle_mel_len(_p(dr, F32P), dr.shape[0], dr.shape[1], ml.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)))
    assert np.array_equal(ml, g["mel_lens"])


@pytest.mark.skipif(not ref_harness.available(), reason="needs /root/reference (build container only)")
def test_oracle_vs_live_reference():
    ds, batch, _ = S.make_workload("tiny_ms", seed=11)
    d = C.write_preprocessed_dir(tempfile.mkdtemp(), ds)
    mc, pc = C.model_config(ds), C.preprocess_config(ds, d)
    sd = S.fastspeech2_state_dict(mc, ds, seed=99)
    model = ref_harness.build_fastspeech2(pc, mc, sd)  # strict=True: the key/shape contract
    with torch.no_grad(), ref_harness.quiet():
        ref = model(batch["speakers"], batch["texts"], batch["src_lens"], batch["max_src_len"], batch["mels"],
                    batch["mel_lens"], batch["max_mel_len"])
        mine = O.fastspeech2_forward(sd, mc, batch)
    for a, b in zip(ref, mine):
        assert a.dtype == b.dtype and a.shape == b.shape
        if a.dtype.is_floating_point:
            assert (a - b).abs().max().item() <= 2e-5
        else:
            assert torch.equal(a, b)
    h = C.hifigan_config()
    vsd = S.hifigan_state_dict(h, seed=5)
    gen = ref_harness.build_generator(h, vsd)
    mel = torch.randn(1, 80, 12, generator=torch.Generator().manual_seed(3))
    with torch.no_grad():
        assert (gen(mel) - O.hifigan_forward(S.fold_weight_norm(vsd), h, mel)).abs().max().item() <= 2e-5


@pytest.mark.skipif(not ref_harness.available(), reason="needs /root/reference (build container only)")
def test_oracle_vs_live_reference_beyond_the_position_table():
    """More than 1000 output frames: the reference's decoder builds its position table on the fly (Models.py:154-162)
    and crops nothing; the oracle must follow it (the GPU test test_edge_shapes_against_the_oracle leans on this)."""
    ds = "BC2013"
    g = torch.Generator().manual_seed(3)
    sl, ml = torch.tensor([170, 90]), torch.tensor([300, 1000])
    L, Tr = 170, 1000
    texts = torch.randint(1, C.N_SRC_VOCAB, (2, L), generator=g) * (torch.arange(L)[None, :] < sl[:, None])
    mels = torch.randn(2, Tr, 80, generator=g) * (torch.arange(Tr)[None, :, None] < ml[:, None, None])
    batch = {"speakers": torch.zeros(2, dtype=torch.int64), "texts": texts, "src_lens": sl, "max_src_len": L,
             "mels": mels, "mel_lens": ml, "max_mel_len": Tr}
    d = C.write_preprocessed_dir(tempfile.mkdtemp(), ds)
    mc, pc = C.model_config(ds), C.preprocess_config(ds, d)
    sd = S.fastspeech2_state_dict(mc, ds)
    model = ref_harness.build_fastspeech2(pc, mc, sd)
    with torch.no_grad(), ref_harness.quiet():
        ref = model(batch["speakers"], batch["texts"], batch["src_lens"], batch["max_src_len"], batch["mels"],
                    batch["mel_lens"], batch["max_mel_len"])
        mine = O.fastspeech2_forward(sd, mc, batch)
    assert ref[0].shape[1] > 1000
    for a, b in zip(ref, mine):
        assert a.dtype == b.dtype and a.shape == b.shape
        if a.dtype.is_floating_point:
            assert (a - b).abs().max().item() <= 5e-5
        else:
            assert torch.equal(a, b)


def test_teacher_forced_oracle_matches_golden(golden):
    """p/e/d targets given (the training / fine-control call)."""
    g = golden("tiny_tf")
    mc = C.model_config("BC2013")
    sd = S.fastspeech2_state_dict(mc, "BC2013")
    batch, tg = S.make_teacher_forced_batch()
    with torch.no_grad():
        out = O.fastspeech2_forward(sd, mc, batch, **tg)
    names = ["mel", "postnet_mel", "pitch", "energy", "log_d", "d_rounded", "src_masks", "mel_masks", "src_lens", "mel_lens"]
    for n, o in zip(names, out):
        if o.dtype.is_floating_point:
            np.testing.assert_allclose(o.numpy(), g[n], atol=2e-5, rtol=0, err_msg=n)
        else:
            assert np.array_equal(o.numpy(), g[n]) and o.numpy().dtype == g[n].dtype, n
