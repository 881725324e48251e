"""Deterministic synthetic weights and batches.

There are no checkpoints and no datasets offline, so every test and the benchmark run on
random-init weights of the reference architecture and on synthetic inputs of the shapes
BASELINE.json names.  Everything is drawn from ``torch.Generator`` objects seeded on the CPU, so
the very same tensors are produced here, on the GPU box, and by oracle/make_golden.py (which
loads them into the real reference with ``strict=True`` - that load is also the check that the
key names and shapes below are the reference's state_dict contract).

Choices that differ from the reference's *default* initialisers, on purpose:
* ``duration_predictor.linear_layer.bias = duration_bias`` (default 2.1) so an utterance of ~100
  phonemes expands to ~800 frames;
* LayerNorm / BatchNorm affine terms, BatchNorm running statistics and the FiLM scalars are
  perturbed away from (1, 0) so a kernel that ignored them could not pass parity;
* HiFi-GAN convolutions are drawn variance-preserving (the default N(0, 0.01) init gives a nearly
  silent wave) and stored in checkpoint form (``weight_g``/``weight_v``).
"""
import math

import numpy as np
import torch

from . import configs as C


def sinusoid_table(n_position, d_hid):
    """float64 sinusoid table cast to float32 - transformer/Models.py:10-30, vectorised."""
    pos = np.arange(n_position, dtype=np.float64)[:, None]
    j = np.arange(d_hid)
    denom = np.power(10000, 2 * (j // 2) / d_hid)[None, :]
    table = pos / denom
    table[:, 0::2] = np.sin(table[:, 0::2])
    table[:, 1::2] = np.cos(table[:, 1::2])
    return torch.from_numpy(table.astype(np.float32))


class _Rng:
    def __init__(self, seed):
        self.g = torch.Generator(device="cpu")
        self.g.manual_seed(int(seed))

    def uniform(self, shape, bound):
        return (torch.rand(shape, generator=self.g, dtype=torch.float32) * 2.0 - 1.0) * bound

    def normal(self, shape, std=1.0, mean=0.0):
        return torch.randn(shape, generator=self.g, dtype=torch.float32) * std + mean

    def randint(self, lo, hi, shape):
        return torch.randint(lo, hi, shape, generator=self.g, dtype=torch.int64)


def _dense(r, sd, name, out_f, in_f, k=None, bias=True):
    fan_in = in_f * (k or 1)
    bound = 1.0 / math.sqrt(fan_in)
    shape = (out_f, in_f) if k is None else (out_f, in_f, k)
    sd[name + ".weight"] = r.uniform(shape, bound)
    if bias:
        sd[name + ".bias"] = r.uniform((out_f,), bound)


def _layer_norm(r, sd, name, n):
    sd[name + ".weight"] = 1.0 + r.uniform((n,), 0.2)
    sd[name + ".bias"] = r.uniform((n,), 0.1)


def _fft_block(r, sd, p, d_model, d_inner, kernel, film):
    for w in ("w_qs", "w_ks", "w_vs"):
        _dense(r, sd, f"{p}.slf_attn.{w}", d_model, d_model)
    _layer_norm(r, sd, f"{p}.slf_attn.layer_norm", d_model)
    _dense(r, sd, f"{p}.slf_attn.fc", d_model, d_model)
    _dense(r, sd, f"{p}.pos_ffn.w_1", d_inner, d_model, kernel[0])
    _dense(r, sd, f"{p}.pos_ffn.w_2", d_model, d_inner, kernel[1])
    _layer_norm(r, sd, f"{p}.pos_ffn.layer_norm", d_model)
    if film:
        sd[f"{p}.film.s_gamma"] = 1.0 + r.uniform((1,), 0.3)
        sd[f"{p}.film.s_beta"] = 1.0 + r.uniform((1,), 0.3)


def _variance_predictor(r, sd, p, d, filt, k):
    _dense(r, sd, f"{p}.conv_layer.conv1d_1.conv", filt, d, k)
    _layer_norm(r, sd, f"{p}.conv_layer.layer_norm_1", filt)
    _dense(r, sd, f"{p}.conv_layer.conv1d_2.conv", filt, filt, k)
    _layer_norm(r, sd, f"{p}.conv_layer.layer_norm_2", filt)
    sd[f"{p}.film.s_gamma"] = 1.0 + r.uniform((1,), 0.3)
    sd[f"{p}.film.s_beta"] = 1.0 + r.uniform((1,), 0.3)
    _dense(r, sd, f"{p}.linear_layer", 1, filt)


def fastspeech2_state_dict(model_config, dataset="BC2013", seed=1234, duration_bias=2.1, n_mel=80):
    """Full FastSpeech2 + ReferenceEncoder state_dict with the reference's keys and shapes."""
    r = _Rng(seed)
    sd = {}
    tr = model_config["transformer"]
    re = model_config["reference_encoder"]
    vp = model_config["variance_predictor"]
    n_pos = model_config["max_seq_len"] + 1
    d = tr["encoder_hidden"]

    emb = r.normal((C.N_SRC_VOCAB, d))
    emb[0].zero_()  # padding_idx = PAD = 0 (transformer/Models.py:56-58)
    sd["encoder.src_word_emb.weight"] = emb
    sd["encoder.position_enc"] = sinusoid_table(n_pos, d).unsqueeze(0)
    for i in range(tr["encoder_layer"]):
        _fft_block(r, sd, f"encoder.layer_stack.{i}", d, tr["conv_filter_size"], tr["conv_kernel_size"], True)

    filt = re["conv_filter_size"]
    sd["reference_encoder.position_enc"] = sinusoid_table(n_pos, filt).unsqueeze(0)
    for i in range(re["conv_layer"]):
        _dense(r, sd, f"reference_encoder.layer_stack.{i}.0.conv", filt, n_mel if i == 0 else filt,
               re["conv_kernel_size"])
        _layer_norm(r, sd, f"reference_encoder.layer_stack.{i}.2", filt)
    dm = re["encoder_hidden"]
    _dense(r, sd, "reference_encoder.fftb_linear.linear", dm, filt, bias=False)
    for i in range(re["encoder_layer"]):
        _fft_block(r, sd, f"reference_encoder.fftb_stack.{i}", dm, filt,
                   [re["conv_kernel_size"], re["conv_kernel_size"]], False)
    _dense(r, sd, "reference_encoder.feature_wise_affine.linear", 2 * dm, dm, bias=False)
    # keep FiLM modulation O(0.3) so (gamma+1) stays well away from 0
    sd["reference_encoder.feature_wise_affine.linear.weight"] *= 2.0

    stats = C.STATS[dataset]
    nb = model_config["variance_embedding"]["n_bins"]
    ve = model_config["variance_embedding"]

    def edges(lo, hi, quant):  # model/modules.py:51-75
        if quant == "log":
            return torch.exp(torch.linspace(np.log(lo), np.log(hi), nb - 1))
        return torch.linspace(lo, hi, nb - 1)

    sd["variance_adaptor.pitch_bins"] = edges(stats["pitch"][0], stats["pitch"][1], ve["pitch_quantization"])
    sd["variance_adaptor.energy_bins"] = edges(stats["energy"][0], stats["energy"][1], ve["energy_quantization"])
    for name in ("duration_predictor", "pitch_predictor", "energy_predictor"):
        _variance_predictor(r, sd, f"variance_adaptor.{name}", d, vp["filter_size"], vp["kernel_size"])
    sd["variance_adaptor.duration_predictor.linear_layer.bias"] = torch.full((1,), float(duration_bias))
    # spread pitch / energy predictions over many buckets instead of two or three
    sd["variance_adaptor.pitch_predictor.linear_layer.weight"] *= 8.0
    sd["variance_adaptor.energy_predictor.linear_layer.weight"] *= 8.0
    if dataset == "LJSpeech_paper":  # un-normalised features: move the predictions into the range of the bins
        sd["variance_adaptor.pitch_predictor.linear_layer.weight"] *= 40.0
        sd["variance_adaptor.pitch_predictor.linear_layer.bias"] = torch.full((1,), 260.0)
        sd["variance_adaptor.energy_predictor.linear_layer.weight"] *= 20.0
        sd["variance_adaptor.energy_predictor.linear_layer.bias"] = torch.full((1,), 120.0)
    sd["variance_adaptor.pitch_embedding.weight"] = r.normal((nb, d))
    sd["variance_adaptor.energy_embedding.weight"] = r.normal((nb, d))

    dd = tr["decoder_hidden"]
    sd["decoder.position_enc"] = sinusoid_table(n_pos, dd).unsqueeze(0)
    for i in range(tr["decoder_layer"]):
        _fft_block(r, sd, f"decoder.layer_stack.{i}", dd, tr["conv_filter_size"], tr["conv_kernel_size"], True)

    _dense(r, sd, "mel_linear", n_mel, dd)
    chans = [n_mel, 512, 512, 512, 512, n_mel]
    for i in range(5):
        p = f"postnet.convolutions.{i}"
        _dense(r, sd, f"{p}.0.conv", chans[i + 1], chans[i], 5)
        sd[f"{p}.1.weight"] = 1.0 + r.uniform((chans[i + 1],), 0.2)
        sd[f"{p}.1.bias"] = r.uniform((chans[i + 1],), 0.1)
        sd[f"{p}.1.running_mean"] = r.uniform((chans[i + 1],), 0.1)
        sd[f"{p}.1.running_var"] = 1.0 + r.uniform((chans[i + 1],), 0.3)
        sd[f"{p}.1.num_batches_tracked"] = torch.tensor(0, dtype=torch.int64)

    if model_config["multi_speaker"]:
        sd["speaker_emb.weight"] = r.normal((C.N_SPEAKERS[dataset], d), std=0.3)
    return sd


def hifigan_state_dict(h=None, seed=4321, gain=1.0):
    """Generator state_dict in checkpoint form (weight_g, weight_v, bias) - hifigan/models.py:113-147."""
    h = h or C.hifigan_config()
    r = _Rng(seed)
    sd = {}

    def wn_conv(name, out_c, in_c, k, transposed=False, g=1.0):
        # weight_norm(dim=0): one gain per slice along dim 0.  For a Conv1d row norm g the output
        # variance is g^2 * var(x) for iid x, so g ~ 1 preserves variance.
        shape = (in_c, out_c, k) if transposed else (out_c, in_c, k)
        v = r.normal(shape)
        sd[name + ".weight_v"] = v
        sd[name + ".weight_g"] = g * (1.0 + r.uniform((shape[0], 1, 1), 0.1))
        sd[name + ".bias"] = r.uniform((out_c,), 0.05)

    c0 = h["upsample_initial_channel"]
    wn_conv("conv_pre", c0, h["num_mels"], 7, g=gain)
    ch = c0
    for i, (u, k) in enumerate(zip(h["upsample_rates"], h["upsample_kernel_sizes"])):
        cin, cout = c0 // (2 ** i), c0 // (2 ** (i + 1))
        wn_conv(f"ups.{i}", cout, cin, k, transposed=True, g=1.4 * gain * math.sqrt(u * cout / cin))
        ch = cout
        for j, (rk, rd) in enumerate(zip(h["resblock_kernel_sizes"], h["resblock_dilation_sizes"])):
            idx = i * len(h["resblock_kernel_sizes"]) + j
            for m in range(len(rd)):
                wn_conv(f"resblocks.{idx}.convs1.{m}", ch, ch, rk, g=0.9 * gain)
                wn_conv(f"resblocks.{idx}.convs2.{m}", ch, ch, rk, g=0.35 * gain)
    wn_conv("conv_post", 1, ch, 7, g=0.5 * gain)
    return sd


def fold_weight_norm(sd):
    """weight = g * v / ||v|| per output row (dim 0) - what remove_weight_norm() leaves behind
    (hifigan/models.py:105-109, 167-174).  Keys already holding a plain ``weight`` pass through."""
    out = {}
    for k, v in sd.items():
        if k.endswith(".weight_v"):
            base = k[: -len(".weight_v")]
            g = sd[base + ".weight_g"]
            norm = v.flatten(1).norm(dim=1).view(-1, *([1] * (v.dim() - 1)))
            out[base + ".weight"] = v * (g / norm)
        elif k.endswith(".weight_g"):
            continue
        else:
            out[k] = v
    return out


# ---------------------------------------------------------------------------------------------
# synthetic batches
# ---------------------------------------------------------------------------------------------
def make_batch(batch, src_len_range, ref_len_range, n_speakers=1, seed=1, n_mel=80, fixed_ref_len=None):
    """Batch tuple in the reference's layout (dataset.py:198-217 minus ids/raw_texts):
    speakers[B] i64, texts[B,L] i64 (0 = PAD beyond src_len), src_lens[B] i64, max_src_len,
    mels[B,Tr,80] f32 (zero beyond mel_len, utils/tools.py:298-316), mel_lens[B] i64, max_mel_len."""
    r = _Rng(seed)
    lo, hi = src_len_range
    src_lens = r.randint(lo, hi + 1, (batch,))
    if fixed_ref_len is not None:
        mel_lens = torch.full((batch,), int(fixed_ref_len), dtype=torch.int64)
    else:
        mel_lens = r.randint(ref_len_range[0], ref_len_range[1] + 1, (batch,))
    L = int(src_lens.max())
    Tr = int(mel_lens.max())
    texts = r.randint(1, C.N_SRC_VOCAB, (batch, L))
    texts = texts * (torch.arange(L)[None, :] < src_lens[:, None])
    mels = r.normal((batch, Tr, n_mel))
    mels = mels * (torch.arange(Tr)[None, :, None] < mel_lens[:, None, None])
    speakers = r.randint(0, n_speakers, (batch,))
    return {
        "speakers": speakers,
        "texts": texts,
        "src_lens": src_lens,
        "max_src_len": L,
        "mels": mels,
        "mel_lens": mel_lens,
        "max_mel_len": Tr,
    }


WORKLOADS = {
    # name: (dataset, batch, src_len_range, ref_len_range, fixed_ref_len, vocoder)
    "tiny": ("BC2013", 3, (5, 12), (9, 24), None, True),
    "tiny_ms": ("LibriTTS", 2, (4, 9), (8, 16), None, True),
    "tiny_fl": ("LJSpeech_paper", 3, (5, 12), (9, 24), None, False),  # frame-level pitch / energy, log pitch bins
    "cfg1": ("LJSpeech", 16, (80, 120), (200, 400), None, False),
    "cfg2": ("BC2013", 64, (80, 120), (800, 800), 800, False),
    "cfg3": ("LibriTTS", 32, (80, 120), (200, 400), None, True),
}


def make_workload(name, seed=1):
    ds, b, sl, rl, fixed, voc = WORKLOADS[name]
    batch = make_batch(b, sl, rl, n_speakers=C.N_SPEAKERS[ds], seed=seed, fixed_ref_len=fixed)
    return ds, batch, voc


def make_teacher_forced_batch(batch=3, seed=7, n_mel=80):
    """Batch + (p_targets, e_targets, d_targets) as the training / fine-control callers pass them
    (train.py:86, notebooks/control.ipynb): integer durations whose sum is the reference mel's length,
    pitch / energy targets in the normalised range of the bins."""
    r = _Rng(seed)
    src_lens = r.randint(5, 13, (batch,))
    L = int(src_lens.max())
    valid = torch.arange(L)[None, :] < src_lens[:, None]
    d_targets = r.randint(0, 10, (batch, L)) * valid
    d_targets[:, 0] += 1  # every utterance at least one frame
    mel_lens = d_targets.sum(dim=1)
    Tr = int(mel_lens.max())
    texts = r.randint(1, C.N_SRC_VOCAB, (batch, L)) * valid
    mels = r.normal((batch, Tr, n_mel)) * (torch.arange(Tr)[None, :, None] < mel_lens[:, None, None])
    p_targets = r.normal((batch, L), std=2.0) * valid
    e_targets = r.normal((batch, L), std=2.0) * valid
    b = {"speakers": torch.zeros(batch, dtype=torch.int64), "texts": texts, "src_lens": src_lens, "max_src_len": L,
         "mels": mels, "mel_lens": mel_lens, "max_mel_len": Tr}
    return b, {"p_targets": p_targets, "e_targets": e_targets, "d_targets": d_targets}


# ---------------------------------------------------------------------------------------------
# synthetic corpora for the sharded sweeps (BASELINE configs[3] and [4])
# ---------------------------------------------------------------------------------------------
CORPORA = {
    # name: (dataset, utterances, src_len_range or None (reference encoder only), ref_len_range, max_batch,
    #        bucket_width, vocoder)
    "cfg4": ("BC2013", 10000, None, (100, 1000), 64, 64, False),
    "cfg5": ("BC2013", 2048, (40, 110), (200, 1000), 32, 16, True),
}


class Corpus:
    """Lengths of every utterance (cheap, drawn at once) and lazily drawn payloads: utterance ``i`` has its own
    generator seeded with (seed, i), so any process can materialise any subset - a rank only ever builds the
    utterances of its own shards - and gets the same tensors as any other process."""

    def __init__(self, name, seed=1, n=None):
        ds, n_all, sl, rl, max_batch, width, voc = CORPORA[name]
        self.name, self.dataset, self.vocoder, self.text = name, ds, voc, sl is not None
        self.max_batch, self.bucket_width, self.seed = max_batch, width, int(seed)
        self.n = n_all if n is None else min(int(n), n_all)
        r = _Rng(seed)
        self.src_lens = (r.randint(sl[0], sl[1] + 1, (n_all,)) if sl is not None else torch.ones(n_all, dtype=torch.int64))
        self.ref_lens = r.randint(rl[0], rl[1] + 1, (n_all,))
        self.speakers = r.randint(0, C.N_SPEAKERS[ds], (n_all,))
        self.src_lens, self.ref_lens, self.speakers = self.src_lens[: self.n], self.ref_lens[: self.n], self.speakers[: self.n]

    def plan(self, n_ranks):
        from . import shard as SH

        return SH.plan_shards(self.src_lens.tolist() if self.text else None, self.ref_lens.tolist(),
                              max_batch=self.max_batch, bucket_width=self.bucket_width, n_ranks=n_ranks,
                              vocoder=self.vocoder, text=self.text)

    def utterance(self, i, n_mel=80):
        """(text [src_len] i64, mel [ref_len, 80] f32) of utterance i."""
        r = _Rng(self.seed * 1000003 + 7919 * (int(i) + 1))
        text = r.randint(1, C.N_SRC_VOCAB, (int(self.src_lens[i]),)) if self.text else torch.ones(1, dtype=torch.int64)
        return text, r.normal((int(self.ref_lens[i]), n_mel))

    def shard_batch(self, shard):
        """One shard collated in the reference's batch layout (dataset.py:198-217)."""
        from . import shard as SH

        texts, mels = {}, {}
        for i in shard.utterances:
            texts[i], mels[i] = self.utterance(i)
        return SH.make_shard_batch(shard, self.speakers, texts, mels)
