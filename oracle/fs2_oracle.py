"""TEST INFRASTRUCTURE ONLY - CPU restatement of the reference's text -> mel -> wave arithmetic.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this file, and only as the checker (or the timed CPU arm); the product path
(speakingstyle_b200/) never does.

What it is: a functional, state_dict-driven restatement of the forward pass of
  model/fastspeech2.py:44-120, model/modules.py:105-165 / 174-201 / 249-259 / 361-406,
  model/blocks.py:54-62, transformer/Models.py:73-101 / 140-170, transformer/Layers.py:23-37 /
  140-148, transformer/SubLayers.py:29-57 / 85-93, transformer/Modules.py:14-23,
  hifigan/models.py:96-103 / 149-165 and utils/model.py:97-115,
written against torch's CPU kernels (the reference has no arithmetic of its own: every op is an
ATen CPU call, SURVEY.md section 8c "third-party arithmetic" - PyTorch, 2.11.0 in this image, 1.7.0 pinned
by the reference's environment.yaml:35).  The integer-valued steps (duration rounding, bucketize,
frame expansion indices) are additionally restated in plain C in oracle/intops.c.

Pinning: parity is PINNED against outputs of the reference itself, run in the build container:
tests/golden/*.npz were produced by oracle/make_golden.py from the real reference modules on the
seeded weights/batches of speakingstyle_b200/synth.py, and tests/test_oracle.py checks this file
against them (and, where /root/reference exists, against the live reference).  The reference
ships no tests, golden vectors or checkpoints of its own.

Everything is computed in ``dtype`` (float32 like the reference, or float64 to measure rounding
margins).  Tensors are channels-last [B, S, C] between ops, as in the reference.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F


# --------------------------------------------------------------------------------------------
# small pieces
# --------------------------------------------------------------------------------------------
def mask_from_lengths(lengths, max_len=None):
    """True = padding (utils/tools.py:110-118)."""
    lengths = torch.as_tensor(lengths)
    if max_len is None:
        max_len = int(lengths.max())
    return torch.arange(int(max_len))[None, :] >= lengths[:, None]


def sinusoid_table(n_position, d_hid):
    """transformer/Models.py:10-30: float64 angles, sin on even / cos on odd columns, cast to f32."""
    pos = np.arange(n_position, dtype=np.float64)[:, None]
    j = np.arange(d_hid)
    ang = pos / np.power(10000, 2 * (j // 2) / d_hid)[None, :]
    ang[:, 0::2] = np.sin(ang[:, 0::2])
    ang[:, 1::2] = np.cos(ang[:, 1::2])
    return torch.from_numpy(ang.astype(np.float32))


def _conv_nlc(x, w, b, padding, dilation=1):
    """Conv1d applied to a channels-last tensor (the reference transposes around every conv)."""
    return F.conv1d(x.transpose(1, 2), w, b, padding=padding, dilation=dilation).transpose(1, 2)


def _ln(x, sd, p):
    return F.layer_norm(x, (x.shape[-1],), sd[p + ".weight"], sd[p + ".bias"], 1e-5)


def film(x, gammas, betas, s_gamma, s_beta):
    """model/blocks.py:54-62."""
    return (s_gamma * gammas + 1.0) * x + s_beta * betas


class _SD(dict):
    """state_dict view cast to the working dtype."""

    def __init__(self, sd, dtype):
        super().__init__({k: (v.to(dtype) if v.is_floating_point() else v) for k, v in sd.items()})


# --------------------------------------------------------------------------------------------
# FFT block
# --------------------------------------------------------------------------------------------
def multi_head_attention(sd, p, x, pad_mask, n_head):
    """transformer/SubLayers.py:29-57 + transformer/Modules.py:14-23 (key-padding mask only)."""
    B, S, D = x.shape
    dk = D // n_head
    q = F.linear(x, sd[p + ".w_qs.weight"], sd[p + ".w_qs.bias"]).view(B, S, n_head, dk).transpose(1, 2)
    k = F.linear(x, sd[p + ".w_ks.weight"], sd[p + ".w_ks.bias"]).view(B, S, n_head, dk).transpose(1, 2)
    v = F.linear(x, sd[p + ".w_vs.weight"], sd[p + ".w_vs.bias"]).view(B, S, n_head, dk).transpose(1, 2)
    score = torch.matmul(q, k.transpose(-1, -2)) / np.power(dk, 0.5)
    score = score.masked_fill(pad_mask[:, None, None, :], -math.inf)
    ctx = torch.matmul(torch.softmax(score, dim=-1), v)  # [B,H,S,dk]
    ctx = ctx.transpose(1, 2).reshape(B, S, D)
    out = F.linear(ctx, sd[p + ".fc.weight"], sd[p + ".fc.bias"])
    return _ln(out + x, sd, p + ".layer_norm")


def pos_ffn(sd, p, x):
    """transformer/SubLayers.py:85-93."""
    w1, w2 = sd[p + ".w_1.weight"], sd[p + ".w_2.weight"]
    h = F.relu(_conv_nlc(x, w1, sd[p + ".w_1.bias"], (w1.shape[2] - 1) // 2))
    h = _conv_nlc(h, w2, sd[p + ".w_2.bias"], (w2.shape[2] - 1) // 2)
    return _ln(h + x, sd, p + ".layer_norm")


def fft_block(sd, p, x, pad_mask, n_head, gammas=None, betas=None):
    """transformer/Layers.py:23-37."""
    keep = (~pad_mask)[..., None].to(x.dtype)
    x = multi_head_attention(sd, p + ".slf_attn", x, pad_mask, n_head) * keep
    x = pos_ffn(sd, p + ".pos_ffn", x)
    if gammas is not None:
        x = film(x, gammas, betas, sd[p + ".film.s_gamma"], sd[p + ".film.s_beta"])
    return x * keep


def _pos_enc(sd, key, n, width):
    table = sd[key][0]
    if n > table.shape[0]:  # eval-time regeneration (Models.py:82-87, 145-152; modules.py:374-380)
        table = sinusoid_table(n, width).to(table.dtype)
    return table[:n]


# --------------------------------------------------------------------------------------------
# modules
# --------------------------------------------------------------------------------------------
def reference_encoder(sd, mc, mels, mel_mask, return_pooled=False):
    """model/modules.py:361-406 -> (gammas, betas) each [B,1,d]."""
    rc = mc["reference_encoder"]
    p = "reference_encoder"
    x = mels
    k = rc["conv_kernel_size"]
    for i in range(rc["conv_layer"]):
        x = _conv_nlc(x, sd[f"{p}.layer_stack.{i}.0.conv.weight"], sd[f"{p}.layer_stack.{i}.0.conv.bias"],
                      (k - 1) // 2)
        x = _ln(F.relu(x), sd, f"{p}.layer_stack.{i}.2")
    x = x.masked_fill(mel_mask[..., None], 0.0)
    Tr = x.shape[1]
    x = x + _pos_enc(sd, p + ".position_enc", Tr, rc["conv_filter_size"])[None]
    x = F.linear(x, sd[p + ".fftb_linear.linear.weight"])
    for i in range(rc["encoder_layer"]):
        x = fft_block(sd, f"{p}.fftb_stack.{i}", x, mel_mask, rc["encoder_head"])
    pooled = x.mean(dim=1, keepdim=True)  # over the PADDED length (modules.py:398)
    gb = F.linear(pooled, sd[p + ".feature_wise_affine.linear.weight"])
    d = rc["encoder_hidden"]
    if return_pooled:
        return gb[..., :d], gb[..., d:], pooled
    return gb[..., :d], gb[..., d:]


def encoder(sd, mc, texts, src_mask, gammas, betas):
    """transformer/Models.py:73-101."""
    tr = mc["transformer"]
    L = texts.shape[1]
    x = sd["encoder.src_word_emb.weight"][texts] + _pos_enc(sd, "encoder.position_enc", L, tr["encoder_hidden"])[None]
    for i in range(tr["encoder_layer"]):
        x = fft_block(sd, f"encoder.layer_stack.{i}", x, src_mask, tr["encoder_head"], gammas, betas)
    return x


def variance_predictor(sd, p, x, mask, gammas=None, betas=None):
    """model/modules.py:249-259 (conv k=3 pad 1 twice, ReLU, LayerNorm; FiLM only if given)."""
    w1 = sd[p + ".conv_layer.conv1d_1.conv.weight"]
    h = _conv_nlc(x, w1, sd[p + ".conv_layer.conv1d_1.conv.bias"], (w1.shape[2] - 1) // 2)
    h = _ln(F.relu(h), sd, p + ".conv_layer.layer_norm_1")
    h = _conv_nlc(h, sd[p + ".conv_layer.conv1d_2.conv.weight"], sd[p + ".conv_layer.conv1d_2.conv.bias"], 1)
    h = _ln(F.relu(h), sd, p + ".conv_layer.layer_norm_2")
    if gammas is not None:
        h = film(h, gammas, betas, sd[p + ".film.s_gamma"], sd[p + ".film.s_beta"])
    out = F.linear(h, sd[p + ".linear_layer.weight"], sd[p + ".linear_layer.bias"]).squeeze(-1)
    return out.masked_fill(mask, 0.0)


def bucketize(values, edges):
    """torch.bucketize(right=False): index = #(edges < v); NaN lands in the last bucket."""
    return torch.bucketize(values.contiguous(), edges.contiguous())


def round_durations(log_d, d_control=1.0):
    """model/modules.py:137-140: clamp(round(exp(logd) - 1) * control, min=0); half-to-even."""
    return torch.clamp(torch.round(torch.exp(log_d) - 1) * d_control, min=0)


def length_regulate(x, durations, max_len=None):
    """model/modules.py:174-201 + utils/tools.py:319-337 without the Python double loop.

    Row i of utterance b is repeated max(int(d[b,i]), 0) times (int() truncates toward zero);
    rows are zero beyond the expanded length; T = max_b sum (or max_len).  Returns
    (out [B,T,C], mel_len [B] i64, mel2phone [B,T] i64 with -1 on padding)."""
    B, L, Cc = x.shape
    reps = torch.clamp(durations.to(torch.float64).trunc().to(torch.int64), min=0)
    ends = torch.cumsum(reps, dim=1)
    mel_len = ends[:, -1] if L > 0 else torch.zeros(B, dtype=torch.int64)
    T = int(max_len) if max_len is not None else (int(mel_len.max()) if B > 0 else 0)
    t = torch.arange(T)[None, :].expand(B, T)
    idx = torch.searchsorted(ends, t.contiguous(), right=True)  # #{j : ends[j] <= t}
    valid = t < mel_len[:, None]
    idx = torch.where(valid, idx, torch.zeros_like(idx))
    out = torch.gather(x, 1, idx[..., None].expand(B, T, Cc)) * valid[..., None].to(x.dtype)
    mel2phone = torch.where(valid, idx, torch.full_like(idx, -1))
    return out, mel_len, mel2phone


def variance_adaptor(sd, mc, x, src_mask, gammas, betas, p_control=1.0, e_control=1.0, d_control=1.0,
                     d_targets=None, max_len=None, p_targets=None, e_targets=None, mel_mask=None,
                     levels=("phoneme_level", "phoneme_level")):
    """model/modules.py:105-165.  ``levels`` = (pitch, energy) feature level of the preprocess config: phoneme_level
    predictors run on the token sequence before the LengthRegulator (modules.py:121-131), frame_level ones on the
    expanded sequence after it, masked with mel_mask (modules.py:146-155)."""
    va = "variance_adaptor"
    log_d = variance_predictor(sd, va + ".duration_predictor", x, src_mask, gammas, betas)
    res = {}

    def embed(which, x, mask, target, control):
        pred = variance_predictor(sd, f"{va}.{which}_predictor", x, mask)
        if target is not None:  # modules.py:85-86: embedding from the target, prediction returned unscaled
            idx = bucketize(target.to(pred.dtype), sd[f"{va}.{which}_bins"])
        else:
            pred = pred * control
            idx = bucketize(pred, sd[f"{va}.{which}_bins"])
        res[which], res[which + "_idx"] = pred, idx
        return x + sd[f"{va}.{which}_embedding.weight"][idx]

    if levels[0] == "phoneme_level":
        x = embed("pitch", x, src_mask, p_targets, p_control)
    if levels[1] == "phoneme_level":
        x = embed("energy", x, src_mask, e_targets, e_control)
    x_pre_lr = x
    if d_targets is not None:
        d_rounded = d_targets
        out, mel_len, mel2phone = length_regulate(x, d_targets, max_len)
    else:
        d_rounded = round_durations(log_d, d_control)
        out, _, mel2phone = length_regulate(x, d_rounded, None)
        mel_len = d_rounded.sum(-1).int()  # int32, from the float sum (modules.py:143)
        mel_mask = mask_from_lengths(mel_len)
    if levels[0] == "frame_level":
        out = embed("pitch", out, mel_mask, p_targets, p_control)
    if levels[1] == "frame_level":
        out = embed("energy", out, mel_mask, e_targets, e_control)
    return {
        "x": out, "pitch": res["pitch"], "energy": res["energy"], "log_d": log_d, "d_rounded": d_rounded,
        "mel_len": mel_len, "mel2phone": mel2phone, "pitch_idx": res["pitch_idx"], "energy_idx": res["energy_idx"],
        "x_pre_lr": x_pre_lr, "mel_mask": mel_mask,
    }


def decoder(sd, mc, x, mel_mask, gammas, betas):
    """transformer/Models.py:140-170."""
    tr = mc["transformer"]
    T = x.shape[1]
    x = x + _pos_enc(sd, "decoder.position_enc", T, tr["decoder_hidden"])[None]
    mel_mask = mel_mask[:, :T]
    for i in range(tr["decoder_layer"]):
        x = fft_block(sd, f"decoder.layer_stack.{i}", x, mel_mask, tr["decoder_head"], gammas, betas)
    return x, mel_mask


def postnet(sd, mel):
    """transformer/Layers.py:140-148: 5 x (conv k5 -> BatchNorm(eval) -> tanh except last)."""
    x = mel.transpose(1, 2)
    n = 5
    for i in range(n):
        p = f"postnet.convolutions.{i}"
        x = F.conv1d(x, sd[p + ".0.conv.weight"], sd[p + ".0.conv.bias"], padding=2)
        x = F.batch_norm(x, sd[p + ".1.running_mean"], sd[p + ".1.running_var"], sd[p + ".1.weight"],
                         sd[p + ".1.bias"], False, 0.0, 1e-5)
        if i < n - 1:
            x = torch.tanh(x)
    return x.transpose(1, 2)


def fastspeech2_forward(sd, mc, batch, p_control=1.0, e_control=1.0, d_control=1.0, dtype=torch.float32,
                        intermediates=False, p_targets=None, e_targets=None, d_targets=None,
                        levels=("phoneme_level", "phoneme_level")):
    """model/fastspeech2.py:44-120 (inference: no targets).  Returns the 10-tuple, or a dict."""
    sd = _SD(sd, dtype)
    texts, src_lens = batch["texts"], batch["src_lens"]
    mels = batch["mels"].to(dtype)
    src_mask = mask_from_lengths(src_lens, batch["max_src_len"])
    ref_mask = mask_from_lengths(batch["mel_lens"], batch["max_mel_len"])
    gammas, betas = reference_encoder(sd, mc, mels, ref_mask)
    enc = enc_plain = encoder(sd, mc, texts, src_mask, gammas, betas)
    if "speaker_emb.weight" in sd and mc["multi_speaker"]:
        enc = enc + sd["speaker_emb.weight"][batch["speakers"]][:, None, :]  # pads included (fastspeech2.py:74-77)
    va = variance_adaptor(sd, mc, enc, src_mask, gammas, betas, p_control, e_control, d_control, d_targets=d_targets,
                          max_len=batch["max_mel_len"] if d_targets is not None else None, p_targets=p_targets,
                          e_targets=e_targets, mel_mask=ref_mask if d_targets is not None else None, levels=levels)
    # teacher forcing keeps the reference mel's mask (fastspeech2.py:62-66); otherwise it is rebuilt from mel_len
    mel_mask = va["mel_mask"]
    dec, mel_mask = decoder(sd, mc, va["x"], mel_mask, gammas, betas)
    mel = F.linear(dec, sd["mel_linear.weight"], sd["mel_linear.bias"])
    post = postnet(sd, mel) + mel
    out = (mel, post, va["pitch"], va["energy"], va["log_d"], va["d_rounded"], src_mask, mel_mask, src_lens,
           va["mel_len"])
    if intermediates:
        return {"out": out, "gammas": gammas, "betas": betas, "encoder": enc_plain, "decoder": dec, **va}
    return out


# --------------------------------------------------------------------------------------------
# HiFi-GAN generator
# --------------------------------------------------------------------------------------------
def hifigan_forward(sd, h, mel_bct, dtype=torch.float32, stages=False):
    """hifigan/models.py:149-165 on folded (plain ``weight``) parameters.  mel_bct is [B,80,T]."""
    sd = _SD(sd, dtype)
    x = F.conv1d(mel_bct.to(dtype), sd["conv_pre.weight"], sd["conv_pre.bias"], padding=3)
    nk = len(h["resblock_kernel_sizes"])
    trace = [x]
    for i, (u, k) in enumerate(zip(h["upsample_rates"], h["upsample_kernel_sizes"])):
        x = F.conv_transpose1d(F.leaky_relu(x, 0.1), sd[f"ups.{i}.weight"], sd[f"ups.{i}.bias"], stride=u,
                               padding=(k - u) // 2)
        acc = None
        for j, (rk, rd) in enumerate(zip(h["resblock_kernel_sizes"], h["resblock_dilation_sizes"])):
            p = f"resblocks.{i * nk + j}"
            y = x
            for m, dil in enumerate(rd):
                t = F.conv1d(F.leaky_relu(y, 0.1), sd[f"{p}.convs1.{m}.weight"], sd[f"{p}.convs1.{m}.bias"],
                             padding=(rk * dil - dil) // 2, dilation=dil)
                t = F.conv1d(F.leaky_relu(t, 0.1), sd[f"{p}.convs2.{m}.weight"], sd[f"{p}.convs2.{m}.bias"],
                             padding=(rk - 1) // 2)
                y = t + y
            acc = y if acc is None else acc + y
        x = acc / nk
        trace.append(x)
    x = F.leaky_relu(x)  # default slope 0.01 (hifigan/models.py:161)
    x = torch.tanh(F.conv1d(x, sd["conv_post.weight"], sd["conv_post.bias"], padding=3))
    return (x, trace) if stages else x


def wave_to_int16(wave_b1t, lengths=None, max_wav_value=32768.0):
    """utils/model.py:105-113: scale, numpy astype(int16) (truncation toward zero), per-utterance trim."""
    w = (wave_b1t.squeeze(1).detach().cpu().numpy() * max_wav_value).astype("int16")
    out = [row for row in w]
    if lengths is not None:
        out = [row[: int(n)] for row, n in zip(out, lengths)]
    return out
