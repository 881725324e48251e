"""Host-side mirror of model/modules.py (VarianceAdaptor, LengthRegulator, VariancePredictor,
ReferenceEncoder) and model/blocks.py (FiLM, LinearNorm) over the C ABI.  Inference only.

Names, constructor arguments, forward signatures, return arity and state_dict keys follow the
reference so that its callers and checkpoints work unchanged; every forward enqueues
speakingstyle_b200.ops calls (hand-written sm_100a kernels) and nothing else.
"""
import json
import os
from collections import OrderedDict

import numpy as np
import torch
import torch.nn as nn

from . import ops
from .synth import sinusoid_table
from .transformer import ConvNorm, Dense, FFTBlock, FiLM, Norm, ParamCache, _PositionTable, film_args, \
    lens_from_mask, pack_conv


def _as_lens(lengths, device):
    return torch.as_tensor(lengths).to(device=device, dtype=torch.int32).contiguous()


def get_mask_from_lengths(lengths, max_len=None):
    """utils/tools.py:110-118 - mask[b,t] = t >= lengths[b] (True = padding).  ``max_len=None``
    costs one host read of max(lengths), as in the reference."""
    if max_len is None:
        max_len = int(torch.max(lengths).item())
    return ops.length_mask(_as_lens(lengths, lengths.device), int(max_len))


class LinearNorm(nn.Module):
    """model/blocks.py:65-78 parameter holder (``.linear.weight``)."""

    def __init__(self, in_features, out_features, bias=False):
        super().__init__()
        self.linear = Dense(in_features, out_features, bias=bias)


class Conv(nn.Module):
    """model/modules.py:262-305 parameter holder (``.conv.weight`` / ``.conv.bias``)."""

    def __init__(self, in_channels, out_channels, kernel_size=1, stride=1, padding=0, dilation=1, bias=True,
                 w_init="linear"):
        super().__init__()
        assert stride == 1 and dilation == 1
        self.kernel_size, self.padding = kernel_size, padding
        self.conv = Dense(in_channels, out_channels, kernel_size, bias=bias)


class VariancePredictor(nn.Module, ParamCache):
    """model/modules.py:204-259: conv-ReLU-LN, conv-ReLU-LN, [FiLM], Linear(->1), mask."""

    def __init__(self, model_config):
        super().__init__()
        self.input_size = model_config["transformer"]["encoder_hidden"]
        self.filter_size = model_config["variance_predictor"]["filter_size"]
        self.kernel = model_config["variance_predictor"]["kernel_size"]
        self.conv_output_size = self.filter_size
        self.conv_layer = nn.Sequential(OrderedDict([
            ("conv1d_1", Conv(self.input_size, self.filter_size, kernel_size=self.kernel,
                              padding=(self.kernel - 1) // 2)),
            ("layer_norm_1", Norm(self.filter_size)),
            ("conv1d_2", Conv(self.filter_size, self.filter_size, kernel_size=self.kernel, padding=1)),
            ("layer_norm_2", Norm(self.filter_size)),
        ]))
        self.film = FiLM()
        self.linear_layer = Dense(self.conv_output_size, 1)

    def _pack(self):
        cl = self.conv_layer
        return {"w1": pack_conv(cl.conv1d_1.conv.weight), "w2": pack_conv(cl.conv1d_2.conv.weight),
                "lin_w": self.linear_layer.weight.reshape(-1).contiguous()}

    def forward(self, encoder_output, mask, gammas=None, betas=None, lens=None):
        pk = self.packed()
        cl = self.conv_layer
        if lens is None and mask is not None:
            lens = lens_from_mask(mask)
        k = self.kernel
        h = ops.conv1d(encoder_output, pk["w1"], cl.conv1d_1.conv.bias, k=k, pad=cl.conv1d_1.padding,
                       act=ops.ACT_RELU)
        # the reference has no mask between the convs; batch-invariant mode zeroes the pads so that the second
        # convolution's halo at an utterance end reads zeros whatever the batch was padded to
        h = ops.layernorm(h, cl.layer_norm_1.weight, cl.layer_norm_1.bias, out=h,
                          lens=lens if getattr(self, "batch_invariant", False) else None)
        h = ops.conv1d(h, pk["w2"], cl.conv1d_2.conv.bias, k=k, pad=cl.conv1d_2.padding, act=ops.ACT_RELU)
        return ops.predictor_head(h, cl.layer_norm_2.weight, cl.layer_norm_2.bias, pk["lin_w"],
                                  self.linear_layer.bias, film=film_args(self.film, gammas, betas), lens=lens)


class LengthRegulator(nn.Module):
    """model/modules.py:168-201.  forward(x, duration, max_len) -> (expanded [B,T,C], mel_len [B] i64).
    ``duration`` may be float (predicted, truncated per token like int()) or integer targets."""

    def forward(self, x, duration, max_len=None, pe=None, return_lens=False):
        dur = duration.to(torch.float32).contiguous()
        _, ends, lens2 = ops.duration_scan(None, 1.0, durations=dur)
        lens_host = lens2.cpu()  # the one host read: the output shape depends on the data
        T = int(max_len) if max_len is not None else int(lens_host[0].max())
        out = ops.length_regulate(x.contiguous(), ends, T, pe=pe)
        mel_len = lens2[0].to(torch.int64)
        if return_lens:
            return out, mel_len, lens2, lens_host
        return out, mel_len

    LR = forward


class VarianceAdaptor(nn.Module):
    """model/modules.py:20-165.  Pitch / energy are predicted per phoneme before the LengthRegulator (``phoneme_level``:
    config/{BC2013,LJSpeech,LibriTTS,AISHELL3}/preprocess.yaml) or per frame after it (``frame_level``:
    config/LJSpeech_paper/preprocess.yaml:26,29; modules.py:146-155)."""

    def __init__(self, preprocess_config, model_config):
        super().__init__()
        self.duration_predictor = VariancePredictor(model_config)
        self.length_regulator = LengthRegulator()
        self.pitch_predictor = VariancePredictor(model_config)
        self.energy_predictor = VariancePredictor(model_config)
        self.pitch_feature_level = preprocess_config["preprocessing"]["pitch"]["feature"]
        self.energy_feature_level = preprocess_config["preprocessing"]["energy"]["feature"]
        for lv in (self.pitch_feature_level, self.energy_feature_level):
            if lv not in ("phoneme_level", "frame_level"):
                raise ValueError(f"pitch / energy feature must be phoneme_level or frame_level, got {lv!r}")
        ve = model_config["variance_embedding"]
        n_bins = ve["n_bins"]
        with open(os.path.join(preprocess_config["path"]["preprocessed_path"], "stats.json")) as f:
            stats = json.load(f)
            pitch_min, pitch_max = stats["pitch"][:2]
            energy_min, energy_max = stats["energy"][:2]

        def edges(lo, hi, quant):
            if quant == "log":
                return torch.exp(torch.linspace(np.log(lo), np.log(hi), n_bins - 1))
            return torch.linspace(lo, hi, n_bins - 1)

        self.pitch_bins = nn.Parameter(edges(pitch_min, pitch_max, ve["pitch_quantization"]), requires_grad=False)
        self.energy_bins = nn.Parameter(edges(energy_min, energy_max, ve["energy_quantization"]),
                                        requires_grad=False)
        d = model_config["transformer"]["encoder_hidden"]
        self.pitch_embedding = nn.Embedding(n_bins, d)
        self.energy_embedding = nn.Embedding(n_bins, d)
        for p in self.parameters():
            p.requires_grad_(False)

    def forward(self, x, src_mask, mel_mask=None, max_len=None, pitch_target=None, energy_target=None,
                duration_target=None, p_control=1.0, e_control=1.0, d_control=1.0, gammas=None, betas=None,
                src_lens=None, decoder_pe=None, return_extra=False, inplace=False):
        lens = src_lens if src_lens is not None else lens_from_mask(src_mask)
        log_d = self.duration_predictor(x, src_mask, gammas, betas, lens=lens)
        if not inplace:
            x = x.clone()  # the reference leaves its input untouched; the embedding adds below are in place
        # With a target the embedding comes from the target and the prediction is returned unscaled
        # (modules.py:85-86, 96-97); without, from control * prediction (88-91, 99-102).
        binv = getattr(self, "batch_invariant", False)

        def embed(which, x, lens_, target, control):
            """get_pitch_embedding / get_energy_embedding + the add (modules.py:83-103, 123-131, 146-155), in place."""
            pred = getattr(self, which + "_predictor")(x, None, lens=lens_)
            bins, table = getattr(self, which + "_bins"), getattr(self, which + "_embedding").weight
            if target is not None:
                idx = ops.bucket_embed_add(target.to(torch.float32).contiguous().clone(), 1.0, bins, table, x,
                                           want_idx=return_extra)
            else:
                idx = ops.bucket_embed_add(pred, control, bins, table, x, want_idx=return_extra)
            if binv:
                ops.mask_rows_(x, lens_)  # the embedding was added at the pads too (modules.py:126,131)
            return pred, idx

        pitch = energy = pitch_idx = energy_idx = None
        if self.pitch_feature_level == "phoneme_level":
            pitch, pitch_idx = embed("pitch", x, lens, pitch_target, p_control)
        if self.energy_feature_level == "phoneme_level":
            energy, energy_idx = embed("energy", x, lens, energy_target, e_control)
        frame_level = "frame_level" in (self.pitch_feature_level, self.energy_feature_level)
        if frame_level:
            # the frame-level predictors read the expanded sequence WITHOUT the decoder's position table, so the
            # LengthRegulator kernel must not fuse that add here; it is applied after the embedding adds below
            lr_pe, late_pe = None, decoder_pe
        else:
            lr_pe, late_pe = decoder_pe, None
        if duration_target is not None:
            # teacher forcing (modules.py:133-135): expand by the given durations to max_len; mel_mask is the caller's
            d_rounded = duration_target
            self.last_mel_len_host = None
            T = int(max_len) if max_len is not None else None
            pe = lr_pe(T) if (callable(lr_pe) and T is not None) else (None if callable(lr_pe) else lr_pe)
            out, mel_len = self.length_regulator(x, duration_target, max_len, pe=pe)
            frame_lens = lens_from_mask(mel_mask) if mel_mask is not None else mel_len.to(torch.int32)
            extra = {"mel2phone": None, "T": out.shape[1], "mel_len_expand": mel_len, "pe_added": pe is not None}
        else:
            d_rounded, ends, lens2 = ops.duration_scan(log_d, d_control)
            lens_host = lens2.cpu()  # one host read (the reference does B*L of them, modules.py:193)
            T = int(lens_host[0].max())
            pe = lr_pe(T) if callable(lr_pe) else lr_pe
            res = ops.length_regulate(x, ends, T, pe=pe, want_mel2phone=return_extra)
            out, m2p = res if return_extra else (res, None)
            mel_len = lens2[1]  # int32, from the float sum (modules.py:143)
            mel_mask = ops.length_mask(mel_len, int(lens_host[1].max()))
            self.last_mel_len_host = lens_host[1]  # already on the host: lets the wave epilogue trim without a second sync
            frame_lens = mel_len
            extra = {"mel2phone": m2p, "T": T, "mel_len_expand": lens2[0], "pe_added": pe is not None}
        if frame_level:
            T = out.shape[1]
            if mel_mask is not None and mel_mask.shape[1] != T:
                # the reference masks the [B, T] predictions with this mask and fails on the shape (modules.py:146-155
                # with d_control != 1, where int(sum d) != sum int(d)); say why instead
                raise RuntimeError(f"frame-level pitch / energy: mel_mask has {mel_mask.shape[1]} frames, the expanded "
                                   f"sequence {T} (non-integer durations); the reference fails here too")
            if self.pitch_feature_level == "frame_level":
                pitch, pitch_idx = embed("pitch", out, frame_lens, pitch_target, p_control)
            if self.energy_feature_level == "frame_level":
                energy, energy_idx = embed("energy", out, frame_lens, energy_target, e_control)
            pe = late_pe(T) if callable(late_pe) else late_pe
            if pe is not None:
                out = ops.add_position(out, pe)
                extra["pe_added"] = True
        extra.update(pitch_idx=pitch_idx, energy_idx=energy_idx)
        ret = (out, pitch, energy, log_d, d_rounded, mel_len, mel_mask)
        return ret + (extra,) if return_extra else ret


class ReferenceEncoder(nn.Module, ParamCache, _PositionTable):
    """model/modules.py:307-406: 3 x (Conv1d k3 -> ReLU -> LayerNorm(1024)), mask, + position table,
    Linear 1024->256 (no bias), 4 FFT blocks (8 heads, conv k [3,3], no FiLM), mean over the padded
    time axis, Linear 256->512 (no bias), split into (gammas, betas) [B,1,256]."""

    def __init__(self, preprocess_config, model_config):
        super().__init__()
        rc = model_config["reference_encoder"]
        self.max_seq_len = model_config["max_seq_len"] + 1
        n_conv_layers = rc["conv_layer"]
        kernel_size = rc["conv_kernel_size"]
        n_head = rc["encoder_head"]
        self.d_model = rc["encoder_hidden"]
        n_mel = preprocess_config["preprocessing"]["mel"]["n_mel_channels"]
        d_k = self.d_model // n_head
        self.filter_size = rc["conv_filter_size"]
        self.kernel_size = kernel_size
        self.layer_stack = nn.ModuleList([
            nn.Sequential(OrderedDict([
                ("0", ConvNorm(n_mel if i == 0 else self.filter_size, self.filter_size, kernel_size=kernel_size,
                               stride=1, padding=(kernel_size - 1) // 2, dilation=1, transform=True)),
                ("2", Norm(self.filter_size)),
            ])) for i in range(n_conv_layers)])
        self.position_enc = nn.Parameter(sinusoid_table(self.max_seq_len, self.filter_size).unsqueeze(0),
                                         requires_grad=False)
        self.fftb_linear = LinearNorm(self.filter_size, self.d_model)
        self.fftb_stack = nn.ModuleList([
            FFTBlock(self.d_model, n_head, d_k, d_k, self.filter_size, [kernel_size, kernel_size],
                     dropout=rc["dropout"], film=False) for _ in range(rc["encoder_layer"])])
        self.feature_wise_affine = LinearNorm(self.d_model, 2 * self.d_model)

    def _pack(self):
        return {
            "convs": [pack_conv(seq[0].conv.weight) for seq in self.layer_stack],
            "lin": pack_conv(self.fftb_linear.linear.weight),
            "affine": pack_conv(self.feature_wise_affine.linear.weight),
        }

    def forward(self, mel, max_len, mask, lens=None, return_pooled=False):
        pk = self.packed()
        if lens is None:
            lens = lens_from_mask(mask)
        B, Tr, _ = mel.shape
        k = self.kernel_size
        x = mel.contiguous()
        n = len(self.layer_stack)
        for i, seq in enumerate(self.layer_stack):
            x = ops.conv1d(x, pk["convs"][i], seq[0].conv.bias, k=k, pad=(k - 1) // 2, act=ops.ACT_RELU)
            last = i == n - 1
            # no mask between the conv layers (modules.py:366-371); mask + position add after the last.
            # Batch-invariant mode masks after every layer (the padded frames' bias would otherwise leak into
            # the last valid frames through the next layer's halo).
            binv = getattr(self, "batch_invariant", False)
            x = ops.layernorm(x, seq[1].weight, seq[1].bias, lens=lens if (last or binv) else None,
                              pe=self._pe(Tr, self.filter_size) if last else None, out=x)
        x = ops.conv1d(x, pk["lin"], None)
        for blk in self.fftb_stack:
            x, _ = blk(x, lens=lens)
        # mean over the PADDED length (modules.py:398); batch-invariant mode: over the utterance's own frames
        pooled = ops.mean_pool(x, lens=lens if getattr(self, "batch_invariant", False) else None)
        gb = ops.conv1d(pooled.unsqueeze(0), pk["affine"], None, backend="ffma")[0]  # [B, 2d]: one tiny GEMV
        d = self.d_model
        gammas = gb[:, :d].unsqueeze(1)
        betas = gb[:, d:].unsqueeze(1)
        if return_pooled:
            return gammas, betas, pooled
        return gammas, betas
