"""Drop-in for model/fastspeech2.py: ``FastSpeech2(preprocess_config, model_config)`` with the
reference's forward signature, 10-tuple return and state_dict keys, running on hand-written
sm_100a kernels through the C ABI.  Inference (eval / no_grad) only; no CPU fallback."""
import json
import os

import torch
import torch.nn as nn

from . import ops
from .modules import ReferenceEncoder, VarianceAdaptor, _as_lens
from .transformer import Decoder, Dense, Encoder, ParamCache, PostNet, pack_conv


class FastSpeech2(nn.Module, ParamCache):
    """model/fastspeech2.py:13-120."""

    def __init__(self, preprocess_config, model_config, precision="fp32", batch_invariant=False):
        super().__init__()
        self.model_config = model_config
        # "fp32": every kernel accumulates in fp32 FMA (mel max-abs <= 1e-3 vs the reference);
        # "bf16": Decoder / mel_linear / PostNet run on tcgen05 with bf16 operands (relative L2 <= 1e-2);
        # durations, bucket ids and mel_lens are computed in fp32 in both modes and are identical.
        if precision not in ("fp32", "bf16"):
            raise ValueError("precision must be 'fp32' or 'bf16'")
        self.precision = precision
        self.encoder = Encoder(model_config)
        self.reference_encoder = ReferenceEncoder(preprocess_config, model_config)
        self.variance_adaptor = VarianceAdaptor(preprocess_config, model_config)
        self.decoder = Decoder(model_config)
        self.mel_linear = Dense(model_config["transformer"]["decoder_hidden"],
                                preprocess_config["preprocessing"]["mel"]["n_mel_channels"])
        self.postnet = PostNet()
        self.speaker_emb = None
        if model_config["multi_speaker"]:
            with open(os.path.join(preprocess_config["path"]["preprocessed_path"], "speakers.json"), "r") as f:
                n_speaker = len(json.load(f))
            self.speaker_emb = nn.Embedding(n_speaker, model_config["transformer"]["encoder_hidden"])
        for p in self.parameters():
            p.requires_grad_(False)
        self.eval()
        self.set_batch_invariant(batch_invariant)

    def set_batch_invariant(self, flag=True):
        """Opt-in mode in which an utterance's outputs do not depend on what it was batched with.
        The reference leaks padding in six places: the reference encoder's conv stack is not masked between
        layers, the second convolution of its FFT blocks ([3, 3] kernels) reads the first one's output at padded
        frames, and its style vector is a mean over the PADDED length (modules.py:366-371, 353, 398); the speaker,
        pitch and energy embeddings are added at padded tokens (fastspeech2.py:74-77, modules.py:126, 131) and the
        predictors' two convolutions are not masked in between; mel_linear's bias sits in the padded frames that
        the PostNet convolves over.  This mode zeroes the pads at each of those points and averages over the valid
        frames.  It DEVIATES from the reference by design (the parity tests run with it off); with it on, running
        an utterance alone or in any batch gives the same integers and the same mel, bit for bit."""
        self.batch_invariant = bool(flag)
        for m in self.modules():
            if m is not self:
                m.batch_invariant = bool(flag)
        return self

    def _pack(self):
        return {"mel_w": pack_conv(self.mel_linear.weight)}

    def packed(self):  # only mel_linear's own parameters key this cache
        srcs = [self.mel_linear.weight]
        key = tuple((t.data_ptr(), t._version) for t in srcs)
        if self.__dict__.get("_pack_key") != key:
            with torch.no_grad():
                self.__dict__["_pack_val"] = self._pack()
            self.__dict__["_pack_key"] = key
        return self.__dict__["_pack_val"]

    @property
    def device(self):
        return self.mel_linear.weight.device

    def train(self, mode=True):
        if mode:
            raise NotImplementedError("speakingstyle_b200.FastSpeech2 is an inference path; training "
                                      "(train.py:45-101) stays with the reference implementation")
        return super().train(False)

    @torch.no_grad()
    def forward(self, speakers, texts, src_lens, max_src_len, mels=None, mel_lens=None, max_mel_len=None,
                p_targets=None, e_targets=None, d_targets=None, p_control=1.0, e_control=1.0, d_control=1.0,
                return_extra=False):
        """model/fastspeech2.py:44-120.  p_control / e_control / d_control: numbers, or tensors with one value per
        phoneme ([B, L], the fine-control use of notebooks/control.ipynb)."""
        dev = self.device
        if dev.type != "cuda":
            raise RuntimeError("speakingstyle_b200.FastSpeech2 needs a CUDA device (.to('cuda')); no CPU fallback")
        # the library launches on cudart's CURRENT device and torch's current stream of that device: make the model's
        # own device current for the duration of the call (a model on cuda:1 with device 0 current would otherwise
        # launch on GPU 0 with GPU 1 pointers)
        with torch.cuda.device(dev):
            return self._forward(speakers, texts, src_lens, max_src_len, mels, mel_lens, max_mel_len, p_targets,
                                 e_targets, d_targets, p_control, e_control, d_control, return_extra)

    def _forward(self, speakers, texts, src_lens, max_src_len, mels, mel_lens, max_mel_len, p_targets, e_targets,
                 d_targets, p_control, e_control, d_control, return_extra):
        dev = self.device
        if mels is None or mel_lens is None:
            # the reference dereferences mel_masks / mask unconditionally (fastspeech2.py:79, modules.py:384)
            raise AttributeError("a reference mel (mels, mel_lens, max_mel_len) is mandatory at inference")
        nb = True
        speakers = torch.as_tensor(speakers).to(dev, torch.int64, non_blocking=nb).contiguous()
        texts = torch.as_tensor(texts).to(dev, torch.int64, non_blocking=nb).contiguous()
        src_lens_in = torch.as_tensor(src_lens).to(dev, non_blocking=nb)
        mels = torch.as_tensor(mels).to(dev, torch.float32, non_blocking=nb).contiguous()
        max_src_len, max_mel_len = int(max_src_len), int(max_mel_len)
        if texts.shape[1] != max_src_len or mels.shape[1] != max_mel_len:
            raise ValueError("texts / mels must be padded to max_src_len / max_mel_len (dataset.py:198-217)")
        src_lens32 = _as_lens(src_lens_in, dev)
        ref_lens32 = _as_lens(torch.as_tensor(mel_lens), dev)

        if p_targets is not None:
            p_targets = torch.as_tensor(p_targets).to(dev, torch.float32, non_blocking=nb)
        if e_targets is not None:
            e_targets = torch.as_tensor(e_targets).to(dev, torch.float32, non_blocking=nb)
        if d_targets is not None:
            d_targets = torch.as_tensor(d_targets).to(dev, non_blocking=nb)
        src_masks = ops.length_mask(src_lens32, max_src_len)
        gammas, betas = self.reference_encoder(mels, max_mel_len, None, lens=ref_lens32)
        x = self.encoder(texts, None, gammas, betas, lens=src_lens32)
        if self.speaker_emb is not None:
            x = ops.add_speaker(x, speakers, self.speaker_emb.weight)  # pads included (fastspeech2.py:74-77)
            if self.batch_invariant:
                ops.mask_rows_(x, src_lens32)

        ref_masks = ops.length_mask(ref_lens32, max_mel_len) if d_targets is not None else None
        (x, p_pred, e_pred, log_d, d_rounded, mel_lens_out, mel_masks, extra) = self.variance_adaptor(
            x, src_masks, ref_masks, max_mel_len, p_targets, e_targets, d_targets, p_control, e_control, d_control,
            gammas, betas, src_lens=src_lens32, decoder_pe=lambda T: self.decoder._pe(T, self.decoder.d_model),
            return_extra=True, inplace=True)
        T = extra["T"]
        if d_targets is not None:
            # teacher forcing: the decoder mask is the reference mel's mask (fastspeech2.py:62-66, modules.py:133-135)
            dec_lens = ref_lens32  # every kernel clamps lengths to its own row count
        else:
            # decoder key mask = mask from mel_len cropped to the expanded length (Models.py:154-162)
            dec_lens = mel_lens_out  # may exceed T with d_control != 1; every kernel clamps lengths to its own row count
            if mel_masks.shape[1] != T:
                mel_masks = ops.length_mask(dec_lens, T)
        inv_lens = dec_lens if self.batch_invariant else None  # zero the padded mel frames (mel_linear's bias)
        if self.precision == "bf16":
            # Decoder, mel_linear and PostNet on the tensor cores (bf16 operands, fp32 accumulation and
            # fp32 residual stream).  Everything before this point decides integers and stays fp32.
            (x, xb), _ = self.decoder(x, mel_masks, gammas, betas, lens=dec_lens, pe_added=True, precision="bf16",
                                      return_bf16=True)
            mel, melb = ops.conv1d_tc(xb, self.packed()["mel_w"], self.mel_linear.bias, act2=ops.ACT_NONE, lens=inv_lens)
            postnet_mel = self.postnet(mel, residual=mel, precision="bf16", xb=melb, lens=inv_lens)
        else:
            # Everything up to here decides integers (durations, bucket ids) and runs on the 3xTF32 kernels.  From the
            # decoder on only the mel is produced: its wide layers take the split-bf16 kernels (ops.fp32_convs()).
            prec = "bf16x3" if ops.fp32_convs() == "bf16x3" else "fp32"
            x, _ = self.decoder(x, mel_masks, gammas, betas, lens=dec_lens, pe_added=True, precision=prec)
            mel = ops.conv1d(x, self.packed()["mel_w"], self.mel_linear.bias, lens=inv_lens)
            postnet_mel = self.postnet(mel, residual=mel, lens=inv_lens, precision=prec)
        out = (mel, postnet_mel, p_pred, e_pred, log_d, d_rounded, src_masks, mel_masks, src_lens_in, mel_lens_out)
        if return_extra:
            extra.update(gammas=gammas, betas=betas)
            return out, extra
        return out
