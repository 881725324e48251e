"""Glue the reference keeps in utils/model.py and utils/tools.py, for this path only."""
import torch

from . import configs
from .fastspeech2 import FastSpeech2
from .hifigan import Generator


def get_model(preprocess_config, model_config, device="cuda", state_dict=None, precision="fp32"):
    """utils/model.py:11-45 without the checkpoint-file handling: build, load, eval, move."""
    model = FastSpeech2(preprocess_config, model_config, precision=precision)
    if state_dict is not None:
        model.load_state_dict(state_dict, strict=True)
    return model.to(device)


def get_vocoder(h=None, device="cuda", state_dict=None, precision="fp32"):
    """utils/model.py:62-94: Generator(h), load ckpt["generator"], eval, remove_weight_norm, move."""
    gen = Generator(h or configs.hifigan_config(), precision=precision)
    if state_dict is not None:
        gen.load_state_dict(state_dict, strict=True)
    gen.remove_weight_norm()
    return gen.to(device)


@torch.no_grad()
def _ragged_pcm(vocoder, mels, n_samples, channels_last, device_out=False, pcm_scale=32768.0):
    """Runs the vocoder with the ragged PCM epilogue writing into pinned host memory; returns one int16 numpy view
    per utterance (no padded wave, no separate D2H copy - the kernel's stores are what crosses PCIe).
    device_out: keep the trimmed waves on the device instead (one flat int16 tensor, returned as per-utterance views,
    no synchronisation) - for callers that post-process on the GPU."""
    if not mels.is_cuda:
        raise RuntimeError("speakingstyle_b200 vocoder needs CUDA tensors; there is no CPU fallback")
    with torch.cuda.device(mels.device):  # launches and the stream synchronisation go to the mel's own device
        return _ragged_pcm_on_device(vocoder, mels, n_samples, channels_last, device_out, pcm_scale)


def _ragged_pcm_on_device(vocoder, mels, n_samples, channels_last, device_out, pcm_scale):
    T = mels.shape[1] if channels_last else mels.shape[2]
    hop = 1
    for u in vocoder.h["upsample_rates"]:
        hop *= u
    n = [max(0, min(int(v), T * hop)) for v in n_samples]
    offs = [0]
    for v in n:
        offs.append(offs[-1] + v)
    dev = mels.device
    if device_out:
        host = torch.empty(max(offs[-1], 1), dtype=torch.int16, device=dev)
    else:
        host = torch.empty(max(offs[-1], 1), dtype=torch.int16, pin_memory=True)  # cached pinned allocator: cheap after the first call
    n_dev = torch.tensor(n, dtype=torch.int32).to(dev, non_blocking=True)
    o_dev = torch.tensor(offs[:-1], dtype=torch.int64).to(dev, non_blocking=True)
    # every layer runs only on the frames that can reach a returned sample (Generator.forward, frame_lens)
    frame_lens = [-(-v // hop) for v in n]
    vocoder(mels, channels_last=channels_last, pcm_out=(n_dev, o_dev, host), frame_lens=frame_lens, pcm_scale=pcm_scale)
    if device_out:
        return [host[offs[i]:offs[i + 1]] for i in range(len(n))]
    torch.cuda.current_stream(dev).synchronize()
    flat = host.numpy()
    return [flat[offs[i]:offs[i + 1]] for i in range(len(n))]


def vocoder_infer(mels, vocoder, model_config=None, preprocess_config=None, lengths=None):
    """utils/model.py:97-115: mels [B,80,T] -> list of int16 numpy arrays trimmed to ``lengths``.
    The x32768 scaling, the int16 truncation and the trim happen in the output kernel, which writes each
    utterance's valid samples straight into pinned host memory; with lengths=None the whole padded
    waves are returned, as the reference does."""
    scale = 32768.0
    if preprocess_config is not None:
        scale = float(preprocess_config["preprocessing"]["audio"]["max_wav_value"])
    B, T = mels.shape[0], mels.shape[2]
    hop = 1
    for u in vocoder.h["upsample_rates"]:
        hop *= u
    n = [T * hop] * B if lengths is None else [int(v) for v in lengths]
    return _ragged_pcm(vocoder, mels, n, channels_last=False, pcm_scale=scale)


@torch.no_grad()
def synthesize_batch(model, vocoder, batch, p_control=1.0, e_control=1.0, d_control=1.0, hop_length=256,
                     device_out=False):
    """One batch of synthesize.py:128-150 + utils/tools.py:183-230 without plots and files:
    host batch dict -> (10-tuple of model outputs, list of int16 waves or None).  The waves are numpy views of pinned
    host memory, trimmed to mel_len * hop; with device_out=True they stay on the device (int16 tensors, no sync)."""
    out = model(batch["speakers"], batch["texts"], batch["src_lens"], batch["max_src_len"], batch["mels"],
                batch["mel_lens"], batch["max_mel_len"], p_control=p_control, e_control=e_control,
                d_control=d_control)
    wavs = None
    if vocoder is not None:
        mel_len = getattr(model.variance_adaptor, "last_mel_len_host", None)  # read once inside forward()
        if mel_len is None or len(mel_len) != out[9].shape[0]:
            mel_len = out[9].cpu()
        wavs = _ragged_pcm(vocoder, out[1], [int(v) * hop_length for v in mel_len.tolist()], channels_last=True,
                           device_out=device_out)
    return out, wavs
