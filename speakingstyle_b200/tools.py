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
def vocoder_infer(mels, vocoder, model_config=None, preprocess_config=None, lengths=None):
    """utils/model.py:97-115: mels [B,80,T] -> list of int16 numpy arrays trimmed to ``lengths``.
    The scaling and the int16 truncation happen in the output kernel; only int16 crosses PCIe."""
    scale = 32768.0
    if preprocess_config is not None:
        scale = float(preprocess_config["preprocessing"]["audio"]["max_wav_value"])
    if scale != 32768.0:
        raise NotImplementedError("max_wav_value other than 32768 is not wired into the PCM epilogue")
    _, pcm = vocoder(mels, pcm=True)
    wavs = [w for w in pcm.cpu().numpy()]
    if lengths is not None:
        wavs = [w[: int(n)] for w, n in zip(wavs, lengths)]
    return wavs


@torch.no_grad()
def synthesize_batch(model, vocoder, batch, p_control=1.0, e_control=1.0, d_control=1.0, hop_length=256):
    """One batch of synthesize.py:128-150 + utils/tools.py:183-230 without plots and files:
    host batch dict -> (10-tuple of model outputs, list of int16 waves or None)."""
    out = model(batch["speakers"], batch["texts"], batch["src_lens"], batch["max_src_len"], batch["mels"],
                batch["mel_lens"], batch["max_mel_len"], p_control=p_control, e_control=e_control,
                d_control=d_control)
    wavs = None
    if vocoder is not None:
        lengths = (out[9] * hop_length).cpu().tolist()
        _, pcm = vocoder(out[1], channels_last=True, pcm=True)
        wavs = [w[:n] for w, n in zip(pcm.cpu().numpy(), lengths)]
    return out, wavs
