"""Configuration dictionaries for the text -> mel -> wave path.

The reference passes three YAML dicts around (synthesize.py:235-240); only two of them are read
by the path (model/fastspeech2.py:16, model/modules.py:23-81, 310-322, transformer/Models.py:36-51).
The values below restate the shipped YAML/JSON files so that tests and the benchmark can run on a
machine where the reference checkout does not exist:

* config/BC2013/model.yaml (the only model.yaml that carries a ``reference_encoder`` block, lines 13-20)
* config/{LJSpeech,LibriTTS}/model.yaml + that block
* config/*/preprocess.yaml keys read on the path (mel.n_mel_channels, pitch/energy.feature,
  audio.max_wav_value, stft.hop_length, path.preprocessed_path)
* hifigan/config.json (generator hyper-parameters, lines 11-15, 20-24)
* preprocessed_data/{LJSpeech,LibriTTS}/stats.json min/max (bin edges, model/modules.py:43-75)

A caller holding the reference's own YAML dicts can pass them unchanged.
"""
import copy
import json
import os

N_SRC_VOCAB = 361  # len(text.symbols.symbols) + 1  (transformer/Models.py:40, text/symbols.py:21-29)

_REFERENCE_ENCODER = {
    "encoder_layer": 4,
    "encoder_head": 8,
    "encoder_hidden": 256,
    "conv_layer": 3,
    "conv_filter_size": 1024,
    "conv_kernel_size": 3,
    "dropout": 0.1,
}

_MODEL_BASE = {
    "transformer": {
        "encoder_layer": 4,
        "encoder_head": 2,
        "encoder_hidden": 256,
        "decoder_layer": 6,
        "decoder_head": 2,
        "decoder_hidden": 256,
        "conv_filter_size": 1024,
        "conv_kernel_size": [9, 1],
        "encoder_dropout": 0.2,
        "decoder_dropout": 0.2,
    },
    "reference_encoder": _REFERENCE_ENCODER,
    "variance_predictor": {"filter_size": 256, "kernel_size": 3, "dropout": 0.5},
    "variance_embedding": {
        "pitch_quantization": "linear",
        "energy_quantization": "linear",
        "n_bins": 256,
    },
    "multi_speaker": False,
    "max_seq_len": 1000,
    "vocoder": {"model": "HiFi-GAN", "speaker": "LJSpeech"},
}

STATS = {
    # preprocessed_data/LJSpeech/stats.json (BC2013 ships none; SURVEY.md 8c uses LJSpeech's)
    "LJSpeech": {
        "pitch": [-2.917079304729967, 11.391254536985784, 207.6309860026605, 46.77559025098988],
        "energy": [-1.431044578552246, 8.184337615966797, 37.32621679053821, 26.044180782835863],
    },
    # preprocessed_data/LibriTTS/stats.json
    "LibriTTS": {
        "pitch": [-2.646310080183867, 11.922013280384945, 163.55966796034886, 61.80669044989039],
        "energy": [-1.248658537864685, 9.75546646118164, 41.65338755249414, 33.35850956918866],
    },
}
STATS["BC2013"] = STATS["LJSpeech"]
# config/LJSpeech_paper (frame-level pitch / energy, log pitch quantisation, no normalisation) ships no stats.json;
# these are plausible UN-normalised ranges (Hz / frame energy) with LJSpeech's mean and std - synthetic, stated as such
STATS["LJSpeech_paper"] = {
    "pitch": [71.0, 795.8, 207.6309860026605, 46.77559025098988],
    "energy": [0.018, 315.0, 37.32621679053821, 26.044180782835863],
}
N_SPEAKERS = {"LJSpeech": 1, "BC2013": 1, "LibriTTS": 904, "LJSpeech_paper": 1}  # len(speakers.json)

HIFIGAN_CONFIG = {
    "resblock": "1",
    "upsample_rates": [8, 8, 2, 2],
    "upsample_kernel_sizes": [16, 16, 4, 4],
    "upsample_initial_channel": 512,
    "resblock_kernel_sizes": [3, 7, 11],
    "resblock_dilation_sizes": [[1, 3, 5], [1, 3, 5], [1, 3, 5]],
    "num_mels": 80,
    "hop_size": 256,
    "sampling_rate": 22050,
}


class AttrDict(dict):
    """Same convenience type as hifigan/__init__.py:4-7 (keys readable as attributes)."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.__dict__ = self


def model_config(dataset="BC2013"):
    cfg = copy.deepcopy(_MODEL_BASE)
    if dataset == "LibriTTS":
        cfg["multi_speaker"] = True
        cfg["vocoder"]["speaker"] = "universal"
    elif dataset == "LJSpeech_paper":  # config/LJSpeech_paper/model.yaml:5,19 (+ the BC2013 reference_encoder block)
        cfg["transformer"]["decoder_layer"] = 4
        cfg["variance_embedding"]["pitch_quantization"] = "log"
    elif dataset not in ("BC2013", "LJSpeech"):
        raise ValueError(f"unknown dataset {dataset!r}")
    return cfg


def preprocess_config(dataset="BC2013", preprocessed_path=None):
    # config/LJSpeech_paper/preprocess.yaml:26-30: frame-level features, not normalised
    level = "frame_level" if dataset == "LJSpeech_paper" else "phoneme_level"
    norm = dataset != "LJSpeech_paper"
    return {
        "dataset": dataset,
        "path": {"preprocessed_path": preprocessed_path or ""},
        "preprocessing": {
            "audio": {"sampling_rate": 22050, "max_wav_value": 32768.0},
            "stft": {"filter_length": 1024, "hop_length": 256, "win_length": 1024},
            "mel": {"n_mel_channels": 80, "mel_fmin": 0, "mel_fmax": 8000},
            "pitch": {"feature": level, "normalization": norm},
            "energy": {"feature": level, "normalization": norm},
        },
    }


def hifigan_config():
    return AttrDict(copy.deepcopy(HIFIGAN_CONFIG))


def write_preprocessed_dir(path, dataset="BC2013"):
    """Create the two JSON files the constructors read (modules.py:43-49, fastspeech2.py:31-38)."""
    os.makedirs(path, exist_ok=True)
    with open(os.path.join(path, "stats.json"), "w") as f:
        json.dump(STATS[dataset], f)
    with open(os.path.join(path, "speakers.json"), "w") as f:
        json.dump({f"spk{i}": i for i in range(N_SPEAKERS[dataset])}, f)
    return path
