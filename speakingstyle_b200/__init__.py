"""speakingstyle_b200 - B200-native (sm_100a) drop-in for the batched text -> mel -> wave inference
path of lordzuko/SpeakingStyle: FastSpeech2 + FiLM ReferenceEncoder + VarianceAdaptor /
LengthRegulator + PostNet + HiFi-GAN Generator, behind the reference's own Python API.

Everything that computes is a hand-written CUDA kernel reached through the C ABI declared in
include/speakingstyle_b200.h (libspeakingstyle_b200.so, loaded with ctypes).  PyTorch provides
device memory, streams and torch.distributed only.  There is no CPU or eager fallback: importing
the model classes works anywhere, running them needs the built library and a CUDA device.
"""
from .audio import STFT, TacotronSTFT, get_mel_from_wav  # noqa: F401
from .configs import AttrDict, hifigan_config, model_config, preprocess_config  # noqa: F401
from .fastspeech2 import FastSpeech2  # noqa: F401
from .hifigan import Generator, ResBlock  # noqa: F401
from .modules import (LengthRegulator, ReferenceEncoder, VarianceAdaptor, VariancePredictor,  # noqa: F401
                      get_mask_from_lengths)
from .transformer import Decoder, Encoder, FFTBlock, MultiHeadAttention, PositionwiseFeedForward, PostNet  # noqa: F401
from .tools import get_model, get_vocoder, synthesize_batch, vocoder_infer  # noqa: F401

__version__ = "0.1.0"
