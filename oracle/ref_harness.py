"""TEST INFRASTRUCTURE ONLY - loads the unmodified reference from /root/reference on the CPU.

Used (a) by oracle/make_golden.py to produce the fixtures under tests/golden/ and (b) by the
``not gpu`` test that pins oracle/fs2_oracle.py against the real reference modules.  It cannot
travel to the GPU box (there is no /root/reference there); nothing on the product path, in the
``-m gpu`` tests, in smoke() or in bench.py imports this file.

How the import is made to work: the reference imports three packages at module
import time that this image does not have (matplotlib: utils/tools.py:7-9; unidecode:
text/cleaners.py:18; inflect: text/numbers.py:3).  None of them is executed by the forward path,
so inert stand-ins are registered in ``sys.modules`` first.  The reference also picks its device
from a module-global (utils/tools.py:15, model/modules.py:17); both are forced to CPU.
"""
import contextlib
import io
import os
import sys
import types

import torch

REFERENCE_ROOT = os.environ.get("SS_REFERENCE_ROOT", "/root/reference")


def available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "model"))


def _install_stubs():
    def stub(name, **attrs):
        if name in sys.modules:
            return sys.modules[name]
        m = types.ModuleType(name)
        for k, v in attrs.items():
            setattr(m, k, v)
        sys.modules[name] = m
        return m

    try:
        import unidecode  # noqa: F401
    except ImportError:
        stub("unidecode", unidecode=lambda s: s)
    try:
        import inflect  # noqa: F401
    except ImportError:
        stub("inflect", engine=lambda: None)
    try:
        import matplotlib  # noqa: F401
    except ImportError:
        mpl = stub("matplotlib", use=lambda *a, **k: None)
        mpl.pyplot = stub("matplotlib.pyplot")
        mpl.pylab = stub("matplotlib.pylab")


def load_audio():
    """The reference's audio.stft module (STFT, TacotronSTFT) with stand-ins for the librosa symbols it imports at
    module level (audio/stft.py:5-6; librosa is not installed here): pad_center / tiny are one-liners, and
    filters.mel is the restatement in oracle/mel_oracle.py - so the filterbank VALUES are unpinned, everything
    else (reflect pad, Fourier-basis convolution, magnitude, log compression, energy) is the reference's own code."""
    if "audio" in _loaded:
        return _loaded["audio"]
    if not available():
        raise RuntimeError(f"reference not found at {REFERENCE_ROOT}")
    import numpy as np
    from oracle import mel_oracle

    def pad_center(data, size, axis=-1, **kw):
        n = data.shape[axis]
        lpad = int((size - n) // 2)
        lengths = [(0, 0)] * data.ndim
        lengths[axis] = (lpad, int(size - n - lpad))
        return np.pad(data, lengths, **kw)

    try:
        import librosa  # noqa: F401
    except ImportError:
        lib = types.ModuleType("librosa")
        util = types.ModuleType("librosa.util")
        util.pad_center, util.tiny = pad_center, (lambda x: np.finfo(np.float32).tiny)
        filters = types.ModuleType("librosa.filters")
        filters.mel = lambda sr, n_fft, n_mels=128, fmin=0.0, fmax=None: mel_oracle.slaney_mel(sr, n_fft, n_mels, fmin, fmax)
        lib.util, lib.filters = util, filters
        sys.modules.update({"librosa": lib, "librosa.util": util, "librosa.filters": filters})
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import torch.cuda
    with contextlib.redirect_stdout(io.StringIO()):
        import audio.stft as ref_stft
    _loaded["audio"] = ref_stft
    return ref_stft


_loaded = {}


def load():
    """Returns a namespace with the reference's FastSpeech2, Generator, AttrDict and helpers."""
    if "ns" in _loaded:
        return _loaded["ns"]
    if not available():
        raise RuntimeError(f"reference not found at {REFERENCE_ROOT}")
    _install_stubs()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    with contextlib.redirect_stdout(io.StringIO()):
        import model as ref_model  # noqa
        import model.modules as ref_modules
        import utils.tools as ref_tools
        import hifigan as ref_hifigan
        import transformer.Models as ref_tmodels
    cpu = torch.device("cpu")
    ref_modules.device = cpu
    ref_tools.device = cpu
    ns = types.SimpleNamespace(
        FastSpeech2=ref_model.FastSpeech2,
        Generator=ref_hifigan.Generator,
        AttrDict=ref_hifigan.AttrDict,
        get_sinusoid_encoding_table=ref_tmodels.get_sinusoid_encoding_table,
        get_mask_from_lengths=ref_tools.get_mask_from_lengths,
        modules=ref_modules,
        tools=ref_tools,
    )
    _loaded["ns"] = ns
    return ns


@contextlib.contextmanager
def quiet():
    """The reference prints from inside forward() (fastspeech2.py:70,73,79,103)."""
    with contextlib.redirect_stdout(io.StringIO()):
        yield


def build_fastspeech2(preprocess_config, model_config, state_dict=None):
    ns = load()
    with quiet():
        m = ns.FastSpeech2(preprocess_config, model_config).eval()
    if state_dict is not None:
        m.load_state_dict(state_dict, strict=True)
    return m


def build_generator(h, state_dict=None):
    """state_dict uses weight_g / weight_v keys (checkpoint form, utils/model.py:81-92)."""
    ns = load()
    import warnings

    with quiet(), warnings.catch_warnings():
        warnings.simplefilter("ignore")
        g = ns.Generator(ns.AttrDict(dict(h)))
        if state_dict is not None:
            g.load_state_dict(state_dict, strict=True)
        g.eval()
        g.remove_weight_norm()
    return g
