"""ctypes binding of include/speakingstyle_b200.h.

The shared library is the product: if it is missing or cannot be loaded this module raises - there
is no eager/PyTorch fallback anywhere in the package.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# SS_LIB_PATH: A/B timing of two builds of the same ABI on one box (debug only; the default is the in-tree build)
LIB_PATH = os.environ.get("SS_LIB_PATH") or os.path.join(_HERE, "lib", "libspeakingstyle_b200.so")

vp, i32, i64, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float

ACT_NONE, ACT_RELU, ACT_TANH, ACT_LRELU = 0, 1, 2, 3


class Conv1dF32Params(C.Structure):
    _fields_ = [
        ("x", vp), ("x_batch_stride", i64), ("x_row_stride", i64), ("w", vp), ("bias", vp), ("residual", vp),
        ("lens", vp), ("y", vp), ("accum", vp), ("y_batch_stride", i64), ("y_row_stride", i64), ("y_offset", i64),
        ("y_batch_elems", i64), ("B", i32), ("S_in", i32), ("S_out", i32), ("Cin", i32), ("Cout", i32), ("k", i32),
        ("dilation", i32), ("pad", i32), ("in_act", i32), ("in_slope", f32), ("act", i32), ("slope", f32),
        ("accum_scale", f32), ("accum_init", i32),
    ]


class Conv1dTcParams(C.Structure):
    _fields_ = [
        ("x", vp), ("x_batch_stride", i64), ("x_row_stride", i64), ("w", vp), ("bias", vp), ("residual", vp),
        ("lens", vp), ("y", vp), ("y_act", vp), ("accum", vp), ("y_batch_stride", i64), ("y_row_stride", i64),
        ("y_offset", i64), ("y_batch_elems", i64), ("B", i32), ("S_in", i32), ("S_out", i32), ("Cin", i32),
        ("Cout", i32), ("k", i32), ("dilation", i32), ("pad", i32), ("act", i32), ("slope", f32), ("act2", i32),
        ("slope2", f32), ("y_dtype", i32), ("residual_dtype", i32), ("accum_scale", f32), ("accum_init", i32),
        ("rows", vp), ("pair_c", i32),
    ]


class ResUnitParams(C.Structure):
    _fields_ = [
        ("x", vp), ("residual", vp), ("w1", vp), ("b1", vp), ("w2", vp), ("b2", vp), ("y", vp), ("y_pair", vp),
        ("accum", vp), ("rows", vp), ("B", i32), ("S", i32), ("C", i32), ("k", i32), ("dilation", i32), ("slope", f32),
        ("act2", i32), ("slope2", f32), ("accum_scale", f32), ("accum_init", i32),
    ]


class Conv1dTf32Params(C.Structure):
    _fields_ = [
        ("x", vp), ("x_batch_stride", i64), ("x_row_stride", i64), ("w_hi", vp), ("w_lo", vp), ("bias", vp),
        ("residual", vp), ("lens", vp), ("y", vp), ("y_act", vp), ("accum", vp), ("y_batch_stride", i64),
        ("y_row_stride", i64), ("y_offset", i64), ("y_batch_elems", i64), ("B", i32), ("S_in", i32), ("S_out", i32),
        ("Cin", i32), ("Cout", i32), ("k", i32), ("dilation", i32), ("pad", i32), ("in_act", i32), ("in_slope", f32),
        ("act", i32), ("slope", f32), ("act2", i32), ("slope2", f32), ("accum_scale", f32), ("accum_init", i32),
        ("k_group", i32), ("rows", vp),
    ]


class LayerNormParams(C.Structure):
    _fields_ = [
        ("x", vp), ("residual", vp), ("ln_w", vp), ("ln_b", vp), ("gamma", vp), ("beta", vp), ("s_gamma", vp),
        ("s_beta", vp), ("lens", vp), ("pe", vp), ("y", vp), ("y_bf16", vp), ("B", i32), ("S", i32), ("C", i32), ("pre_relu", i32),
        ("bf16_pair", i32),
    ]


class PredictorHeadParams(C.Structure):
    _fields_ = [
        ("h", vp), ("ln_w", vp), ("ln_b", vp), ("gamma", vp), ("beta", vp), ("s_gamma", vp), ("s_beta", vp),
        ("lin_w", vp), ("lin_b", vp), ("lens", vp), ("pred", vp), ("B", i32), ("L", i32), ("C", i32),
    ]


_SIGNATURES = {
    "ss_status_string": (C.c_char_p, [C.c_int]),
    "ss_last_error": (C.c_char_p, []),
    "ss_version": (C.c_int, []),
    "ss_launch_count": (i64, []),
    "ss_launch_count_reset": (None, []),
    "ss_embed_posenc": (C.c_int, [vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp]),
    "ss_add_posenc": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, C.c_int, vp]),
    "ss_conv1d_f32": (C.c_int, [C.POINTER(Conv1dF32Params), vp]),
    "ss_conv1d_tc": (C.c_int, [C.POINTER(Conv1dTcParams), vp]),
    "ss_conv1d_tf32x3": (C.c_int, [C.POINTER(Conv1dTf32Params), vp]),
    "ss_conv1d_tf32x3_ts": (C.c_int, [C.POINTER(Conv1dTf32Params), vp]),
    "ss_conv1d_tf32x3_ts2": (C.c_int, [C.POINTER(Conv1dTf32Params), vp]),
    "ss_conv1d_tf32x3_panel": (C.c_int, [C.POINTER(Conv1dTf32Params), vp]),
    "ss_conv1d_tf32x3_panel_supported": (C.c_int, [C.POINTER(Conv1dTf32Params)]),
    "ss_cast_bf16": (C.c_int, [vp, vp, i64, C.c_int, f32, vp]),
    "ss_conv1d_bf16x3": (C.c_int, [C.POINTER(Conv1dTcParams), vp]),
    "ss_conv1d_bf16_panel": (C.c_int, [C.POINTER(Conv1dTcParams), vp]),
    "ss_resunit_bf16x3": (C.c_int, [C.POINTER(ResUnitParams), vp]),
    "ss_resunit_bf16x3_supported": (C.c_int, [C.c_int, C.c_int, C.c_int]),
    "ss_conv1d_bf16x3_plan": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int]),
    "ss_split_bf16": (C.c_int, [vp, vp, i64, C.c_int, C.c_int, f32, vp]),
    "ss_layernorm": (C.c_int, [C.POINTER(LayerNormParams), vp]),
    "ss_attention_f32": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp]),
    "ss_attention_tc": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp]),
    "ss_predictor_head": (C.c_int, [C.POINTER(PredictorHeadParams), vp]),
    "ss_bucket_embed_add": (C.c_int, [vp, f32, vp, C.c_int, vp, vp, vp, C.c_int, C.c_int, C.c_int, vp]),
    "ss_bucket_embed_add_v": (C.c_int, [vp, vp, vp, C.c_int, vp, vp, vp, C.c_int, C.c_int, C.c_int, vp]),
    "ss_duration_scan": (C.c_int, [vp, f32, vp, vp, vp, vp, C.c_int, C.c_int, vp]),
    "ss_duration_scan_v": (C.c_int, [vp, vp, vp, vp, vp, vp, C.c_int, C.c_int, vp]),
    "ss_length_regulate": (C.c_int, [vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp]),
    "ss_length_regulate_noalloc": (C.c_int, [vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp]),
    "ss_length_mask": (C.c_int, [vp, vp, C.c_int, C.c_int, vp]),
    "ss_mean_pool": (C.c_int, [vp, vp, C.c_int, C.c_int, C.c_int, vp]),
    "ss_mean_pool_lens": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, C.c_int, vp]),
    "ss_mask_rows": (C.c_int, [vp, vp, C.c_int, C.c_int, C.c_int, vp]),
    "ss_add_speaker": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp]),
    "ss_hifigan_post": (C.c_int, [vp, vp, vp, vp, vp, f32, f32, C.c_int, C.c_int, C.c_int, C.c_int, vp]),
    "ss_reflect_pad": (C.c_int, [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp]),
    "ss_stft_magnitude": (C.c_int, [vp, vp, vp, C.c_int64, C.c_int, C.c_int, C.c_int, vp]),
    "ss_range_flag": (C.c_int, [vp, i64, f32, f32, vp, vp]),
    "ss_log_clamp": (C.c_int, [vp, C.c_int64, f32, vp]),
    "ss_hifigan_post_trim": (C.c_int, [vp, vp, vp, vp, vp, vp, f32, f32, C.c_int, C.c_int, C.c_int, C.c_int, vp]),
}

_lib = None


def exported_names():
    return sorted(_SIGNATURES)


def lib():
    """The loaded C-ABI library.  Raises (loudly) when the CUDA extension has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m speakingstyle_b200.build` "
                "(or __graft_entry__.build()).  speakingstyle_b200 has no CPU or eager fallback.")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError if the header and the library diverge
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


class SSError(RuntimeError):
    pass


def check(status):
    if status != 0:
        h = lib()
        raise SSError(f"{h.ss_status_string(status).decode()} ({status}): {h.ss_last_error().decode()}")
