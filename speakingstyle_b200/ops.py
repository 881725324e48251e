"""Tensor-level wrappers over the C ABI (include/speakingstyle_b200.h).

torch is used for device memory and streams only: every function here unwraps tensors to raw
device pointers and enqueues ONE library call on torch's current stream.  Nothing in this file
computes with torch operators.
"""
import ctypes as C
import os

import torch

from . import _lib
from ._lib import ACT_LRELU, ACT_NONE, ACT_RELU, ACT_TANH, check  # noqa: F401


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def _stream():
    """The current CUDA stream of the current device as a raw handle.  torch.cuda.current_stream() builds a Python
    Stream object per call (~15 us, more than a small kernel runs); the raw getter is a plain C call."""
    if _raw_stream is not None:
        return C.c_void_p(_raw_stream(torch.cuda.current_device()))
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _chk(t, dtype, name):
    if t is None:
        return
    if not t.is_cuda:
        raise _lib.SSError(f"{name}: expected a CUDA tensor (speakingstyle_b200 has no CPU path)")
    if t.dtype != dtype:
        raise _lib.SSError(f"{name}: expected {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise _lib.SSError(f"{name}: expected a contiguous tensor")


class KernelTimer:
    """Optional per-call CUDA-event timing of one kernel family, used by bench.py to measure the
    dominant kernel's average launch duration inside the timed region (events are recorded on the
    stream the kernel is launched on).  Costs two event records per call while installed.
    ``ctx`` is whatever the caller sets before a group of launches (bench.py: the index of the batch being
    processed); it is stored with every record so the caller can attribute launches afterwards."""

    def __init__(self):
        self.records = []  # (start_event, end_event, work, tag, ctx, meta)
        self.ctx = None

    def run(self, work, fn, tag=None, meta=None):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        status = fn()
        b.record()
        self.records.append((a, b, work, tag, self.ctx, meta))
        return status

    def summary(self):
        """(calls, total_ms, total_work) - call after a synchronize."""
        ms = sum(r[0].elapsed_time(r[1]) for r in self.records)
        return len(self.records), ms, sum(r[2] for r in self.records)

    def rows(self):
        """[(ms, work, tag, ctx, meta)] per launch - call after a synchronize.
        meta = (kind, B, S_in, Cin, Cout, k, dilation) of the convolution."""
        return [(a.elapsed_time(b), w, tag, ctx, meta) for a, b, w, tag, ctx, meta in self.records]

    def by_tag(self):
        """{tag: (calls, ms, work)} sorted by time, for the per-shape breakdown bench.py can print."""
        agg = {}
        for a, b, w, tag, _, _ in self.records:
            c, ms, ww = agg.get(tag, (0, 0.0, 0.0))
            agg[tag] = (c + 1, ms + a.elapsed_time(b), ww + w)
        return dict(sorted(agg.items(), key=lambda kv: -kv[1][1]))


conv_timer = None  # set to a KernelTimer to time every ss_conv1d_f32 launch; work = algorithmic FLOPs


def _tag(kind, p):
    return f"{kind} B={p.B} S={p.S_in} Cin={p.Cin} Cout={p.Cout} k={p.k} dil={p.dilation}"


def _meta(kind, p):
    return (kind, int(p.B), int(p.S_in), int(p.Cin), int(p.Cout), int(p.k), int(p.dilation))


def _run_conv(p, flops):
    fn = lambda: _lib.lib().ss_conv1d_f32(C.byref(p), _stream())  # noqa: E731
    check(conv_timer.run(flops, fn, _tag("f32", p), _meta("f32", p)) if conv_timer is not None else fn())


def launch_count():
    return int(_lib.lib().ss_launch_count())


def launch_count_reset():
    _lib.lib().ss_launch_count_reset()


# ---------------------------------------------------------------------------------------------
def embed_posenc(texts, emb, pe):
    _chk(texts, torch.int64, "texts"); _chk(emb, torch.float32, "emb"); _chk(pe, torch.float32, "pe")
    B, L = texts.shape
    V, D = emb.shape
    if pe.shape[0] < L or pe.shape[1] != D:
        raise _lib.SSError("embed_posenc: position table too small")
    out = torch.empty((B, L, D), dtype=torch.float32, device=texts.device)
    check(_lib.lib().ss_embed_posenc(_ptr(texts), _ptr(emb), _ptr(pe), _ptr(out), B, L, D, V, _stream()))
    return out


def add_position(x, pe):
    _chk(x, torch.float32, "x"); _chk(pe, torch.float32, "pe")
    B, S, Cc = x.shape
    if pe.shape[0] < S or pe.shape[1] != Cc:
        raise _lib.SSError("add_position: position table too small")
    out = torch.empty_like(x)
    check(_lib.lib().ss_add_posenc(_ptr(x), _ptr(pe), _ptr(out), B, S, Cc, _stream()))
    return out


fp32_backend = "tf32x3"  # "tf32x3": 3xTF32 tcgen05 kernel (default); "ffma": the FP32-pipe kernel
# experiment knobs (environment overrides are for A/B timing only; the defaults are what the tests pin)
tf32_k_group = int(os.environ.get("SS_TF32_KGROUP", "0"))  # 0 = library default (accumulator restart every 8 chunks)
# "auto": A operand in tensor memory - the CTA-pair kernel ("ts2") for long reductions, the single-CTA one ("ts") for
# short ones (they are bit-identical; the pair form is faster from 12 chunks up and slower on k = 1, K = 256 layers);
# "ss": split operands in shared memory (the first version, kept as a cross-check)
tf32_variant = os.environ.get("SS_TF32_VARIANT", "auto")
# narrow layers (Cin, Cout <= 64, k >= 2: the HiFi-GAN 64- / 32-channel stages) go to the panel kernel in "auto"
tf32_panel = os.environ.get("SS_TF32_PANEL", "1") != "0"
# ... including those whose weights stream through a ring (64 channels, k > 3); "0": those stay on the per-tap kernel
tf32_panel_streamed = os.environ.get("SS_TF32_PANEL_STREAMED", "1") != "0"


def tf32_round(t):
    """Nearest TF32 value (10 explicit mantissa bits), ties away from zero in magnitude, as an fp32 tensor."""
    return ((t.contiguous().view(torch.int32) + 0x1000) & ~0x1FFF).view(torch.float32)


class ConvW:
    """One convolution weight in the layouts the kernels want, built lazily from the canonical
    [k, Cout, Cin] fp32 tensor: .ffma [k*Cin, Cout] (ss_conv1d_f32), .tf32 = (hi, lo) [k, Cout, Cin]
    (ss_conv1d_tf32x3), .bf16 [k, Cout, Cin] (ss_conv1d_tc)."""

    def __init__(self, kcc):
        self.kcc = kcc.contiguous()
        self.k, self.cout, self.cin = self.kcc.shape
        self._ffma = self._tf32 = self._bf16 = self._bf16x3 = self._bf16p = None

    @property
    def ffma(self):
        if self._ffma is None:
            self._ffma = self.kcc.permute(0, 2, 1).reshape(self.k * self.cin, self.cout).contiguous()
        return self._ffma

    @property
    def tf32(self):
        if self._tf32 is None:
            hi = tf32_round(self.kcc)
            self._tf32 = (hi, tf32_round(self.kcc - hi))
        return self._tf32

    @property
    def bf16(self):
        if self._bf16 is None:
            self._bf16 = self.kcc.to(torch.bfloat16)
        return self._bf16


    def _panel_tiles(self, w, bn):
        """[k, Cout, Cin] -> the weight tiles of the panel kernels (layout: include/speakingstyle_b200.h), one number per
        value: [tiles, bn rows, 64]."""
        if self.cin == 32:  # two taps per 128-byte row
            npair = (self.k + 1) // 2
            wp = torch.zeros(2 * npair, self.cout, 32, dtype=w.dtype, device=w.device)
            wp[: self.k] = w
            return wp.view(npair, 2, self.cout // 32, 32, 32).permute(2, 0, 3, 1, 4).reshape(-1, 32, 64)
        nch = self.cin // 64
        return w.view(self.k, self.cout // bn, bn, nch, 64).permute(1, 3, 0, 2, 4).reshape(-1, bn, 64)

    @property
    def bf16x3(self):
        """Split-bf16 weight tiles for ss_conv1d_bf16x3 ([W_hi rows ; W_lo rows] per tile), or None when the layer
        shape is not handled by that kernel."""
        if self._bf16x3 is None:
            bn = _lib.lib().ss_conv1d_bf16x3_plan(self.cin, self.cout, self.k, 1)
            if bn == 0:
                self._bf16x3 = False
            else:
                hi = self.kcc.to(torch.bfloat16)
                lo = (self.kcc - hi.float()).to(torch.bfloat16)
                self._bf16x3 = torch.cat([self._panel_tiles(hi, bn), self._panel_tiles(lo, bn)], dim=1).contiguous()
        return None if self._bf16x3 is False else self._bf16x3

    @property
    def bf16p(self):
        """bf16 weight tiles for ss_conv1d_bf16_panel, or None when the layer shape is not handled by that kernel."""
        if self._bf16p is None:
            bn = _lib.lib().ss_conv1d_bf16x3_plan(self.cin, self.cout, self.k, 1)
            self._bf16p = False if bn == 0 else self._panel_tiles(self.kcc.to(torch.bfloat16), bn).contiguous()
        return None if self._bf16p is False else self._bf16p


def pack_conv(weight):
    """nn.Conv1d weight [Cout, Cin, k] or nn.Linear weight [out, in] -> ConvW."""
    if weight.dim() == 2:
        weight = weight.unsqueeze(-1)
    return ConvW(weight.permute(2, 0, 1))


def pack_conv_transpose(weight, stride):
    """nn.ConvTranspose1d weight [Cin, Cout, k] (k % stride == 0) -> ConvW of the phase form: taps = k/stride,
    stride*Cout GEMM columns ordered (phase r, co); tap j' reads input row s-(taps-1-j') with kernel index
    r + (taps-1-j')*stride."""
    cin, cout, k = weight.shape
    taps = k // stride
    w4 = weight.reshape(cin, cout, taps, stride).permute(2, 3, 1, 0).flip(0)  # [j', r, co, ci]
    return ConvW(w4.reshape(taps, stride * cout, cin))


def _tf32_call(x, W, bias, residual, lens, y, y_act, accum, *, B, S_in, S_out, y_batch_stride, y_row_stride,
               y_offset, y_batch_elems, dilation, pad, in_act, in_slope, act, slope, act2, slope2, accum_scale,
               accum_init, flops, variant=None, rows=None):
    hi, lo = W.tf32
    p = _lib.Conv1dTf32Params(
        x=x.data_ptr(), x_batch_stride=S_in * W.cin, x_row_stride=W.cin, w_hi=hi.data_ptr(), w_lo=lo.data_ptr(),
        bias=None if bias is None else bias.data_ptr(), residual=None if residual is None else residual.data_ptr(),
        lens=None if lens is None else lens.data_ptr(), y=None if y is None else y.data_ptr(),
        y_act=None if y_act is None else y_act.data_ptr(), accum=None if accum is None else accum.data_ptr(),
        y_batch_stride=y_batch_stride, y_row_stride=y_row_stride, y_offset=y_offset, y_batch_elems=y_batch_elems,
        B=B, S_in=S_in, S_out=S_out, Cin=W.cin, Cout=W.cout, k=W.k, dilation=dilation, pad=pad, in_act=in_act,
        in_slope=in_slope, act=act, slope=slope, act2=act2 or 0, slope2=slope2, accum_scale=accum_scale,
        accum_init=int(accum_init), k_group=tf32_k_group, rows=None if rows is None else rows.data_ptr())
    v = variant or tf32_variant
    if v == "auto":
        # the panel kernel takes the narrow layers: 1 = weights of all taps resident in shared memory (every 32-channel
        # layer, 64-channel layers with k = 3), 2 = weights streamed once per tile (64 channels, k = 7 / 11)
        mode = (_lib.lib().ss_conv1d_tf32x3_panel_supported(C.byref(p))
                if (tf32_panel and W.cin <= 64 and W.cout <= 64 and W.k >= 2) else 0)
        if mode == 1 or (mode == 2 and tf32_panel_streamed):
            v = "panel"
        else:
            # CTA pair from 12 chunks up (128-channel k = 3 layers: 5-13 % faster than the single-CTA kernel,
            # scripts/hifigan_ncu.py with SS_TF32_VARIANT=ts2); slower on the 8-chunk layers (K = 256, k = 1)
            v = "ts2" if W.k * ((W.cin + 31) // 32) >= 12 else "ts"
            # ragged schedule on short sequences (FFT blocks: a few hundred frames per utterance): the pair kernel
            # schedules 256-row units, so an utterance of 534 frames costs 768 rows against 640 with 128-row tiles -
            # more than the pair form gains
            if v == "ts2" and rows is not None and S_out <= 2048:
                v = "ts"
    entry = {"ts": _lib.lib().ss_conv1d_tf32x3_ts, "ts2": _lib.lib().ss_conv1d_tf32x3_ts2,
             "ss": _lib.lib().ss_conv1d_tf32x3, "panel": _lib.lib().ss_conv1d_tf32x3_panel}[v]
    fn = lambda: entry(C.byref(p), _stream())  # noqa: E731
    check(conv_timer.run(flops, fn, _tag("tf32x3", p), _meta("tf32x3", p)) if conv_timer is not None else fn())


def conv1d(x, w, bias=None, *, k=None, dilation=1, pad=0, act=ACT_NONE, slope=0.0, in_act=ACT_NONE, in_slope=0.0,
           residual=None, lens=None, out=None, accum=None, accum_scale=1.0, accum_init=False, write_y=True,
           backend=None, s_out=None, rows=None):
    """fp32-accurate convolution: x [B,S,Cin] f32, w a ConvW -> y [B,S,Cout] f32.
    backend (default ops.fp32_backend): "tf32x3" = ss_conv1d_tf32x3, "ffma" = ss_conv1d_f32.
    s_out: output length when it is not S (a "valid" convolution: pad=0, s_out = S - (k-1)*dilation); 3xTF32 only.
    rows: int32 [B] on the device - ragged schedule, only rows t < rows[b] are wanted (the tensor-memory 3xTF32
    kernels skip whole tiles past that; the cross-check backends compute every row, which gives the same wanted rows)."""
    _chk(x, torch.float32, "x"); _chk(bias, torch.float32, "bias"); _chk(rows, torch.int32, "rows")
    _chk(residual, torch.float32, "residual"); _chk(lens, torch.int32, "lens"); _chk(accum, torch.float32, "accum")
    B, S, Cin = x.shape
    if not isinstance(w, ConvW):
        raise _lib.SSError("conv1d: weight must be a ConvW (ops.pack_conv)")
    if w.cin != Cin or (k is not None and k != w.k):
        raise _lib.SSError(f"conv1d: weight is [k={w.k}, Cout={w.cout}, Cin={w.cin}] but x has Cin={Cin}, k={k}")
    k, Cout = w.k, w.cout
    So = S if s_out is None else int(s_out)
    be = backend or fp32_backend
    if So != S and (not be.startswith("tf32x3") or So < 1):
        raise _lib.SSError(f"conv1d: s_out={So} needs the 3xTF32 backend and s_out >= 1")
    y = None
    if write_y:
        y = out if out is not None else torch.empty((B, So, Cout), dtype=torch.float32, device=x.device)
        _chk(y, torch.float32, "out")
    if be.startswith("tf32x3"):
        _tf32_call(x, w, bias, residual, lens, y, None, accum, variant=be[7:] or None, B=B, S_in=S, S_out=So,
                   y_batch_stride=So * Cout,
                   y_row_stride=Cout, y_offset=0, y_batch_elems=So * Cout, dilation=dilation, pad=pad, in_act=in_act,
                   in_slope=in_slope, act=act, slope=slope, act2=None, slope2=0.0, accum_scale=accum_scale,
                   accum_init=accum_init, flops=2.0 * B * So * Cout * k * Cin, rows=rows)
        return y
    w = w.ffma
    p = _lib.Conv1dF32Params(
        x=x.data_ptr(), x_batch_stride=S * Cin, x_row_stride=Cin, w=w.data_ptr(),
        bias=None if bias is None else bias.data_ptr(), residual=None if residual is None else residual.data_ptr(),
        lens=None if lens is None else lens.data_ptr(), y=None if y is None else y.data_ptr(),
        accum=None if accum is None else accum.data_ptr(), y_batch_stride=S * Cout, y_row_stride=Cout, y_offset=0,
        y_batch_elems=S * Cout, B=B, S_in=S, S_out=S, Cin=Cin, Cout=Cout, k=k, dilation=dilation, pad=pad,
        in_act=in_act, in_slope=in_slope, act=act, slope=slope, accum_scale=accum_scale, accum_init=int(accum_init))
    _run_conv(p, 2.0 * B * S * Cout * k * Cin)
    return y


def conv_transpose1d(x, w, bias, *, cout, stride, pad, in_act=ACT_NONE, in_slope=0.0, backend=None, rows=None):
    """ConvTranspose1d in phase form: x [B,S,Cin] f32, w = pack_conv_transpose(weight, stride), bias tiled to
    [stride*cout].  Output [B, T_out, cout] with T_out = (S-1)*stride - 2*pad + taps*stride.
    rows (ragged schedule) counts rows of the phase-form GEMM, i.e. INPUT positions: row s writes output samples
    [stride*s - pad, stride*(s+1) - pad)."""
    _chk(x, torch.float32, "x"); _chk(bias, torch.float32, "bias"); _chk(rows, torch.int32, "rows")
    B, S, Cin = x.shape
    taps = w.k
    k_full = taps * stride
    T_out = (S - 1) * stride - 2 * pad + k_full
    S_out = (T_out - 1 + pad) // stride + 1
    y = torch.empty((B, T_out, cout), dtype=torch.float32, device=x.device)
    be = backend or fp32_backend
    if be.startswith("tf32x3"):
        _tf32_call(x, w, bias, None, None, y, None, None, variant=be[7:] or None, B=B, S_in=S, S_out=S_out,
                   y_batch_stride=T_out * cout,
                   y_row_stride=stride * cout, y_offset=-pad * cout, y_batch_elems=T_out * cout, dilation=1,
                   pad=taps - 1, in_act=in_act, in_slope=in_slope, act=ACT_NONE, slope=0.0, act2=None, slope2=0.0,
                   accum_scale=0.0, accum_init=False, flops=2.0 * B * T_out * cout * taps * Cin, rows=rows)
        return y
    w = w.ffma
    p = _lib.Conv1dF32Params(
        x=x.data_ptr(), x_batch_stride=S * Cin, x_row_stride=Cin, w=w.data_ptr(), bias=bias.data_ptr(), residual=None,
        lens=None, y=y.data_ptr(), accum=None, y_batch_stride=T_out * cout, y_row_stride=stride * cout,
        y_offset=-pad * cout, y_batch_elems=T_out * cout, B=B, S_in=S, S_out=S_out, Cin=Cin, Cout=stride * cout,
        k=taps, dilation=1, pad=taps - 1, in_act=in_act, in_slope=in_slope, act=ACT_NONE, slope=0.0,
        accum_scale=0.0, accum_init=0)
    _run_conv(p, 2.0 * B * T_out * cout * taps * Cin)
    return y


F32, BF16 = 0, 1
_DT = {torch.float32: F32, torch.bfloat16: BF16}


def cast_bf16(x, act=ACT_NONE, slope=0.0):
    """bf16(act(x)) - the entry into the tensor-core path."""
    _chk(x, torch.float32, "x")
    out = torch.empty(x.shape, dtype=torch.bfloat16, device=x.device)
    check(_lib.lib().ss_cast_bf16(_ptr(x), _ptr(out), x.numel(), act, float(slope), _stream()))
    return out


def pack_conv_tc(weight):
    """[Cout, Cin, k] or [out, in] fp32 -> bf16 [k, Cout, Cin] (tap-major, Cin contiguous) for ss_conv1d_tc."""
    return pack_conv(weight).bf16


def bf16_panel_enabled():
    """bf16 mode: ss_conv1d_bf16_panel for the layer shapes it takes (default), or always ss_conv1d_tc
    (environment variable SS_BF16_PANEL=0)."""
    return os.environ.get("SS_BF16_PANEL", "1") != "0"


def _tc_call(p, flops, panel=False):
    f = _lib.lib().ss_conv1d_bf16_panel if panel else _lib.lib().ss_conv1d_tc
    kind = "tcp" if panel else "tc"
    fn = lambda: f(C.byref(p), _stream())  # noqa: E731
    check(conv_timer.run(flops, fn, _tag(kind, p), _meta(kind, p)) if conv_timer is not None else fn())


def conv1d_tc(x, w, bias=None, *, dilation=1, pad=0, act=ACT_NONE, slope=0.0, residual=None, lens=None,
              out_dtype=torch.float32, want_y=True, out=None, act2=None, slope2=0.0, accum=None, accum_scale=1.0,
              accum_init=False, rows=None):
    """Tensor-core convolution.  x [B,S,Cin] bf16, w [k,Cout,Cin] bf16 (pack_conv_tc).
    Returns (y or None, y_act or None); y_act (bf16) is produced when act2 is not None."""
    panel = None  # ss_conv1d_bf16_panel (activations fetched once per tile) where the layer shape allows
    if isinstance(w, ConvW):
        if bf16_panel_enabled() and dilation * (w.k - 1) <= 128:
            panel = w.bf16p
        k, Cout, cin_w = w.k, w.cout, w.cin
        w = panel if panel is not None else w.bf16
    else:
        k, Cout, cin_w = w.shape
    _chk(x, torch.bfloat16, "x"); _chk(w, torch.bfloat16, "w"); _chk(bias, torch.float32, "bias")
    _chk(lens, torch.int32, "lens"); _chk(accum, torch.float32, "accum")
    B, S, Cin = x.shape
    if cin_w != Cin:
        raise _lib.SSError(f"conv1d_tc: weight Cin {cin_w} != input Cin {Cin}")
    y = None
    if want_y:
        y = out if out is not None else torch.empty((B, S, Cout), dtype=out_dtype, device=x.device)
    y_act = torch.empty((B, S, Cout), dtype=torch.bfloat16, device=x.device) if act2 is not None else None
    p = _lib.Conv1dTcParams(
        x=x.data_ptr(), x_batch_stride=S * Cin, x_row_stride=Cin, w=w.data_ptr(),
        bias=None if bias is None else bias.data_ptr(), residual=None if residual is None else residual.data_ptr(),
        lens=None if lens is None else lens.data_ptr(), y=None if y is None else y.data_ptr(),
        y_act=None if y_act is None else y_act.data_ptr(), accum=None if accum is None else accum.data_ptr(),
        y_batch_stride=S * Cout, y_row_stride=Cout, y_offset=0, y_batch_elems=S * Cout, B=B, S_in=S, S_out=S, Cin=Cin,
        Cout=Cout, k=k, dilation=dilation, pad=pad, act=act, slope=slope, act2=act2 or 0, slope2=slope2,
        y_dtype=_DT[y.dtype] if y is not None else F32, residual_dtype=_DT[residual.dtype] if residual is not None else F32,
        accum_scale=accum_scale, accum_init=int(accum_init), rows=None if rows is None else rows.data_ptr())
    _tc_call(p, 2.0 * B * S * Cout * k * Cin, panel is not None)
    return y, y_act


def x3_ok(w):
    """True when ss_conv1d_bf16x3 handles this ConvW's shape."""
    return isinstance(w, ConvW) and w.bf16x3 is not None


def fp32_convs():
    """Which kernels the fp32 mode uses where no integer decision is downstream (HiFi-GAN, FastSpeech2's decoder and
    PostNet): "bf16x3" (split-bf16, default) or "tf32x3" (environment variable SS_FP32_CONVS)."""
    v = os.environ.get("SS_FP32_CONVS", "bf16x3")
    if v not in ("bf16x3", "tf32x3"):
        raise ValueError("SS_FP32_CONVS must be 'bf16x3' or 'tf32x3'")
    return v


def split_bf16(x, act=ACT_NONE, slope=0.0):
    """act(x) [.., C] fp32 -> split-bf16 pair rows [.., 2, C]: hi = bf16(v), lo = bf16(v - hi) (ss_conv1d_bf16x3's input)."""
    _chk(x, torch.float32, "x")
    Cn = x.shape[-1]
    out = torch.empty(x.shape[:-1] + (2, Cn), dtype=torch.bfloat16, device=x.device)
    check(_lib.lib().ss_split_bf16(_ptr(x), _ptr(out), x.numel() // Cn, Cn, act, float(slope), _stream()))
    return out


def conv1d_x3(x, w, bias=None, *, dilation=1, pad=0, act=ACT_NONE, slope=0.0, residual=None, lens=None, want_y=True,
              out=None, act2=None, slope2=0.0, accum=None, accum_scale=1.0, accum_init=False, rows=None, out_pair=None):
    """Split-bf16 convolution (ss_conv1d_bf16x3).  x: pair rows [B,S,2,Cin] bf16 (split_bf16 or a previous y_pair);
    w: ConvW.  Returns (y fp32 or None, y_pair [B,S,2,Cout] bf16 of act2(v) or None)."""
    wt = w.bf16x3
    if wt is None:
        raise _lib.SSError(f"conv1d_x3: layer Cin={w.cin} Cout={w.cout} k={w.k} is not handled by ss_conv1d_bf16x3")
    _chk(x, torch.bfloat16, "x"); _chk(bias, torch.float32, "bias")
    _chk(lens, torch.int32, "lens"); _chk(accum, torch.float32, "accum"); _chk(residual, torch.float32, "residual")
    B, S, two, Cin = x.shape
    if two != 2 or Cin != w.cin:
        raise _lib.SSError(f"conv1d_x3: x must be pair rows [B,S,2,{w.cin}], got {tuple(x.shape)}")
    Cout, k = w.cout, w.k
    y = None
    if want_y:
        y = out if out is not None else torch.empty((B, S, Cout), dtype=torch.float32, device=x.device)
    y_pair = None
    if act2 is not None:
        y_pair = out_pair if out_pair is not None else torch.empty((B, S, 2, Cout), dtype=torch.bfloat16, device=x.device)
    p = _lib.Conv1dTcParams(
        x=x.data_ptr(), x_batch_stride=S * 2 * Cin, x_row_stride=2 * Cin, w=wt.data_ptr(),
        bias=None if bias is None else bias.data_ptr(), residual=None if residual is None else residual.data_ptr(),
        lens=None if lens is None else lens.data_ptr(), y=None if y is None else y.data_ptr(),
        y_act=None if y_pair is None else y_pair.data_ptr(), accum=None if accum is None else accum.data_ptr(),
        y_batch_stride=S * Cout, y_row_stride=Cout, y_offset=0, y_batch_elems=S * Cout, B=B, S_in=S, S_out=S, Cin=Cin,
        Cout=Cout, k=k, dilation=dilation, pad=pad, act=act, slope=slope, act2=act2 or 0, slope2=slope2,
        y_dtype=F32, residual_dtype=F32, accum_scale=accum_scale, accum_init=int(accum_init),
        rows=None if rows is None else rows.data_ptr())
    fn = lambda: _lib.lib().ss_conv1d_bf16x3(C.byref(p), _stream())  # noqa: E731
    flops = 2.0 * B * S * Cout * k * Cin
    check(conv_timer.run(flops, fn, _tag("bf16x3", p), _meta("bf16x3", p)) if conv_timer is not None else fn())
    return y, y_pair


def resunit_ok(w1, w2, dilation):
    """True when ss_resunit_bf16x3 takes this c1 / c2 pair (HiFi-GAN's 32-channel stage)."""
    return (isinstance(w1, ConvW) and isinstance(w2, ConvW) and w1.cin == w1.cout == w2.cin == w2.cout and w1.k == w2.k
            and os.environ.get("SS_RESUNIT", "1") != "0"
            and _lib.lib().ss_resunit_bf16x3_supported(w1.cin, w1.k, dilation) == 1 and w1.bf16x3 is not None)


def resunit_x3(x, residual, w1, b1, w2, b2, *, dilation, slope, want_y=True, out=None, act2=None, slope2=0.0, accum=None,
               accum_scale=1.0, accum_init=False, rows=None, out_pair=None):
    """One ResBlock unit y' = residual + c2(leaky_relu(c1(x))) in one launch (ss_resunit_bf16x3).  x: pair rows
    [B,S,2,C] of leaky_relu(y); residual [B,S,C] f32.  Returns (y' or None, pair rows of act2(y') or None)."""
    _chk(x, torch.bfloat16, "x"); _chk(residual, torch.float32, "residual"); _chk(accum, torch.float32, "accum")
    _chk(b1, torch.float32, "b1"); _chk(b2, torch.float32, "b2")
    B, S, two, Cn = x.shape
    if two != 2 or Cn != w1.cin:
        raise _lib.SSError(f"resunit_x3: x must be pair rows [B,S,2,{w1.cin}], got {tuple(x.shape)}")
    y = None
    if want_y:
        y = out if out is not None else torch.empty((B, S, Cn), dtype=torch.float32, device=x.device)
    y_pair = None
    if act2 is not None:
        y_pair = out_pair if out_pair is not None else torch.empty((B, S, 2, Cn), dtype=torch.bfloat16, device=x.device)
    p = _lib.ResUnitParams(
        x=x.data_ptr(), residual=None if residual is None else residual.data_ptr(), w1=w1.bf16x3.data_ptr(),
        b1=b1.data_ptr(), w2=w2.bf16x3.data_ptr(), b2=b2.data_ptr(), y=None if y is None else y.data_ptr(),
        y_pair=None if y_pair is None else y_pair.data_ptr(), accum=None if accum is None else accum.data_ptr(),
        rows=None if rows is None else rows.data_ptr(), B=B, S=S, C=Cn, k=w1.k, dilation=dilation, slope=slope,
        act2=act2 or 0, slope2=slope2, accum_scale=accum_scale, accum_init=int(accum_init))
    fn = lambda: _lib.lib().ss_resunit_bf16x3(C.byref(p), _stream())  # noqa: E731
    flops = 2.0 * 2.0 * B * S * Cn * w1.k * Cn  # both convolutions
    meta = ("resunit", B, S, Cn, Cn, w1.k, dilation)
    tag = f"resunit B={B} S={S} Cin={Cn} Cout={Cn} k={w1.k} dil={dilation}"
    check(conv_timer.run(flops, fn, tag, meta) if conv_timer is not None else fn())
    return y, y_pair


def conv_transpose1d_x3(x, w, bias, *, cout, stride, pad, act2=None, slope2=0.0, want_y=True, rows=None):
    """ConvTranspose1d in phase form on the split-bf16 kernel.  x: pair rows [B,S,2,Cin] bf16; w: ConvW from
    pack_conv_transpose; bias tiled to [stride*cout].  Returns (y f32 [B,T_out,cout] or None, pair rows
    [B,T_out,2,cout] of act2(y) or None)."""
    wt = w.bf16x3
    if wt is None:
        raise _lib.SSError(f"conv_transpose1d_x3: layer Cin={w.cin} Cout={w.cout} is not handled by ss_conv1d_bf16x3")
    _chk(x, torch.bfloat16, "x"); _chk(bias, torch.float32, "bias")
    B, S, two, Cin = x.shape
    if two != 2 or Cin != w.cin or w.cout != stride * cout:
        raise _lib.SSError(f"conv_transpose1d_x3: x {tuple(x.shape)} / weight [{w.k},{w.cout},{w.cin}] mismatch")
    taps = w.k
    T_out = (S - 1) * stride - 2 * pad + taps * stride
    S_out = (T_out - 1 + pad) // stride + 1
    y = torch.empty((B, T_out, cout), dtype=torch.float32, device=x.device) if want_y else None
    y_pair = torch.empty((B, T_out, 2, cout), dtype=torch.bfloat16, device=x.device) if act2 is not None else None
    p = _lib.Conv1dTcParams(
        x=x.data_ptr(), x_batch_stride=S * 2 * Cin, x_row_stride=2 * Cin, w=wt.data_ptr(), bias=bias.data_ptr(),
        residual=None, lens=None, y=None if y is None else y.data_ptr(),
        y_act=None if y_pair is None else y_pair.data_ptr(), accum=None, y_batch_stride=T_out * cout,
        y_row_stride=stride * cout, y_offset=-pad * cout, y_batch_elems=T_out * cout, B=B, S_in=S, S_out=S_out, Cin=Cin,
        Cout=stride * cout, k=taps, dilation=1, pad=taps - 1, act=ACT_NONE, slope=0.0, act2=act2 or 0, slope2=slope2,
        y_dtype=F32, residual_dtype=F32, accum_scale=0.0, accum_init=0, rows=None if rows is None else rows.data_ptr(),
        pair_c=cout)
    fn = lambda: _lib.lib().ss_conv1d_bf16x3(C.byref(p), _stream())  # noqa: E731
    flops = 2.0 * B * T_out * cout * taps * Cin
    check(conv_timer.run(flops, fn, _tag("bf16x3", p), _meta("bf16x3", p)) if conv_timer is not None else fn())
    return y, y_pair


def conv_transpose1d_tc(x, w, bias, *, cout, stride, pad, act2=None, slope2=0.0, want_y=True, rows=None):
    """ConvTranspose1d in phase form on tensor cores.  x [B,S,Cin] bf16; w [taps, stride*cout, Cin] bf16;
    bias tiled to [stride*cout] f32.  Returns (y f32 [B,T_out,cout] or None, y_act bf16 or None)."""
    panel = None
    if isinstance(w, ConvW):
        if bf16_panel_enabled():
            panel = w.bf16p
        taps = w.k
        w = panel if panel is not None else w.bf16
    else:
        taps = w.shape[0]
    _chk(x, torch.bfloat16, "x"); _chk(w, torch.bfloat16, "w"); _chk(bias, torch.float32, "bias")
    B, S, Cin = x.shape
    T_out = (S - 1) * stride - 2 * pad + taps * stride
    S_out = (T_out - 1 + pad) // stride + 1
    y = torch.empty((B, T_out, cout), dtype=torch.float32, device=x.device) if want_y else None
    y_act = torch.empty((B, T_out, cout), dtype=torch.bfloat16, device=x.device) if act2 is not None else None
    p = _lib.Conv1dTcParams(
        x=x.data_ptr(), x_batch_stride=S * Cin, x_row_stride=Cin, w=w.data_ptr(), bias=bias.data_ptr(), residual=None,
        lens=None, y=None if y is None else y.data_ptr(), y_act=None if y_act is None else y_act.data_ptr(), accum=None,
        y_batch_stride=T_out * cout, y_row_stride=stride * cout, y_offset=-pad * cout, y_batch_elems=T_out * cout, B=B,
        S_in=S, S_out=S_out, Cin=Cin, Cout=stride * cout, k=taps, dilation=1, pad=taps - 1, act=ACT_NONE, slope=0.0,
        act2=act2 or 0, slope2=slope2, y_dtype=F32, residual_dtype=F32, accum_scale=0.0, accum_init=0,
        rows=None if rows is None else rows.data_ptr())
    _tc_call(p, 2.0 * B * T_out * cout * taps * Cin, panel is not None)
    return y, y_act


def layernorm(x, ln_w, ln_b, *, residual=None, film=None, lens=None, pe=None, pre_relu=False, out=None,
              want_bf16=False, want_pair=False):
    """film = (gamma [B,C], beta [B,C], s_gamma [1], s_beta [1]) or None.  With want_bf16 returns (y, y_bf16); with
    want_pair (y, split-bf16 pair rows [B,S,2,C])."""
    _chk(x, torch.float32, "x"); _chk(residual, torch.float32, "residual"); _chk(lens, torch.int32, "lens")
    _chk(pe, torch.float32, "pe"); _chk(ln_w, torch.float32, "ln_w"); _chk(ln_b, torch.float32, "ln_b")
    B, S, Cc = x.shape
    if pe is not None and (pe.shape[0] < S or pe.shape[1] != Cc):
        raise _lib.SSError("layernorm: position table too small")
    y = out if out is not None else torch.empty_like(x)
    yb = torch.empty(x.shape, dtype=torch.bfloat16, device=x.device) if want_bf16 else None
    if want_pair:
        yb = torch.empty((B, S, 2, Cc), dtype=torch.bfloat16, device=x.device)
    g = be = sg = sb = None
    if film is not None:
        g, be, sg, sb = film
        for t, n in ((g, "gamma"), (be, "beta"), (sg, "s_gamma"), (sb, "s_beta")):
            _chk(t, torch.float32, n)
    p = _lib.LayerNormParams(
        x=x.data_ptr(), residual=None if residual is None else residual.data_ptr(), ln_w=ln_w.data_ptr(),
        ln_b=ln_b.data_ptr(), gamma=None if g is None else g.data_ptr(), beta=None if be is None else be.data_ptr(),
        s_gamma=None if sg is None else sg.data_ptr(), s_beta=None if sb is None else sb.data_ptr(),
        lens=None if lens is None else lens.data_ptr(), pe=None if pe is None else pe.data_ptr(), y=y.data_ptr(),
        y_bf16=None if yb is None else yb.data_ptr(), B=B, S=S, C=Cc, pre_relu=int(pre_relu), bf16_pair=int(want_pair))
    check(_lib.lib().ss_layernorm(C.byref(p), _stream()))
    return (y, yb) if (want_bf16 or want_pair) else y


def attention(qkv, lens, n_head, backend=None):
    """Fused masked-softmax attention.  backend "tc" = ss_attention_tc (tcgen05, head sizes 32 / 128), "simt" =
    ss_attention_f32; default: tensor cores whenever the head size allows and ops.fp32_backend is "tf32x3"."""
    _chk(qkv, torch.float32, "qkv"); _chk(lens, torch.int32, "lens")
    B, S, D3 = qkv.shape
    D = D3 // 3
    dk = D // n_head
    ctx = torch.empty((B, S, D), dtype=torch.float32, device=qkv.device)
    if backend is None:
        backend = "tc" if (dk in (32, 128) and fp32_backend == "tf32x3") else "simt"
    fn = _lib.lib().ss_attention_tc if backend == "tc" else _lib.lib().ss_attention_f32
    check(fn(_ptr(qkv), _ptr(lens), _ptr(ctx), B, S, n_head, dk, _stream()))
    return ctx


def predictor_head(h, ln_w, ln_b, lin_w, lin_b, *, film=None, lens=None):
    _chk(h, torch.float32, "h"); _chk(lens, torch.int32, "lens")
    B, L, Cc = h.shape
    pred = torch.empty((B, L), dtype=torch.float32, device=h.device)
    g = be = sg = sb = None
    if film is not None:
        g, be, sg, sb = film
    p = _lib.PredictorHeadParams(
        h=h.data_ptr(), ln_w=ln_w.data_ptr(), ln_b=ln_b.data_ptr(), gamma=None if g is None else g.data_ptr(),
        beta=None if be is None else be.data_ptr(), s_gamma=None if sg is None else sg.data_ptr(),
        s_beta=None if sb is None else sb.data_ptr(), lin_w=lin_w.data_ptr(), lin_b=lin_b.data_ptr(),
        lens=None if lens is None else lens.data_ptr(), pred=pred.data_ptr(), B=B, L=L, C=Cc)
    check(_lib.lib().ss_predictor_head(C.byref(p), _stream()))
    return pred


def _control_vec(control, shape, device):
    """A control given as a tensor / array (one value per token, notebooks/control.ipynb) -> fp32 [B, L] on the device;
    None for a plain number."""
    if isinstance(control, (int, float)):
        return None
    c = torch.as_tensor(control)
    if c.numel() == 1:
        return None
    c = c.to(device, torch.float32)
    if tuple(c.shape) != tuple(shape):
        c = c.expand(shape)  # [L] or [B, 1] broadcast as the reference's `prediction * control` would
    return c.contiguous()


def _control_scalar(control):
    return float(control) if isinstance(control, (int, float)) else float(torch.as_tensor(control).reshape(-1)[0])


def bucket_embed_add(pred, control, edges, table, x, want_idx=False):
    """In place: pred *= control; x += table[bucketize(pred, edges)].  Returns idx (int32) if asked.
    control: a number, or a tensor with one value per token ([B, L], model/modules.py:88,99)."""
    _chk(pred, torch.float32, "pred"); _chk(edges, torch.float32, "edges"); _chk(table, torch.float32, "table")
    _chk(x, torch.float32, "x")
    B, L, Cc = x.shape
    idx = torch.empty((B, L), dtype=torch.int32, device=x.device) if want_idx else None
    cv = _control_vec(control, (B, L), x.device)
    if cv is not None:
        check(_lib.lib().ss_bucket_embed_add_v(_ptr(pred), _ptr(cv), _ptr(edges), edges.numel(), _ptr(table), _ptr(x),
                                               _ptr(idx), B, L, Cc, _stream()))
    else:
        check(_lib.lib().ss_bucket_embed_add(_ptr(pred), _control_scalar(control), _ptr(edges), edges.numel(), _ptr(table),
                                             _ptr(x), _ptr(idx), B, L, Cc, _stream()))
    return idx


def duration_scan(log_d, d_control, durations=None):
    """Returns (d_rounded f32 [B,L], ends i32 [B,L], lens2 i32 [2,B] = (expanded length, int(float sum)))."""
    src = log_d if log_d is not None else durations
    _chk(src, torch.float32, "durations")
    B, L = src.shape
    d_rounded = torch.empty_like(src) if log_d is not None else durations
    ends = torch.empty((B, L), dtype=torch.int32, device=src.device)
    lens2 = torch.empty((2, B), dtype=torch.int32, device=src.device)
    cv = _control_vec(d_control, (B, L), src.device) if log_d is not None else None
    if cv is not None:
        check(_lib.lib().ss_duration_scan_v(_ptr(log_d), _ptr(cv), _ptr(d_rounded), _ptr(ends), _ptr(lens2[0]),
                                            _ptr(lens2[1]), B, L, _stream()))
    else:
        check(_lib.lib().ss_duration_scan(_ptr(log_d), _control_scalar(d_control), _ptr(d_rounded), _ptr(ends),
                                          _ptr(lens2[0]), _ptr(lens2[1]), B, L, _stream()))
    return d_rounded, ends, lens2


def length_regulate(x, ends, T, pe=None, want_mel2phone=False):
    _chk(x, torch.float32, "x"); _chk(ends, torch.int32, "ends"); _chk(pe, torch.float32, "pe")
    B, L, Cc = x.shape
    if pe is not None and (pe.shape[0] < T or pe.shape[1] != Cc):
        raise _lib.SSError("length_regulate: position table too small")
    out = torch.empty((B, T, Cc), dtype=torch.float32, device=x.device)
    m2p = torch.empty((B, T), dtype=torch.int32, device=x.device) if want_mel2phone else None
    check(_lib.lib().ss_length_regulate(_ptr(x), _ptr(ends), _ptr(pe), _ptr(out), _ptr(m2p), B, L, int(T), Cc,
                                        _stream()))
    return (out, m2p) if want_mel2phone else out


def length_mask(lens, T):
    _chk(lens, torch.int32, "lens")
    B = lens.shape[0]
    mask = torch.empty((B, int(T)), dtype=torch.bool, device=lens.device)
    check(_lib.lib().ss_length_mask(_ptr(lens), _ptr(mask), B, int(T), _stream()))
    return mask


def mean_pool(x, lens=None):
    """Mean over the padded length (the reference, modules.py:398); with lens: over each utterance's own frames."""
    _chk(x, torch.float32, "x"); _chk(lens, torch.int32, "lens")
    B, S, Cc = x.shape
    out = torch.empty((B, Cc), dtype=torch.float32, device=x.device)
    if lens is None:
        check(_lib.lib().ss_mean_pool(_ptr(x), _ptr(out), B, S, Cc, _stream()))
    else:
        check(_lib.lib().ss_mean_pool_lens(_ptr(x), _ptr(lens), _ptr(out), B, S, Cc, _stream()))
    return out


def mask_rows_(x, lens):
    """x[b, t, :] = 0 for t >= lens[b], in place (batch-invariant mode)."""
    _chk(x, torch.float32, "x"); _chk(lens, torch.int32, "lens")
    B, S, Cc = x.shape
    check(_lib.lib().ss_mask_rows(_ptr(x), _ptr(lens), B, S, Cc, _stream()))
    return x


def add_speaker(x, ids, table):
    _chk(x, torch.float32, "x"); _chk(ids, torch.int64, "speakers"); _chk(table, torch.float32, "speaker table")
    B, L, Cc = x.shape
    check(_lib.lib().ss_add_speaker(_ptr(x), _ptr(ids), _ptr(table), B, L, Cc, table.shape[0], _stream()))
    return x


def hifigan_post(x, w, bias, *, k, in_slope=0.01, want_wave=True, want_pcm=False, pcm_scale=32768.0):
    _chk(x, torch.float32, "x"); _chk(w, torch.float32, "w")
    B, T, Cc = x.shape
    wave = torch.empty((B, T), dtype=torch.float32, device=x.device) if want_wave else None
    pcm = torch.empty((B, T), dtype=torch.int16, device=x.device) if want_pcm else None
    check(_lib.lib().ss_hifigan_post(_ptr(x), _ptr(w), _ptr(bias), _ptr(wave), _ptr(pcm), float(pcm_scale),
                                     float(in_slope), B, T, Cc, k, _stream()))
    return wave, pcm


def hifigan_post_trim(x, w, bias, n_samples, out_offset, pcm_out, *, k, in_slope=0.01, pcm_scale=32768.0):
    """Ragged PCM epilogue (ss_hifigan_post_trim): utterance b's first n_samples[b] int16 samples go to
    pcm_out[out_offset[b]:].  n_samples [B] int32 / out_offset [B] int64 on the device; pcm_out is a flat int16
    tensor on the device or in page-locked host memory (written by the kernel over PCIe, no separate copy)."""
    _chk(x, torch.float32, "x"); _chk(w, torch.float32, "w")
    _chk(n_samples, torch.int32, "n_samples"); _chk(out_offset, torch.int64, "out_offset")
    if pcm_out.dtype != torch.int16 or not pcm_out.is_contiguous() or not (pcm_out.is_cuda or pcm_out.is_pinned()):
        raise _lib.SSError("hifigan_post_trim: pcm_out must be a contiguous int16 tensor on the device or in pinned host memory")
    B, T, Cc = x.shape
    check(_lib.lib().ss_hifigan_post_trim(_ptr(x), _ptr(w), _ptr(bias), _ptr(n_samples), _ptr(out_offset),
                                          pcm_out.data_ptr(), float(pcm_scale), float(in_slope), B, T, Cc, k, _stream()))
    return pcm_out


def reflect_pad(y, half, n_out):
    """ss_reflect_pad: y [B,T] f32 -> [B,n_out], out[b,i] = y[b, reflect(i - half)] (F.pad mode="reflect")."""
    _chk(y, torch.float32, "y")
    B, T = y.shape
    out = torch.empty((B, n_out), dtype=torch.float32, device=y.device)
    check(_lib.lib().ss_reflect_pad(_ptr(y), _ptr(out), B, T, int(half), int(n_out), _stream()))
    return out


def stft_magnitude(spec, n_bins, want_energy=True):
    """ss_stft_magnitude: spec [B,F,>=2*n_bins] (re | im | pad) -> (mag [B,F,ceil4(n_bins)] zero-padded, energy [B,F])."""
    _chk(spec, torch.float32, "spec")
    B, F, W = spec.shape
    ms = (n_bins + 3) // 4 * 4
    mag = torch.empty((B, F, ms), dtype=torch.float32, device=spec.device)
    energy = torch.empty((B, F), dtype=torch.float32, device=spec.device) if want_energy else None
    check(_lib.lib().ss_stft_magnitude(_ptr(spec), _ptr(mag), _ptr(energy), B * F, int(n_bins), W, ms, _stream()))
    return mag, energy


def out_of_range(y, lo=-1.0, hi=1.0):
    """ss_range_flag: True if any element of y is outside [lo, hi] or NaN (one pass, one 4-byte read-back)."""
    _chk(y, torch.float32, "y")
    flag = torch.zeros(1, dtype=torch.int32, device=y.device)
    check(_lib.lib().ss_range_flag(_ptr(y), y.numel(), float(lo), float(hi), _ptr(flag), _stream()))
    return bool(flag.item())


def log_clamp_(x, clip=1e-5):
    """ss_log_clamp: x = log(max(x, clip)) in place."""
    _chk(x, torch.float32, "x")
    check(_lib.lib().ss_log_clamp(_ptr(x), x.numel(), float(clip), _stream()))
    return x
