"""Drop-in for hifigan/models.py ``Generator`` (HiFi-GAN V1) over the C ABI.  Inference only.

``Generator(h).forward(mel [B,80,T]) -> wave [B,1,256*T]`` with the reference's attribute names
(conv_pre, ups, resblocks, conv_post), ``remove_weight_norm()`` and a ``state_dict`` that accepts
either checkpoint form (``weight_g``/``weight_v``, utils/model.py:81-92) or folded ``weight`` keys.

Kernel mapping:
  conv_pre                  -> ss_conv1d (k=7)
  leaky_relu + ConvTranspose -> ss_conv1d in phase form: k/stride taps, stride*C_out GEMM columns
                               whose rows interleave into the up-sampled sequence; lrelu on load
  ResBlock conv pairs       -> ss_conv1d with lrelu on load; c1's epilogue needs nothing else, c2's
                               epilogue adds the residual; the last c2 of each ResBlock also
                               accumulates (1/3) into the stage mean (xs / num_kernels)
  leaky_relu(0.01) + conv_post + tanh -> ss_hifigan_post (optionally straight to int16 PCM)
"""
import torch
import torch.nn as nn

from . import ops
from .configs import AttrDict  # noqa: F401  (same helper as hifigan/__init__.py)
from .transformer import ParamCache, pack_conv

LRELU_SLOPE = 0.1


def get_padding(kernel_size, dilation=1):
    return int((kernel_size * dilation - dilation) / 2)


class WNConv(nn.Module):
    """Parameter holder for weight_norm(Conv1d) / weight_norm(ConvTranspose1d): holds
    (weight_g, weight_v, bias) until remove_weight_norm(), then (weight, bias)."""

    def __init__(self, shape, out_channels, std=0.01):
        super().__init__()
        v = torch.randn(shape) * std
        self.weight_g = nn.Parameter(v.flatten(1).norm(dim=1).view(-1, 1, 1), requires_grad=False)
        self.weight_v = nn.Parameter(v, requires_grad=False)
        self.bias = nn.Parameter(torch.zeros(out_channels), requires_grad=False)

    @property
    def normalized(self):
        return "weight_g" in self._parameters

    def folded_weight(self):
        if not self.normalized:
            return self.weight
        v = self.weight_v
        return v * (self.weight_g / v.flatten(1).norm(dim=1).view(-1, 1, 1))

    def remove_weight_norm(self):
        if self.normalized:
            w = self.folded_weight().detach().clone()
            del self._parameters["weight_g"], self._parameters["weight_v"]
            self.weight = nn.Parameter(w, requires_grad=False)
            ParamCache.epoch += 1  # Parameter objects replaced: cached source lists are stale

    def _load_from_state_dict(self, state_dict, prefix, *args, **kwargs):
        has_plain = prefix + "weight" in state_dict
        has_wn = prefix + "weight_v" in state_dict
        if self.normalized and has_plain and not has_wn:
            # folded checkpoint into a weight-normed module: g = ||w||, v = w reproduces w exactly
            w = state_dict.pop(prefix + "weight")
            state_dict[prefix + "weight_v"] = w
            state_dict[prefix + "weight_g"] = w.flatten(1).norm(dim=1).view(-1, 1, 1)
        elif not self.normalized and has_wn and not has_plain:
            v = state_dict.pop(prefix + "weight_v")
            g = state_dict.pop(prefix + "weight_g")
            state_dict[prefix + "weight"] = v * (g / v.flatten(1).norm(dim=1).view(-1, 1, 1))
        return super()._load_from_state_dict(state_dict, prefix, *args, **kwargs)


class ResBlock(nn.Module):
    """hifigan/models.py:20-109."""

    def __init__(self, h, channels, kernel_size=3, dilation=(1, 3, 5)):
        super().__init__()
        self.h = h
        self.channels, self.kernel_size, self.dilation = channels, kernel_size, tuple(dilation)
        self.convs1 = nn.ModuleList([WNConv((channels, channels, kernel_size), channels) for _ in dilation])
        self.convs2 = nn.ModuleList([WNConv((channels, channels, kernel_size), channels) for _ in dilation])

    def remove_weight_norm(self):
        for m in list(self.convs1) + list(self.convs2):
            m.remove_weight_norm()


class Generator(nn.Module, ParamCache):
    """hifigan/models.py:112-174."""

    def __init__(self, h, precision="fp32"):
        super().__init__()
        self.h = h
        if precision not in ("fp32", "bf16"):
            raise ValueError("precision must be 'fp32' or 'bf16'")
        self.precision = precision  # "bf16": tcgen05 convolutions (wave relative L2 <= 1e-2)
        # fp32 mode: "bf16x3" = split-bf16 pairs, three bf16 tensor-core products per multiply (ss_conv1d_bf16x3; wave
        # within ~2e-5 of the reference), "tf32x3" = the 3xTF32 kernels (environment variable SS_FP32_CONVS)
        self.fp32_convs = ops.fp32_convs()
        self.num_kernels = len(h["resblock_kernel_sizes"])
        self.num_upsamples = len(h["upsample_rates"])
        c0 = h["upsample_initial_channel"]
        self.conv_pre = WNConv((c0, h.get("num_mels", 80), 7), c0)
        self.ups = nn.ModuleList()
        for i, (u, k) in enumerate(zip(h["upsample_rates"], h["upsample_kernel_sizes"])):
            if k % u != 0:
                raise NotImplementedError("ConvTranspose1d phase form needs kernel_size % stride == 0")
            self.ups.append(WNConv((c0 // (2 ** i), c0 // (2 ** (i + 1)), k), c0 // (2 ** (i + 1))))
        self.resblocks = nn.ModuleList()
        ch = c0
        for i in range(len(self.ups)):
            ch = c0 // (2 ** (i + 1))
            for k, d in zip(h["resblock_kernel_sizes"], h["resblock_dilation_sizes"]):
                self.resblocks.append(ResBlock(h, ch, k, d))
        self.conv_post = WNConv((1, ch, 7), 1)
        self.eval()

    def remove_weight_norm(self):
        for m in self.ups:
            m.remove_weight_norm()
        for m in self.resblocks:
            m.remove_weight_norm()
        self.conv_pre.remove_weight_norm()
        self.conv_post.remove_weight_norm()

    def train(self, mode=True):
        if mode:
            raise NotImplementedError("speakingstyle_b200 Generator is inference-only")
        return super().train(False)

    # ---------------------------------------------------------------------------------------
    def _pack(self):
        h = self.h
        pk = {"pre": pack_conv(self.conv_pre.folded_weight()), "ups": [], "res": []}
        for i, u in enumerate(h["upsample_rates"]):
            pk["ups"].append((ops.pack_conv_transpose(self.ups[i].folded_weight(), u),
                              self.ups[i].bias.repeat(u).contiguous()))
        for rb in self.resblocks:
            pk["res"].append(([pack_conv(c.folded_weight()) for c in rb.convs1],
                              [pack_conv(c.folded_weight()) for c in rb.convs2]))
        wp = self.conv_post.folded_weight()  # [1, C, k] -> [k*C]
        pk["post"] = wp[0].t().contiguous().reshape(-1)
        # every ResBlock layer must be a shape ss_conv1d_bf16x3 handles, or the whole generator stays on 3xTF32
        pk["x3"] = all(w.bf16x3 is not None for w1s, w2s in pk["res"] for w in w1s + w2s)
        return pk

    # ---------------------------------------------------------------------------------------
    def _halo_by_rate(self):
        """{rows per frame: mel frames} - how far to the RIGHT of an utterance's last frame the layers running at that rate
        must still be computed so that every sample of the utterance is exact: the sum of the right-hand halos of all
        layers DOWNSTREAM of them (conv: dilation*(k-1)/2 rows; transposed conv: one input row), each divided by that
        layer's rows per frame - counting, for the layers of a stage, the whole stage (its first convolution has the
        other eleven behind it) - rounded up, plus one.  utils/model.py:109-113 trims wave b to mel_len[b]*hop, so
        nothing further right is ever observed.  Deep stages need little: 2-3 frames at 256 / 128 / 64 rows per
        frame, 11 at 8 rows per frame, 15 for conv_pre."""
        h = self.h
        rates = [1]
        for u in h["upsample_rates"]:
            rates.append(rates[-1] * u)
        need = 3.0 / rates[-1]  # conv_post, k = 7
        halo = {}
        for i in reversed(range(len(h["upsample_rates"]))):
            rb = max((k - 1) // 2 * (sum(d) + len(d)) for k, d in zip(h["resblock_kernel_sizes"], h["resblock_dilation_sizes"]))
            need += rb / rates[i + 1]
            halo[rates[i + 1]] = int(need) + 2
            need += 1.0 / rates[i]  # ConvTranspose1d: output t reads input rows <= (t + pad) / stride
        halo[1] = int(need) + 2  # conv_pre's output
        return halo

    def halo_frames(self):
        """The largest of the per-rate halos (that of conv_pre's output): mel frames past an utterance's end that can
        still reach one of its samples."""
        return max(self._halo_by_rate().values())

    def _ragged_rows(self, frame_lens, T, dev):
        """Per-layer row limits of the ragged schedule for utterances of ``frame_lens`` valid mel frames in a batch padded
        to T: {rate: int32 [B]} for the layers running at ``rate`` rows per frame, {("up", i): int32 [B]} for the
        phase-form GEMM of ups[i] (rows = input positions).  One small host -> device copy."""
        halo = self._halo_by_rate()
        rates = [1]
        for u in self.h["upsample_rates"]:
            rates.append(rates[-1] * u)
        keep = {r: [min(T, max(0, int(n)) + halo[r]) for n in frame_lens] for r in rates}
        # phase-form GEMM of ups[i]: row s writes output samples [u*s - pad, u*(s+1) - pad), so the samples of frame
        # f-1 of its OUTPUT stage need rows <= f*rate (the kernel clamps to its own row count)
        table = [[f * r for f in keep[r]] for r in rates] + \
                [[f * rates[i] + 1 for f in keep[rates[i + 1]]] for i in range(len(rates) - 1)]
        t = torch.tensor(table, dtype=torch.int32).pin_memory().to(dev, non_blocking=True)
        rows = {r: t[j] for j, r in enumerate(rates)}
        rows.update({("up", i): t[len(rates) + i] for i in range(len(rates) - 1)})
        return rows

    def _post(self, xs, pk, pcm, pcm_out, pcm_scale=32768.0):
        """Output head.  pcm_out = (n_samples [B] int32 dev, out_offset [B] int64 dev, flat int16 buffer): ragged PCM
        straight into that buffer (ss_hifigan_post_trim) instead of the padded float wave."""
        if pcm_out is not None:
            return ops.hifigan_post_trim(xs, pk["post"], self.conv_post.bias, *pcm_out, k=7, in_slope=0.01,
                                         pcm_scale=pcm_scale)
        wave, pcm16 = ops.hifigan_post(xs, pk["post"], self.conv_post.bias, k=7, in_slope=0.01, want_wave=True,
                                       want_pcm=pcm, pcm_scale=pcm_scale)
        wave = wave.unsqueeze(1)
        return (wave, pcm16) if pcm else wave

    def _forward_bf16(self, x, pcm, pcm_out=None, rows=None, pcm_scale=32768.0):
        """Tensor-core path.  Convolution operands are bf16; the residual stream of every ResBlock and
        the stage mean stay fp32.  Each epilogue also emits the leaky-ReLU'd bf16 copy that the next
        convolution consumes (the activation cannot sit between TMA and the MMA)."""
        h, pk, nk = self.h, self.packed(), self.num_kernels
        L = ops.ACT_LRELU
        rate = 1
        xa = ops.conv1d_tc(ops.cast_bf16(x), pk["pre"], self.conv_pre.bias, pad=3, want_y=False, act2=L,
                           slope2=LRELU_SLOPE, rows=rows and rows[rate])[1]
        n_up = self.num_upsamples
        xs = None
        for i, (u, k) in enumerate(zip(h["upsample_rates"], h["upsample_kernel_sizes"])):
            x32, xa = ops.conv_transpose1d_tc(xa, pk["ups"][i][0], pk["ups"][i][1], cout=self.ups[i].bias.numel(),
                                              stride=u, pad=(k - u) // 2, act2=L, slope2=LRELU_SLOPE,
                                              rows=rows and rows[("up", i)])
            rate *= u
            rw = rows and rows[rate]
            xs, ybuf = torch.empty_like(x32), torch.empty_like(x32)
            for j in range(nk):
                rb = self.resblocks[i * nk + j]
                w1s, w2s = pk["res"][i * nk + j]
                y32, ya = x32, xa
                nd, rk = len(rb.dilation), rb.kernel_size
                for m, dil in enumerate(rb.dilation):
                    ta = ops.conv1d_tc(ya, w1s[m], rb.convs1[m].bias, dilation=dil, pad=get_padding(rk, dil),
                                       want_y=False, act2=L, slope2=LRELU_SLOPE, rows=rw)[1]
                    if m < nd - 1:
                        y32, ya = ops.conv1d_tc(ta, w2s[m], rb.convs2[m].bias, pad=get_padding(rk, 1), residual=y32,
                                                out=ybuf, act2=L, slope2=LRELU_SLOPE, rows=rw)
                    else:
                        last_rb = j == nk - 1
                        feeds_ups = last_rb and i < n_up - 1  # the next ConvTranspose reads lrelu(mean)
                        nxt = ops.conv1d_tc(ta, w2s[m], rb.convs2[m].bias, pad=get_padding(rk, 1), residual=y32,
                                            want_y=False, accum=xs, accum_scale=1.0 / nk, accum_init=(j == 0),
                                            act2=L if feeds_ups else None, slope2=LRELU_SLOPE, rows=rw)[1]
            xa = nxt
        return self._post(xs, pk, pcm, pcm_out, pcm_scale)

    def _forward_x3(self, x, pcm, pcm_out=None, rows=None, pcm_scale=32768.0):
        """fp32 mode on the split-bf16 kernel (ss_conv1d_bf16x3): transposed convolutions (phase form) and ResBlocks.
        The residual stream y and the stage mean xs are fp32; every convolution reads its input as split-bf16 PAIR
        rows of leaky_relu(.) written by the epilogue that produced it (c1 -> c2: hifigan/models.py:98-101; c2 -> next
        c1: the residual sum; last c2 of a stage -> next transposed convolution: the stage mean).  conv_pre (80 input
        channels) stays on the 3xTF32 kernel, and so does any transposed convolution whose shape the split-bf16
        kernel does not take.  Dataflow as in _forward_bf16."""
        h, pk, nk = self.h, self.packed(), self.num_kernels
        L = ops.ACT_LRELU
        rate = 1
        x = ops.conv1d(x, pk["pre"], self.conv_pre.bias, k=7, pad=3, rows=rows and rows[rate])
        n_up = self.num_upsamples
        xa = None  # pair rows of leaky_relu(x) when the next transposed convolution runs on the split-bf16 kernel
        for i, (u, k) in enumerate(zip(h["upsample_rates"], h["upsample_kernel_sizes"])):
            wu, bu = pk["ups"][i]
            if wu.bf16x3 is not None:
                if xa is None:
                    xa = ops.split_bf16(x, act=L, slope=LRELU_SLOPE)
                x, xa = ops.conv_transpose1d_x3(xa, wu, bu, cout=self.ups[i].bias.numel(), stride=u, pad=(k - u) // 2,
                                                act2=L, slope2=LRELU_SLOPE, rows=rows and rows[("up", i)])
            else:
                x = ops.conv_transpose1d(x, wu, bu, cout=self.ups[i].bias.numel(), stride=u, pad=(k - u) // 2,
                                         in_act=L, in_slope=LRELU_SLOPE, rows=rows and rows[("up", i)])
                xa = ops.split_bf16(x, act=L, slope=LRELU_SLOPE)
            # x: the stage input (fp32 residual stream), xa: its leaky-ReLU'd pair, read by the first c1 of all ResBlocks
            rate *= u
            rw = rows and rows[rate]
            xs, ybuf = torch.empty_like(x), torch.empty_like(x)
            tbuf, yabufs = torch.empty_like(xa), (torch.empty_like(xa), torch.empty_like(xa))
            nxt = None
            for j in range(nk):
                rb = self.resblocks[i * nk + j]
                w1s, w2s = pk["res"][i * nk + j]
                y, ya = x, xa
                nd, rk = len(rb.dilation), rb.kernel_size
                for m, dil in enumerate(rb.dilation):
                    last = m == nd - 1
                    # the last ResBlock's last epilogue also emits leaky_relu(mean) as the pair the next transposed
                    # convolution reads (hifigan/models.py:152-153), when that one runs on the split-bf16 kernel
                    feeds = last and j == nk - 1 and i < n_up - 1 and pk["ups"][i + 1][0].bf16x3 is not None
                    tail = dict(act2=L if (not last or feeds) else None, slope2=LRELU_SLOPE, rows=rw,
                                out_pair=yabufs[m & 1] if (not last or feeds) else None)  # never the pair being read
                    if last:  # x_j = c2(..) + y goes straight into the stage mean
                        tail.update(want_y=False, accum=xs, accum_scale=1.0 / nk, accum_init=(j == 0))
                    else:
                        tail.update(out=ybuf)
                    if ops.resunit_ok(w1s[m], w2s[m], dil):  # 32-channel stage: c1 -> c2 in one launch
                        y_new, ya_new = ops.resunit_x3(ya, y, w1s[m], rb.convs1[m].bias, w2s[m], rb.convs2[m].bias,
                                                       dilation=dil, slope=LRELU_SLOPE, **tail)
                    else:
                        ta = ops.conv1d_x3(ya, w1s[m], rb.convs1[m].bias, dilation=dil, pad=get_padding(rk, dil),
                                           want_y=False, act2=L, slope2=LRELU_SLOPE, rows=rw, out_pair=tbuf)[1]
                        y_new, ya_new = ops.conv1d_x3(ta, w2s[m], rb.convs2[m].bias, pad=get_padding(rk, 1), residual=y,
                                                      **tail)
                    if last:
                        nxt = ya_new
                    else:
                        y, ya = y_new, ya_new
            x, xa = xs, nxt
        return self._post(x, pk, pcm, pcm_out, pcm_scale)

    @torch.no_grad()
    def forward(self, x, channels_last=False, pcm=False, pcm_out=None, frame_lens=None, pcm_scale=32768.0):
        """x: mel [B,80,T] (reference layout) or [B,T,80] with channels_last=True.
        Returns wave [B,1,256*T] float32; with pcm=True returns (wave, int16 [B,256*T]).
        pcm_out = (n_samples [B] int32, out_offset [B] int64, flat int16 buffer - device or pinned host): the ragged
        PCM epilogue of SURVEY 8(f1); returns that buffer and produces no padded wave.
        pcm_scale: the int16 full scale (preprocess.yaml audio.max_wav_value, utils/model.py:105-108).
        frame_lens (host ints, one per utterance; only with pcm_out): the utterances' valid mel frames.  The caller
        will only ever look at samples < frame_lens[b]*hop (utils/model.py:109-113), so every layer is run on the rows
        that can reach those samples - frame_lens[b] + halo_frames() frames' worth - instead of on the whole padded
        batch.  The samples that are written are bit-identical to the padded computation."""
        if frame_lens is not None and pcm_out is None:
            raise ValueError("frame_lens (ragged schedule) needs pcm_out: the padded wave would hold uncomputed samples")
        if not x.is_cuda:
            raise RuntimeError("speakingstyle_b200 Generator needs CUDA tensors; there is no CPU fallback")
        with torch.cuda.device(x.device):  # launches go to the tensor's own device, whatever the current one is
            return self._forward(x, channels_last, pcm, pcm_out, frame_lens, pcm_scale)

    def _forward(self, x, channels_last, pcm, pcm_out, frame_lens, pcm_scale):
        if not x.is_cuda:
            raise RuntimeError("speakingstyle_b200 Generator needs CUDA tensors; there is no CPU fallback")
        h = self.h
        pk = self.packed()
        if not channels_last:
            x = x.transpose(1, 2)
        x = x.to(torch.float32).contiguous()
        rows = self._ragged_rows(frame_lens, x.shape[1], x.device) if frame_lens is not None else None
        if self.precision == "bf16":
            return self._forward_bf16(x, pcm, pcm_out, rows, pcm_scale)
        if self.fp32_convs == "bf16x3" and pk["x3"]:
            return self._forward_x3(x, pcm, pcm_out, rows, pcm_scale)
        rate = 1
        x = ops.conv1d(x, pk["pre"], self.conv_pre.bias, k=7, pad=3, rows=rows and rows[rate])
        nk = self.num_kernels
        for i, (u, k) in enumerate(zip(h["upsample_rates"], h["upsample_kernel_sizes"])):
            wu, bu = pk["ups"][i]
            x = ops.conv_transpose1d(x, wu, bu, cout=self.ups[i].bias.numel(), stride=u, pad=(k - u) // 2,
                                     in_act=ops.ACT_LRELU, in_slope=LRELU_SLOPE, rows=rows and rows[("up", i)])
            rate *= u
            rw = rows and rows[rate]
            # four buffers per stage: x (stage input, read by all three ResBlocks), xs (their mean),
            # tbuf (output of c1) and ybuf (the running residual stream of one ResBlock).  c2 may write
            # in place over its residual operand (same element, same thread) but never over its conv input.
            xs, tbuf, ybuf = torch.empty_like(x), torch.empty_like(x), torch.empty_like(x)
            for j in range(nk):
                rb = self.resblocks[i * nk + j]
                w1s, w2s = pk["res"][i * nk + j]
                y = x
                nd = len(rb.dilation)
                rk = rb.kernel_size
                for m, dil in enumerate(rb.dilation):
                    # c1's output is only ever read through the leaky-ReLU in front of c2 (hifigan/models.py:98-101), so c1's
                    # epilogue applies it once per element and c2 takes its input as is - the same fp32 operation on the
                    # same value, done once instead of once per tap in c2's operand split
                    t = ops.conv1d(y, w1s[m], rb.convs1[m].bias, k=rk, dilation=dil, pad=get_padding(rk, dil),
                                   in_act=ops.ACT_LRELU, in_slope=LRELU_SLOPE, act=ops.ACT_LRELU, slope=LRELU_SLOPE,
                                   out=tbuf, rows=rw)
                    if m < nd - 1:
                        y = ops.conv1d(t, w2s[m], rb.convs2[m].bias, k=rk, pad=get_padding(rk, 1), residual=y, out=ybuf,
                                       rows=rw)
                    else:  # last pair: x_j = c2(..) + y goes straight into the stage mean
                        ops.conv1d(t, w2s[m], rb.convs2[m].bias, k=rk, pad=get_padding(rk, 1), residual=y, write_y=False,
                                   accum=xs, accum_scale=1.0 / nk, accum_init=(j == 0), rows=rw)
            x = xs
        return self._post(x, pk, pcm, pcm_out, pcm_scale)
