"""Host-side mirror of the reference's ``transformer`` package over the C ABI.

Same class names, constructor arguments, attribute names and ``state_dict`` keys as
transformer/{Models,Layers,SubLayers}.py, so a checkpoint of the reference loads unchanged.  The
sub-modules only HOLD parameters (no torch operator is ever applied to them); ``forward`` enqueues
calls of speakingstyle_b200.ops, which are the hand-written sm_100a kernels.  Inference only.
"""
import math

import torch
import torch.nn as nn

from . import ops
from .configs import N_SRC_VOCAB
from .synth import sinusoid_table

PAD = 0  # transformer/Constants.py:1


class ParamCache:
    """Kernel-layout copies of parameters, rebuilt when a source parameter changes."""

    epoch = 0  # bumped whenever some module replaces Parameter objects

    def packed(self):
        # the walk over parameters / buffers is done once (it costs more host time than a small kernel runs); in-place
        # updates (load_state_dict, .to()) are still seen through data_ptr / _version.  Code that REPLACES Parameter
        # objects (remove_weight_norm) bumps ParamCache.epoch, which invalidates every cached list.
        srcs = self.__dict__.get("_pack_srcs")
        if srcs is None or self.__dict__.get("_pack_epoch") != ParamCache.epoch:
            srcs = [p for p in self.parameters(recurse=True)] + [b for b in self.buffers(recurse=True)]
            self.__dict__["_pack_srcs"] = srcs
            self.__dict__["_pack_epoch"] = ParamCache.epoch
        key = tuple((t.data_ptr(), t._version) for t in srcs)
        if self.__dict__.get("_pack_key") != key:
            with torch.no_grad():
                self.__dict__["_pack_val"] = self._pack()
            self.__dict__["_pack_key"] = key
        return self.__dict__["_pack_val"]


pack_conv = ops.pack_conv  # nn.Conv1d / nn.Linear weight -> ops.ConvW (all kernel layouts, built lazily)


class Dense(nn.Module):
    """Parameter holder with nn.Linear's / nn.Conv1d's names, shapes and default initialiser."""

    def __init__(self, in_f, out_f, kernel_size=None, bias=True):
        super().__init__()
        shape = (out_f, in_f) if kernel_size is None else (out_f, in_f, kernel_size)
        bound = 1.0 / math.sqrt(in_f * (kernel_size or 1))
        self.weight = nn.Parameter(torch.empty(shape).uniform_(-bound, bound), requires_grad=False)
        if bias:
            self.bias = nn.Parameter(torch.empty(out_f).uniform_(-bound, bound), requires_grad=False)
        else:
            self.register_parameter("bias", None)

    def forward(self, *a, **k):
        raise RuntimeError("parameter holder: compute goes through speakingstyle_b200.ops")


class Norm(nn.Module):
    """nn.LayerNorm parameter holder."""

    def __init__(self, n):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(n), requires_grad=False)
        self.bias = nn.Parameter(torch.zeros(n), requires_grad=False)

    def forward(self, *a, **k):
        raise RuntimeError("parameter holder: compute goes through speakingstyle_b200.ops")


class FiLM(nn.Module):
    """model/blocks.py:43-62 - the two learnable scales; the modulation is fused into ss_layernorm."""

    def __init__(self):
        super().__init__()
        self.s_gamma = nn.Parameter(torch.ones(1), requires_grad=False)
        self.s_beta = nn.Parameter(torch.ones(1), requires_grad=False)


def film_args(film, gammas, betas):
    if film is None or gammas is None or betas is None:
        return None
    B = gammas.shape[0]
    return (gammas.reshape(B, -1).contiguous(), betas.reshape(B, -1).contiguous(), film.s_gamma, film.s_beta)


def lens_from_mask(mask):
    """Key-padding masks on this path are always prefix masks (utils/tools.py:110-118), so a
    [B,S] bool mask (True = pad) is equivalent to per-row lengths, which is what the kernels take."""
    return (mask.shape[1] - mask.sum(dim=1)).to(torch.int32)


class MultiHeadAttention(nn.Module, ParamCache):
    """transformer/SubLayers.py:8-57."""

    def __init__(self, n_head, d_model, d_k, d_v, dropout=0.1):
        super().__init__()
        assert d_k == d_v and n_head * d_k == d_model
        self.n_head, self.d_k, self.d_v = n_head, d_k, d_v
        self.w_qs = Dense(d_model, n_head * d_k)
        self.w_ks = Dense(d_model, n_head * d_k)
        self.w_vs = Dense(d_model, n_head * d_v)
        self.layer_norm = Norm(d_model)
        self.fc = Dense(n_head * d_v, d_model)

    def _pack(self):
        wqkv = torch.cat([self.w_qs.weight, self.w_ks.weight, self.w_vs.weight], dim=0)  # [3D, D]
        return {
            "wqkv": pack_conv(wqkv),
            "bqkv": torch.cat([self.w_qs.bias, self.w_ks.bias, self.w_vs.bias]).contiguous(),
            "wfc": pack_conv(self.fc.weight),
        }

    def forward(self, x, lens, mask_rows=True, xb=None, xp=None):
        """x [B,S,D] -> LayerNorm(fc(attention) + x), rows >= lens zeroed (Layers.py:27-28).
        With ``xb`` (bf16 copy of x) the two projections run on the tensor cores and the result is
        returned as (fp32, bf16).  With ``xp`` (split-bf16 pair rows of x) they run on ss_conv1d_bf16x3 (fp32-class
        accuracy) and the result is (fp32, pair rows)."""
        pk = self.packed()
        ln = self.layer_norm
        # Ragged schedule: the attention kernel reads only Q / K / V rows below ceil128(lens) and writes zeros for the
        # padded queries, and the LayerNorm below overwrites the padded rows with zeros without reading them - so the
        # two k = 1 projections skip the tiles that hold nothing but padding (SubLayers.py:39-55 computes them and
        # Layers.py:27-28 throws them away).
        rows = lens if mask_rows else None
        if xb is not None:
            qkv, _ = ops.conv1d_tc(xb, pk["wqkv"], pk["bqkv"], rows=rows)
            ctx = ops.attention(qkv, lens, self.n_head)
            y, _ = ops.conv1d_tc(ops.cast_bf16(ctx), pk["wfc"], self.fc.bias, residual=x, rows=rows)
            return ops.layernorm(y, ln.weight, ln.bias, lens=lens if mask_rows else None, out=y, want_bf16=True)
        if xp is not None:
            qkv, _ = ops.conv1d_x3(xp, pk["wqkv"], pk["bqkv"], rows=rows)
            ctx = ops.attention(qkv, lens, self.n_head)
            y, _ = ops.conv1d_x3(ops.split_bf16(ctx), pk["wfc"], self.fc.bias, residual=x, rows=rows)
            return ops.layernorm(y, ln.weight, ln.bias, lens=lens if mask_rows else None, out=y, want_pair=True)
        qkv = ops.conv1d(x, pk["wqkv"], pk["bqkv"], rows=rows)
        ctx = ops.attention(qkv, lens, self.n_head)
        y = ops.conv1d(ctx, pk["wfc"], self.fc.bias, residual=x, rows=rows)
        return ops.layernorm(y, ln.weight, ln.bias, lens=lens if mask_rows else None, out=y)


class PositionwiseFeedForward(nn.Module, ParamCache):
    """transformer/SubLayers.py:60-93."""

    def __init__(self, d_in, d_hid, kernel_size, dropout=0.1):
        super().__init__()
        self.kernel_size = list(kernel_size)
        self.w_1 = Dense(d_in, d_hid, kernel_size[0])
        self.w_2 = Dense(d_hid, d_in, kernel_size[1])
        self.layer_norm = Norm(d_in)

    def _pack(self):
        return {"w1": pack_conv(self.w_1.weight), "w2": pack_conv(self.w_2.weight)}

    def forward(self, x, lens=None, film=None, xb=None, xp=None):
        pk = self.packed()
        k1, k2 = self.kernel_size
        ln = self.layer_norm
        # Batch-invariant mode: when w_2 has a halo (the reference encoder's [3, 3] kernels) it reads w_1's output at
        # padded frames, which a lone utterance would see as zero padding - mask the intermediate there.
        hl = lens if (getattr(self, "batch_invariant", False) and k2 > 1) else None
        # Ragged schedule (encoder / decoder blocks, w_2 with k = 1): the block's input is zero on the padded frames
        # (the attention sub-layer masked it), w_2 reads no neighbour rows, and the LayerNorm below zeroes the padded
        # rows without reading them - tiles of pure padding are skipped.  Not for the reference encoder's [3, 3]
        # blocks, whose w_2 reads w_1's (non-zero) output on padded frames (SubLayers.py:88-91).
        rows = lens if (lens is not None and k2 == 1) else None
        if xb is not None:  # tensor-core path: the 1024-channel intermediate only ever exists in bf16
            h, _ = ops.conv1d_tc(xb, pk["w1"], self.w_1.bias, pad=(k1 - 1) // 2, act=ops.ACT_RELU, lens=hl,
                                 out_dtype=torch.bfloat16, rows=rows)
            y, _ = ops.conv1d_tc(h, pk["w2"], self.w_2.bias, pad=(k2 - 1) // 2, residual=x, rows=rows)
            return ops.layernorm(y, ln.weight, ln.bias, film=film, lens=lens, out=y, want_bf16=True)
        if xp is not None:  # split-bf16 path: the 1024-channel intermediate only ever exists as pair rows of relu(.)
            _, hp = ops.conv1d_x3(xp, pk["w1"], self.w_1.bias, pad=(k1 - 1) // 2, want_y=False, act2=ops.ACT_RELU, lens=hl,
                                  rows=rows)
            y, _ = ops.conv1d_x3(hp, pk["w2"], self.w_2.bias, pad=(k2 - 1) // 2, residual=x, rows=rows)
            return ops.layernorm(y, ln.weight, ln.bias, film=film, lens=lens, out=y, want_pair=True)
        h = ops.conv1d(x, pk["w1"], self.w_1.bias, k=k1, pad=(k1 - 1) // 2, act=ops.ACT_RELU, lens=hl, rows=rows)
        y = ops.conv1d(h, pk["w2"], self.w_2.bias, k=k2, pad=(k2 - 1) // 2, residual=x, rows=rows)
        return ops.layernorm(y, ln.weight, ln.bias, film=film, lens=lens, out=y)


class FFTBlock(nn.Module):
    """transformer/Layers.py:11-37."""

    def __init__(self, d_model, n_head, d_k, d_v, d_inner, kernel_size, dropout=0.1, film=True):
        super().__init__()
        self.slf_attn = MultiHeadAttention(n_head, d_model, d_k, d_v, dropout=dropout)
        self.pos_ffn = PositionwiseFeedForward(d_model, d_inner, kernel_size, dropout=dropout)
        if film:
            self.film = FiLM()

    def forward(self, enc_input, gammas=None, betas=None, mask=None, slf_attn_mask=None, lens=None, xb=None, xp=None):
        """Returns (output, None): the attention probabilities the reference returns here are
        dropped by every caller (Models.py:98-101) and are never materialised.
        With ``xb`` (bf16 copy of the input) the block runs its projections and convolutions on the
        tensor cores and returns ((fp32, bf16), None)."""
        if lens is None:
            lens = lens_from_mask(mask)
        film = film_args(getattr(self, "film", None), gammas, betas)
        if xb is not None:
            x, xb = self.slf_attn(enc_input, lens, xb=xb)
            return self.pos_ffn(x, lens=lens, film=film, xb=xb), None
        if xp is not None:  # split-bf16 pair rows: ((fp32, pair), None)
            x, xp = self.slf_attn(enc_input, lens, xp=xp)
            return self.pos_ffn(x, lens=lens, film=film, xp=xp), None
        x = self.slf_attn(enc_input, lens)
        x = self.pos_ffn(x, lens=lens, film=film)
        return x, None


class _PositionTable:
    """position_enc parameter + on-demand longer table (Models.py:82-87,145-152; modules.py:374-380)."""

    def _pe(self, n, width):
        table = self.position_enc[0]
        if n <= table.shape[0]:
            return table
        cache = self.__dict__.setdefault("_long_pe", {})
        if n not in cache:
            cache.clear()
            cache[n] = sinusoid_table(n, width).to(table.device)
        return cache[n]


class Encoder(nn.Module, _PositionTable):
    """transformer/Models.py:33-101."""

    def __init__(self, config):
        super().__init__()
        tr = config["transformer"]
        n_position = config["max_seq_len"] + 1
        d = tr["encoder_hidden"]
        self.max_seq_len = config["max_seq_len"]
        self.d_model = d
        self.src_word_emb = nn.Embedding(N_SRC_VOCAB, d, padding_idx=PAD)
        self.src_word_emb.weight.requires_grad_(False)
        self.position_enc = nn.Parameter(sinusoid_table(n_position, d).unsqueeze(0), requires_grad=False)
        self.layer_stack = nn.ModuleList([
            FFTBlock(d, tr["encoder_head"], d // tr["encoder_head"], d // tr["encoder_head"], tr["conv_filter_size"],
                     tr["conv_kernel_size"], dropout=tr["encoder_dropout"]) for _ in range(tr["encoder_layer"])])

    def forward(self, src_seq, mask, gammas=None, betas=None, return_attns=False, lens=None):
        if lens is None:
            lens = lens_from_mask(mask)
        x = ops.embed_posenc(src_seq.contiguous(), self.src_word_emb.weight, self._pe(src_seq.shape[1], self.d_model))
        for layer in self.layer_stack:
            x, _ = layer(x, gammas=gammas, betas=betas, lens=lens)
        return x


class Decoder(nn.Module, _PositionTable):
    """transformer/Models.py:104-170.  ``forward`` takes the sequence with the position table
    ALREADY added when ``pe_added`` (the LengthRegulator kernel writes x + PE in its one pass)."""

    def __init__(self, config):
        super().__init__()
        tr = config["transformer"]
        n_position = config["max_seq_len"] + 1
        d = tr["decoder_hidden"]
        self.max_seq_len = config["max_seq_len"]
        self.d_model = d
        self.position_enc = nn.Parameter(sinusoid_table(n_position, d).unsqueeze(0), requires_grad=False)
        self.layer_stack = nn.ModuleList([
            FFTBlock(d, tr["decoder_head"], d // tr["decoder_head"], d // tr["decoder_head"], tr["conv_filter_size"],
                     tr["conv_kernel_size"], dropout=tr["decoder_dropout"]) for _ in range(tr["decoder_layer"])])

    def forward(self, enc_seq, mask, gammas=None, betas=None, return_attns=False, lens=None, pe_added=False,
                precision="fp32", return_bf16=False):
        T = enc_seq.shape[1]
        mask = mask[:, :T]
        if lens is None:
            lens = lens_from_mask(mask)
        x = enc_seq
        if not pe_added:
            x = ops.add_position(enc_seq, self._pe(T, self.d_model))
        if precision == "bf16":
            xb = ops.cast_bf16(x)
            for layer in self.layer_stack:
                (x, xb), _ = layer(x, gammas=gammas, betas=betas, lens=lens, xb=xb)
            return ((x, xb), mask) if return_bf16 else (x, mask)
        if precision == "bf16x3" and all(ops.x3_ok(w) for layer in self.layer_stack for w in (
                layer.slf_attn.packed()["wqkv"], layer.slf_attn.packed()["wfc"], layer.pos_ffn.packed()["w1"],
                layer.pos_ffn.packed()["w2"])):
            # fp32 mode, split-bf16 kernels: nothing downstream of the decoder decides an integer, and the mel stays
            # within ~1e-5 of the 3xTF32 path (bar 1e-3)
            xp = ops.split_bf16(x)
            for layer in self.layer_stack:
                (x, xp), _ = layer(x, gammas=gammas, betas=betas, lens=lens, xp=xp)
            return x, mask
        for layer in self.layer_stack:
            x, _ = layer(x, gammas=gammas, betas=betas, lens=lens)
        return x, mask


class ConvNorm(nn.Module):
    """transformer/Layers.py:39-77 parameter holder (``.conv.weight`` / ``.conv.bias``)."""

    def __init__(self, in_channels, out_channels, kernel_size=1, stride=1, padding=None, dilation=1, bias=True,
                 w_init_gain="linear", transform=False):
        super().__init__()
        if padding is None:
            assert kernel_size % 2 == 1
            padding = int(dilation * (kernel_size - 1) / 2)
        assert stride == 1
        self.kernel_size, self.padding, self.dilation = kernel_size, padding, dilation
        self.conv = Dense(in_channels, out_channels, kernel_size, bias=bias)
        self.transform = transform


class _BatchNormHolder(nn.Module):
    def __init__(self, n):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(n), requires_grad=False)
        self.bias = nn.Parameter(torch.zeros(n), requires_grad=False)
        self.register_buffer("running_mean", torch.zeros(n))
        self.register_buffer("running_var", torch.ones(n))
        self.register_buffer("num_batches_tracked", torch.tensor(0, dtype=torch.long))
        self.eps = 1e-5


class PostNet(nn.Module, ParamCache):
    """transformer/Layers.py:80-148.  BatchNorm1d (eval) is folded into each convolution's weight
    and bias when the parameters are packed; tanh is the convolution kernel's epilogue."""

    def __init__(self, n_mel_channels=80, postnet_embedding_dim=512, postnet_kernel_size=5, postnet_n_convolutions=5):
        super().__init__()
        self.kernel_size = postnet_kernel_size
        chans = [n_mel_channels] + [postnet_embedding_dim] * (postnet_n_convolutions - 1) + [n_mel_channels]
        self.convolutions = nn.ModuleList()
        for i in range(postnet_n_convolutions):
            self.convolutions.append(nn.Sequential(
                ConvNorm(chans[i], chans[i + 1], kernel_size=postnet_kernel_size, stride=1,
                         padding=int((postnet_kernel_size - 1) / 2), dilation=1),
                _BatchNormHolder(chans[i + 1])))

    def _pack(self):
        out = []
        for conv, bn in self.convolutions:
            scale = bn.weight / torch.sqrt(bn.running_var + bn.eps)
            w = conv.conv.weight * scale[:, None, None]
            b = (conv.conv.bias - bn.running_mean) * scale + bn.bias
            out.append((pack_conv(w), b.contiguous()))
        return out

    def forward(self, x, residual=None, precision="fp32", xb=None, lens=None):
        """x [B,T,n_mel] -> postnet(x) (+ residual fused into the last convolution if given).
        precision="bf16": tensor-core convolutions, bf16 activations between layers (xb = bf16 copy of x).
        lens (batch-invariant mode): rows >= lens[b] are zeroed after every layer."""
        pk = self.packed()
        k = self.kernel_size
        n = len(pk)
        if precision == "bf16":
            h = xb if xb is not None else ops.cast_bf16(x)
            for i, (wtc, b) in enumerate(pk):
                last = i == n - 1
                h, _ = ops.conv1d_tc(h, wtc, b, pad=(k - 1) // 2, act=ops.ACT_NONE if last else ops.ACT_TANH,
                                     residual=residual if last else None, lens=lens,
                                     out_dtype=torch.float32 if last else torch.bfloat16)
            return h
        xp = None
        for i, (w, b) in enumerate(pk):
            last = i == n - 1
            act = ops.ACT_NONE if last else ops.ACT_TANH
            if precision == "bf16x3" and ops.x3_ok(w):
                # split-bf16 kernel for the 512 -> 512 layers; the next layer takes pair rows of tanh(.) if it runs on
                # this kernel too, else fp32
                if xp is None:
                    xp = ops.split_bf16(x)
                nxt_x3 = not last and ops.x3_ok(pk[i + 1][0])
                x, xp = ops.conv1d_x3(xp, w, b, pad=(k - 1) // 2, act=act, residual=residual if last else None, lens=lens,
                                      want_y=not nxt_x3, act2=ops.ACT_NONE if nxt_x3 else None)
            else:
                x = ops.conv1d(x, w, b, k=k, pad=(k - 1) // 2, act=act, residual=residual if last else None, lens=lens)
                xp = None
        return x
