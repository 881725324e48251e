"""Per-kernel parity through the C ABI (speakingstyle_b200.ops = ctypes calls) against the oracle.

Integer / index outputs: bit-exact.  fp32 outputs: max-abs tolerance written at each check
(north_star: fp32 max-abs <= 1e-3 end to end; single kernels are held to 1e-4 or tighter)."""
import ctypes
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from oracle import fs2_oracle as O  # noqa: E402
from oracle import intops  # noqa: E402

DEV = "cuda:0"
F32P = ctypes.POINTER(ctypes.c_float)
I64P = ctypes.POINTER(ctypes.c_int64)


def rnd(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(*shape, generator=g) * scale


@pytest.fixture(scope="module")
def ops():
    from speakingstyle_b200 import ops as o

    return o


def pack(w):
    from speakingstyle_b200 import ops

    return ops.pack_conv(w.to(DEV))


def lrelu(x, s):
    return F.leaky_relu(x, s)


@pytest.mark.parametrize("B,S,Cin,Cout,k,dil", [
    (2, 37, 80, 1024, 3, 1),     # reference-encoder first conv
    (3, 150, 256, 1024, 9, 1),   # FFT block w_1
    (2, 130, 1024, 256, 1, 1),   # FFT block w_2
    (2, 70, 256, 768, 1, 1),     # fused QKV projection
    (2, 300, 64, 64, 11, 5),     # HiFi-GAN ResBlock, dilated
    (1, 257, 32, 32, 7, 3),
    (2, 90, 512, 80, 5, 1),      # PostNet last layer (Cout not a multiple of the tile)
    (1, 1, 256, 512, 1, 1),      # single row (feature_wise_affine)
    (2, 5, 256, 256, 3, 1),      # sequence shorter than the kernel's halo
])
@pytest.mark.parametrize("backend", ["ffma", "tf32x3_ss", "tf32x3_ts"])
def test_conv1d_matches_torch(ops, backend, B, S, Cin, Cout, k, dil):
    """Both fp32-accurate kernels: the FP32-pipe one and the 3xTF32 tcgen05 one."""
    x, w, b = rnd(B, S, Cin, seed=1), rnd(Cout, Cin, k, seed=2, scale=1 / math.sqrt(Cin * k)), rnd(Cout, seed=3)
    pad = (k - 1) // 2 * dil
    ref = F.conv1d(x.transpose(1, 2).double(), w.double(), b.double(), padding=pad, dilation=dil).transpose(1, 2)
    y = ops.conv1d(x.to(DEV), pack(w), b.to(DEV), k=k, dilation=dil, pad=pad, backend=backend)
    err = (y.cpu().double() - ref).abs()
    assert err.max().item() <= 2e-5, (err.max().item(), err.pow(2).mean().sqrt().item())


def test_tf32x3_is_at_least_as_accurate_as_fp32_fma(ops):
    """The claim the integer-deciding chain rests on: on a long reduction
    (K = 3 x 1024) the 3xTF32 tensor-core result is no further from the fp64 truth than the
    FP32-pipe kernel's."""
    B, S, Cin, Cout, k = 2, 400, 1024, 1024, 3
    x, w = rnd(B, S, Cin, seed=51), rnd(Cout, Cin, k, seed=52, scale=1 / math.sqrt(Cin * k))
    ref = F.conv1d(x.transpose(1, 2).double(), w.double(), None, padding=1).transpose(1, 2)
    W = pack(w)
    e = {}
    for backend in ("ffma", "tf32x3_ss", "tf32x3_ts"):
        y = ops.conv1d(x.to(DEV), W, None, pad=1, backend=backend).cpu().double()
        e[backend] = ((y - ref).pow(2).mean().sqrt().item(), (y - ref).abs().max().item())
    import os
    d = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    if os.path.isdir(d):
        with open(os.path.join(d, "parity_report.txt"), "a") as f:
            f.write(f"conv K=3072 error vs fp64 (rms, max): {e}\n")
    for be in ("tf32x3_ss", "tf32x3_ts"):
        assert e[be][0] <= 1.5 * e["ffma"][0] and e[be][1] <= 1e-5, be


@pytest.mark.parametrize("backend", ["ffma", "tf32x3_ss", "tf32x3_ts"])
def test_conv1d_epilogues(ops, backend, monkeypatch):
    monkeypatch.setattr(ops, "fp32_backend", backend)
    B, S, Cin, Cout, k = 3, 77, 64, 128, 3
    x, w, b = rnd(B, S, Cin, seed=4), rnd(Cout, Cin, k, seed=5, scale=0.1), rnd(Cout, seed=6)
    res = rnd(B, S, Cout, seed=7)
    lens = torch.tensor([77, 40, 1], dtype=torch.int32)
    conv = lambda inp: F.conv1d(inp.transpose(1, 2), w, b, padding=1).transpose(1, 2)  # noqa: E731
    keep = (torch.arange(S)[None, :] < lens[:, None])[..., None]
    xd, wd, bd, resd, lensd = x.to(DEV), pack(w), b.to(DEV), res.to(DEV), lens.to(DEV)
    # relu + mask
    y = ops.conv1d(xd, wd, bd, k=k, pad=1, act=ops.ACT_RELU, lens=lensd)
    assert (y.cpu() - F.relu(conv(x)) * keep).abs().max().item() <= 2e-5
    # tanh + residual
    y = ops.conv1d(xd, wd, bd, k=k, pad=1, act=ops.ACT_TANH, residual=resd)
    assert (y.cpu() - (torch.tanh(conv(x)) + res)).abs().max().item() <= 2e-5
    # lrelu on load + residual, in place over the residual buffer
    buf = resd.clone()
    y = ops.conv1d(xd, wd, bd, k=k, pad=1, in_act=ops.ACT_LRELU, in_slope=0.1, residual=buf, out=buf)
    assert (y.cpu() - (conv(lrelu(x, 0.1)) + res)).abs().max().item() <= 2e-5
    # accumulate into a running mean without writing y
    acc = torch.full((B, S, Cout), 7.0, device=DEV)
    ops.conv1d(xd, wd, bd, k=k, pad=1, write_y=False, accum=acc, accum_scale=0.5, accum_init=True)
    ops.conv1d(xd, wd, bd, k=k, pad=1, act=ops.ACT_LRELU, slope=0.2, write_y=False, accum=acc, accum_scale=0.25)
    want = 0.5 * conv(x) + 0.25 * lrelu(conv(x), 0.2)
    assert (acc.cpu() - want).abs().max().item() <= 2e-5


@pytest.mark.parametrize("B,S,Cin,Cout,k,dil", [
    (7, 50, 64, 128, 9, 1),     # several utterances inside one 128-row tile; halo wider than the gaps
    (5, 129, 256, 256, 1, 1),   # one row past a tile: every tile straddles
    (3, 300, 64, 80, 11, 5),    # dilated taps reaching 25 frames across an utterance end; Cout not a tile multiple
    (64, 119, 256, 256, 3, 1),  # the encoder's shape
    (9, 1, 256, 256, 1, 1),     # S = 1
])
def test_conv1d_flat_tiling_across_utterances(ops, B, S, Cin, Cout, k, dil):
    """The default 3xTF32 kernel tiles stride-1 same-length layers over the flattened B*S rows:
    taps must not leak across utterance ends, and bias / relu / residual / row mask must land on the right
    (utterance, frame) of tiles that straddle.  Checked against torch per utterance and against the kernel's own
    per-utterance tiling (one launch per utterance)."""
    x, w, b = rnd(B, S, Cin, seed=61), rnd(Cout, Cin, k, seed=62, scale=1 / math.sqrt(Cin * k)), rnd(Cout, seed=63)
    res = rnd(B, S, Cout, seed=64)
    lens = torch.tensor([(S * (i + 1)) // B if i % 2 else S for i in range(B)], dtype=torch.int32)
    pad = (k - 1) // 2 * dil
    conv = F.conv1d(x.transpose(1, 2).double(), w.double(), b.double(), padding=pad, dilation=dil).transpose(1, 2)
    keep = (torch.arange(S)[None, :] < lens[:, None])[..., None]
    want = (F.relu(conv) + res.double()) * keep
    W = pack(w)
    y = ops.conv1d(x.to(DEV), W, b.to(DEV), k=k, dilation=dil, pad=pad, act=ops.ACT_RELU, residual=res.to(DEV),
                   lens=lens.to(DEV), backend="tf32x3_ts")
    assert (y.cpu().double() - want).abs().max().item() <= 2e-5
    # one launch per utterance (B = 1 turns flat tiling off): per-utterance tiling must agree to the last bit
    y2 = torch.cat([ops.conv1d(x[i:i + 1].to(DEV), W, b.to(DEV), k=k, dilation=dil, pad=pad, act=ops.ACT_RELU,
                               residual=res[i:i + 1].to(DEV), lens=lens[i:i + 1].to(DEV), backend="tf32x3_ts")
                    for i in range(B)])
    assert torch.equal(y, y2)


@pytest.mark.parametrize("B,S,Cin,Cout,k,dil", [(3, 150, 256, 1024, 9, 1), (7, 50, 64, 128, 9, 1), (2, 300, 64, 80, 11, 5),
                                                 (5, 257, 256, 256, 1, 1), (1, 700, 128, 384, 3, 1), (2, 37, 80, 64, 3, 1),
                                                 (2, 37, 80, 1024, 3, 1), (1, 1, 256, 512, 1, 1), (4, 256, 64, 128, 5, 1)])
def test_cta_pair_conv_is_bit_identical_to_single_cta(ops, B, S, Cin, Cout, k, dil):
    """The cta_group::2 form (two CTAs per 256 x 128 tile) does the same sums in the same order."""
    x, w, b = rnd(B, S, Cin, seed=71), rnd(Cout, Cin, k, seed=72, scale=1 / math.sqrt(Cin * k)), rnd(Cout, seed=73)
    res = rnd(B, S, Cout, seed=74)
    lens = torch.tensor([max(1, S - 13 * i) for i in range(B)], dtype=torch.int32)
    pad = (k - 1) // 2 * dil
    W = pack(w)
    kw = dict(k=k, dilation=dil, pad=pad, act=ops.ACT_LRELU, slope=0.1, residual=res.to(DEV), lens=lens.to(DEV))
    y1 = ops.conv1d(x.to(DEV), W, b.to(DEV), backend="tf32x3_ts", **kw)
    y2 = ops.conv1d(x.to(DEV), W, b.to(DEV), backend="tf32x3_ts2", **kw)
    assert torch.equal(y1, y2)



@pytest.mark.parametrize("B,S,Cin,Cout,k,dil", [
    (3, 1000, 32, 32, 11, 5),   # weights resident, halo 25 each side, three panels
    (3, 1000, 32, 32, 3, 1),
    (2, 777, 64, 64, 3, 5),     # 64 channels, S not a multiple of the tile
    (2, 777, 64, 64, 11, 5),    # weights streamed (22 tiles of 16 KB do not fit)
    (20, 300, 64, 64, 7, 1),    # streamed, more tiles than CTAs: weight ring wrap-around across tiles
    (2, 640, 64, 64, 3, 3),
    (2, 900, 64, 32, 7, 3),     # Cin != Cout
    (1, 50, 32, 64, 2, 1),      # shorter than a tile, even tap count
    (40, 300, 32, 32, 7, 1),    # more tiles than CTAs: panel / weight ring wrap-around
])
def test_panel_conv_matches_the_per_tap_kernel(ops, B, S, Cin, Cout, k, dil):
    """ss_conv1d_tf32x3_panel (activation panel fetched once per tile, shifted rows per tap, resident weights, hi*W_hi
    and hi*W_lo as one N = 2*BN instruction) against ss_conv1d_tf32x3_ts: same products, the correction terms summed
    per accumulator-restart group instead of per tile - equal to ~1 ulp - with every epilogue HiFi-GAN uses."""
    g = torch.Generator().manual_seed(12)
    x = torch.randn(B, S, Cin, generator=g).to(DEV)
    W = ops.pack_conv(torch.randn(Cout, Cin, k, generator=g).to(DEV) / math.sqrt(Cin * k))
    bias = torch.randn(Cout, generator=g).to(DEV)
    res = torch.randn(B, S, Cout, generator=g).to(DEV)
    pad = (k - 1) * dil // 2
    kw = dict(k=k, dilation=dil, pad=pad, in_act=ops.ACT_LRELU, in_slope=0.1)
    for extra in (dict(), dict(residual=res), dict(residual=res, act=ops.ACT_RELU)):
        a = ops.conv1d(x, W, bias, backend="tf32x3_ts", **kw, **extra)
        b = ops.conv1d(x, W, bias, backend="tf32x3_panel", **kw, **extra)
        assert (a - b).abs().max().item() <= 4e-6, extra
    acc_a = torch.full((B, S, Cout), 0.25, device=DEV)
    acc_b = acc_a.clone()
    ops.conv1d(x, W, bias, backend="tf32x3_ts", residual=res, write_y=False, accum=acc_a, accum_scale=1 / 3, **kw)
    ops.conv1d(x, W, bias, backend="tf32x3_panel", residual=res, write_y=False, accum=acc_b, accum_scale=1 / 3, **kw)
    assert (acc_a - acc_b).abs().max().item() <= 4e-6
    # ragged schedule on the panel kernel
    rows_host = ([S, 0, 1, S // 2 + 3] * 10)[:B]
    rows = torch.tensor(rows_host, dtype=torch.int32, device=DEV)
    full = ops.conv1d(x, W, bias, backend="tf32x3_panel", residual=res, **kw)
    out = torch.full((B, S, Cout), 7.5, device=DEV)
    ops.conv1d(x, W, bias, backend="tf32x3_panel", residual=res, out=out, rows=rows, **kw)
    for b_, r in enumerate(rows_host):
        assert torch.equal(out[b_, :r], full[b_, :r]), (b_, r)
        assert bool((out[b_, -(-r // 128) * 128:] == 7.5).all()), (b_, r)
    # and against fp64 directly
    ref = torch.nn.functional.conv1d(torch.nn.functional.leaky_relu(x.double(), 0.1).transpose(1, 2),
                                     W.kcc.permute(1, 2, 0).double(), bias.double(), padding=pad, dilation=dil).transpose(1, 2)
    if ref.shape[1] == S:
        assert (full.double() - res.double() - ref).abs().max().item() <= 2e-5


@pytest.mark.parametrize("variant,B,S,Cin,Cout,k,dil", [
    ("ts", 5, 1000, 32, 32, 11, 5),    # narrow tile (8 split warps, two correction accumulators)
    ("ts", 4, 700, 64, 64, 3, 1),
    ("ts", 3, 900, 256, 256, 1, 1),    # 96-column short-reduction form
    ("ts2", 4, 1200, 128, 128, 7, 3),  # CTA pair: 256-row scheduling units
    ("ts2", 3, 515, 256, 1024, 9, 1),
])
def test_conv1d_ragged_schedule(ops, variant, B, S, Cin, Cout, k, dil):
    """rows[b]: tiles past an utterance's wanted rows are skipped; the wanted rows are bit-identical to the full
    computation and memory past the last computed tile is left alone."""
    g = torch.Generator().manual_seed(11)
    x = torch.randn(B, S, Cin, generator=g).to(DEV)
    W = ops.pack_conv(torch.randn(Cout, Cin, k, generator=g).to(DEV) / math.sqrt(Cin * k))
    bias = torch.randn(Cout, generator=g).to(DEV)
    res = torch.randn(B, S, Cout, generator=g).to(DEV)
    rows_host = [S, 0, 1, S // 2 + 3, 130][:B]
    rows = torch.tensor(rows_host, dtype=torch.int32, device=DEV)
    pad = (k - 1) // 2 * dil
    kw = dict(k=k, dilation=dil, pad=pad, in_act=ops.ACT_LRELU, in_slope=0.1, residual=res, backend="tf32x3_" + variant)
    full = ops.conv1d(x, W, bias, **kw)
    out = torch.full((B, S, Cout), 7.5, device=DEV)
    ops.conv1d(x, W, bias, out=out, rows=rows, **kw)
    unit = 256 if variant == "ts2" and Cout > 64 else 128
    for b, r in enumerate(rows_host):
        assert torch.equal(out[b, :r], full[b, :r]), (b, r)
        done = -(-r // unit) * unit  # the scheduling unit holding the last wanted row is computed whole
        assert bool((out[b, done:] == 7.5).all()), (b, r)
    # running-mean epilogue (HiFi-GAN MRF) on the same schedule
    acc_full = torch.zeros(B, S, Cout, device=DEV)
    acc = torch.zeros(B, S, Cout, device=DEV)
    ops.conv1d(x, W, bias, write_y=False, accum=acc_full, accum_scale=1 / 3, accum_init=False, **kw)
    ops.conv1d(x, W, bias, write_y=False, accum=acc, accum_scale=1 / 3, accum_init=False, rows=rows, **kw)
    for b, r in enumerate(rows_host):
        assert torch.equal(acc[b, :r], acc_full[b, :r]), (b, r)


@pytest.mark.parametrize("Cin,Cout,k,u,S", [(512, 256, 16, 8, 13), (128, 64, 4, 2, 50), (64, 32, 4, 2, 129)])
def test_conv_transpose_phase_form(ops, Cin, Cout, k, u, S):
    from speakingstyle_b200 import configs as C
    from speakingstyle_b200.hifigan import Generator

    B = 2
    x, w, b = rnd(B, S, Cin, seed=8), rnd(Cin, Cout, k, seed=9, scale=1 / math.sqrt(Cin * k / u)), rnd(Cout, seed=10)
    pad = (k - u) // 2
    ref = F.conv_transpose1d(lrelu(x, 0.1).transpose(1, 2), w, b, stride=u, padding=pad).transpose(1, 2)
    for backend in ("ffma", "tf32x3_ss", "tf32x3_ts"):
        y = ops.conv_transpose1d(x.to(DEV), ops.pack_conv_transpose(w.to(DEV), u), b.repeat(u).to(DEV), cout=Cout,
                                 stride=u, pad=pad, in_act=ops.ACT_LRELU, in_slope=0.1, backend=backend)
        assert y.shape == ref.shape
        assert (y.cpu() - ref).abs().max().item() <= 2e-5, backend


@pytest.mark.parametrize("C", [256, 1024])
def test_layernorm_variants(ops, C):
    B, S = 3, 45
    x, res = rnd(B, S, C, seed=11, scale=2.0), rnd(B, S, C, seed=12)
    w, b = 1 + 0.2 * rnd(C, seed=13), 0.1 * rnd(C, seed=14)
    gam, bet = 0.3 * rnd(B, C, seed=15), 0.3 * rnd(B, C, seed=16)
    sg, sb = torch.tensor([1.2]), torch.tensor([0.7])
    pe = rnd(S + 3, C, seed=17)
    lens = torch.tensor([45, 20, 0], dtype=torch.int32)
    keep = (torch.arange(S)[None, :] < lens[:, None])[..., None].float()
    d = lambda t: t.to(DEV)  # noqa: E731
    # plain
    y = ops.layernorm(d(x), d(w), d(b))
    assert (y.cpu() - F.layer_norm(x, (C,), w, b)).abs().max().item() <= 1e-5
    # residual + FiLM + mask (FFT block tail)
    y = ops.layernorm(d(x), d(w), d(b), residual=d(res), film=(d(gam), d(bet), d(sg), d(sb)), lens=d(lens))
    want = O.film(F.layer_norm(x + res, (C,), w, b), gam[:, None], bet[:, None], sg, sb) * keep
    assert (y.cpu() - want).abs().max().item() <= 1e-5
    # relu first, mask, then position add (reference-encoder conv stack tail), in place
    buf = d(x).clone()
    y = ops.layernorm(buf, d(w), d(b), pre_relu=True, lens=d(lens), pe=d(pe), out=buf)
    want = F.layer_norm(F.relu(x), (C,), w, b) * keep + pe[:S][None]
    assert (y.cpu() - want).abs().max().item() <= 1e-5


@pytest.mark.parametrize("H,dk,S,lens", [(2, 128, 150, [150, 64, 1]), (8, 32, 203, [203, 129, 65]),
                                          (8, 32, 800, [800, 513, 64, 1]), (3, 32, 127, [127, 128 - 64, 5]),
                                          (2, 128, 30, [30, 7, 12]), (4, 64, 70, [70, 70, 3]),
                                          (2, 128, 696, [696, 400, 33])])
@pytest.mark.parametrize("backend", ["simt", "tc"])
def test_attention_matches_oracle_math(ops, backend, H, dk, S, lens):
    if backend == "tc" and dk not in (32, 128):
        pytest.skip("the tcgen05 attention kernels cover head sizes 32 and 128")
    B, D = len(lens), H * dk
    qkv = rnd(B, S, 3 * D, seed=18)
    lens_t = torch.tensor(lens, dtype=torch.int32)
    q, k, v = [t.view(B, S, H, dk).transpose(1, 2).double() for t in qkv.split(D, dim=-1)]
    score = q @ k.transpose(-1, -2) / np.power(dk, 0.5)
    pad = torch.arange(S)[None, :] >= lens_t[:, None]
    score = score.masked_fill(pad[:, None, None, :], -math.inf)
    ref = (torch.softmax(score, -1) @ v).transpose(1, 2).reshape(B, S, D)
    ctx = ops.attention(qkv.to(DEV), lens_t.to(DEV), H, backend=backend).cpu().double()
    valid = ~pad
    assert (ctx - ref)[valid].abs().max().item() <= 2e-5
    assert ctx[pad].abs().max().item() == 0.0  # padded query rows are written as zeros


def test_predictor_head_and_bucket_embed(ops):
    B, L, C = 3, 41, 256
    h = F.relu(rnd(B, L, C, seed=19))
    w, b = 1 + 0.2 * rnd(C, seed=20), 0.1 * rnd(C, seed=21)
    lw, lb = rnd(C, seed=22, scale=0.3), torch.tensor([0.4])
    gam, bet = 0.3 * rnd(B, C, seed=23), 0.3 * rnd(B, C, seed=24)
    sg, sb = torch.tensor([0.9]), torch.tensor([1.1])
    lens = torch.tensor([41, 17, 5], dtype=torch.int32)
    pad = torch.arange(L)[None, :] >= lens[:, None]
    d = lambda t: t.to(DEV)  # noqa: E731
    ln = F.layer_norm(h, (C,), w, b)
    want = (O.film(ln, gam[:, None], bet[:, None], sg, sb) @ lw + lb).masked_fill(pad, 0.0)
    got = ops.predictor_head(d(h), d(w), d(b), d(lw), d(lb), film=(d(gam), d(bet), d(sg), d(sb)), lens=d(lens))
    assert (got.cpu() - want).abs().max().item() <= 2e-5
    want2 = (ln @ lw + lb).masked_fill(pad, 0.0)
    got2 = ops.predictor_head(d(h), d(w), d(b), d(lw), d(lb), lens=d(lens))
    assert (got2.cpu() - want2).abs().max().item() <= 2e-5

    # bucketize + embedding add: indices bit-exact, including exact edges / out of range / NaN
    edges = torch.linspace(-2.9, 11.4, 255)
    pred = torch.cat([rnd(100, seed=25, scale=4.0), edges[::9], torch.tensor([-50.0, 99.0, float("nan")])])
    n = pred.numel()
    pred = pred.view(1, n).contiguous()
    table = rnd(256, C, seed=26)
    x = rnd(1, n, C, seed=27)
    control = 1.25
    want_idx = torch.bucketize(pred * control, edges)
    xd, pd = d(x).clone(), d(pred).clone()
    idx = ops.bucket_embed_add(pd, control, d(edges), d(table), xd, want_idx=True)
    assert torch.equal(idx.cpu().long(), want_idx)
    assert torch.equal(xd.cpu(), x + table[want_idx])  # one fp32 add per element: exact
    assert torch.equal(pd.cpu()[~pred.isnan()], (pred * control)[~pred.isnan()])


def test_length_regulator_bit_exact(ops, golden):
    g = golden("length_regulator")
    x = torch.from_numpy(g["x"])
    for dur, want, want_len, max_len in ((g["d"], g["out"], g["mel_len"], None), (g["d"], g["out_fix"], g["mel_len_fix"], 24),
                                         (g["d_int"].astype(np.float32), g["out_int"], g["mel_len_int"], None)):
        d_t = torch.from_numpy(np.ascontiguousarray(dur)).to(DEV)
        _, ends, lens2 = ops.duration_scan(None, 1.0, durations=d_t)
        assert np.array_equal(lens2[0].cpu().numpy(), want_len)
        T = int(lens2[0].max()) if max_len is None else max_len
        out, m2p = ops.length_regulate(x.to(DEV), ends, T, want_mel2phone=True)
        assert np.array_equal(out.cpu().numpy(), want)
        _, _, m2p_ref = O.length_regulate(x, torch.from_numpy(np.ascontiguousarray(dur)), max_len)
        assert np.array_equal(m2p.cpu().numpy(), m2p_ref.numpy())
    # masks
    lens = torch.from_numpy(g["lens"]).to(torch.int32).to(DEV)
    assert np.array_equal(ops.length_mask(lens, 7).cpu().numpy(), g["mask"])
    assert np.array_equal(ops.length_mask(lens, 9).cpu().numpy(), g["mask9"])


def test_length_regulator_full_size_properties(ops):
    """BASELINE-sized case (B=64, L=120, C=256, ~700 frames): against the C oracle on a sample of
    rows, plus size-independent properties (lengths = sum of repeats; every output row equals its
    source row; padding rows are exactly the position table)."""
    B, L, C = 64, 120, 256
    g = torch.Generator().manual_seed(31)
    x = torch.randn(B, L, C, generator=g)
    log_d = torch.randn(B, L, generator=g) * 0.35 + 2.0
    src_lens = torch.randint(80, 121, (B,), generator=g)
    log_d = log_d.masked_fill(torch.arange(L)[None, :] >= src_lens[:, None], 0.0)  # predictor output is masked
    # keep clear of .5 boundaries so libm expf vs CUDA expf cannot matter for this test
    frac = (torch.exp(log_d) - 1) % 1.0
    log_d = torch.where((frac - 0.5).abs() < 1e-3, log_d + 0.01, log_d)
    d_rounded, ends, lens2 = ops.duration_scan(log_d.to(DEV), 1.0)
    ld = np.ascontiguousarray(log_d.numpy().reshape(-1))
    want_d = np.empty_like(ld)
    intops().ss_oracle_round_durations(ld.ctypes.data_as(F32P), 1.0, ld.size, want_d.ctypes.data_as(F32P))
    assert np.array_equal(d_rounded.cpu().numpy().reshape(-1), want_d)
    reps = d_rounded.cpu().clamp(min=0).long()
    assert torch.equal(ends.cpu().long(), reps.cumsum(1))
    assert torch.equal(lens2[0].cpu().long(), reps.sum(1))
    assert torch.equal(lens2[1].cpu(), d_rounded.cpu().sum(-1).int())
    T = int(lens2[0].max())
    pe = torch.randn(T + 5, C, generator=g)
    out, m2p = ops.length_regulate(x.to(DEV), ends, T, pe=pe.to(DEV), want_mel2phone=True)
    out, m2p = out.cpu(), m2p.cpu().long()
    xs = np.ascontiguousarray(x.numpy())
    dn = np.ascontiguousarray(d_rounded.cpu().numpy())
    ml = np.zeros(B, dtype=np.int64)
    ref = np.empty((B, T, C), dtype=np.float32)
    ref_m2p = np.empty((B, T), dtype=np.int64)
    intops().ss_oracle_length_regulate(xs.ctypes.data_as(F32P), dn.ctypes.data_as(F32P), B, L, C, -1,
                                       ref.ctypes.data_as(F32P), ml.ctypes.data_as(I64P), ref_m2p.ctypes.data_as(I64P))
    assert np.array_equal(m2p.numpy(), ref_m2p)
    assert torch.equal(out, torch.from_numpy(ref) + pe[:T][None])
    # without the position add the copy is verbatim
    out2 = ops.length_regulate(x.to(DEV), ends, T).cpu()
    assert np.array_equal(out2.numpy(), ref)


def test_small_rowwise_kernels(ops):
    B, L, D, V = 3, 19, 256, 361
    texts = torch.randint(0, V, (B, L), generator=torch.Generator().manual_seed(40))
    emb, pe = rnd(V, D, seed=41), rnd(L + 2, D, seed=42)
    got = ops.embed_posenc(texts.to(DEV), emb.to(DEV), pe.to(DEV))
    assert torch.equal(got.cpu(), emb[texts] + pe[:L][None])
    x = rnd(B, L, D, seed=43)
    assert torch.equal(ops.add_position(x.to(DEV), pe.to(DEV)).cpu(), x + pe[:L][None])
    got = ops.mean_pool(x.to(DEV)).cpu()
    assert (got - x.mean(dim=1)).abs().max().item() <= 1e-6
    spk = torch.tensor([2, 0, 1])
    table = rnd(3, D, seed=44)
    xd = x.to(DEV).clone()
    ops.add_speaker(xd, spk.to(DEV), table.to(DEV))
    assert torch.equal(xd.cpu(), x + table[spk][:, None, :])


def test_hifigan_post_and_pcm(ops):
    B, T, C, k = 2, 700, 32, 7
    x = rnd(B, T, C, seed=45)
    w, b = rnd(1, C, k, seed=46, scale=0.4), torch.tensor([0.05])
    ref = torch.tanh(F.conv1d(F.leaky_relu(x, 0.01).transpose(1, 2), w, b, padding=3)).squeeze(1)
    wave, pcm = ops.hifigan_post(x.to(DEV), w[0].t().contiguous().reshape(-1).to(DEV), b.to(DEV), k=k, want_pcm=True)
    assert (wave.cpu() - ref).abs().max().item() <= 1e-5
    # the int16 epilogue is exact given the kernel's own float wave
    flat = np.ascontiguousarray(wave.cpu().numpy().reshape(-1))
    want = np.empty(flat.size, dtype=np.int16)
    intops().ss_oracle_wave_int16(flat.ctypes.data_as(F32P), flat.size, 32768.0,
                                  want.ctypes.data_as(ctypes.POINTER(ctypes.c_int16)))
    assert np.array_equal(pcm.cpu().numpy().reshape(-1), want)


@pytest.mark.parametrize("where", ["device", "pinned_host"])
def test_hifigan_post_trim_writes_ragged_pcm(ops, where):
    """each utterance's first n_samples int16 samples, back to back, and nothing else - into device
    memory or straight into page-locked host memory.  Bit-exact against the padded PCM epilogue."""
    B, T, C, k = 5, 1000, 32, 7
    x = rnd(B, T, C, seed=47).to(DEV)
    w, b = rnd(1, C, k, seed=48, scale=0.4)[0].t().contiguous().reshape(-1).to(DEV), torch.tensor([0.05]).to(DEV)
    _, pcm = ops.hifigan_post(x, w, b, k=k, want_pcm=True)
    n = [1000, 0, 257, 999, 1]          # full, empty, one past a CTA tile, ...
    slack = [3, 0, 5, 0, 2]             # gaps between utterances must stay untouched
    offs, pos = [], 0
    for ni, sl in zip(n, slack):
        offs.append(pos)
        pos += ni + sl
    buf = torch.full((pos + 7,), -12345, dtype=torch.int16)
    buf = buf.to(DEV) if where == "device" else buf.pin_memory()
    out = ops.hifigan_post_trim(x, w, b, torch.tensor(n, dtype=torch.int32, device=DEV),
                                torch.tensor(offs, dtype=torch.int64, device=DEV), buf, k=k)
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    want = np.full(pos + 7, -12345, dtype=np.int16)
    for i in range(B):
        want[offs[i]:offs[i] + n[i]] = pcm[i, :n[i]].cpu().numpy()
    assert np.array_equal(got, want)
    # n_samples beyond T is clamped
    big = torch.zeros(T + 4, dtype=torch.int16, device=DEV)
    ops.hifigan_post_trim(x[:1], w, b, torch.tensor([T + 100], dtype=torch.int32, device=DEV),
                          torch.zeros(1, dtype=torch.int64, device=DEV), big, k=k)
    assert torch.equal(big[:T], pcm[0]) and int(big[T:].abs().sum()) == 0


def test_errors_cross_the_abi_as_exceptions(ops):
    from speakingstyle_b200._lib import SSError

    with pytest.raises(SSError, match="CUDA tensor"):
        ops.mean_pool(torch.zeros(1, 2, 4))
    with pytest.raises(SSError, match="multiples of 4"):
        ops.conv1d(torch.zeros(1, 4, 6, device=DEV), ops.pack_conv(torch.zeros(8, 6, device=DEV)))
    with pytest.raises(SSError, match="dk="):
        ops.attention(torch.zeros(1, 4, 3 * 48, device=DEV), torch.ones(1, dtype=torch.int32, device=DEV), 1)
