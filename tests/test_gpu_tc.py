"""tcgen05 / TMA convolution kernel (ss_conv1d_tc) through the C ABI against torch on the SAME
bf16-rounded operands: with exactly representable inputs the only difference left is the fp32
accumulation order, so fp32 outputs are held to 2e-4 and bf16 outputs to one bf16 rounding."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def rnd(*shape, seed=0, scale=1.0):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed)) * scale


def bf(x):
    return x.to(torch.bfloat16).float()


@pytest.fixture(scope="module")
def ops():
    from speakingstyle_b200 import ops as o

    return o


def ref_conv(x, w, b, k, dil):
    pad = (k - 1) // 2 * dil
    return F.conv1d(x.transpose(1, 2).double(), w.double(), None if b is None else b.double(), padding=pad,
                    dilation=dil).transpose(1, 2)


@pytest.mark.parametrize("B,S,Cin,Cout,k,dil", [
    (2, 300, 256, 1024, 9, 1),   # FFT w_1 (BN=256, 36 K-chunks)
    (2, 130, 1024, 256, 1, 1),   # FFT w_2
    (8, 700, 256, 1024, 3, 1),   # > 148 tiles: persistent loop + both accumulator buffers
    (2, 100, 80, 512, 7, 1),     # conv_pre: Cin not a multiple of the 64-channel chunk (TMA zero fill)
    (3, 90, 512, 80, 5, 1),      # PostNet last layer: Cout < BN
    (2, 500, 128, 128, 7, 3),    # BN=128, dilated
    (2, 777, 64, 64, 11, 5),     # BN=64, dilation 5, halo 25
    (1, 1000, 32, 32, 3, 1),     # BN=32, Cin half a chunk
    (2, 5, 256, 256, 3, 1),      # shorter than one tile and than the halo
])
def test_conv1d_tc_matches_torch(ops, B, S, Cin, Cout, k, dil):
    x = bf(rnd(B, S, Cin, seed=1))
    w = bf(rnd(Cout, Cin, k, seed=2, scale=1 / math.sqrt(Cin * k)))
    b = rnd(Cout, seed=3)
    ref = ref_conv(x, w, b, k, dil)
    y, _ = ops.conv1d_tc(x.to(DEV).to(torch.bfloat16), ops.pack_conv_tc(w.to(DEV)), b.to(DEV), dilation=dil,
                         pad=(k - 1) // 2 * dil)
    err = (y.cpu().double() - ref).abs().max().item()
    assert err <= 2e-4, err


def test_conv1d_tc_epilogues(ops):
    B, S, Cin, Cout, k = 3, 200, 64, 128, 3
    x, w, b = bf(rnd(B, S, Cin, seed=4)), bf(rnd(Cout, Cin, k, seed=5, scale=0.1)), rnd(Cout, seed=6)
    res = rnd(B, S, Cout, seed=7)
    lens = torch.tensor([200, 77, 1], dtype=torch.int32)
    keep = (torch.arange(S)[None, :] < lens[:, None])[..., None]
    conv = ref_conv(x, w, b, k, 1).float()
    xd, wd, bd = x.to(DEV).to(torch.bfloat16), ops.pack_conv_tc(w.to(DEV)), b.to(DEV)
    # relu + mask, bf16 output
    y, _ = ops.conv1d_tc(xd, wd, bd, pad=1, act=ops.ACT_RELU, lens=lens.to(DEV), out_dtype=torch.bfloat16)
    want = F.relu(conv) * keep
    assert y.dtype == torch.bfloat16
    assert (y.float().cpu() - want).abs().max().item() <= 2e-4 + want.abs().max().item() * 2 ** -8
    # tanh + fp32 residual, fp32 y plus activated bf16 copy
    y, ya = ops.conv1d_tc(xd, wd, bd, pad=1, act=ops.ACT_TANH, residual=res.to(DEV), act2=ops.ACT_LRELU, slope2=0.1)
    want = torch.tanh(conv) + res
    assert (y.cpu() - want).abs().max().item() <= 2e-4
    assert torch.equal(ya.cpu(), F.leaky_relu(y.cpu(), 0.1).to(torch.bfloat16))
    # bf16 residual
    y, _ = ops.conv1d_tc(xd, wd, bd, pad=1, residual=res.to(DEV).to(torch.bfloat16))
    assert (y.cpu() - (conv + bf(res))).abs().max().item() <= 2e-4
    # running mean: two accumulations, no y; y_act follows the accumulated value
    acc = torch.full((B, S, Cout), 3.0, device=DEV)
    ops.conv1d_tc(xd, wd, bd, pad=1, want_y=False, accum=acc, accum_scale=0.5, accum_init=True)
    _, ya = ops.conv1d_tc(xd, wd, bd, pad=1, act=ops.ACT_LRELU, slope=0.2, want_y=False, accum=acc, accum_scale=0.25,
                          act2=ops.ACT_NONE)
    want = 0.5 * conv + 0.25 * F.leaky_relu(conv, 0.2)
    assert (acc.cpu() - want).abs().max().item() <= 2e-4
    assert torch.equal(ya.cpu(), acc.cpu().to(torch.bfloat16))


@pytest.mark.parametrize("Cin,Cout,k,u,S", [(512, 256, 16, 8, 50), (256, 128, 16, 8, 333), (128, 64, 4, 2, 1000),
                                             (64, 32, 4, 2, 129)])
def test_conv_transpose_tc(ops, Cin, Cout, k, u, S):
    B = 2
    x = bf(F.leaky_relu(rnd(B, S, Cin, seed=8), 0.1))
    w = bf(rnd(Cin, Cout, k, seed=9, scale=1 / math.sqrt(Cin * k / u)))
    b = rnd(Cout, seed=10)
    pad = (k - u) // 2
    ref = F.conv_transpose1d(x.transpose(1, 2).double(), w.double(), b.double(), stride=u, padding=pad).transpose(1, 2)
    taps = k // u
    w4 = w.reshape(Cin, Cout, taps, u).permute(2, 3, 1, 0).flip(0).reshape(taps, u * Cout, Cin).contiguous()
    y, ya = ops.conv_transpose1d_tc(x.to(DEV).to(torch.bfloat16), w4.to(DEV).to(torch.bfloat16), b.repeat(u).to(DEV),
                                    cout=Cout, stride=u, pad=pad, act2=ops.ACT_LRELU, slope2=0.1)
    assert y.shape == ref.shape
    assert (y.cpu().double() - ref).abs().max().item() <= 2e-4
    assert torch.equal(ya.cpu(), F.leaky_relu(y.cpu(), 0.1).to(torch.bfloat16))


def test_cast_bf16(ops):
    x = rnd(3, 50, 80, seed=11)
    assert torch.equal(ops.cast_bf16(x.to(DEV)).cpu(), x.to(torch.bfloat16))
    assert torch.equal(ops.cast_bf16(x.to(DEV), ops.ACT_LRELU, 0.1).cpu(), F.leaky_relu(x, 0.1).to(torch.bfloat16))


@pytest.mark.parametrize("B,S,Cin,Cout,k,dil", [(5, 1500, 32, 32, 11, 1), (4, 900, 64, 64, 7, 3), (3, 700, 128, 128, 3, 5),
                                                (3, 400, 256, 256, 11, 1)])
def test_conv1d_tc_ragged_schedule(ops, B, S, Cin, Cout, k, dil):
    """rows[b] on the bf16 kernel (scheduling unit = 256 / BN row tiles): wanted rows bit-identical, the rest untouched."""
    x = rnd(B, S, Cin, seed=21).to(DEV).to(torch.bfloat16)
    w = ops.pack_conv_tc(rnd(Cout, Cin, k, seed=22, scale=1 / math.sqrt(Cin * k)).to(DEV))
    b = rnd(Cout, seed=23).to(DEV)
    res = rnd(B, S, Cout, seed=24).to(DEV)
    rows_host = [S, 0, 1, S // 2 + 3, 130][:B]
    rows = torch.tensor(rows_host, dtype=torch.int32, device=DEV)
    pad = (k - 1) // 2 * dil
    kw = dict(dilation=dil, pad=pad, residual=res, act2=ops.ACT_LRELU, slope2=0.1)
    y_full, a_full = ops.conv1d_tc(x, w, b, **kw)
    out = torch.full((B, S, Cout), 7.5, device=DEV)
    y, a = ops.conv1d_tc(x, w, b, out=out, rows=rows, **kw)
    bn = 256 if Cout >= 256 else 128 if Cout > 64 else 64 if Cout > 32 else 32
    unit = 128 * max(1, 256 // bn)
    for i, r in enumerate(rows_host):
        assert torch.equal(y[i, :r], y_full[i, :r]) and torch.equal(a[i, :r], a_full[i, :r]), (i, r)
        done = min(S, -(-r // unit) * unit)
        assert bool((out[i, done:] == 7.5).all()), (i, r)


@pytest.mark.parametrize("B,S,C,k,dil", [(3, 300, 32, 3, 1), (2, 1000, 32, 11, 5), (3, 500, 64, 7, 3), (2, 900, 64, 11, 1),
                                         (2, 700, 128, 11, 5), (3, 260, 256, 3, 1), (2, 300, 256, 7, 1)])
def test_bf16_panel_kernel_matches_conv1d_tc(ops, B, S, C, k, dil):
    """ss_conv1d_bf16_panel (what conv1d_tc dispatches to when it is given a ConvW of a shape that kernel takes) against
    ss_conv1d_tc on the same bf16 operands: same products, fp32 accumulation in a different order."""
    x = rnd(B, S, C, seed=21).to(DEV).to(torch.bfloat16)
    w = rnd(C, C, k, seed=22, scale=1.0 / math.sqrt(C * k)).to(DEV)
    b = rnd(C, seed=23).to(DEV)
    res = rnd(B, S, C, seed=24).to(DEV)
    W = ops.pack_conv(w)
    assert W.bf16p is not None
    pad = (k - 1) // 2 * dil
    y0, a0 = ops.conv1d_tc(x, W.bf16, b, dilation=dil, pad=pad, residual=res, act2=ops.ACT_LRELU, slope2=0.1)
    y1, a1 = ops.conv1d_tc(x, W, b, dilation=dil, pad=pad, residual=res, act2=ops.ACT_LRELU, slope2=0.1)
    assert (y0 - y1).abs().max().item() <= 2e-5 * y0.abs().max().item()
    assert (a0.float() - a1.float()).abs().max().item() <= 2.0 ** -7 * y0.abs().max().item()
    rows = torch.tensor([S, 129, 0][:B], dtype=torch.int32, device=DEV)
    out = torch.full_like(y1, -5.0)
    ops.conv1d_tc(x, W, b, dilation=dil, pad=pad, residual=res, rows=rows, out=out)
    for i, r in enumerate(rows.tolist()):
        done = min(S, (r + 127) // 128 * 128)
        assert torch.equal(out[i, :r], y1[i, :r]) and bool((out[i, done:] == -5.0).all())


@pytest.mark.parametrize("Cin,Cout,k,u,S", [(512, 256, 16, 8, 50), (128, 64, 4, 2, 1000), (64, 32, 4, 2, 2001)])
def test_bf16_panel_conv_transpose(ops, Cin, Cout, k, u, S):
    x = rnd(2, S, Cin, seed=25).to(DEV).to(torch.bfloat16)
    w = rnd(Cin, Cout, k, seed=26, scale=1.0 / math.sqrt(Cin * k / u)).to(DEV)
    b = rnd(Cout, seed=27).to(DEV).repeat(u).contiguous()
    W = ops.pack_conv_transpose(w, u)
    assert W.bf16p is not None
    y0, a0 = ops.conv_transpose1d_tc(x, W.bf16, b, cout=Cout, stride=u, pad=(k - u) // 2, act2=ops.ACT_LRELU, slope2=0.1)
    y1, a1 = ops.conv_transpose1d_tc(x, W, b, cout=Cout, stride=u, pad=(k - u) // 2, act2=ops.ACT_LRELU, slope2=0.1)
    assert (y0 - y1).abs().max().item() <= 2e-5 * y0.abs().max().item()
    assert (a0.float() - a1.float()).abs().max().item() <= 2.0 ** -7 * y0.abs().max().item()
