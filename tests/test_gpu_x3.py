"""Split-bf16 ("bf16x3") convolution kernel (ss_conv1d_bf16x3) through the C ABI against an fp64 torch convolution of the
fp32 operands.  Three bf16 products per multiply leave ~2^-16 relative per product, so outputs are held to 3e-5 of
the output scale (the 3xTF32 kernels: 1e-6; a plain bf16 convolution: 1e-2)."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
TOL = 3e-5


def rnd(*shape, seed=0, scale=1.0):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed)) * scale


@pytest.fixture(scope="module")
def ops():
    from speakingstyle_b200 import ops as o

    return o


def ref_conv(x, w, b, k, dil):
    pad = (k - 1) // 2 * dil
    return F.conv1d(x.transpose(1, 2).double(), w.double(), None if b is None else b.double(), padding=pad,
                    dilation=dil).transpose(1, 2)


def unpair(p):
    return p[..., 0, :].double() + p[..., 1, :].double()


def test_split_bf16_pairs(ops):
    x = rnd(3, 50, 64, seed=1, scale=3.0).to(DEV)
    p = ops.split_bf16(x)
    assert p.shape == (3, 50, 2, 64) and p.dtype == torch.bfloat16
    assert torch.equal(p[..., 0, :], x.to(torch.bfloat16))
    assert (unpair(p) - x.double()).abs().max().item() <= 2.0 ** -17 * x.abs().max().item()
    q = ops.split_bf16(x, act=ops.ACT_LRELU, slope=0.1)
    assert (unpair(q) - F.leaky_relu(x, 0.1).double()).abs().max().item() <= 2.0 ** -17 * x.abs().max().item()


@pytest.mark.parametrize("B,S,C,k,dil", [
    (3, 300, 32, 3, 1), (2, 1000, 32, 11, 5), (5, 257, 32, 7, 3), (40, 1500, 32, 11, 1),   # narrow: two taps per weight row
    (3, 300, 64, 3, 1), (2, 900, 64, 11, 1), (3, 500, 64, 7, 5),                          # 64 channels: resident / ring
    (3, 300, 128, 3, 1), (2, 700, 128, 11, 5), (12, 1300, 128, 7, 1),                      # two chunks, weight ring
    (2, 400, 256, 7, 3), (2, 300, 256, 11, 5), (3, 260, 256, 3, 1),                        # two column tiles
])
def test_conv1d_x3_matches_fp64(ops, B, S, C, k, dil):
    x = rnd(B, S, C, seed=2).to(DEV)
    w = rnd(C, C, k, seed=3, scale=1.0 / math.sqrt(C * k)).to(DEV)
    b = rnd(C, seed=4).to(DEV)
    W = ops.pack_conv(w)
    assert W.bf16x3 is not None
    pad = (k - 1) // 2 * dil
    y, yp = ops.conv1d_x3(ops.split_bf16(x), W, b, dilation=dil, pad=pad, act2=ops.ACT_LRELU, slope2=0.1)
    ref = ref_conv(x, w, b, k, dil)
    scale = ref.abs().max().item()
    assert (y.double() - ref).abs().max().item() <= TOL * scale
    assert (unpair(yp) - F.leaky_relu(y, 0.1).double()).abs().max().item() <= 2.0 ** -16 * scale
    torch.cuda.synchronize()


def test_conv1d_x3_epilogues(ops):
    B, S, C, k = 3, 400, 64, 7
    x = rnd(B, S, C, seed=5).to(DEV)
    w = rnd(C, C, k, seed=6, scale=1.0 / math.sqrt(C * k)).to(DEV)
    b = rnd(C, seed=7).to(DEV)
    res = rnd(B, S, C, seed=8).to(DEV)
    lens = torch.tensor([400, 130, 0], dtype=torch.int32, device=DEV)
    W = ops.pack_conv(w)
    xp = ops.split_bf16(x)
    ref = ref_conv(x, w, b, k, 1)
    mask = (torch.arange(S, device=DEV)[None, :] < lens[:, None])[..., None]
    # activation, residual, row mask, in-place over the residual
    out = res.clone()
    y, _ = ops.conv1d_x3(xp, W, b, pad=3, act=ops.ACT_LRELU, slope=0.1, residual=out, lens=lens, out=out)
    want = (F.leaky_relu(ref, 0.1) + res.double()) * mask
    assert y.data_ptr() == out.data_ptr()
    assert (y.double() - want).abs().max().item() <= TOL * want.abs().max().item()
    # running mean: init, then accumulate; the pair copy follows the accumulated value
    acc = torch.full((B, S, C), 7.0, device=DEV)
    ops.conv1d_x3(xp, W, b, pad=3, want_y=False, accum=acc, accum_scale=0.5, accum_init=True)
    _, yp = ops.conv1d_x3(xp, W, b, pad=3, residual=res, want_y=False, accum=acc, accum_scale=0.25, act2=ops.ACT_LRELU,
                          slope2=0.1)
    want = 0.5 * ref + 0.25 * (ref + res.double())
    assert (acc.double() - want).abs().max().item() <= TOL * want.abs().max().item()
    assert (unpair(yp) - F.leaky_relu(acc, 0.1).double()).abs().max().item() <= 2.0 ** -16 * want.abs().max().item()


@pytest.mark.parametrize("B,S,C,k,dil", [(5, 1500, 32, 11, 1), (4, 900, 64, 7, 3), (3, 700, 128, 3, 5), (6, 520, 256, 11, 1)])
def test_conv1d_x3_ragged_schedule(ops, B, S, C, k, dil):
    """rows[b] wanted rows per utterance: those rows are bit-identical to the full computation, rows of tiles past them
    are left untouched."""
    x = rnd(B, S, C, seed=9).to(DEV)
    w = rnd(C, C, k, seed=10, scale=1.0 / math.sqrt(C * k)).to(DEV)
    b = rnd(C, seed=11).to(DEV)
    W = ops.pack_conv(w)
    xp = ops.split_bf16(x)
    pad = (k - 1) // 2 * dil
    full, fullp = ops.conv1d_x3(xp, W, b, dilation=dil, pad=pad, act2=ops.ACT_LRELU, slope2=0.1)
    rows = torch.tensor([S, 1, 0, 129, S - 1, 300][:B], dtype=torch.int32, device=DEV)
    out = torch.full((B, S, C), -5.0, device=DEV)
    outp = torch.full((B, S, 2, C), -5.0, device=DEV, dtype=torch.bfloat16)
    ops.conv1d_x3(xp, W, b, dilation=dil, pad=pad, act2=ops.ACT_LRELU, slope2=0.1, rows=rows, out=out, out_pair=outp)
    for i, r in enumerate(rows.tolist()):
        done = min(S, (r + 127) // 128 * 128)
        assert torch.equal(out[i, :r], full[i, :r]) and torch.equal(outp[i, :r], fullp[i, :r])
        assert bool((out[i, done:] == -5.0).all()) and bool((outp[i, done:] == -5.0).all())


@pytest.mark.parametrize("Cin,Cout,k,u,S", [(512, 256, 16, 8, 50), (256, 128, 16, 8, 333), (128, 64, 4, 2, 1000),
                                            (64, 32, 4, 2, 2001)])
def test_conv_transpose_x3(ops, Cin, Cout, k, u, S):
    """HiFi-GAN's four up-sampling layers (hifigan/models.py:124-127) in phase form, pair rows in and out."""
    B = 2
    x = rnd(B, S, Cin, seed=12).to(DEV)
    w = rnd(Cin, Cout, k, seed=13, scale=1.0 / math.sqrt(Cin * k / u)).to(DEV)
    b = rnd(Cout, seed=14).to(DEV)
    W = ops.pack_conv_transpose(w, u)
    assert W.bf16x3 is not None
    pad = (k - u) // 2
    y, yp = ops.conv_transpose1d_x3(ops.split_bf16(x), W, b.repeat(u).contiguous(), cout=Cout, stride=u, pad=pad,
                                    act2=ops.ACT_LRELU, slope2=0.1)
    ref = F.conv_transpose1d(x.transpose(1, 2).double(), w.double(), b.double(), stride=u, padding=pad).transpose(1, 2)
    assert y.shape == ref.shape and yp.shape == (B, ref.shape[1], 2, Cout)
    scale = ref.abs().max().item()
    assert (y.double() - ref).abs().max().item() <= TOL * scale
    assert (unpair(yp) - F.leaky_relu(y, 0.1).double()).abs().max().item() <= 2.0 ** -16 * scale
    # ragged: rows of the phase-form GEMM (input positions)
    rows = torch.tensor([S // 3, S], dtype=torch.int32, device=DEV)
    y2, _ = ops.conv_transpose1d_x3(ops.split_bf16(x), W, b.repeat(u).contiguous(), cout=Cout, stride=u, pad=pad,
                                    act2=ops.ACT_LRELU, slope2=0.1, rows=rows)
    n = (S // 3 - 1) * u - pad  # samples fully determined by GEMM rows < S // 3
    assert torch.equal(y2[0, :n], y[0, :n]) and torch.equal(y2[1], y[1])


def test_conv1d_x3_plan_rejects_other_shapes(ops):
    assert ops.pack_conv(torch.zeros(80, 80, 3, device=DEV)).bf16x3 is None
    assert ops.pack_conv(torch.zeros(32, 64, 3, device=DEV)).bf16x3 is None  # [Cout, Cin, k]: Cin 64 -> Cout 32


@pytest.mark.parametrize("C,k,dil,S", [(32, 3, 1, 1000), (32, 3, 5, 126), (32, 7, 3, 777), (32, 11, 1, 2500), (32, 11, 5, 118),
                                       (32, 11, 3, 4000), (64, 3, 1, 1000), (64, 3, 3, 126), (64, 3, 5, 5000)])
def test_resunit_is_bit_identical_to_two_launches(ops, C, k, dil, S):
    """ss_resunit_bf16x3 (c1 -> leaky ReLU -> c2 + residual in one launch, intermediate in shared memory) against the
    two ss_conv1d_bf16x3 launches it replaces: same MMAs in the same order, same epilogue arithmetic."""
    B = 3
    L = ops.ACT_LRELU
    y = rnd(B, S, C, seed=31).to(DEV)
    xp = ops.split_bf16(y, act=L, slope=0.1)
    W1 = ops.pack_conv(rnd(C, C, k, seed=32, scale=1.0 / math.sqrt(C * k)).to(DEV))
    W2 = ops.pack_conv(rnd(C, C, k, seed=33, scale=1.0 / math.sqrt(C * k)).to(DEV))
    b1, b2 = rnd(C, seed=34).to(DEV), rnd(C, seed=35).to(DEV)
    assert ops.resunit_ok(W1, W2, dil)
    _, tp = ops.conv1d_x3(xp, W1, b1, dilation=dil, pad=(k - 1) // 2 * dil, want_y=False, act2=L, slope2=0.1)
    y2, yp2 = ops.conv1d_x3(tp, W2, b2, pad=(k - 1) // 2, residual=y, act2=L, slope2=0.1)
    y1, yp1 = ops.resunit_x3(xp, y, W1, b1, W2, b2, dilation=dil, slope=0.1, act2=L, slope2=0.1)
    torch.cuda.synchronize()
    assert torch.equal(y1, y2) and torch.equal(yp1, yp2)
    # running-mean form, in place over the residual, ragged rows
    acc2 = torch.full((B, S, C), 0.25, device=DEV)
    acc1 = acc2.clone()
    ops.conv1d_x3(tp, W2, b2, pad=(k - 1) // 2, residual=y, want_y=False, accum=acc2, accum_scale=1.0 / 3)
    ops.resunit_x3(xp, y, W1, b1, W2, b2, dilation=dil, slope=0.1, want_y=False, accum=acc1, accum_scale=1.0 / 3)
    assert torch.equal(acc1, acc2)
    rows = torch.tensor([S, min(S, 130), 0], dtype=torch.int32, device=DEV)
    out = torch.full((B, S, C), -5.0, device=DEV)
    outp = torch.full((B, S, 2, C), -5.0, device=DEV, dtype=torch.bfloat16)
    ops.resunit_x3(xp, y, W1, b1, W2, b2, dilation=dil, slope=0.1, act2=L, slope2=0.1, rows=rows, out=out, out_pair=outp)
    TM = 128 - (k - 1)
    for i, r in enumerate(rows.tolist()):
        done = min(S, (r + TM - 1) // TM * TM)
        assert torch.equal(out[i, :r], y2[i, :r]) and torch.equal(outp[i, :r], yp2[i, :r])
        assert bool((out[i, done:] == -5.0).all()) and bool((outp[i, done:] == -5.0).all())
