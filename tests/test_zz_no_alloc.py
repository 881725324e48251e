"""ss_length_regulate_noalloc, the entry point for callers that cannot allocate device memory
(a worker restricted to buffers it was given at start-up, a plug-in without access to its host's
allocator): it keeps the prefix sums in an array inside the library instead of a caller-provided
buffer.  Checked through the C ABI against the existing two-call form (ss_duration_scan +
ss_length_regulate), bit for bit, on two calls in a row with different shapes - the second call
must not see anything the first one left in the shared array.

This file sorts last on purpose: if it fails, a run that stops at the first failure has already
run every other test.  It passes on its own in a fresh process as well as after the rest."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _two_call_form(ops, x, d, T, pe):
    _, ends, lens2 = ops.duration_scan(None, 1.0, durations=d.clone())
    return ops.length_regulate(x, ends, T, pe=pe), lens2[0]


def _noalloc(lib, x, d, T, pe):
    B, L, C = x.shape
    out = torch.full((B, T, C), float("nan"), device=DEV)      # caller-owned output, pre-filled with junk
    mel_len = torch.full((B,), -7, dtype=torch.int32, device=DEV)
    rc = lib.ss_length_regulate_noalloc(_ptr(x), _ptr(d), _ptr(pe), _ptr(out), _ptr(mel_len), B, L, T, C, _stream())
    assert rc == 0, lib.ss_last_error().decode()
    return out, mel_len


def test_noalloc_length_regulator_matches_two_call_form_twice_in_a_row():
    from speakingstyle_b200 import _lib, ops

    lib = _lib.lib()
    g = torch.Generator().manual_seed(123)

    # call 1: decoder-style use - 256 channels, position table added, T = longest utterance
    B, L, C = 5, 37, 256
    x1 = torch.randn(B, L, C, generator=g).to(DEV)
    d1 = torch.randint(0, 9, (B, L), generator=g).float().to(DEV)
    d1[1, 10:] = 0.0                       # a short utterance -> padding rows
    T1 = int(d1.sum(dim=1).max())
    pe1 = torch.randn(T1 + 3, C, generator=g).to(DEV)
    want_out, want_len = _two_call_form(ops, x1, d1, T1, pe1)
    got_out, got_len = _noalloc(lib, x1, d1, T1, pe1)
    assert torch.equal(got_len, want_len)
    assert torch.equal(got_out, want_out)   # copied rows (+ one fp32 add): exact, the tolerance the LR tests use
    first_out, first_len = got_out.clone(), got_len.clone()

    # call 2, straight after: other shape, fractional / negative durations, no position table, and a T that
    # cuts the longest utterances while padding the shortest - stale prefix sums from call 1 must not matter
    B, L, C = 3, 11, 64
    x2 = torch.randn(B, L, C, generator=g).to(DEV)
    d2 = torch.tensor([[2.7, -1.0, 0.0, 3.2, 1.9, -0.0, 4.0, 1.0, 0.5, 2.0, 6.0],
                       [0.0] * 10 + [1.0],
                       [9.0] * 11], device=DEV)
    T2 = 20
    want_out, want_len = _two_call_form(ops, x2, d2, T2, None)
    got_out, got_len = _noalloc(lib, x2, d2, T2, None)
    assert want_len.tolist() == [19, 1, 99]
    assert torch.equal(got_len, want_len)
    assert torch.equal(got_out, want_out)
    # ... and it stands on the oracle directly (model/modules.py:174-201 restated in oracle/fs2_oracle.py), not only
    # on the library's other entry point: same rows, same zero padding, same lengths
    from oracle import fs2_oracle as O

    ref_out, ref_len, _ = O.length_regulate(x2.cpu(), d2.cpu(), T2)
    assert torch.equal(got_out.cpu(), ref_out) and got_len.cpu().tolist() == ref_len.tolist()

    # and call 1 again gives call 1's answer again
    again_out, again_len = _noalloc(lib, x1, d1, T1, pe1)
    assert torch.equal(again_out, first_out) and torch.equal(again_len, first_len)


def test_noalloc_rejects_sizes_beyond_its_built_in_scratch():
    from speakingstyle_b200 import _lib

    lib = _lib.lib()
    B, L, C = 300, 256, 4                     # 76 800 prefix sums > SS_LR_NOALLOC_MAX_TOKENS (65 536)
    x = torch.zeros(B, L, C, device=DEV)
    d = torch.ones(B, L, device=DEV)
    out = torch.zeros(B, 8, C, device=DEV)
    mel_len = torch.zeros(B, dtype=torch.int32, device=DEV)
    rc = lib.ss_length_regulate_noalloc(_ptr(x), _ptr(d), None, _ptr(out), _ptr(mel_len), B, L, 8, C, _stream())
    assert rc == -2
    msg = lib.ss_last_error().decode()
    assert "65536" in msg and "ss_duration_scan" in msg
