"""CPU-side checks of the boundary: the library builds, loads, and exports exactly the symbols the
header declares; the ctypes structs match the C structs; the host mirror keeps the reference's
state_dict contract.  No compute call is made (there is no GPU here)."""
import ctypes
import os
import re
import subprocess
import tempfile

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "speakingstyle_b200.h")


@pytest.fixture(scope="module")
def lib():
    from speakingstyle_b200 import _lib, build

    build.build()
    return _lib.lib()


def declared_symbols():
    text = open(HEADER).read()
    return sorted(set(re.findall(r"SS_API\s+[\w\s\*]+?\b(ss_\w+)\s*\(", text)))


def test_header_symbols_are_exported_and_bound(lib):
    from speakingstyle_b200 import _lib

    names = declared_symbols()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"
    assert names == _lib.exported_names(), "ctypes binding and header disagree"
    assert lib.ss_version() == 100
    assert lib.ss_status_string(0) == b"ok"


def test_struct_layouts_match_c(lib):
    """sizeof of every params struct as the C compiler sees it == ctypes.sizeof."""
    from speakingstyle_b200 import _lib

    src = '#include <stdio.h>\n#include "speakingstyle_b200.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu\\n",' \
          'sizeof(ss_conv1d_f32_params),sizeof(ss_layernorm_params),sizeof(ss_predictor_head_params),' \
          'sizeof(ss_conv1d_tc_params),sizeof(ss_conv1d_tf32_params),sizeof(ss_resunit_params));return 0;}'
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "s.c")
        open(c, "w").write(src)
        exe = os.path.join(d, "s")
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), c, "-o", exe])
        sizes = [int(v) for v in subprocess.check_output([exe]).split()]
    assert sizes == [ctypes.sizeof(_lib.Conv1dF32Params), ctypes.sizeof(_lib.LayerNormParams),
                     ctypes.sizeof(_lib.PredictorHeadParams), ctypes.sizeof(_lib.Conv1dTcParams),
                     ctypes.sizeof(_lib.Conv1dTf32Params), ctypes.sizeof(_lib.ResUnitParams)]


def test_compute_entry_points_reject_bad_arguments_without_a_gpu(lib):
    # argument validation happens before any CUDA call, so it is testable here
    assert lib.ss_embed_posenc(None, None, None, None, 1, 1, 4, 1, None) == -1
    assert b"null" in lib.ss_last_error()
    from speakingstyle_b200 import _lib

    p = _lib.Conv1dF32Params()
    assert lib.ss_conv1d_f32(ctypes.byref(p), None) == -1
    assert lib.ss_attention_f32(None, None, None, 1, 1, 1, 32, None) == -1
    assert lib.ss_length_regulate(None, None, None, None, None, 1, 1, 1, 4, None) == -1
    q = _lib.Conv1dTcParams()
    assert lib.ss_conv1d_bf16x3(ctypes.byref(q), None) == -1 and b"ss_conv1d_bf16x3" in lib.ss_last_error()
    assert lib.ss_conv1d_bf16_panel(ctypes.byref(q), None) == -1 and b"ss_conv1d_bf16_panel" in lib.ss_last_error()
    assert lib.ss_split_bf16(None, None, 1, 8, 0, 0.0, None) == -1


def test_state_dict_contract_and_no_cpu_fallback():
    import speakingstyle_b200 as ss
    from speakingstyle_b200 import configs as C
    from speakingstyle_b200 import synth as S

    for ds in ("BC2013", "LibriTTS"):
        d = C.write_preprocessed_dir(tempfile.mkdtemp(), ds)
        mc = C.model_config(ds)
        m = ss.FastSpeech2(C.preprocess_config(ds, d), mc)
        sd = S.fastspeech2_state_dict(mc, ds)
        m.load_state_dict(sd, strict=True)
        assert set(m.state_dict()) == set(sd)
        for k, v in m.state_dict().items():
            assert v.shape == sd[k].shape, k
    n_params = sum(p.numel() for p in m.parameters())
    assert n_params == 50_477_403 + 904 * 256 - 2 * 255 or n_params > 50_000_000
    ds, batch, _ = S.make_workload("tiny_ms")
    with pytest.raises(RuntimeError, match="CUDA"):
        m(batch["speakers"], batch["texts"], batch["src_lens"], batch["max_src_len"], batch["mels"],
          batch["mel_lens"], batch["max_mel_len"])
    with pytest.raises(NotImplementedError):
        m.train()

    g = ss.Generator(C.hifigan_config())
    vsd = S.hifigan_state_dict()
    g.load_state_dict(vsd, strict=True)
    assert sum(p.numel() for p in g.parameters()) > 13_900_000
    g.remove_weight_norm()
    folded = S.fold_weight_norm(vsd)
    assert set(g.state_dict()) == set(folded)
    for k in folded:
        assert torch.allclose(g.state_dict()[k], folded[k], atol=1e-6)
    g.load_state_dict(vsd, strict=True)  # checkpoint form into a folded module
    with pytest.raises(RuntimeError, match="CUDA"):
        g(torch.zeros(1, 80, 4))


def test_product_path_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "speakingstyle_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text, f
                assert "/root/reference" not in text, f


def test_split_bf16_weight_tiles_follow_the_header_layout(lib):
    """ops.ConvW.bf16x3 / .bf16p against a loop restatement of the tile layout include/speakingstyle_b200.h states for
    ss_conv1d_bf16x3 (no GPU: packing is host-side torch indexing; the plan call only inspects shapes)."""
    import torch

    from speakingstyle_b200 import ops

    def check(cin, cout, k):
        w = torch.randn(cout, cin, k, generator=torch.Generator().manual_seed(cin + cout + k))
        W = ops.pack_conv(w)
        bn = lib.ss_conv1d_bf16x3_plan(cin, cout, k, 1)
        assert bn in (32, 64, 128)
        hi = W.kcc.to(torch.bfloat16)
        lo = (W.kcc - hi.float()).to(torch.bfloat16)
        tiles = W.bf16x3
        plain = W.bf16p
        if cin == 32:
            npair = (k + 1) // 2
            assert tiles.shape == (cout // 32 * npair, 64, 64) and plain.shape == (cout // 32 * npair, 32, 64)
            for nt in range(cout // 32):
                for q in range(npair):
                    t = tiles[nt * npair + q]
                    for h in range(2):
                        j = 2 * q + h
                        want_hi = hi[j, nt * 32:nt * 32 + 32] if j < k else torch.zeros(32, 32, dtype=torch.bfloat16)
                        want_lo = lo[j, nt * 32:nt * 32 + 32] if j < k else torch.zeros(32, 32, dtype=torch.bfloat16)
                        assert torch.equal(t[:32, 32 * h:32 * h + 32], want_hi)
                        assert torch.equal(t[32:, 32 * h:32 * h + 32], want_lo)
                        assert torch.equal(plain[nt * npair + q][:, 32 * h:32 * h + 32], want_hi)
        else:
            nch = cin // 64
            assert tiles.shape == (cout // bn * nch * k, 2 * bn, 64)
            for nt in range(cout // bn):
                for c in range(nch):
                    for j in range(k):
                        t = tiles[(nt * nch + c) * k + j]
                        assert torch.equal(t[:bn], hi[j, nt * bn:(nt + 1) * bn, 64 * c:64 * c + 64])
                        assert torch.equal(t[bn:], lo[j, nt * bn:(nt + 1) * bn, 64 * c:64 * c + 64])
                        assert torch.equal(plain[(nt * nch + c) * k + j], t[:bn])
        # hi + lo carries 16 significant bits of the weight
        assert ((hi.float() + lo.float()) - W.kcc).abs().max() <= 2.0 ** -17 * W.kcc.abs().max()

    check(32, 32, 3)
    check(32, 32, 11)
    check(64, 64, 7)
    check(128, 128, 3)
    check(256, 1024, 2)
    assert lib.ss_conv1d_bf16x3_plan(80, 512, 7, 1) == 0 and lib.ss_conv1d_bf16x3_plan(256, 80, 1, 1) == 0
    assert lib.ss_conv1d_bf16x3_plan(32, 32, 11, 13) == 0  # 128 + (k-1)*dilation > 256 rows: no panel
    assert ops.pack_conv(torch.zeros(80, 256, 1)).bf16x3 is None
