"""Shard planner and the N>1 path on CPU: world_size-2 gloo processes, each running its own shards
through the CPU oracle (test infrastructure standing in for the GPU model), no data-path collective."""
import os
import socket
import tempfile

import pytest
import torch
import torch.multiprocessing as mp

from speakingstyle_b200 import configs as C
from speakingstyle_b200 import shard as SH
from speakingstyle_b200 import synth as S


def _lens(n, seed, lo, hi):
    return torch.randint(lo, hi + 1, (n,), generator=torch.Generator().manual_seed(seed)).tolist()


def test_plan_is_an_exact_cover_and_deterministic():
    src, ref = _lens(2048, 1, 40, 110), _lens(2048, 2, 200, 1000)   # BASELINE cfg5 lengths
    p1 = SH.plan_shards(src, ref, max_batch=32, bucket_width=16, n_ranks=8)
    p2 = SH.plan_shards(src, ref, max_batch=32, bucket_width=16, n_ranks=8)
    ids = sorted(i for s in p1.shards for i in s.utterances)
    assert ids == list(range(2048))
    assert [(s.utterances, s.rank) for s in p1.shards] == [(s.utterances, s.rank) for s in p2.shards]
    assert all(1 <= len(s.utterances) <= 32 for s in p1.shards)
    for s in p1.shards:  # padding waste bounded by the bucket width
        assert s.max_src_len - min(src[i] for i in s.utterances) < 16
    assert sorted({s.rank for s in p1.shards}) == list(range(8))
    assert p1.imbalance() < 1.05, p1.imbalance()   # LPT over ~70 shards per rank


def test_reference_sweep_plan_cfg4():
    ref = _lens(10000, 3, 100, 1000)   # BASELINE cfg4: 10k mels, bucket width 64, batch <= 64
    for n in (1, 2, 4, 8):
        p = SH.plan_shards(None, ref, max_batch=64, bucket_width=64, n_ranks=n, text=False)
        assert sorted(i for s in p.shards for i in s.utterances) == list(range(10000))
        assert all(s.max_ref_len - min(ref[i] for i in s.utterances) < 64 for s in p.shards)
        assert p.imbalance() < 1.03
    with pytest.raises(ValueError):
        SH.plan_shards([1, 2], [1], n_ranks=1)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _utterances(n, seed):
    g = torch.Generator().manual_seed(seed)
    texts = [torch.randint(1, C.N_SRC_VOCAB, (int(torch.randint(4, 12, (1,), generator=g)),), generator=g) for _ in range(n)]
    mels = [torch.randn(int(torch.randint(8, 40, (1,), generator=g)), 80, generator=g) for _ in range(n)]
    return [0] * n, texts, mels


def _oracle_compute(sd, mc):
    from oracle import fs2_oracle as O

    def compute(batch):
        with torch.no_grad():
            out = O.fastspeech2_forward(sd, mc, batch)
        return {"postnet_mel": out[1], "d_rounded": out[5], "mel_len": out[9]}

    return compute


def _worker(rank, world, port, outdir):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(2)
    mc = C.model_config("BC2013")
    sd = S.fastspeech2_state_dict(mc, "BC2013")
    speakers, texts, mels = _utterances(10, seed=5)
    plan = SH.plan_shards([t.numel() for t in texts], [m.shape[0] for m in mels], max_batch=3, bucket_width=4,
                          n_ranks=world, vocoder=False)
    res = SH.run_rank(plan, rank, _oracle_compute(sd, mc), speakers, texts, mels)
    frames = float(sum(int(r["mel_len"]) for r in res.values()))
    tot = SH.gather_counters({"frames": frames, "utterances": float(len(res))})
    torch.save({"res": res, "tot": tot, "plan": [(s.utterances, s.rank) for s in plan.shards]},
               os.path.join(outdir, f"rank{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_gloo_match_single_process():
    world = 2
    outdir = tempfile.mkdtemp()
    mp.spawn(_worker, args=(world, _free_port(), outdir), nprocs=world, join=True)
    r = [torch.load(os.path.join(outdir, f"rank{k}.pt"), weights_only=False) for k in range(world)]
    assert r[0]["plan"] == r[1]["plan"]                      # same plan computed independently
    assert r[0]["tot"] == r[1]["tot"] and r[0]["tot"]["utterances"] == 10.0
    merged = {**r[0]["res"], **r[1]["res"]}
    assert sorted(merged) == list(range(10)) and not (set(r[0]["res"]) & set(r[1]["res"]))
    # single process, same plan: identical results (a shard is a batch; ranks do not interact)
    mc = C.model_config("BC2013")
    sd = S.fastspeech2_state_dict(mc, "BC2013")
    speakers, texts, mels = _utterances(10, seed=5)
    plan = SH.plan_shards([t.numel() for t in texts], [m.shape[0] for m in mels], max_batch=3, bucket_width=4,
                          n_ranks=1, vocoder=False)
    assert [s.utterances for s in plan.shards] == [u for u, _ in r[0]["plan"]]
    torch.set_num_threads(2)
    single = SH.run_rank(plan, 0, _oracle_compute(sd, mc), speakers, texts, mels)
    for uid in range(10):
        assert torch.equal(single[uid]["d_rounded"], merged[uid]["d_rounded"])
        assert int(single[uid]["mel_len"]) == int(merged[uid]["mel_len"])
        assert torch.equal(single[uid]["postnet_mel"], merged[uid]["postnet_mel"])
    assert r[0]["tot"]["frames"] == float(sum(int(v["mel_len"]) for v in single.values()))


def _dynamic_worker(rank, world, port, outdir):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(2)
    mc = C.model_config("BC2013")
    sd = S.fastspeech2_state_dict(mc, "BC2013")
    speakers, texts, mels = _utterances(10, seed=5)
    plan = SH.plan_shards([t.numel() for t in texts], [m.shape[0] for m in mels], max_batch=3, bucket_width=4,
                          n_ranks=world, vocoder=False)
    queue = SH.ShardQueue()
    out = {}
    for sweep in (1, 2):  # two sweeps, one counter each
        out[sweep] = SH.run_rank_dynamic(plan, queue, sweep, _oracle_compute(sd, mc), speakers, texts, mels)
    torch.save(out, os.path.join(outdir, f"dyn{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_dynamic_schedule_two_ranks_claim_every_shard_exactly_once():
    """ShardQueue: ranks claim the planned shards from a shared counter; whichever rank runs a shard, the shard is the
    planned batch, so the union over ranks equals the static / single-process result bit for bit."""
    world = 2
    outdir = tempfile.mkdtemp()
    mp.spawn(_dynamic_worker, args=(world, _free_port(), outdir), nprocs=world, join=True)
    r = [torch.load(os.path.join(outdir, f"dyn{k}.pt"), weights_only=False) for k in range(world)]
    mc = C.model_config("BC2013")
    sd = S.fastspeech2_state_dict(mc, "BC2013")
    speakers, texts, mels = _utterances(10, seed=5)
    plan = SH.plan_shards([t.numel() for t in texts], [m.shape[0] for m in mels], max_batch=3, bucket_width=4,
                          n_ranks=1, vocoder=False)
    torch.set_num_threads(2)
    single = SH.run_rank(plan, 0, _oracle_compute(sd, mc), speakers, texts, mels)
    for sweep in (1, 2):
        a, b = r[0][sweep], r[1][sweep]
        assert not (set(a) & set(b)) and sorted({**a, **b}) == list(range(10))
        merged = {**a, **b}
        for uid in range(10):
            assert torch.equal(single[uid]["d_rounded"], merged[uid]["d_rounded"])
            assert torch.equal(single[uid]["postnet_mel"], merged[uid]["postnet_mel"])
    order = SH.dynamic_order(plan)
    assert [s.cost for s in order] == sorted((s.cost for s in plan.shards), reverse=True)
