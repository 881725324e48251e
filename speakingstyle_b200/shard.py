"""Length-bucketed shard planner and per-rank driver for multi-GPU batched synthesis.

The path shards by utterance: every utterance is independent, weights are replicated on every GPU
and there is NO collective on the data path.  What needs care is that the
reference's outputs depend on how a batch is padded (mean-pool over the padded reference length,
embeddings added on padding - SURVEY.md section 0), so a *shard is a reference batch*: the plan (which
utterances, in which order, form each batch) is part of the result's definition.  The planner is
therefore a pure, deterministic function of the lengths - every rank computes the same plan from
the same inputs without talking to the others, and an oracle can be run on the very same shards.

  plan = plan_shards(src_lens, ref_lens, max_batch=32, bucket_width=64, n_ranks=8)
  for shard in plan.shards_of(rank): ...             # one process per GPU

torch.distributed is used only off the data path (gathering counters for a report).
"""
from dataclasses import dataclass, field
from typing import Dict, List, Sequence

import torch

# Relative per-unit costs (algorithmic MFLOP): encoder+predictors per phoneme, reference
# encoder per reference frame, decoder+postnet per output frame, vocoder per output frame.
COST_PHONEME, COST_REF_FRAME, COST_FRAME, COST_VOC_FRAME = 25.8, 31.6, 48.2, 614.1
EXPECTED_FRAMES_PER_PHONEME = 5.5   # what the synthetic weights of synth.py produce; a real model's own mean belongs here
VOCODER_HALO_FRAMES = 15            # hifigan.Generator.halo_frames(): frames computed beyond each utterance's own
FIXED_COST = 6.0e5                  # per shard: ~210 launches and one host read, ~3 ms, in the same MFLOP units


@dataclass
class Shard:
    index: int
    utterances: List[int]          # utterance ids, in batch order
    max_src_len: int
    max_ref_len: int
    cost: float
    rank: int = -1


@dataclass
class ShardPlan:
    shards: List[Shard]
    n_ranks: int
    rank_cost: List[float] = field(default_factory=list)

    def shards_of(self, rank: int) -> List[Shard]:
        return [s for s in self.shards if s.rank == rank]

    def imbalance(self) -> float:
        """max rank cost / mean rank cost (1.0 = perfect)."""
        mean = sum(self.rank_cost) / max(1, len(self.rank_cost))
        return max(self.rank_cost) / mean if mean > 0 else 1.0


def shard_cost(n, max_src, max_ref, vocoder=True, text=True, sum_src=None):
    """Work of one batch as the GPU executes it: the reference encoder, the encoder and the predictors run on the
    PADDED batch (n x max length); the decoder's FFT blocks and the vocoder run a ragged schedule, so
    their work follows the utterances' own (expected) frame counts plus the vocoder's halo; mel_linear / PostNet are
    padded again.  ``sum_src`` = sum of the utterances' phoneme counts (default: n * max_src, the padded bound)."""
    if sum_src is None:
        sum_src = n * max_src
    c = FIXED_COST + n * max_ref * COST_REF_FRAME
    if text:
        padded_frames = n * max_src * EXPECTED_FRAMES_PER_PHONEME
        own_frames = sum_src * EXPECTED_FRAMES_PER_PHONEME
        c += n * max_src * COST_PHONEME + own_frames * (COST_FRAME - 8.7) + padded_frames * 8.7  # 8.7: mel_linear + PostNet
        if vocoder:
            c += (own_frames + n * VOCODER_HALO_FRAMES) * COST_VOC_FRAME
    return c


def plan_shards(src_lens: Sequence[int], ref_lens: Sequence[int], max_batch: int = 32, bucket_width: int = 64,
                n_ranks: int = 1, vocoder: bool = True, text: bool = True, sort_key: str = "auto") -> ShardPlan:
    """Bucket utterances by length, cut buckets into batches of <= max_batch, assign batches to ranks
    longest-processing-time-first.  Deterministic: ties are broken by utterance id."""
    n = len(ref_lens)
    src = [int(x) for x in src_lens] if text else [0] * n
    ref = [int(x) for x in ref_lens]
    if text and len(src) != n:
        raise ValueError("src_lens and ref_lens must have the same length")
    if max_batch < 1 or bucket_width < 1 or n_ranks < 1:
        raise ValueError("max_batch, bucket_width and n_ranks must be positive")
    if sort_key == "auto":
        sort_key = "src" if text else "ref"
    primary = src if sort_key == "src" else ref
    # bucket on the primary length (padding waste <= bucket_width), order inside a bucket by the other
    order = sorted(range(n), key=lambda i: (primary[i] // bucket_width, (ref if sort_key == "src" else src)[i], i))
    shards: List[Shard] = []
    cur: List[int] = []
    cur_bucket = None
    for i in order + [None]:
        b = None if i is None else primary[i] // bucket_width
        if cur and (i is None or b != cur_bucket or len(cur) == max_batch):
            ms = max(src[j] for j in cur) if text else 0
            mr = max(ref[j] for j in cur)
            shards.append(Shard(len(shards), cur, ms, mr,
                                shard_cost(len(cur), ms, mr, vocoder, text, sum(src[j] for j in cur) if text else None)))
            cur = []
        if i is not None:
            cur.append(i)
            cur_bucket = b
    # LPT assignment
    rank_cost = [0.0] * n_ranks
    for s in sorted(shards, key=lambda s: (-s.cost, s.index)):
        r = min(range(n_ranks), key=lambda k: (rank_cost[k], k))
        s.rank = r
        rank_cost[r] += s.cost
    _refine(shards, rank_cost)
    return ShardPlan(shards, n_ranks, rank_cost)


def _refine(shards: List[Shard], rank_cost: List[float], max_rounds: int = 200) -> None:
    """Deterministic local search after LPT: move one shard from the most loaded rank to another, or swap one of its
    shards with one of another rank's, whenever that lowers the larger of the two loads; stop when nothing does.
    With ~8 shards per rank LPT alone leaves ~2 % between the extremes; this brings it to a few tenths."""
    n = len(rank_cost)
    if n < 2:
        return
    for _ in range(max_rounds):
        hi = max(range(n), key=lambda k: (rank_cost[k], -k))
        best = None  # (new larger load, tie-breakers..., action)
        mine = [s for s in shards if s.rank == hi]
        for lo in range(n):
            if lo == hi:
                continue
            theirs = [s for s in shards if s.rank == lo]
            for a in mine:
                new_hi, new_lo = rank_cost[hi] - a.cost, rank_cost[lo] + a.cost  # move a
                cand = (max(new_hi, new_lo), a.index, -1, lo)
                if cand[0] < rank_cost[hi] - 1e-9 and (best is None or cand < best):
                    best = cand
                for b in theirs:  # swap a <-> b
                    new_hi, new_lo = rank_cost[hi] - a.cost + b.cost, rank_cost[lo] + a.cost - b.cost
                    cand = (max(new_hi, new_lo), a.index, b.index, lo)
                    if cand[0] < rank_cost[hi] - 1e-9 and (best is None or cand < best):
                        best = cand
        if best is None:
            return
        _, ai, bi, lo = best
        a = next(s for s in shards if s.index == ai)
        rank_cost[hi] -= a.cost
        rank_cost[lo] += a.cost
        a.rank = lo
        if bi >= 0:
            b = next(s for s in shards if s.index == bi)
            rank_cost[lo] -= b.cost
            rank_cost[hi] += b.cost
            b.rank = hi


def make_shard_batch(shard: Shard, speakers, texts: Sequence[torch.Tensor], mels: Sequence[torch.Tensor]) -> Dict:
    """Collate one shard into the reference's batch layout (dataset.py:198-217): zero padding, PAD id 0."""
    ids = shard.utterances
    B = len(ids)
    src_lens = torch.tensor([int(texts[i].numel()) for i in ids], dtype=torch.int64)
    mel_lens = torch.tensor([int(mels[i].shape[0]) for i in ids], dtype=torch.int64)
    L, Tr = int(src_lens.max()), int(mel_lens.max())
    n_mel = mels[ids[0]].shape[1]
    t = torch.zeros((B, L), dtype=torch.int64)
    m = torch.zeros((B, Tr, n_mel), dtype=torch.float32)
    for r, i in enumerate(ids):
        t[r, : src_lens[r]] = texts[i]
        m[r, : mel_lens[r]] = mels[i]
    spk = torch.tensor([int(speakers[i]) for i in ids], dtype=torch.int64)
    return {"speakers": spk, "texts": t, "src_lens": src_lens, "max_src_len": L, "mels": m, "mel_lens": mel_lens,
            "max_mel_len": Tr}


def run_rank(plan: ShardPlan, rank: int, compute, speakers, texts, mels) -> Dict[int, Dict]:
    """Process this rank's shards with ``compute(batch) -> dict of per-batch tensors / lists`` and
    return {utterance id: per-utterance result}.  ``compute`` is the GPU model on a GPU box."""
    out: Dict[int, Dict] = {}
    for shard in plan.shards_of(rank):
        batch = make_shard_batch(shard, speakers, texts, mels)
        res = compute(batch)
        for row, uid in enumerate(shard.utterances):
            out[uid] = {k: v[row] for k, v in res.items()}
    return out


def dynamic_order(plan: ShardPlan) -> List[Shard]:
    """The planned shards, most expensive first: the order in which ranks claim them under the dynamic schedule."""
    return sorted(plan.shards, key=lambda s: (-s.cost, s.index))


class ShardQueue:
    """Dynamic schedule for one node: ranks claim positions of ``dynamic_order(plan)`` from one counter per sweep in the
    process group's key-value store (``TCPStore.add`` is atomic).  Host-side control plane only - no GPU data moves
    and nothing is added to the data path; the shards themselves stay exactly as planned (so results are unchanged and
    an oracle can still be run on the very same batches), only WHICH GPU runs a shard is decided at run time.  That
    absorbs what a static assignment cannot see: GPUs of one box differ by a few per cent under their power caps."""

    def __init__(self, store=None, prefix="ss_shard_queue"):
        if store is None:
            import torch.distributed as dist

            store = dist.distributed_c10d._get_default_store()
        self.store, self.prefix = store, prefix

    def claim(self, sweep: int) -> int:
        """Next unclaimed position of sweep number ``sweep`` (every rank must number its sweeps the same way)."""
        return int(self.store.add(f"{self.prefix}_{sweep}", 1)) - 1

    def positions(self, sweep: int, n: int):
        while True:
            i = self.claim(sweep)
            if i >= n:
                return
            yield i


def run_rank_dynamic(plan: ShardPlan, queue: ShardQueue, sweep: int, compute, speakers, texts, mels) -> Dict[int, Dict]:
    """Like run_rank, but this rank processes whichever planned shards it manages to claim."""
    order = dynamic_order(plan)
    out: Dict[int, Dict] = {}
    for pos in queue.positions(sweep, len(order)):
        shard = order[pos]
        res = compute(make_shard_batch(shard, speakers, texts, mels))
        for row, uid in enumerate(shard.utterances):
            out[uid] = {k: v[row] for k, v in res.items()}
    return out


def gather_counters(local: Dict[str, float], group=None) -> Dict[str, float]:
    """Sum scalar counters (frames, seconds ...) over ranks - off the data path, for the report only."""
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        return dict(local)
    keys = sorted(local)
    t = torch.tensor([float(local[k]) for k in keys], dtype=torch.float64)
    if dist.get_backend(group) == "nccl":
        t = t.cuda()
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return {k: float(v) for k, v in zip(keys, t.cpu().tolist())}
