#!/usr/bin/env python
"""Benchmark of the batched text -> mel -> wave path.  Contract: see DESIGN.md "Measurement".

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload auto|cfg1..cfg5]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workloads are BASELINE.json's configs.  Two kinds:

* single batch (cfg1, cfg2, cfg3): a step is ONE pass of the hot path over one synthetic batch.  Rank r runs its own
  batch (seed 1+r, weights replicated): utterances are independent, no collective on the data path, weak scaling.
* sharded sweep (cfg4, cfg5): the corpus (10 000 reference mels / 2048 utterances) is cut into length-bucketed shards
  by ``shard.plan_shards`` (a pure function of the lengths: every rank computes the same plan), each rank materialises
  and runs only its own shards, and a step is ONE pass of a rank over all of its shards.  Total work is fixed as N
  grows: strong scaling.  There is still no collective on the data path; the time of a step is the max over ranks.

``--workload auto`` (the default) is cfg3 - FastSpeech2 + ReferenceEncoder + HiFi-GAN, text -> int16 wave, the largest
single-GPU config - at N = 1 and the cfg5 sweep at N > 1.  One JSON line is printed by rank 0.

--impl reference times the CPU restatement of the reference (oracle/, kind "port": the reference is pure Python and
cannot travel to the GPU box) on the host cores, on a bounded, fixed sample of the same workload (the same sample
the GPU arm's ``cpu_baseline`` leg uses).
"""
import argparse
import json
import os
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

from speakingstyle_b200 import configs as C  # noqa: E402
from speakingstyle_b200 import synth as S  # noqa: E402

HOP, SR = 256, 22050
UPS = (1, 8, 64, 128, 256)  # HiFi-GAN rows per mel frame at conv_pre / the four stages
WORKLOAD_DESC = {
    "cfg1": "BASELINE configs[0]: LJSpeech(+ref-enc) B=16, L~U[80,120], Tr~U[200,400], text->mel",
    "cfg2": "BASELINE configs[1]: BC2013 FastSpeech2+ReferenceEncoder B=64, L~U[80,120], Tr=800, text+ref->mel",
    "cfg3": "BASELINE configs[2]: LibriTTS multi-speaker FastSpeech2 + HiFi-GAN B=32, text->22.05kHz int16 wave",
    "cfg4": "BASELINE configs[3]: ReferenceEncoder sweep over 10000 reference mels, Tr~U[100,1000], length-bucketed "
            "(width 64, batch<=64), sharded by shard.plan_shards",
    "cfg5": "BASELINE configs[4]: 2048 utterances, L~U[40,110], Tr~U[200,1000], FastSpeech2+ReferenceEncoder+HiFi-GAN "
            "-> int16 wave, length-bucketed shards (width 16, batch<=32) planned by shard.plan_shards",
}
SWEEPS = ("cfg4", "cfg5")
# the fixed CPU sample of each workload (cpu_baseline leg and --impl reference use the same one)
CPU_SAMPLE = {"cfg1": 16, "cfg2": 64, "cfg3": 4, "cfg4": 64, "cfg5": 4}


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"hbm_gbs": p["hbm_gbs"], "tflops_burst": p["bf16_tflops"],
                "tflops_sustained": p.get("bf16_tflops_sustained", p["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "tflops_burst": 1590.0, "tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons during the timed region (pynvml, 10 ms period)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()

    def run(self):
        try:
            import pynvml as nv

            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {
                getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
                getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
                getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
                getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            }
            get = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
            while not self._stop_evt.is_set():
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                mask = get(h)
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
                time.sleep(0.01)
        except Exception as e:  # pragma: no cover
            self.reasons.add(f"sampler_error:{type(e).__name__}")

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def pin(batch):
    return {k: (v.pin_memory() if torch.is_tensor(v) else v) for k, v in batch.items()}


def to_dev(batch, dev):
    return {k: (v.to(dev, non_blocking=True) if torch.is_tensor(v) else v) for k, v in batch.items()}


def call(model, b):
    return model(b["speakers"], b["texts"], b["src_lens"], b["max_src_len"], b["mels"], b["mel_lens"], b["max_mel_len"])


def sub_batch(batch, n):
    """First n utterances, re-padded (a shard is a batch as composed)."""
    L, Tr = int(batch["src_lens"][:n].max()), int(batch["mel_lens"][:n].max())
    return {"speakers": batch["speakers"][:n], "texts": batch["texts"][:n, :L].contiguous(),
            "src_lens": batch["src_lens"][:n], "max_src_len": L, "mels": batch["mels"][:n, :Tr].contiguous(),
            "mel_lens": batch["mel_lens"][:n], "max_mel_len": Tr}


def resolve_workload(name, world):
    if name != "auto":
        return name
    return "cfg3" if world == 1 else "cfg5"


# ---------------------------------------------------------------------------------------------
# the work of one rank: a list of batches (one for the single-batch workloads, its shards for a sweep)
# ---------------------------------------------------------------------------------------------
def rank_batches(workload, rank, world, limit=None, all_shards=False):
    """(dataset, [batch, ...], vocoder?, text?, plan-or-None).  ``limit`` caps the corpus of a sweep (utterances).
    all_shards (dynamic schedule): every rank holds every shard, most expensive first."""
    if workload in SWEEPS:
        corpus = S.Corpus(workload, seed=1, n=limit)
        plan = corpus.plan(world)
        from speakingstyle_b200 import shard as SH

        mine = SH.dynamic_order(plan) if all_shards else plan.shards_of(rank)
        return corpus.dataset, [corpus.shard_batch(s) for s in mine], corpus.vocoder, corpus.text, plan
    ds, batch, voc = S.make_workload(workload, seed=1 + rank)
    return ds, [batch], voc, True, None


# ---------------------------------------------------------------------------------------------
# CPU arm (oracle port of the reference) - also the cpu_baseline leg of the GPU arm
# ---------------------------------------------------------------------------------------------
def cpu_sample(workload):
    """The fixed CPU sample of a workload: (state dict, model config, [batches], vocoder state, h, text?, description)."""
    n = CPU_SAMPLE[workload]
    if workload == "cfg4":
        corpus = S.Corpus("cfg4", seed=1, n=n)
        batches = [corpus.shard_batch(s) for s in corpus.plan(1).shards]
        ds, voc, text = corpus.dataset, False, False
        what = f"the first {n} reference mels of the cfg4 corpus, planned into {len(batches)} shards"
    elif workload == "cfg5":
        corpus = S.Corpus("cfg5", seed=1)
        shard0 = corpus.plan(1).shards[0]
        batches = [sub_batch(corpus.shard_batch(shard0), n)]
        ds, voc, text = corpus.dataset, True, True
        what = f"the first {n} utterances of shard 0 of the cfg5 plan"
    else:
        ds, batch, voc = S.make_workload(workload, seed=1)
        B = batch["texts"].shape[0]
        batches = [sub_batch(batch, min(n, B))]
        text = True
        what = f"the first {min(n, B)} utterances of the {B}-utterance {workload} batch"
    mc = C.model_config(ds)
    sd = S.fastspeech2_state_dict(mc, ds)
    h = C.hifigan_config() if voc else None
    vsd = S.fold_weight_norm(S.hifigan_state_dict(h)) if voc else None
    return sd, mc, batches, vsd, h, text, what


def cpu_pass(sd, mc, batches, vsd, h, text):
    """One pass of the oracle over the sample; returns the frames it produced (reference frames for cfg4)."""
    from oracle import fs2_oracle as O

    frames = 0
    with torch.no_grad():
        for b in batches:
            if not text:
                O.reference_encoder(sd, mc, b["mels"], O.mask_from_lengths(b["mel_lens"], b["max_mel_len"]))
                frames += int(b["mel_lens"].sum())
                continue
            out = O.fastspeech2_forward(sd, mc, b)
            if vsd is not None:
                O.hifigan_forward(vsd, h, out[1].transpose(1, 2))
            frames += int(out[9].sum())
    return frames


def cpu_baseline(workload):
    """1 warm-up pass, then 3 timed passes over the fixed sample, all host threads."""
    torch.set_num_threads(os.cpu_count() or 1)
    sd, mc, batches, vsd, h, text, what = cpu_sample(workload)
    cpu_pass(sd, mc, batches, vsd, h, text)
    t0 = time.perf_counter()
    frames = sum(cpu_pass(sd, mc, batches, vsd, h, text) for _ in range(3))
    dt = time.perf_counter() - t0
    return {"value": frames / dt, "unit": "mel-frames/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"{what}; 3 timed passes after 1 warm-up, torch {torch.__version__} CPU fp32, oracle/fs2_oracle.py"}


def run_reference_arm(args, rank, world):
    if rank != 0:
        return
    workload = resolve_workload(args.workload, max(world, args.gpus))
    torch.set_num_threads(os.cpu_count() or 1)
    sd, mc, batches, vsd, h, text, what = cpu_sample(workload)
    for _ in range(args.warmup):
        cpu_pass(sd, mc, batches, vsd, h, text)
    t0 = time.perf_counter()
    frames = 0
    for _ in range(args.steps):
        frames += cpu_pass(sd, mc, batches, vsd, h, text)
    dt = time.perf_counter() - t0
    value = frames / dt
    sample = f"{what} per step, oracle port of the reference (torch {torch.__version__} CPU fp32)"
    voc = vsd is not None
    line = {
        "impl": "reference", "metric": "mel_frames_per_s", "value": value, "unit": "mel-frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
        "scaling": "strong" if workload in SWEEPS else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload, "desc": WORKLOAD_DESC[workload], "vocoder": voc},
        "cpu_baseline": {"value": value, "unit": "mel-frames/s", "cores": torch.get_num_threads(), "kind": "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": "mel-frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    if voc:
        line["audio_seconds_per_s"] = value * HOP / SR
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------
def lr_roofline(dev, pk):
    """LengthRegulator kernel alone at a BASELINE cfg5-like shard size, L2 flushed between launches."""
    from speakingstyle_b200 import ops

    B, L, Cc = 1024, 100, 256  # ~1 GB of output: large enough that launch latency is not what is measured
    g = torch.Generator().manual_seed(3)
    x = torch.randn(B, L, Cc, generator=g).to(dev)
    d = torch.randint(4, 12, (B, L), generator=g).float().to(dev)
    _, ends, lens2 = ops.duration_scan(None, 1.0, durations=d)
    T = int(lens2[0].max())
    pe = torch.randn(T, Cc, generator=g).to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    times = []
    for i in range(8):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        ops.length_regulate(x, ends, T, pe=pe)
        b.record()
        torch.cuda.synchronize()
        if i >= 3:
            times.append(a.elapsed_time(b))
    ms = sum(times) / len(times)
    nbytes = 4 * Cc * (B * L + B * T) + 4 * B * L + 4 * B
    ach = nbytes / (ms * 1e-3) / 1e9
    traffic = None
    try:  # DRAM bytes of this very launch shape from the committed ncu capture
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        traffic = tj.get(f"length_regulate_kernel B={B} L={L} T={T} C={Cc}", {}).get("bytes")
    except Exception:
        pass
    return {"kernel": "length_regulate_kernel", "bound": "hbm", "achieved": ach, "peak": pk["hbm_gbs"], "unit": "GB/s",
            "frac": ach / pk["hbm_gbs"], "traffic": traffic, "algorithmic_bytes": nbytes,
            "shape": f"B={B} L={L} T={T} C={Cc}", "ms": ms,
            "peak_source": pk["source"]}


class Runner:
    """Model(s) of one workload on one device plus this rank's batches; ``step(src)`` = one pass over them."""

    def __init__(self, workload, precision, dev, rank=0, world=1, limit=None, queue=None):
        import speakingstyle_b200 as ss

        self.ss, self.dev, self.workload, self.precision = ss, dev, workload, precision
        # queue (dynamic schedule of a sweep on several GPUs): a callable returning the next unclaimed position in the
        # cost-ordered shard list, shared by all ranks - see shard_queue()
        self.queue = queue if (queue is not None and workload in SWEEPS and world > 1) else None
        self.rank, self.world = rank, world
        ds, batches, voc, text, plan = rank_batches(workload, rank, world, limit, all_shards=self.queue is not None)
        self.text, self.plan, self.vocoder = text, plan, voc
        d = C.write_preprocessed_dir(tempfile.mkdtemp(), ds)
        mc, pc = C.model_config(ds), C.preprocess_config(ds, d)
        self.model = ss.get_model(pc, mc, dev, S.fastspeech2_state_dict(mc, ds), precision=precision)
        self.gen = ss.get_vocoder(C.hifigan_config(), dev, S.hifigan_state_dict(), precision=precision) if voc else None
        self.host = [pin(b) for b in batches]
        self.resident = [to_dev(b, dev) for b in batches]
        self.info = [None] * len(batches)  # per batch: padded / valid sizes, filled by the first pass
        self.timer = None

    def _one(self, b, i, e2e):
        """One batch.  Returns (frames as a device tensor or int, bytes read back to the host)."""
        if self.timer is not None:
            self.timer.ctx = i
        if not self.text:  # cfg4: ReferenceEncoder alone -> (gammas, betas, pooled)
            lens = b["mel_lens"].to(self.dev, torch.int32, non_blocking=True)
            ga, be, pooled = self.model.reference_encoder(b["mels"].to(self.dev, non_blocking=True), b["max_mel_len"],
                                                          None, lens=lens, return_pooled=True)
            d2h = 0
            if e2e:
                for t in (ga, be, pooled):
                    t.to("cpu")
                    d2h += t.numel() * 4
            return None, d2h, None
        if self.gen is not None:
            # public API (tools.synthesize_batch), the reference's synthesize.py:128-150 + vocoder_infer: waves trimmed to
            # mel_len * hop.  e2e: H2D inside forward(), the ragged PCM epilogue writes the int16 waves straight into
            # pinned host memory (it synchronises the stream before returning them).  Device-resident: same call, the
            # trimmed int16 waves stay in HBM.
            o, waves = self.ss.synthesize_batch(self.model, self.gen, b, device_out=not e2e)
            d2h = 2 * sum(len(x) for x in waves) if e2e else 0
        else:
            o = call(self.model, b)  # with host inputs the H2D of the pinned tensors happens inside forward()
            d2h = 0
        if e2e:
            o[1].to("cpu")
            o[9].to("cpu")
            d2h += o[1].numel() * 4 + o[9].numel() * 4
        return o, d2h, None

    def _positions(self):
        """Which of this rank's batches a step runs, in order.  Static schedule: all of them.  Dynamic: the first pass
        runs every shard once (it fills the per-shard sizes every rank needs for its bookkeeping); from then on each
        rank claims the next unclaimed shard of the cost-ordered list until the list is exhausted."""
        if self.queue is None or any(i is None for i in self.info):
            yield from range(len(self.host))
            return
        self.sweeps = getattr(self, "sweeps", 0) + 1
        while True:
            i = self.queue(self.sweeps)
            if i >= len(self.host):
                return
            yield i

    def step(self, e2e=False):
        d2h = 0
        src = self.host if e2e else self.resident
        self.ran = []
        for i in self._positions():
            b = src[i]
            self.ran.append(i)
            o, nb, _ = self._one(b, i, e2e)
            d2h += nb
            if self.info[i] is None:
                hb = self.host[i]
                inf = {"B": int(hb["mels"].shape[0]), "L": int(hb["max_src_len"]), "Tr": int(hb["max_mel_len"]),
                       "ref_frames": int(hb["mel_lens"].sum()), "tokens": int(hb["src_lens"].sum()), "T": 0, "frames": 0}
                if o is not None:
                    inf["T"] = int(o[1].shape[1])
                    inf["frames"] = int(o[9].clamp(max=inf["T"]).sum())
                self.info[i] = inf
        return d2h

    # totals of this rank over one step (dynamic schedule: every rank holds every shard, so these are the whole job's)
    def frames(self):
        return sum((i["frames"] if self.text else i["ref_frames"]) for i in self.info)

    def h2d_bytes(self):
        return sum(v.numel() * v.element_size() for b in self.host for v in b.values() if torch.is_tensor(v))

    def valid_fraction(self, ctx, S_in):
        """Share of a launch's rows that are valid positions (SURVEY 8d counts algorithmic work on those only)."""
        inf = self.info[ctx] if ctx is not None and ctx < len(self.info) else None
        if inf is None:
            return 1.0
        T = inf["T"]
        if T and S_in % T == 0 and S_in // T in UPS:
            return inf["frames"] / (inf["B"] * T)
        if S_in == inf["L"] and self.text:
            return inf["tokens"] / (inf["B"] * inf["L"])
        if S_in == inf["Tr"]:
            return inf["ref_frames"] / (inf["B"] * inf["Tr"])
        return 1.0


def shard_queue():
    """Dynamic schedule of a sweep (speakingstyle_b200.shard.ShardQueue): ranks claim the planned shards, most expensive
    first, from one counter per sweep in the process group's store.  None when no store is available."""
    from speakingstyle_b200 import shard as SH

    try:
        q = SH.ShardQueue(prefix="ss_bench_sweep")
    except Exception:
        return None
    return q.claim


def time_steps(runner, steps, flush, barrier, e2e):
    """K steps, each bracketed by CUDA events on the launching stream, L2 flushed in between (outside the pair)."""
    barrier()
    evs, d2h = [], 0
    for _ in range(steps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        d2h = runner.step(e2e=e2e)
        b.record()
        evs.append((a, b))
    barrier()
    return sum(a.elapsed_time(b) for a, b in evs), d2h


def quick_measure(workload, precision, dev, steps=10, warmup=3, limit=None):
    """Secondary numbers for the same JSON line: device-resident and end-to-end throughput of another
    workload / precision mode, same timing rules as the headline (events per step, L2 flushed)."""
    r = Runner(workload, precision, dev, limit=limit)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for _ in range(warmup):
        r.step()
    torch.cuda.synchronize()
    frames = r.frames()
    # host enqueue time of a step (no synchronisation inside): what a launch-bound batch is limited by
    t0 = time.perf_counter()
    r.step()
    host_ms = 1e3 * (time.perf_counter() - t0)
    torch.cuda.synchronize()
    ms, _ = time_steps(r, steps, flush, torch.cuda.synchronize, e2e=False)
    e_ms, d2h = time_steps(r, steps, flush, torch.cuda.synchronize, e2e=True)
    res = {"workload": workload, "desc": WORKLOAD_DESC[workload], "precision_mode": precision, "vocoder": r.vocoder,
           "mel_frames_per_s": frames * steps / (ms * 1e-3), "ms_per_step": ms / steps,
           "e2e_mel_frames_per_s": frames * steps / (e_ms * 1e-3), "e2e_ms_per_step": e_ms / steps,
           "host_enqueue_ms_per_step": host_ms, "valid_frames_per_step": frames, "steps": steps,
           "h2d_bytes_per_step": r.h2d_bytes(), "d2h_bytes_per_step": d2h}
    if r.vocoder:
        res["audio_seconds_per_s"] = res["mel_frames_per_s"] * HOP / SR
        res["e2e_audio_seconds_per_s"] = res["e2e_mel_frames_per_s"] * HOP / SR
    if r.plan is not None:
        res["corpus_limit"] = limit
        res["utterances"] = sum(len(s_.utterances) for s_ in r.plan.shards)
        res["shards"] = len(r.host)
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="auto", choices=["auto"] + sorted(WORKLOAD_DESC),
                    help="auto: cfg3 on one GPU, the cfg5 sharded sweep on several")
    ap.add_argument("--precision", default="fp32", choices=["fp32", "bf16"],
                    help="fp32: all kernels fp32-accurate (mel / wave max-abs <= 1e-3); bf16: Decoder/PostNet/HiFi-GAN "
                         "on tcgen05 with bf16 operands (relative L2 <= 1e-2); integers identical in both")
    ap.add_argument("--limit", type=int, default=None, help="sweeps: use only the first LIMIT utterances of the corpus")
    ap.add_argument("--schedule", default="dynamic", choices=["dynamic", "static"],
                    help="sweeps on several GPUs: dynamic = ranks claim the planned shards from a shared counter (default); "
                         "static = each rank runs the shards shard.plan_shards assigned to it")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the secondary workloads / bf16 mode block")
    ap.add_argument("--conv-breakdown", action="store_true", help="print per-shape conv timings to stderr")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return

    import torch.distributed as dist

    from speakingstyle_b200 import ops

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: speakingstyle_b200 has no CPU path "
                         "(use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    workload = resolve_workload(args.workload, world)
    sweep = workload in SWEEPS

    queue = shard_queue() if (sweep and world > 1 and args.schedule == "dynamic") else None
    runner = Runner(workload, args.precision, dev, rank, world, args.limit, queue=queue)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    for _ in range(args.warmup):
        runner.step()
    torch.cuda.synchronize()
    frames = runner.frames()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing (value) ----
    sampler = ClockSampler(local_rank)
    sampler.start()
    timer = ops.KernelTimer()
    ops.conv_timer = runner.timer = timer
    ops.launch_count_reset()
    wall0 = time.perf_counter()
    dev_ms, _ = time_steps(runner, args.steps, flush, barrier, e2e=False)
    wall = time.perf_counter() - wall0
    launches = ops.launch_count()
    ops.conv_timer = runner.timer = None
    clocks = sampler.stop()

    # ---- end-to-end timing through the public API with host buffers ----
    e2e_steps = args.steps if not sweep else min(args.steps, 5)
    e2e_ms, d2h = time_steps(runner, e2e_steps, flush, barrier, e2e=True)
    e2e_ms *= args.steps / e2e_steps  # per-step time is what is compared; scale to K steps for the max below
    h2d = runner.h2d_bytes()

    # ---- max over ranks ----
    t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device=dev)
    f = torch.tensor([frames, h2d, d2h], dtype=torch.float64, device=dev)
    per_rank = [t.clone() for _ in range(world)]
    if world > 1:
        dist.all_gather(per_rank, t)
        if runner.queue is None:
            dist.all_reduce(f, op=dist.ReduceOp.SUM)
        else:  # every rank holds every shard: frames / h2d are already the whole job's; d2h is what THIS rank read back
            d = f[2:3].clone()
            dist.all_reduce(d, op=dist.ReduceOp.SUM)
            f[2] = d[0]
    else:
        per_rank = [t]
    dev_ms_max = max(x[0].item() for x in per_rank)
    e2e_ms_max = max(x[1].item() for x in per_rank)
    total_frames, h2d_all, d2h_all = f.tolist()

    if rank == 0:
        pk = peaks()
        value = total_frames * args.steps / (dev_ms_max * 1e-3)
        e2e_value = total_frames * args.steps / (e2e_ms_max * 1e-3)
        # per conv family (kind, Cin, Cout, k, dilation): launches, ms, FLOPs as launched (padded rows included) and on
        # valid positions only (what SURVEY 8d counts)
        fam = {}
        tot_ms = tot_valid = tot_padded = 0.0
        rows = timer.rows()
        for ms, w, tag, ctx, meta in rows:
            kind, B_, S_in, Cin, Cout, k, dil = meta
            fr = runner.valid_fraction(ctx, S_in)
            key = (kind, Cin, Cout, k, dil)
            c, m, wv, wp = fam.get(key, (0, 0.0, 0.0, 0.0))
            fam[key] = (c + 1, m + ms, wv + w * fr, wp + w)
            tot_ms += ms
            tot_valid += w * fr
            tot_padded += w
        fam = dict(sorted(fam.items(), key=lambda kv: -kv[1][1]))
        if args.conv_breakdown:
            for key, (c, ms, wv, wp) in fam.items():
                sys.stderr.write(f"{ms / args.steps:9.3f} ms/step {c // args.steps:5d} calls/step "
                                 f"{wv / (ms * 1e-3) / 1e12:8.1f} TFLOP/s valid ({wp / (ms * 1e-3) / 1e12:7.1f} padded)  "
                                 f"{key[0]} Cin={key[1]} Cout={key[2]} k={key[3]} dil={key[4]}\n")
            for tag, (c, ms, w) in timer.by_tag().items():
                sys.stderr.write(f"{ms / args.steps:9.3f} ms/step {c // args.steps:4d} calls/step "
                                 f"{w / (ms * 1e-3) / 1e12:8.1f} TFLOP/s  {tag}\n")
            sys.stderr.write(f"step {dev_ms / args.steps:.3f} ms, convs {tot_ms / args.steps:.3f} ms\n")
        ach = tot_valid / (tot_ms * 1e-3) / 1e12 if tot_ms > 0 else 0.0
        roof_all = {"kernel": "all conv / linear launches of the step (3xTF32 tcgen05 kernels"
                              + (" + bf16 tcgen05 kernel)" if args.precision == "bf16" else
                                 " + split-bf16 bf16x3 tcgen05 kernel for the HiFi-GAN ResBlocks)"),
                    "bound": "tensor", "achieved": ach, "achieved_padded": tot_padded / (tot_ms * 1e-3) / 1e12 if tot_ms else 0.0,
                    "peak": pk["tflops_sustained"], "unit": "TFLOP/s", "frac": ach / pk["tflops_sustained"],
                    "launches": len(rows), "avg_launch_ms": tot_ms / max(len(rows), 1), "share_of_step": tot_ms / dev_ms}
        # the dominant kernel = the conv family with the largest share of the step (per-launch figures, as the
        # contract asks: algorithmic FLOPs of one launch / its average CUDA-event duration inside the timed region)
        key, (c, ms, wv, wp) = next(iter(fam.items()))
        d_ach = wv / (ms * 1e-3) / 1e12
        kname = {"tf32x3": "3xTF32 tcgen05 implicit-GEMM conv (fp32-accurate; conv1d_tf32_ts / _ts2 / hifigan kernels)",
                 "bf16x3": "conv1d_bf16x3_kernel (split-bf16 tcgen05 implicit-GEMM conv: 3 bf16 MMAs per algorithmic MAC, "
                           "fp32-class accuracy; HiFi-GAN ResBlocks of the fp32 mode)",
                 "resunit": "resunit32_kernel (split-bf16 c1 -> leaky ReLU -> c2 + residual of one HiFi-GAN ResBlock unit, "
                            "intermediate in shared memory)",
                 "tc": "conv1d_tc_kernel (bf16 tcgen05 implicit-GEMM conv)",
                 "tcp": "conv1d_bf16x3_kernel<SPLIT=false> (bf16 tcgen05 implicit-GEMM conv, panel feeding)", "f32": "conv1d_f32_kernel (FP32 pipe)"}[key[0]]
        roof = {"kernel": kname, "shape": f"Cin={key[1]} Cout={key[2]} k={key[3]} dil={key[4]}", "bound": "tensor",
                "achieved": d_ach, "achieved_padded": wp / (ms * 1e-3) / 1e12, "peak": pk["tflops_sustained"],
                "unit": "TFLOP/s", "frac": d_ach / pk["tflops_sustained"], "traffic": None, "launches": c,
                "avg_launch_ms": ms / max(c, 1), "flops_per_launch": wv / max(c, 1),
                "flops_per_launch_padded": wp / max(c, 1), "share_of_step": ms / dev_ms,
                "peak_source": pk["source"] + " bf16 dense (sustained)",
                "note": "achieved counts FLOPs on valid positions only (SURVEY 8d); achieved_padded counts every row the "
                        "launch computed.  frac divides by the measured bf16 peak as the contract asks.  A 3xTF32 kernel "
                        "issues 3 TF32 MMAs per algorithmic MAC and the TF32 pipe retires 2048 MAC/clk/SM (half of bf16), "
                        "so its own ceiling at the SM clock sampled during the run is 148*2048*2*f/3; a split-bf16 "
                        "(bf16x3) kernel issues 3 bf16 MMAs per algorithmic MAC: ceiling = measured bf16 peak / 3"}
        if key[0] == "tf32x3" and clocks.get("sm_mhz"):
            ceil = 148 * 2048 * 2 * clocks["sm_mhz"] * 1e6 / 3.0 / 1e12
            roof["tf32x3_ceiling_at_sampled_clock"] = ceil
            roof["frac_of_3xtf32_ceiling"] = d_ach / ceil
        if key[0] in ("bf16x3", "resunit"):
            roof["bf16x3_ceiling"] = pk["tflops_sustained"] / 3.0
            roof["frac_of_bf16x3_ceiling"] = d_ach / roof["bf16x3_ceiling"]
        try:  # DRAM bytes of the same conv family from the committed ncu --set full capture (profiles/traffic.json)
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
            v_ = tj.get(f"{key[0]} Cin={key[1]} Cout={key[2]} k={key[3]} dil={key[4]}")
            if v_:
                # the capture is one fixed-size launch with every row computed; the launches timed here are ragged
                # and of several sizes, so the captured bytes are scaled by FLOPs to "per launch like achieved"
                scale = roof["flops_per_launch"] / v_["flops"]
                roof["traffic"] = v_["bytes"] * scale
                roof["algorithmic_bytes"] = v_["algorithmic_bytes"] * scale
                roof["traffic_capture"] = {"bytes": v_["bytes"], "algorithmic_bytes": v_["algorithmic_bytes"],
                                           "shape": v_["shape"], "mix": v_["mix"], "source": v_["source"]}
        except Exception:
            pass
        inf0 = runner.info[0]
        cfg = {"workload": workload, "desc": WORKLOAD_DESC[workload], "vocoder": bool(runner.vocoder),
               "precision_mode": args.precision, "valid_frames_per_step": total_frames,
               "l2": "256 MiB buffer written between timed iterations", "wall_s": wall}
        if sweep:
            plan = runner.plan
            cfg.update(utterances=sum(len(s.utterances) for s in plan.shards), shards=len(plan.shards),
                       schedule="dynamic (shared counter, planned shards, most expensive first)" if runner.queue else "static (shard.plan_shards assignment)",
                       shards_rank0_last_sweep=len(runner.ran), planner_imbalance=plan.imbalance(),
                       per_rank_ms_per_step=[x[0].item() / args.steps for x in per_rank],
                       # no communication on the data path: one GPU would need the sum of the ranks' times, so
                       # mean / max over ranks IS the strong-scaling efficiency of this run
                       balance_efficiency=sum(x[0].item() for x in per_rank) / (world * dev_ms_max))
        else:
            cfg.update(batch_per_gpu=inf0["B"], padded_frames=inf0["T"])
        line = {
            "metric": "mel_frames_per_s", "value": value, "unit": "mel-frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_ms_max / args.steps, "higher_is_better": True,
            "scaling": "strong" if sweep else "weak",
            "vs_baseline": None, "dtype": "f32" if args.precision == "fp32" else "bf16", "data": "synthetic",
            "config": cfg,
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "mel-frames/s", "h2d_bytes_per_step": h2d_all, "d2h_bytes_per_step": d2h_all,
                    "ms_per_step": e2e_ms_max / args.steps, "steps": e2e_steps},
            "gpu_launches": launches,
            "roofline": roof,
            "roofline_all_convs": roof_all,
        }
        if runner.vocoder:
            line["audio_seconds_per_s"] = value * HOP / SR
            line["e2e"]["audio_seconds_per_s"] = e2e_value * HOP / SR
        if not runner.text:
            line["config"]["frames_are"] = "reference-mel frames encoded (ReferenceEncoder only)"
        try:
            line["roofline_length_regulator"] = lr_roofline(dev, pk)
        except Exception as e:  # pragma: no cover
            line["roofline_length_regulator"] = {"error": str(e)}
        if world == 1 and not args.no_extra:
            del runner  # free the headline model before building the secondary ones
            torch.cuda.empty_cache()
            extra = []
            other = "bf16" if args.precision == "fp32" else "fp32"
            # (workload, precision, corpus limit, steps, warm-up).  The cfg5 entry is the WHOLE 2048-utterance sweep on this
            # one GPU - the same workload `--gpus N` runs - so that strong scaling can be read against a one-GPU figure
            # of the same corpus (3 timed sweeps of ~4 s each)
            jobs = [(workload, other, None, 10, 3)]
            jobs += [(w, "fp32", lim, st, wu) for w, lim, st, wu in (("cfg2", None, 10, 3), ("cfg1", None, 10, 3),
                                                                     ("cfg5", None, 3, 3), ("cfg4", 1024, 10, 3))
                     if w != workload]
            for wl, pr, lim, st, wu in jobs:
                try:
                    extra.append(quick_measure(wl, pr, dev, steps=st, warmup=wu, limit=lim))
                except Exception as e:  # pragma: no cover
                    extra.append({"workload": wl, "precision_mode": pr, "error": str(e)})
                torch.cuda.empty_cache()
            line["extra"] = extra
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(workload)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
