"""Is the cfg2 step limited by the host enqueue rate?  Compares the wall time the Python call takes to return
(no sync) with the device time of the same step, with and without the per-conv event timer."""
import sys, tempfile, time, torch
sys.path.insert(0, ".")
import speakingstyle_b200 as ss
from speakingstyle_b200 import configs as C, ops, synth as S
dev = torch.device("cuda", 0)
ds, batch, voc = S.make_workload(sys.argv[1] if len(sys.argv) > 1 else "cfg2", seed=1)
d = C.write_preprocessed_dir(tempfile.mkdtemp(), ds)
mc, pc = C.model_config(ds), C.preprocess_config(ds, d)
model = ss.get_model(pc, mc, dev, S.fastspeech2_state_dict(mc, ds), precision=sys.argv[2] if len(sys.argv) > 2 else "fp32")
gen = ss.get_vocoder(C.hifigan_config(), dev, S.hifigan_state_dict(), precision=sys.argv[2] if len(sys.argv) > 2 else "fp32") if voc else None
b = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in batch.items()}
def step():
    out = model(b["speakers"], b["texts"], b["src_lens"], b["max_src_len"], b["mels"], b["mel_lens"], b["max_mel_len"])
    if gen is not None:
        gen(out[1], channels_last=True, pcm=True)
for _ in range(3): step()
torch.cuda.synchronize()
for label, timer in (("no timer", None), ("conv timer", ops.KernelTimer())):
    ops.conv_timer = timer
    host, devt = [], []
    for _ in range(5):
        torch.cuda.synchronize()
        a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ops.launch_count_reset()
        t0 = time.perf_counter(); a.record(); step(); e.record(); t1 = time.perf_counter()
        torch.cuda.synchronize()
        host.append((t1 - t0) * 1e3); devt.append(a.elapsed_time(e))
    print(f"{label:10s}: host enqueue {sorted(host)[2]:.2f} ms, device {sorted(devt)[2]:.2f} ms, launches {ops.launch_count()}")
ops.conv_timer = None
