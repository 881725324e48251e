"""cProfile of the host side of one cfg2 step (where do the ~100 us per launch go?)."""
import cProfile, pstats, sys, tempfile, torch
sys.path.insert(0, ".")
import speakingstyle_b200 as ss
from speakingstyle_b200 import configs as C, synth as S
dev = torch.device("cuda", 0)
ds, batch, voc = S.make_workload("cfg2", seed=1)
d = C.write_preprocessed_dir(tempfile.mkdtemp(), ds)
mc, pc = C.model_config(ds), C.preprocess_config(ds, d)
model = ss.get_model(pc, mc, dev, S.fastspeech2_state_dict(mc, ds))
b = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in batch.items()}
def step():
    return model(b["speakers"], b["texts"], b["src_lens"], b["max_src_len"], b["mels"], b["mel_lens"], b["max_mel_len"])
for _ in range(3): step()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(5): step()
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr); st.sort_stats("tottime").print_stats(28)
