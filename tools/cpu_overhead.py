import sys, time, torch
sys.path.insert(0, "/root/repo")
from onepose_b200 import synth
from onepose_b200.matcher import GATsSuperGlue
m = GATsSuperGlue(dict(synth.DEFAULT_HPARAMS)).cuda().eval(); m.load_state_dict(synth.make_weights(0, damped=True), strict=True)
obj = synth.make_object(0, 8192, 8)
fr = [synth.make_frame(obj, f, 2048) for f in range(64)]
d = {k: v.cuda() for k, v in synth.pack_batch(obj, fr).items()}; d["query_intrinsic"] = torch.as_tensor(fr[0].K).cuda()
m.pnp_overlap = True
for _ in range(3): m(d)
torch.cuda.synchronize()
t = time.perf_counter()
for _ in range(10): m(d)
t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
print(f"CPU enqueue time per step {1e3 * (t1 - t) / 10:.2f} ms; GPU-complete per step {1e3 * (t2 - t) / 10:.2f} ms")
