"""Measurement of SURVEY.md 8(f) rows f1 / f2 at the headline shape: object-DB build (reference CPU restatement vs GPU) and the
device-side sequence loop (frames/s including upload of every frame from pinned host memory and the evaluator)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from onepose_b200 import objectdb, synth
from onepose_b200.evaluator import run_sequence
from onepose_b200.matcher import GATsSuperGlue
from oracle import objectdb_oracle as orc

N3, L, N2, F = 8192, 8, 2048, 256
rs = np.random.RandomState(0)
idxs = rs.randint(1, 13, size=N3)
clt = rs.standard_normal((int(idxs.sum()), 256)).astype(np.float32)
sc = rs.rand(clt.shape[0], 1).astype(np.float32)
t0 = time.perf_counter(); np.random.seed(1)
avg_ref = orc.mean_descriptors(clt, idxs)
d_ref, _ = orc.pad_features3d_random(avg_ref.T, orc.mean_descriptors(sc, idxs), N3)
l_ref, _ = orc.build_features3d_leaves(clt.T, sc, idxs, N3, L)
t_cpu = time.perf_counter() - t0
clt_dev, sc_dev = torch.from_numpy(clt).cuda(), torch.from_numpy(sc).cuda()
for it in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter(); np.random.seed(1)
    avg = objectdb.mean_descriptors(clt_dev, idxs); avs = objectdb.mean_scores(sc_dev, idxs)
    db = objectdb.ObjectDB.from_arrays(rs.rand(N3, 3).astype(np.float32), avg.t(), avs, clt_dev.t(), sc_dev, idxs, L, "cuda")
    torch.cuda.synchronize(); t_gpu = time.perf_counter() - t0
print(f"DB build N3={N3} L={L} ({clt.shape[0]} collected leaves): CPU restatement {t_cpu * 1e3:.0f} ms, GPU {t_gpu * 1e3:.1f} ms "
      f"(host index plan included), identical: {np.array_equal(db.descriptors2d_db.cpu().numpy(), l_ref)}")

obj = synth.make_object(0, N3, L)
frames = [synth.make_frame(obj, f, N2) for f in range(64)]
q = synth.pack_batch(obj, frames)
db = objectdb.ObjectDB(torch.from_numpy(obj.keypoints3d).cuda(), q["descriptors3d_db"][0].t().contiguous().cuda(),
                       q["descriptors2d_db"][0].t().contiguous().cuda(), L)
m = GATsSuperGlue(dict(synth.DEFAULT_HPARAMS)).cuda().eval()
m.load_state_dict(synth.make_weights(0, damped=True), strict=True)
kp = q["keypoints2d"].repeat(F // 64, 1, 1).pin_memory(); ds = q["descriptors2d_query"].repeat(F // 64, 1, 1).pin_memory()
gt = np.tile(np.stack([f.pose_gt for f in frames]), (F // 64, 1, 1))
run_sequence(m, db, kp[:64], ds[:64], frames[0].K, gt[:64])
torch.cuda.synchronize(); t0 = time.perf_counter()
res = run_sequence(m, db, kp, ds, frames[0].K, gt)
torch.cuda.synchronize(); dt = time.perf_counter() - t0
print(f"run_sequence: {F} frames in {dt * 1e3:.1f} ms = {F / dt:.0f} frames/s (upload + match + PnP + evaluator); eval {res['eval']}")
