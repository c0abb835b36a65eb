"""Golden fixture for SURVEY.md 8(f) row f3, from the UNMODIFIED reference SuperPoint run in the build container.

    python tests/golden/make_golden_superpoint.py        (needs /root/reference; CPU only)

superpoint.npz holds
* stage_*  : outputs of the reference's module-level functions simple_nms / remove_borders / top_k_keypoints /
             sample_descriptors (src/models/extractors/SuperPoint/superpoint.py:47-113) on the seeded arrays of
             ``oracle.superpoint_oracle.sp_inputs`` (only outputs are stored).  sample_descriptors is run twice: as this
             container's torch 2.11 runs it (the version test at :106 picks align_corners=False) and with
             ``torch.__version__`` reading "1.8.0", the reference's pinned version (align_corners=True);
* net_*    : ``SuperPoint.forward`` (:172-246) with ``torch.manual_seed(7)`` initial weights on ``sp_image``; the detector
             logits and the dense descriptor map are captured by forward hooks on convPb / convDb so that the post-network
             stages can be checked on exactly the reference's intermediate values.  Two configs: the SfM one of
             src/sfm/extract_features.py:9-17 and one with max_keypoints=40 (top-k path).
"""
import os
import sys

sys.dont_write_bytecode = True
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, "/root/reference")
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import numpy as np
import torch

from oracle import superpoint_oracle as spo

SFM_CONF = {'descriptor_dim': 256, 'nms_radius': 3, 'max_keypoints': 4096, 'keypoints_threshold': 0.6}   # extract_features.py:9-17
TOPK_CONF = {'nms_radius': 4, 'max_keypoints': 40, 'keypoint_threshold': 0.016, 'remove_borders': 6}


def main():
    import src.models.extractors.SuperPoint.superpoint as ref

    torch.set_num_threads(1)
    out = {}
    # ---- stage level
    logits, dense = spo.sp_inputs()
    scores = spo.softmax_shuffle(logits)                       # the oracle's map; the stages below are the reference's
    for r in (0, 1, 3, 4):
        out[f"stage_nms_r{r}"] = ref.simple_nms(torch.from_numpy(scores), r).numpy()
    nms = torch.from_numpy(out["stage_nms_r3"])[0]
    H, W = nms.shape
    for name, (thr, border, mk) in {"a": (0.005, 4, -1), "b": (0.05, 0, 25), "c": (0.005, 10, 60)}.items():
        k = torch.nonzero(nms > thr)
        s = nms[tuple(k.t())]
        k, s = ref.remove_borders(k, s, border, H, W)
        if mk >= 0:
            k, s = ref.top_k_keypoints(k, s, mk)
        out[f"stage_sel_{name}_kpts"] = torch.flip(k, [1]).float().numpy()
        out[f"stage_sel_{name}_scores"] = s.numpy()
    kp = torch.from_numpy(out["stage_sel_a_kpts"])
    nd = torch.nn.functional.normalize(torch.from_numpy(dense), p=2, dim=1)
    out["stage_desc_ac0"] = ref.sample_descriptors(kp[None].clone(), nd, 8)[0].numpy()
    assert int(torch.__version__[2]) <= 2
    real = torch.__version__
    torch.__version__ = "1.8.0"
    try:
        out["stage_desc_ac1"] = ref.sample_descriptors(kp[None].clone(), nd, 8)[0].numpy()
    finally:
        torch.__version__ = real

    # ---- whole network
    img = torch.from_numpy(spo.sp_image())
    for tag, conf in (("sfm", SFM_CONF), ("topk", TOPK_CONF)):
        torch.manual_seed(7)
        model = ref.SuperPoint(conf).eval()
        cap = {}
        h1 = model.convPb.register_forward_hook(lambda m, i, o: cap.__setitem__("logits", o.detach().clone()))
        h2 = model.convDb.register_forward_hook(lambda m, i, o: cap.__setitem__("dense", o.detach().clone()))
        with torch.no_grad():
            pred = model(img)
        h1.remove(); h2.remove()
        out["net_logits"] = cap["logits"].numpy()             # same weights and image for both configs
        out["net_dense"] = cap["dense"].numpy()
        out[f"net_{tag}_kpts"] = pred["keypoints"][0].numpy()
        out[f"net_{tag}_scores"] = pred["scores"][0].numpy()
        out[f"net_{tag}_desc"] = pred["descriptors"][0].numpy()
        print(tag, "keypoints:", pred["keypoints"][0].shape[0])
    out["torch_version"] = np.array(real)
    np.savez_compressed(os.path.join(HERE, "superpoint.npz"), **out)
    print("wrote superpoint.npz", os.path.getsize(os.path.join(HERE, "superpoint.npz")))


if __name__ == "__main__":
    main()
