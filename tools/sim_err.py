import sys, os, torch
sys.path.insert(0, "/root/repo")
from tests import helpers
from tests.test_gpu_matcher import _run_sim
for shape in [(1, 2048, 8192), (2, 300, 500)]:
    B, N2, N3 = shape
    g = torch.Generator().manual_seed(N2 * 7 + N3)
    m2 = torch.nn.functional.normalize(torch.randn(B, N2, 256, generator=g), dim=2)
    m3 = torch.nn.functional.normalize(torch.randn(B, N3, 256, generator=g), dim=2)
    k = min(N2, N3) // 2
    m3[:, :k] = torch.nn.functional.normalize(m2[:, :k] + 0.05 * torch.randn(B, k, 256, generator=g), dim=2)
    tc = _run_sim("op_sim_dual_softmax", m2.cuda(), m3.cuda())
    s = torch.matmul(m2.double(), m3.double().transpose(1, 2)) / 0.07
    ref = torch.softmax(s, 1) * torch.softmax(s, 2)
    print(os.environ.get("OPB_SIM_MMA"), shape, "conf rel err", float(helpers.conf_rel_err(tc[0].cpu(), ref)))
