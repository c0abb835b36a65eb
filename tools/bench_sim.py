"""Micro-benchmark of op_sim_dual_softmax at the headline shape (CUDA events; used for kernel experiments)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from onepose_b200._lib import call, lib, ptr, stream_ptr

B, N2, N3 = int(os.environ.get("B", 32)), 2048, 8192
g = torch.Generator().manual_seed(0)
m2 = torch.nn.functional.normalize(torch.randn(B, N2, 256, generator=g), dim=2).cuda()
m3 = torch.nn.functional.normalize(torch.randn(B, N3, 256, generator=g), dim=2).cuda()
conf = torch.empty(B, N2, N3, device="cuda")
i0 = torch.empty(B, N2, dtype=torch.int64, device="cuda"); i1 = torch.empty(B, N3, dtype=torch.int64, device="cuda")
s0 = torch.empty(B, N2, device="cuda"); s1 = torch.empty(B, N3, device="cuda")
nbytes = lib.op_sim_dual_softmax_workspace_bytes(B, N2, N3)
ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
def run():
    call("op_sim_dual_softmax", ptr(m2), ptr(m3), B, N2, N3, 0.07, 0.2, ptr(conf), ptr(i0), ptr(i1), ptr(s0), ptr(s1), ptr(ws), nbytes, stream_ptr())
for _ in range(3): run()
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(10): run()
e.record(); torch.cuda.synchronize()
ms = s.elapsed_time(e) / 10
alg = B * (4 * 256 * (N2 + N3) + 4 * N2 * N3 + 24 * (N2 + N3))
print(f"dbg={os.environ.get('OPB_SIM_DBG','0')} B={B}: {ms:.3f} ms/call, {alg / ms / 1e6:.0f} GB/s algorithmic")
