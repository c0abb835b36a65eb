"""Micro-benchmark of op_linear_tc in the matcher's configurations at the headline shape (CUDA events).
   env: B (64), N (8192), OPB_LINEAR_CG, OPB_LIN_DBG"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from onepose_b200._lib import call, lib, ptr, stream_ptr

B, N = int(os.environ.get("B", 64)), int(os.environ.get("N", 8192))
dev = "cuda"
h = lambda *s: (torch.randn(*s, device=dev) * 8).half()
parts = lib.op_linear_tc_stat_parts(N)
x_p = (h(B, N, 256), h(B, N, 256)); q_p = (h(B, N, 256), h(B, N, 256)); hn_p = (h(B, N, 512), h(B, N, 512))
W256 = (h(256, 256), h(256, 256)); W0a = (h(512, 256), h(512, 256)); P = (h(B, 512, 256), h(B, 512, 256)); W1s = (h(B, 256, 512), h(B, 256, 512))
bias = torch.randn(512, device=dev); cb = torch.randn(B, 512, device=dev); ksum = torch.rand(B, 256, device=dev) + 1
x = torch.randn(B, N, 256, device=dev)
eye_p = ((64.0 * torch.eye(256, device=dev)).half(), torch.zeros(256, 256, device=dev).half())
st256 = torch.empty(B, parts, 256, 2, device=dev); st512 = torch.empty(B, parts, 512, 2, device=dev)


def lin(A1, W, bias, Cout, *, A2=None, W2=None, C=None, Cp=None, flags=0, elu=0, stats=None, ksum=None):
    K1 = A1[0].shape[-1]; K2 = 0 if A2 is None else A2[0].shape[-1]
    call("op_linear_tc", ptr(A1[0]), ptr(A1[1]), K1, ptr(A2[0]) if A2 else None, ptr(A2[1]) if A2 else None, K2, ptr(W[0]), ptr(W[1]),
         ptr(W2[0]) if W2 else None, ptr(W2[1]) if W2 else None, ptr(bias), ptr(C), ptr(Cp[0]) if Cp else None, ptr(Cp[1]) if Cp else None,
         B, N, Cout, flags, elu, ptr(stats), ptr(ksum), float(N), stream_ptr())


cases = {
    "k    (ELU1|COLSUM, planes)": (lambda: lin(x_p, W256, bias, 256, Cp=q_p, flags=1 | 16, elu=256, stats=st256), 2.0 * B * N * 256 * 256),
    "q    (ELU1|COLSUM|ZSCALE, planes)": (lambda: lin(x_p, W256, bias, 256, Cp=q_p, flags=1 | 16 | 32, elu=256, stats=st256, ksum=ksum), 2.0 * B * N * 256 * 256),
    "mlp0 (BIASB|CSTATS_RELU|W2B, planes)": (lambda: lin(x_p, W0a, cb, 512, A2=q_p, W2=P, Cp=hn_p, flags=128 | 256 | 512, stats=st512), 2.0 * B * N * 512 * 512),
    "mlp1 (RESIDUAL|COLSUM|WB, fp32+planes)": (lambda: lin(hn_p, W1s, bias, 256, C=x, Cp=x_p, flags=2 | 16 | 64, stats=st256), 2.0 * B * N * 512 * 256),
    "mlp1p (COLSUM|WB|W2_IDENTITY, planes, in place)": (lambda: lin(hn_p, W1s, bias, 256, A2=x_p, W2=eye_p, Cp=x_p, flags=16 | 64 | 1024 | 8192, stats=st256), 2.0 * B * N * 512 * 256),
    "final (fp32)": (lambda: lin(x_p, W256, bias, 256, C=x), 2.0 * B * N * 256 * 256),
    "plain planes": (lambda: lin(x_p, W256, bias, 256, Cp=q_p), 2.0 * B * N * 256 * 256),
}
print(f"B={B} N={N} cg={os.environ.get('OPB_LINEAR_CG', '2')} dbg={os.environ.get('OPB_LIN_DBG', '0')} pf={os.environ.get('OPB_LIN_PF', 'default')}")
for name, (fn, flops) in cases.items():
    for _ in range(2): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(5): fn()
    e.record(); torch.cuda.synchronize()
    ms = s.elapsed_time(e) / 5
    tiles = B * ((N + 127) // 128) * (2 if "mlp0" in name else 1)
    print(f"  {name:42s} {ms * 1e3:8.1f} us  {flops / ms / 1e9:7.1f} TF/s  {ms * 1e3 / (tiles / 148):6.2f} us/tile/SM", flush=True)
