"""op_kv_reduce_static: the KV reduction of linear attention with its scratch in a static device array (no allocation).

Collected last on purpose.  The entry point has no counterpart in the reference;
it is checked through the C ABI against the existing ``op_kv_reduce`` (caller-owned
workspace): same kernels, same summation order, so the repo's bar for such a pair applies -- identical bits -- and against a
float64 restatement of GATs_SuperGlue.py:93-97 with the tolerance of the other fp32 kernel tests.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(B, M, seed):
    g = torch.Generator().manual_seed(seed)
    kv = torch.randn(B, M, 512, generator=g)
    kv[:, :, :256] = torch.nn.functional.elu(kv[:, :, :256]) + 1           # K' = elu(K) + 1 (:91); V as it comes
    return kv.cuda().contiguous()


def _with_workspace(kv):
    from onepose_b200._lib import call, lib, ptr, stream_ptr
    B, M, _ = kv.shape
    KV = torch.empty(B, 4, 64, 64, device="cuda"); Ks = torch.empty(B, 4, 64, device="cuda")
    nbytes = lib.op_kv_reduce_workspace_bytes(B, M)
    ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    call("op_kv_reduce", ptr(kv), ptr(KV), ptr(Ks), B, M, ptr(ws), nbytes, stream_ptr())
    torch.cuda.synchronize()
    return KV, Ks


def _static(kv, KV, Ks):
    from onepose_b200._lib import call, ptr, stream_ptr
    B, M, _ = kv.shape
    call("op_kv_reduce_static", ptr(kv), ptr(KV), ptr(Ks), B, M, stream_ptr())


def _float64(kv):
    B, M, _ = kv.shape
    k = kv[:, :, :256].double().view(B, M, 4, 64); v = kv[:, :, 256:].double().view(B, M, 4, 64) / M
    return torch.einsum("bmhd,bmhq->bhdq", k, v), k.sum(1)


def test_static_scratch_entry_point_equals_workspace_entry_point_on_calls_in_a_row():
    from onepose_b200._lib import lib
    assert lib.op_kv_reduce_static_max_bytes() == 8 << 20
    # large (3 frames at the headline M, 6.4 of the 8 MiB), then a small ragged shape with other data (stale partials of the first
    # call sit behind it in the shared array), then the large one again
    cases = [_inputs(3, 8192, 1), _inputs(2, 1000, 2), _inputs(3, 8192, 3), _inputs(1, 255, 4)]
    outs = []
    for kv in cases:                                       # calls in a row on one stream, no synchronisation in between
        B = kv.shape[0]
        KV = torch.full((B, 4, 64, 64), float("nan"), device="cuda"); Ks = torch.full((B, 4, 64), float("nan"), device="cuda")
        _static(kv, KV, Ks)
        outs.append((KV, Ks))
    torch.cuda.synchronize()
    for kv, (KV, Ks) in zip(cases, outs):
        KV_ref, Ks_ref = _with_workspace(kv)
        assert torch.equal(KV, KV_ref) and torch.equal(Ks, Ks_ref)
        KV64, Ks64 = _float64(kv)
        torch.testing.assert_close(KV.double(), KV64, rtol=2e-5, atol=2e-5)
        torch.testing.assert_close(Ks.double(), Ks64, rtol=2e-5, atol=2e-5)


def test_static_scratch_entry_point_allocates_nothing_and_rejects_oversize():
    from onepose_b200._lib import lib, ptr, stream_ptr
    kv = _inputs(2, 4096, 5)
    KV = torch.empty(2, 4, 64, 64, device="cuda"); Ks = torch.empty(2, 4, 64, device="cuda")
    _static(kv, KV, Ks)                                    # first use on this device: symbol address resolved, kernels loaded
    torch.cuda.synchronize()
    first = (KV.clone(), Ks.clone())
    free0, alloc0, reserved0 = torch.cuda.mem_get_info()[0], torch.cuda.memory_allocated(), torch.cuda.memory_reserved()
    for _ in range(2):
        _static(kv, KV, Ks)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0            # neither cudaMalloc nor torch's allocator was touched
    assert torch.cuda.memory_allocated() == alloc0 and torch.cuda.memory_reserved() == reserved0
    assert torch.equal(KV, first[0]) and torch.equal(Ks, first[1])
    # beyond the static array: a clear error, nothing launched, outputs untouched
    B, M = 4, 8192
    assert lib.op_kv_reduce_workspace_bytes(B, M) > lib.op_kv_reduce_static_max_bytes()
    big = torch.empty(1, device="cuda")                    # never dereferenced: the size check comes first
    keep = KV.clone()
    rc = lib.op_kv_reduce_static(ptr(big), ptr(KV), ptr(Ks), B, M, stream_ptr())
    assert rc == -3                                        # OPB_ERR_WORKSPACE
    msg = lib.op_last_error().decode()
    assert "op_kv_reduce_static" in msg and "8388608" in msg and "op_kv_reduce" in msg
    torch.cuda.synchronize()
    assert torch.equal(KV, keep)
    assert lib.op_kv_reduce_static(None, ptr(KV), ptr(Ks), 1, 16, stream_ptr()) == -1      # OPB_ERR_ARG


def test_static_scratch_entry_point_under_cuda_graph_capture():
    """What the entry point is for: inside a capture nothing may allocate.  The address was resolved by the calls above (or by the
    warm-up here when this test runs alone)."""
    kv = _inputs(2, 2048, 6)
    KV = torch.empty(2, 4, 64, 64, device="cuda"); Ks = torch.empty(2, 4, 64, device="cuda")
    _static(kv, KV, Ks)
    torch.cuda.synchronize()
    ref = (KV.clone(), Ks.clone())
    KV.zero_(); Ks.zero_()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        _static(kv, KV, Ks)
    for _ in range(2):
        KV.zero_(); Ks.zero_()
        g.replay()
        torch.cuda.synchronize()
        assert torch.equal(KV, ref[0]) and torch.equal(Ks, ref[1])
