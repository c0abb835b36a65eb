"""ctypes binding of ``libonepose_b200.so`` -- the C ABI declared in ``include/onepose_b200.h``.

There is no CPU or PyTorch fallback: if the CUDA library is missing or fails to load, importing this
module raises.  (``torch`` is imported first so the CUDA runtime the library links against is the one
already in the process.)
"""
from __future__ import annotations

import ctypes as C
import os

import torch  # noqa: F401  (loads libcudart before our library)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libonepose_b200.so")

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build the sm_100a CUDA library first "
        "(python -m onepose_b200.build, or __graft_entry__.build()). onepose_b200 has no CPU fallback.")

lib = C.CDLL(LIB_PATH)

vp, i32, i64, f32, f64, sz = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_double, C.c_size_t

_SIGS = {
    "op_last_error": (C.c_char_p, []),
    "op_abi_version": (i32, []),
    "op_channel_major_to_point_major": (i32, [vp, vp, i32, i32, i32, i64, vp]),
    "op_channel_major_to_planes": (i32, [vp, i64, vp, vp, vp, i32, i32, vp]),
    "op_channel_major_to_planes_ragged": (i32, [vp, i64, vp, vp, vp, i32, i32, vp, vp]),
    "op_gats_leaf_aggregate": (i32, [vp, i64, vp, vp, vp, vp, vp, i32, i32, i32, f32, i32, i32, vp]),
    "op_linear": (i32, [vp, i32, vp, i32, vp, vp, vp, i32, i32, i32, i32, i32, vp, vp, vp, vp]),
    "op_linear_row_tiles": (i32, [i32]),
    "op_instnorm_finalize": (i32, [vp, i32, i32, i32, i32, f32, vp, vp, vp]),
    "op_kv_reduce_workspace_bytes": (sz, [i32, i32]),
    "op_kv_reduce": (i32, [vp, vp, vp, i32, i32, vp, sz, vp]),
    "op_kv_reduce_static_max_bytes": (sz, []),
    "op_kv_reduce_static": (i32, [vp, vp, vp, i32, i32, vp]),
    "op_attn_apply": (i32, [vp, vp, vp, vp, i32, i32, i32, vp]),
    "op_split_planes": (i32, [vp, vp, vp, i64, vp]),
    "op_norm_relu_split": (i32, [vp, vp, vp, vp, vp, i32, i32, i32, vp]),
    "op_gats_planes_parts": (i32, [i32]),
    "op_gats_leaf_aggregate_planes": (i32, [vp, i64, vp, vp, vp, vp, vp, i32, i32, i32, f32, i32, i32, vp]),
    "op_gather_rows": (i32, [vp, vp, i64, i64, vp, vp]),
    "op_segment_mean": (i32, [vp, vp, i64, i32, vp, vp]),
    "op_pose_errors": (i32, [vp, vp, i32, vp, vp, vp]),
    "op_sp_softmax_shuffle": (i32, [vp, i32, i32, i32, i32, vp, vp]),
    "op_conv_tc": (i32, [vp, vp, i32, i32, i32, i32, vp, vp, i32, vp, i32, i32, i32, i32, vp, vp, vp, i32, vp]),
    "op_conv1a_planes": (i32, [vp, vp, vp, i32, i32, i32, vp, vp, vp]),
    "op_snavely_residual": (i32, [vp, vp, vp, vp, vp, i64, vp, vp]),
    "op_project_points": (i32, [vp, vp, vp, i64, vp, vp, vp, vp, vp]),
    "op_triangulate_points": (i32, [vp, vp, vp, vp, i64, vp, vp]),
    "op_ba_workspace_bytes": (sz, [i32, i32, i32]),
    "op_ba_lm": (i32, [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, f64, vp, sz, vp, vp]),
    "op_pool2_planes": (i32, [vp, vp, i32, i32, i32, i32, vp, vp, vp]),
    "op_sp_nms_workspace_bytes": (sz, [i32, i32, i32]),
    "op_sp_nms": (i32, [vp, i32, i32, i32, i32, vp, vp, sz, vp]),
    "op_sp_select": (i32, [vp, i32, i32, i32, f32, i32, i32, i32, vp, vp, vp, vp, vp]),
    "op_sp_desc_prepare": (i32, [vp, i32, i32, vp, vp]),
    "op_sp_sample": (i32, [vp, vp, vp, i32, i32, i32, i32, i32, i32, vp, vp]),
    "op_linear_tc": (i32, [vp, vp, i32, vp, vp, i32, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, vp, f32, vp]),
    "op_linear_tc_ragged": (i32, [vp, vp, i32, vp, vp, i32, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, vp, f32, vp, vp]),
    "op_linear_tc_stat_parts": (i32, [i32]),
    "op_colsum_partial": (i32, [vp, i32, i32, i32, vp, vp]),
    "op_mlp_center_bias": (i32, [vp, i32, vp, i32, vp, vp, vp, i32, i32, i32, vp, vp]),
    "op_mlp_center_bias_ragged": (i32, [vp, i32, vp, i32, vp, vp, vp, i32, i32, i32, vp, vp, vp]),
    "op_scale_weight_planes": (i32, [vp, vp, i32, i32, i32, vp, vp, vp]),
    "op_instnorm_finalize_tc": (i32, [vp, vp, i32, i32, i32, i32, f32, vp, vp, vp]),
    "op_instnorm_scale_weight_planes": (i32, [vp, i32, i32, i32, f32, vp, i32, i32, vp, vp, vp, vp]),
    "op_instnorm_scale_weight_planes_ragged": (i32, [vp, i32, i32, i32, vp, f32, vp, i32, i32, vp, vp, vp, vp]),
    "op_atb_chunks": (i32, [i32]),
    "op_atb_tc": (i32, [vp, vp, vp, vp, i32, i32, vp, vp]),
    "op_kv_post": (i32, [vp, i32, vp, i32, vp, vp, vp, i32, i32, f32, vp, vp, vp, vp, vp]),
    "op_attn_apply_planes": (i32, [vp, vp, vp, vp, vp, i32, i32, i32, vp]),
    "op_l2_normalize_rows": (i32, [vp, vp, i64, i32, vp]),
    "op_sim_dual_softmax_workspace_bytes": (sz, [i32, i32, i32]),
    "op_sim_dual_softmax": (i32, [vp, vp, i32, i32, i32, f32, f32, vp, vp, vp, vp, vp, vp, sz, vp]),
    "op_sim_dual_softmax_simt": (i32, [vp, vp, i32, i32, i32, f32, f32, vp, vp, vp, vp, vp, vp, sz, vp]),
    "op_sim_dual_softmax_rows": (i32, [vp, vp, i32, i32, i32, f32, f32, vp, vp, vp, vp, vp, vp, sz, vp]),
    "op_sim_dual_softmax_planes": (i32, [vp, vp, vp, vp, i32, i32, i32, f32, f32, vp, vp, vp, vp, vp, vp, sz, vp]),
    "op_sim_dual_softmax_planes_ragged": (i32, [vp, vp, vp, vp, i32, i32, i32, vp, f32, f32, vp, vp, vp, vp, vp, vp, sz, vp]),
    "op_gather_matches": (i32, [vp, vp, vp, vp, i64, i32, i32, i32, f64, vp, vp, vp, vp, vp, vp]),
    "op_pnp_workspace_bytes": (sz, [i32, i32, i32]),
    "op_pnp_cv_samples": (i32, [vp, i32, i32, vp, vp, vp]),
    "op_pnp_refine_lm": (i32, [vp, vp, vp, i32, i32, vp, vp, vp, f64, i32, vp, vp, vp]),
    "op_pnp_ransac": (i32, [vp, vp, vp, i32, i32, vp, vp, vp, i32, f64, f64, i32, f64,
                            vp, vp, vp, vp, vp, vp, vp, sz, vp]),
}
EXPORTS = tuple(_SIGS)

for _name, (_res, _args) in _SIGS.items():
    _fn = getattr(lib, _name)          # AttributeError here = the library does not match the header
    _fn.restype = _res
    _fn.argtypes = _args


class OpError(RuntimeError):
    pass


def check(rc, what=""):
    if rc != 0:
        raise OpError(f"{what} failed (code {rc}): {lib.op_last_error().decode()}")


def ptr(t):
    """Device pointer of a tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()


def stream_ptr():
    return torch.cuda.current_stream().cuda_stream


# kernels launched by each entry point (for the bench's gpu_launches claim)
LAUNCHES = {
    "op_atb_tc": 1, "op_kv_post": 1, "op_split_planes": 1, "op_norm_relu_split": 1, "op_linear_tc": 1, "op_instnorm_finalize_tc": 1, "op_attn_apply_planes": 1,
    "op_channel_major_to_point_major": 1, "op_channel_major_to_planes": 1, "op_gats_leaf_aggregate": 1, "op_linear": 1, "op_instnorm_finalize": 1,
    "op_kv_reduce": 2, "op_kv_reduce_static": 2, "op_attn_apply": 1, "op_l2_normalize_rows": 1, "op_sim_dual_softmax": 9, "op_sim_dual_softmax_rows": 9, "op_sim_dual_softmax_planes": 7, "op_sim_dual_softmax_planes_ragged": 7, "op_linear_tc_ragged": 1, "op_channel_major_to_planes_ragged": 1, "op_sim_dual_softmax_simt": 7,
    "op_gather_matches": 1, "op_pnp_cv_samples": 1, "op_pnp_ransac": 25,
}


class Timeline:
    """Optional per-call CUDA-event timing on the launching stream (used by bench.py only)."""

    def __init__(self):
        self.enabled = False
        self.records = []        # (name, tag, start_event, end_event)
        self.launches = 0
        self.tag = ""
        self.work = 0.0          # algorithmic flops / bytes of the next call (set by the host code)
        self.bytes = 0.0         # tensor-bound calls: the operand / result bytes of the next call as well (its HBM side)
        self.byte_totals = {}    # key -> bytes accumulated over the recorded calls

    def reset(self):
        self.records.clear()
        self.byte_totals.clear()
        self.launches = 0

    def summary(self):
        """name -> (total ms, calls, total work); events must have completed (synchronize first)."""
        out = {}
        for name, tag, s, e, work in self.records:
            key = name if not tag else f"{name}:{tag}"
            ms, n, w = out.get(key, (0.0, 0, 0.0))
            out[key] = (ms + s.elapsed_time(e), n + 1, w + work)
        return out


timeline = Timeline()


def call(name, *args, tag=None):
    """Invoke an ABI entry point, raise on a non-zero return code, optionally time it with CUDA events."""
    fn = getattr(lib, name)
    if timeline.enabled:
        s = torch.cuda.Event(enable_timing=True)
        e = torch.cuda.Event(enable_timing=True)
        s.record()
        rc = fn(*args)
        e.record()
        tg = tag if tag is not None else timeline.tag
        timeline.records.append((name, tg, s, e, timeline.work))
        if timeline.bytes:
            key = name if not tg else f"{name}:{tg}"
            timeline.byte_totals[key] = timeline.byte_totals.get(key, 0.0) + timeline.bytes
        timeline.launches += LAUNCHES.get(name, 1)
    else:
        rc = fn(*args)
    timeline.work = 0.0
    timeline.bytes = 0.0
    check(rc, name)
