"""ctypes binding of libipoke_inn.so (include/ipoke_inn.h).  No fallback: a missing library is an error."""
import ctypes
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "libipoke_inn.so")

PREC_FP32_SPLIT = 0
PREC_BF16 = 1
PRECISIONS = {"fp32": PREC_FP32_SPLIT, "bf16": PREC_BF16}

# every symbol include/ipoke_inn.h declares
SYMBOLS = [
    "ipoke_inn_version", "ipoke_inn_last_error_string", "ipoke_inn_plan_create", "ipoke_inn_plan_destroy",
    "ipoke_inn_packed_bytes", "ipoke_inn_workspace_bytes", "ipoke_inn_num_linears", "ipoke_inn_pack",
    "ipoke_inn_forward", "ipoke_inn_reverse", "ipoke_inn_forward_host", "ipoke_inn_reverse_host",
    "ipoke_inn_actnorm_stats", "ipoke_inn_actnorm_finalize", "ipoke_inn_flow_loss",
    "ipoke_inn_last_launch_count", "ipoke_inn_debug_layout", "ipoke_inn_debug_trace",
    "ipoke_inn_profile_enable", "ipoke_inn_profile_read", "ipoke_inn_set_option",
    "ipoke_inn_saved_bytes", "ipoke_inn_forward_train", "ipoke_inn_backward", "ipoke_inn_profile_read_n",
    "ipoke_inn_adam_chunk_elems", "ipoke_inn_adam_step", "ipoke_inn_reverse_draws", "ipoke_inn_forward_draws",
    "ipoke_inn_param_layout", "ipoke_inn_pack_small", "ipoke_inn_adam_shard_step", "ipoke_inn_master_to_packed",
    "ipoke_inn_grad_to_bf16", "ipoke_inn_flow_loss_ref", "ipoke_inn_flow_loss_ws",
    "ipoke_inn_flow_loss_scratch_bytes", "ipoke_inn_flow_loss_static",
]


class InnDesc(ctypes.Structure):
    _fields_ = [("C", ctypes.c_int32), ("E", ctypes.c_int32), ("H", ctypes.c_int32), ("depth", ctypes.c_int32),
                ("n_flows", ctypes.c_int32), ("conditional", ctypes.c_int32), ("precision", ctypes.c_int32),
                ("alpha", ctypes.c_float), ("mlp_slope", ctypes.c_float)]


_lib = None


def lib():
    """Load the shared library once; raise loudly if it was not built (python __graft_entry__.py build)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). There is no CPU or eager fallback for the INN path.")
    L = ctypes.CDLL(LIB_PATH)
    vp, i32, f32, f64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_double
    L.ipoke_inn_version.restype = i32
    L.ipoke_inn_last_error_string.restype = ctypes.c_char_p
    L.ipoke_inn_last_error_string.argtypes = [i32]
    L.ipoke_inn_plan_create.restype = i32
    L.ipoke_inn_plan_create.argtypes = [ctypes.POINTER(InnDesc), i32, ctypes.POINTER(vp)]
    L.ipoke_inn_plan_destroy.restype = None
    L.ipoke_inn_plan_destroy.argtypes = [vp]
    L.ipoke_inn_packed_bytes.restype = ctypes.c_size_t
    L.ipoke_inn_packed_bytes.argtypes = [vp]
    L.ipoke_inn_workspace_bytes.restype = ctypes.c_size_t
    L.ipoke_inn_workspace_bytes.argtypes = [vp, i32, i32]
    L.ipoke_inn_num_linears.restype = i32
    L.ipoke_inn_num_linears.argtypes = [vp]
    L.ipoke_inn_pack.restype = i32
    L.ipoke_inn_pack.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp]
    L.ipoke_inn_forward.restype = i32
    L.ipoke_inn_forward.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, vp, vp]
    L.ipoke_inn_reverse.restype = i32
    L.ipoke_inn_reverse.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, vp, vp]
    L.ipoke_inn_reverse_draws.restype = i32
    L.ipoke_inn_reverse_draws.argtypes = [vp, vp, vp, vp, vp, i32, i32, vp, vp]
    L.ipoke_inn_forward_draws.restype = i32
    L.ipoke_inn_forward_draws.argtypes = [vp, vp, vp, vp, vp, vp, i32, i32, vp, vp]
    L.ipoke_inn_forward_host.restype = i32
    L.ipoke_inn_forward_host.argtypes = [vp, vp, vp, vp, vp, vp, i32, vp, vp]
    L.ipoke_inn_reverse_host.restype = i32
    L.ipoke_inn_reverse_host.argtypes = [vp, vp, vp, vp, vp, i32, vp, vp]
    L.ipoke_inn_actnorm_stats.restype = i32
    L.ipoke_inn_actnorm_stats.argtypes = [vp, i32, i32, vp, vp]
    L.ipoke_inn_actnorm_finalize.restype = i32
    L.ipoke_inn_actnorm_finalize.argtypes = [vp, f64, i32, vp, vp, vp]
    L.ipoke_inn_flow_loss.restype = i32
    L.ipoke_inn_flow_loss.argtypes = [vp, vp, i32, i32, f32, f32, vp, vp, vp, vp]
    L.ipoke_inn_last_launch_count.restype = i32
    L.ipoke_inn_last_launch_count.argtypes = [vp]
    L.ipoke_inn_debug_layout.restype = i32
    L.ipoke_inn_debug_layout.argtypes = [vp, i32, vp, i32]
    L.ipoke_inn_debug_trace.restype = i32
    L.ipoke_inn_debug_trace.argtypes = [vp, vp, i32]
    L.ipoke_inn_set_option.restype = i32
    L.ipoke_inn_set_option.argtypes = [vp, ctypes.c_char_p, i32]
    L.ipoke_inn_profile_enable.restype = i32
    L.ipoke_inn_profile_enable.argtypes = [vp, i32]
    L.ipoke_inn_profile_read.restype = i32
    L.ipoke_inn_profile_read.argtypes = [vp, vp, vp]
    L.ipoke_inn_saved_bytes.restype = ctypes.c_size_t
    L.ipoke_inn_saved_bytes.argtypes = [vp, i32]
    L.ipoke_inn_forward_train.restype = i32
    L.ipoke_inn_forward_train.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, i32, vp, vp, vp]
    L.ipoke_inn_backward.restype = i32
    L.ipoke_inn_backward.argtypes = [vp] * 14 + [i32, i32, i32, vp, vp]
    L.ipoke_inn_profile_read_n.restype = i32
    L.ipoke_inn_profile_read_n.argtypes = [vp, vp, vp, i32]
    L.ipoke_inn_adam_chunk_elems.restype = i32
    L.ipoke_inn_adam_chunk_elems.argtypes = []
    L.ipoke_inn_adam_step.restype = i32
    L.ipoke_inn_adam_step.argtypes = [vp, vp, vp, i32, f32, f32, f32, f32, f32, i32, i32, f32, vp]
    i64 = ctypes.c_longlong
    L.ipoke_inn_flow_loss_scratch_bytes.restype = i32
    L.ipoke_inn_flow_loss_scratch_bytes.argtypes = []
    L.ipoke_inn_flow_loss_ws.restype = i32
    L.ipoke_inn_flow_loss_ws.argtypes = [vp, vp, vp, i32, i32, f32, f32, vp, vp, vp, vp, vp]
    L.ipoke_inn_flow_loss_ref.restype = i32
    L.ipoke_inn_flow_loss_ref.argtypes = [vp, vp, vp, i32, i32, f32, f32, vp, vp, vp, vp]
    L.ipoke_inn_flow_loss_static.restype = i32
    L.ipoke_inn_flow_loss_static.argtypes = [vp, vp, vp, i32, i32, f32, f32, vp, vp, vp, vp]
    L.ipoke_inn_param_layout.restype = i32
    L.ipoke_inn_param_layout.argtypes = [vp, vp, i32]
    L.ipoke_inn_pack_small.restype = i32
    L.ipoke_inn_pack_small.argtypes = [vp, vp, vp]
    L.ipoke_inn_adam_shard_step.restype = i32
    L.ipoke_inn_adam_shard_step.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, i64, f32, f32, f32, f32, f32, i32, i32, f32, vp]
    L.ipoke_inn_master_to_packed.restype = i32
    L.ipoke_inn_master_to_packed.argtypes = [vp, vp, vp, i64, vp]
    L.ipoke_inn_grad_to_bf16.restype = i32
    L.ipoke_inn_grad_to_bf16.argtypes = [vp, vp, i64, f32, vp]
    _lib = L
    return L


def check(code, what=""):
    if code != 0:
        msg = lib().ipoke_inn_last_error_string(int(code)).decode()
        raise RuntimeError(f"{what or 'ipoke_inn'} failed ({code}): {msg}")
