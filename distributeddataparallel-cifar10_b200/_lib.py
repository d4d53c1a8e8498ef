"""ctypes binding of libnrd.so (include/nrd.h).  No CPU fallback: a missing library is an error."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libnrd.so")

NRD_F32, NRD_BF16 = 0, 1
LOSS_RING = 8          # NRD_LOSS_RING (include/nrd.h)
CONV_CUDA_CORE, CONV_TCGEN05, CONV_TCGEN05_UNFUSED = 0, 1, 2
OPT_FUSE_BN_BWD, OPT_PERSIST_FWD, OPT_TRACE, OPT_PERSIST_BWD, OPT_TC_STEM, OPT_RESERVE_SMS, OPT_PDL, OPT_HEAD_TAIL, OPT_FUSE_POOL, OPT_FUSE_STEM, OPT_ONE_COPY, OPT_TIMELINE, OPT_FUSE_STEM_BWD, OPT_STREAM = range(14)


class NrdError(RuntimeError):
    pass


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: the CUDA extension is the product and there is no CPU fallback. "
            "Build it with `python -c 'import __graft_entry__ as g; g.build()'` (needs nvcc).")
    return C.CDLL(LIB_PATH)


lib = _load()
vp, i32, i64, u64, u32, f32, sz = C.c_void_p, C.c_int, C.c_int64, C.c_uint64, C.c_uint32, C.c_float, C.c_size_t

_PROTOS = {
    "nrd_last_error": (C.c_char_p, []),
    "nrd_version": (i32, []),
    "nrd_param_offsets": (i32, [i32, C.POINTER(i64)]),
    "nrd_workspace_bytes": (sz, [i32, i32, i32, i32]),
    "nrd_plan_create": (i32, [C.POINTER(vp), i32, i32, i32, i32, i32, vp, sz, vp, vp, vp, vp]),
    "nrd_plan_destroy": (i32, [vp]),
    "nrd_plan_set_conv_impl": (i32, [vp, i32]),
    "nrd_plan_conv_impl": (i32, [vp]),
    "nrd_plan_set_option": (i32, [vp, i32, i32]),
    "nrd_plan_status": (i32, [vp, C.POINTER(u32), i32, vp]),
    "nrd_plan_set_bn_sync": (i32, [vp, vp, f32, f32]),
    "nrd_exchange_bytes": (sz, [i32, i32]),
    "nrd_exchange_alloc": (i32, [i32, sz, C.POINTER(vp), vp]),
    "nrd_exchange_open": (i32, [i32, vp, C.POINTER(vp)]),
    "nrd_exchange_release": (i32, [vp, i32]),
    "nrd_plan_set_exchange": (i32, [vp, i32, i32, C.POINTER(vp)]),
    "nrd_plan_tensor": (i32, [vp, i32, i32, C.POINTER(vp), C.POINTER(sz)]),
    "nrd_pack_weights": (i32, [vp, vp]),
    "nrd_forward": (i32, [vp, vp, vp, i32, vp]),
    "nrd_forward_loss": (i32, [vp, vp, vp, f32, vp, vp, vp, vp, vp]),
    "nrd_cross_entropy": (i32, [vp, vp, i32, f32, vp, vp, vp, vp]),
    "nrd_cross_entropy_static": (i32, [vp, vp, i32, f32, vp, vp, vp]),
    "nrd_backward": (i32, [vp, vp, vp]),
    "nrd_backward_head": (i32, [vp, vp, vp]),
    "nrd_backward_body": (i32, [vp, vp, vp]),
    "nrd_sgd_update": (i32, [vp, vp, sz, f32, f32, vp]),
    "nrd_sgd_update_pack": (i32, [vp, f32, f32, vp]),
    "nrd_plan_set_fused_update": (i32, [vp, f32, f32]),
    "nrd_plan_set_loss_mirror": (i32, [vp, vp, vp]),
    "nrd_train_step": (i32, [vp, vp, vp, f32, vp, vp, vp]),
    "nrd_gen_batch": (i32, [vp, vp, i32, u64, u32, u64, vp]),
    "nrd_sgd_momentum_update": (i32, [vp, vp, vp, sz, f32, f32, i32, vp]),
    "nrd_bce_with_logits": (i32, [vp, vp, sz, f32, vp, vp, vp]),
    "nrd_gather_normalize": (i32, [vp, vp, C.c_longlong, vp, i32, C.POINTER(f32), C.POINTER(f32), vp, vp, vp]),
    "nrd_pack_conv_weight": (i32, [vp, i32, vp, vp, vp]),
    "nrd_conv3x3_fwd": (i32, [vp, vp, vp, i32, i32, i32, vp, C.POINTER(i32), vp, vp]),
    "nrd_wgrad_partial_bytes": (sz, [i32, i32]),
    "nrd_conv3x3_wgrad": (i32, [vp, vp, vp, i32, i32, i32, vp, vp]),
    "nrd_bn_finalize": (i32, [vp, i32, i32, i32, vp, vp, vp, vp, vp, vp, vp]),
    "nrd_bn_relu_add": (i32, [vp, vp, vp, vp, i32, i32, i32, vp]),
    "nrd_bn_bwd_scratch_bytes": (sz, [i32, i32]),
    "nrd_bn_relu_bwd": (i32, [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, vp, vp]),
    "nrd_stem_fwd": (i32, [vp, vp, vp, vp, vp, i32, i32, i32, vp]),
    "nrd_stem_bwd_partial_bytes": (sz, [i32, i32]),
    "nrd_stem_bwd": (i32, [vp, vp, vp, vp, vp, vp, i32, i32, i32, vp, vp]),
    "nrd_maxpool_fwd": (i32, [vp, vp, vp, i32, i32, i32, vp]),
    "nrd_maxpool_bwd": (i32, [vp, vp, vp, i32, i32, i32, vp]),
    "nrd_launch_count": (C.c_ulonglong, []),
    "nrd_plan_run_op": (i32, [vp, i32, i32, vp]),
    "nrd_plan_timeline": (i32, [vp, i32, C.POINTER(f32), C.POINTER(C.c_char_p)]),
}
for _n, (_res, _args) in _PROTOS.items():
    _f = getattr(lib, _n)          # AttributeError here = header and library out of sync
    _f.restype, _f.argtypes = _res, _args

EXPORTS = tuple(_PROTOS)


def last_error() -> str:
    return lib.nrd_last_error().decode()


def check(rc: int, what: str = ""):
    if rc != 0:
        raise NrdError(f"{what or 'libnrd'} failed ({rc}): {last_error()}")


def param_offsets(n_chans: int):
    arr = (i64 * 10)()
    check(lib.nrd_param_offsets(n_chans, arr), "nrd_param_offsets")
    return list(arr)
