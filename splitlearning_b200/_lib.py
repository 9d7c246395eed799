"""ctypes binding of libsplitnn.so (include/splitnn.h).  Fails loudly: there is no CPU path.

The library is built in-tree by `splitlearning_b200/csrc/build.sh` (or
`__graft_entry__.build()`); loading never triggers a build.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsplitnn.so")

SN_OPT_NONE, SN_OPT_SGD_MOMENTUM, SN_OPT_ADAM = 0, 1, 2

_vp, _i, _i64, _u64, _f, _d, _sz = C.c_void_p, C.c_int, C.c_int64, C.c_uint64, C.c_float, C.c_double, C.c_size_t
_pp = C.POINTER(C.c_void_p)

# name -> (restype, argtypes); mirrors include/splitnn.h one to one
SIGNATURES = {
    "sn_version": (_i, []),
    "sn_last_error": (C.c_char_p, []),
    "sn_launch_count": (_u64, []),
    "sn_set_matmul_precision": (_i, [_i]),
    "sn_get_matmul_precision": (_i, []),
    "sn_set_skinny_tensor_forward": (_i, [_i]),
    "sn_get_skinny_tensor_forward": (_i, []),
    "sn_set_persistent_tail": (_i, [_i]),
    "sn_get_persistent_tail": (_i, []),
    "sn_set_bwd_tiled_max": (_i, [_i]),
    "sn_get_bwd_tiled_max": (_i, []),
    "sn_conv_pool_forward": (_i, [_vp, _vp, _vp, _i, _vp, _vp]),
    "sn_conv_weights_copy": (_i, [_vp, _vp, _vp, _vp, _vp]),
    "sn_conv_pool_backward_scratch_bytes": (_sz, [_i]),
    "sn_conv_pool_backward": (_i, [_vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp]),
    "sn_linear_forward_scratch_bytes": (_sz, [_i, _i, _i]),
    "sn_linear_forward": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _vp, _f, _vp, _vp, _sz, _vp]),
    "sn_linear_forward_noalloc_max_scratch_bytes": (_sz, []),
    "sn_linear_forward_noalloc": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _vp, _f, _vp, _vp]),
    "sn_cross_entropy": (_i, [_vp, _vp, _i, _i, _vp, _vp, _vp]),
    "sn_eval_counts": (_i, [_vp, _vp, _i, _i, _i64, _vp, _vp]),
    "sn_bob_create": (_i, [_pp, _i, C.POINTER(_i), _i, _i, _d, _i, _d, _d, _d, _d, _d, _d, _i, _i, _pp, _pp]),
    "sn_bob_destroy": (_i, [_vp]),
    "sn_bob_reset_opt_slot": (_i, [_vp, _i, _vp]),
    "sn_bob_opt_state": (_i, [_vp, _i, _i, _i, _pp, C.POINTER(_sz)]),
    "sn_bob_get_opt_step": (_i, [_vp, _i, C.POINTER(_i64)]),
    "sn_bob_set_opt_step": (_i, [_vp, _i, _i64]),
    "sn_bob_set_lr": (_i, [_vp, _d]),
    "sn_bob_set_ce_heads": (_i, [_vp, _i]),
    "sn_bob_set_stream_rows": (_i, [_vp, _i, _sz, C.POINTER(_i)]),
    "sn_bob_profile": (_i, [_vp, _i]),
    "sn_bob_profile_read": (_i, [_vp, _i, _i, C.POINTER(_d), C.POINTER(_i)]),
    "sn_bob_train_step": (_i, [_vp, _i, _vp, _vp, _i, _pp, _u64, _u64, _vp, _vp, _vp, _vp]),
    "sn_bob_forward": (_i, [_vp, _vp, _i, _i, _pp, _u64, _u64, _vp, _vp]),
    "sn_bob_backward_update": (_i, [_vp, _i, _vp, _vp, _i, _vp, _vp]),
    "sn_bob_server_epoch": (_i, [_vp, _i, _vp, _vp, C.POINTER(_i64), _i, _pp, _u64, _u64, _vp, _vp]),
    "sn_alice_create": (_i, [_pp, _i, _i, _d, _d, _d, _d, _d, _d, _i, _vp, _vp]),
    "sn_alice_destroy": (_i, [_vp]),
    "sn_alice_reset_opt": (_i, [_vp, _vp]),
    "sn_alice_opt_state": (_i, [_vp, _i, _i, _pp, C.POINTER(_sz)]),
    "sn_alice_get_opt_step": (_i, [_vp, C.POINTER(_i64)]),
    "sn_alice_set_opt_step": (_i, [_vp, _i64]),
    "sn_alice_forward": (_i, [_vp, _vp, _i, _vp, _vp]),
    "sn_alice_forward_ld": (_i, [_vp, _vp, _i, _vp, _sz, _vp]),
    "sn_alice_forward_host": (_i, [_vp, _vp, _vp, _i, _vp, _vp, _vp, _vp]),
    "sn_vanilla_step_host": (_i, [_vp, _vp, _i, _vp, _vp, _i, _u64, _u64, _vp, _vp, _vp, _vp, _vp, _vp]),
    "sn_alice_backward_update": (_i, [_vp, _vp, _vp, _i, _vp]),
    "sn_alice_local_step": (_i, [_vp, _vp, _vp, _i, _vp, _vp]),
    "sn_alice_local_epoch": (_i, [_vp, _vp, _vp, _i64, _i, _vp, _vp]),
    "sn_vanilla_step": (_i, [_vp, _vp, _i, _vp, _vp, _i, _u64, _u64, _vp, _vp, _vp, _vp]),
    "sn_p2p_enable_all": (_i, [_i]),
    "sn_p2p_copy": (_i, [_vp, _i, _vp, _i, _sz, _vp]),
    "sn_ipc_alloc": (_i, [_sz, _pp, _vp]),
    "sn_ipc_open": (_i, [_vp, _pp]),
    "sn_ipc_close": (_i, [_vp]),
    "sn_ipc_free": (_i, [_vp]),
    "sn_stream_signal": (_i, [_vp, C.c_uint32, _vp]),
    "sn_stream_wait": (_i, [_vp, C.c_uint32, _vp]),
    "sn_flag_signal": (_i, [_vp, C.c_uint32, _vp, _vp, _sz, _vp]),
    "sn_flag_wait": (_i, [_vp, C.c_uint32, _vp]),
    "sn_flag_wait_many": (_i, [_vp, _i, _i, C.c_uint32, _vp]),
    "sn_shard_create": (_i, [_pp, _i, _i, _i, _i, _i, _i, _d, _i, _d, _d, _d, _d, _d, _d, _vp, _vp]),
    "sn_shard_destroy": (_i, [_vp]),
    "sn_shard_reset_opt": (_i, [_vp, _vp]),
    "sn_shard_opt_state": (_i, [_vp, _i, _i, _pp, C.POINTER(_sz)]),
    "sn_shard_get_opt_step": (_i, [_vp, C.POINTER(_i64)]),
    "sn_shard_set_opt_step": (_i, [_vp, _i64]),
    "sn_shard_forward": (_i, [_vp, _vp, _i, _i, _vp, _u64, _u64, _vp, _i, _vp]),
    "sn_shard_backward_update": (_i, [_vp, _vp, _vp, _i, _i, _vp, _i, _vp]),
    "sn_tp_server_epoch": (_i, [_vp, _vp, _i, _i, _vp, _vp, C.POINTER(_i64), _i, _pp, _u64, _u64, _vp, _vp, _vp, _vp,
                                _vp, _vp, C.c_uint32, _vp, _vp]),
    "sn_flag_signal_many": (_i, [_vp, _i, C.c_uint32, _vp]),
    "sn_flag_status": (_i, []),
    "sn_flag_status_clear": (_i, []),
    "sn_lda_partition": (_i, [_vp, _i64, _i, _i, _d, _vp, C.POINTER(_i), C.POINTER(_i), C.POINTER(_d), _vp, _vp]),
    "sn_py_random_shuffle": (_i, [_vp, _i64, _vp, C.POINTER(_i)]),
}

_lib = None


class SplitNNError(RuntimeError):
    pass


def load():
    """Load libsplitnn.so and declare every prototype.  Raises if the library is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise SplitNNError(
            f"{LIB_PATH} is missing: build it with splitlearning_b200/csrc/build.sh "
            "(or __graft_entry__.build()). There is no CPU fallback for the split-NN hot path.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError here = header/library mismatch
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise SplitNNError(load().sn_last_error().decode("utf-8", "replace"))


def ptr(t):
    """Device (or host) address of a torch tensor, or None."""
    return None if t is None else C.c_void_p(t.data_ptr())


def ptr_array(tensors):
    """Host array of pointers (void*[n]) for a list of tensors; None entries become NULL."""
    arr = (C.c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = None if t is None else t.data_ptr()
    return arr
