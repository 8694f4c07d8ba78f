"""ctypes binding of libnjucloud_b200.so -- the only way the Python host side reaches compute.

There is deliberately no fallback: if the library is missing ``lib()`` raises, and ``Context``
raises when no B200 is visible (nc_init fails)."""
import ctypes as C
import os

from .build import library_path

NC_OK = 0
NC_EINVAL, NC_ECUDA, NC_ENOMEM, NC_EUNSUPPORTED, NC_ECONFIG, NC_ECV = -1, -2, -3, -4, -5, -6

_lib = None

c_u8p = C.c_void_p
c_f32p = C.c_void_p
c_i32p = C.c_void_p

_SIGS = {
    "nc_version": (C.c_int, []),
    "nc_init": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "nc_destroy": (C.c_int, [C.c_void_p]),
    "nc_last_error": (C.c_char_p, [C.c_void_p]),
    "nc_launch_count": (C.c_int64, [C.c_void_p]),
    "nc_jpeg_info": (C.c_int, [C.c_void_p, C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "nc_jpeg_entropy_decode": (C.c_int, [C.c_void_p, C.c_size_t, C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]),
    "nc_jpeg_decode": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int64, C.c_int,
                                 C.POINTER(C.c_int64), C.c_void_p]),
    "nc_pp_resize_gray_norm": (C.c_int, [C.c_void_p, c_u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int,
                                         c_u8p, c_f32p, C.c_void_p]),
    "nc_pp_resize_cubic": (C.c_int, [C.c_void_p, c_u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int, c_u8p, C.c_void_p]),
    "nc_pp_resize_gray_norm_host": (C.c_int, [C.c_void_p, c_u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int,
                                              c_u8p, c_f32p, C.POINTER(C.c_int), C.c_void_p]),
    "nc_pp_predict_chain": (C.c_int, [C.c_void_p, c_u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int, C.c_int,
                                      c_f32p, c_u8p, c_i32p, C.c_void_p]),
    "nc_pp_both_chains": (C.c_int, [C.c_void_p, c_u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int, C.c_int,
                                    c_u8p, c_f32p, c_f32p, c_u8p, c_i32p, c_i32p, C.c_void_p]),
    "nc_pp_op": (C.c_int, [C.c_void_p, C.c_int, c_u8p, c_u8p, C.c_int, C.c_double, C.c_double, C.c_int,
                           C.c_void_p, C.c_int64, C.c_void_p]),
    "nc_pp_cubic_table": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p]),
    "nc_pp_cubic_table_f32": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p]),
    "nc_pp_pil_table": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]),
    "nc_pp_gaussian_q8": (C.c_int, [C.c_int, C.c_void_p]),
    "nc_crc32c": (C.c_uint32, [C.c_void_p, C.c_size_t, C.c_uint32]),
    "nc_pp_normalise": (C.c_int, [C.c_void_p, c_u8p, C.c_int64, c_f32p, C.c_void_p]),
    "nc_gather_rows": (C.c_int, [C.c_void_p, c_f32p, C.c_int64, C.c_int, C.c_void_p, C.c_int, c_f32p, C.c_void_p]),
    "nc_gather_rows_status": (C.c_int, [C.c_void_p, C.c_void_p]),
    "nc_matmul": (C.c_int, [C.c_void_p, c_f32p, c_f32p, c_f32p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, c_f32p, C.c_void_p]),
    "nc_dense_forward_noalloc": (C.c_int, [C.c_void_p, c_f32p, c_f32p, c_f32p, C.c_int, C.c_int, C.c_int, c_f32p, C.c_void_p]),
    "nc_net_create": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64,
                                C.POINTER(C.c_void_p)]),
    "nc_net_destroy": (C.c_int, [C.c_void_p]),
    "nc_net_param_count": (C.c_int64, [C.c_void_p]),
    "nc_net_tensor_count": (C.c_int, [C.c_void_p]),
    "nc_net_tensor_info": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64),
                                     C.POINTER(C.c_int), C.POINTER(C.c_int64 * 4)]),
    "nc_net_params": (C.c_void_p, [C.c_void_p]),
    "nc_net_grads": (C.c_void_p, [C.c_void_p]),
    "nc_net_set_params": (C.c_int, [C.c_void_p, C.c_void_p]),
    "nc_net_get_params": (C.c_int, [C.c_void_p, C.c_void_p]),
    "nc_net_set_math": (C.c_int, [C.c_void_p, C.c_int]),
    "nc_net_forward": (C.c_int, [C.c_void_p, c_f32p, C.c_int, c_f32p, C.c_void_p]),
    "nc_net_predict": (C.c_int, [C.c_void_p, c_f32p, C.c_int, c_i32p, C.c_void_p]),
    "nc_net_backward": (C.c_int, [C.c_void_p, c_f32p, c_f32p, C.c_int, C.c_int, C.c_float, c_f32p, c_i32p, C.c_void_p]),
    "nc_net_apply": (C.c_int, [C.c_void_p, C.c_int, C.c_float, C.c_void_p]),
    "nc_net_train_step": (C.c_int, [C.c_void_p, c_f32p, c_f32p, C.c_int, C.c_int, C.c_int, C.c_float, c_f32p, c_i32p,
                                    C.c_void_p]),
    "nc_net_global_step": (C.c_int64, [C.c_void_p]),
    "nc_net_advance_step": (C.c_int, [C.c_void_p, C.c_int64]),
    "nc_net_profile": (C.c_int, [C.c_void_p, C.c_int]),
    "nc_net_profile_read": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_char_p, C.c_size_t, C.POINTER(C.c_int)]),
    "nc_net_attach_comm": (C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int, C.c_int]),
    "nc_net_arena_generation": (C.c_int64, [C.c_void_p]),
    "nc_net_reserve": (C.c_int, [C.c_void_p, C.c_int]),
    "nc_nccl_unique_id": (C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p]),
    "nc_net_attach_nccl": (C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int, C.c_int]),
    "nc_net_allreduce_grads": (C.c_int, [C.c_void_p, C.c_void_p]),
    "nc_net_detach_nccl": (C.c_int, [C.c_void_p]),
    "nc_net_p2p_handles": (C.c_int, [C.c_void_p, C.c_void_p]),
    "nc_net_p2p_attach": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "nc_net_p2p_status": (C.c_int, [C.c_void_p, C.POINTER(C.c_int)]),
    "nc_net_state_slots": (C.c_int, [C.c_void_p, C.c_int]),
    "nc_net_get_state": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "nc_net_set_state": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int64]),
}


def exported_symbols():
    """Every symbol include/njucloud_b200.h declares (kept in sync by tests/test_abi.py)."""
    return sorted(_SIGS)


def lib():
    global _lib
    if _lib is None:
        path = library_path()
        if not os.path.exists(path):
            raise RuntimeError(
                "libnjucloud_b200.so is not built (%s). Run `python -c 'import __graft_entry__ as g; g.build()'`; "
                "there is no CPU fallback." % path)
        L = C.CDLL(path)
        for name, (res, args) in _SIGS.items():
            fn = getattr(L, name)   # AttributeError here = the library does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


class NcError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libnjucloud_b200 error %d: %s" % (code, msg))
        self.code = code
        self.msg = msg


def check(ctx_handle, rc):
    if rc != NC_OK:
        msg = lib().nc_last_error(ctx_handle)
        msg = msg.decode("utf-8", "replace") if msg else ""
        if msg.startswith("IndexError"):
            raise IndexError(msg)
        if rc in (NC_EINVAL, NC_ECONFIG):
            raise ValueError(msg)
        raise NcError(rc, msg)
