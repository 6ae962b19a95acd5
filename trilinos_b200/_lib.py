"""ctypes loader for libb200belos.so (the C-ABI in include/b200_belos.h).

The product path has no CPU fallback: if the shared library is missing the import fails
loudly, and every compute entry point returns B200_ERR_NO_DEVICE when no GPU is present.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libb200belos.so")
SOLVER_LIB_PATH = os.path.join(_HERE, "lib", "libb200belos_solver.so")


class B200Error(RuntimeError):
    """Non-zero status from the C-ABI (mirrors Belos::EpetraMultiVecFailure / EpetraOpFailure,
    belos/epetra/src/BelosEpetraAdapter.hpp:97-106)."""

    def __init__(self, code, msg):
        super().__init__(f"[b200 rc={code}] {msg}")
        self.code = code


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C trilinos_b200/csrc` (there is no CPU fallback for the B200 path)")
    return C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)


lib = _load()

c_i64 = C.c_int64
c_i32 = C.c_int32
c_dbl = C.c_double
vp = C.c_void_p
P = C.POINTER

_SIGS = {
    "b200_last_error": (C.c_char_p, []),
    "b200_last_status": (C.c_int, []),
    "b200_version": (C.c_char_p, []),
    "b200_ctx_create": (C.c_int, [C.c_int, P(vp)]),
    "b200_ctx_destroy": (C.c_int, [vp]),
    "b200_ctx_sync": (C.c_int, [vp]),
    "b200_comm_unique_id": (C.c_int, [vp]),
    "b200_ctx_comm_init": (C.c_int, [vp, vp, C.c_int, C.c_int]),
    "b200_ctx_rank": (C.c_int, [vp, P(C.c_int), P(C.c_int)]),
    "b200_ctx_set_seed": (C.c_int, [vp, C.c_uint]),
    "b200_ctx_counters": (C.c_int, [vp, P(c_i64), P(c_i64)]),
    "b200_ctx_set_fusion": (C.c_int, [vp, C.c_int]),
    "b200_ctx_fusion_stats": (C.c_int, [vp, P(c_i64)]),
    "b200_ctx_deferral_stats": (C.c_int, [vp, P(c_i64), P(c_i64)]),
    "b200_ctx_spec_gram_stats": (C.c_int, [vp, P(c_i64), P(c_i64)]),
    "b200_ctx_findbasis_stats": (C.c_int, [vp, P(c_i64)]),
    "b200_ctx_answered_bytes": (C.c_int, [vp, P(c_dbl), C.c_int]),
    "b200_ctx_timing_enable": (C.c_int, [vp, C.c_int]),
    "b200_ctx_timing_read": (C.c_int, [vp, P(c_dbl), P(c_i64), P(c_dbl), C.c_int]),
    "b200_ctx_stream": (vp, [vp]),
    "b200_ctx_flush_l2": (C.c_int, [vp]),
    "b200_mv_create": (C.c_int, [vp, c_i64, c_i64, c_i64, C.c_int, C.c_int, P(vp)]),
    "b200_mv_destroy": (C.c_int, [vp]),
    "b200_mv_clone": (C.c_int, [vp, C.c_int, P(vp)]),
    "b200_mv_clone_copy": (C.c_int, [vp, P(C.c_int), C.c_int, P(vp)]),
    "b200_mv_clone_view": (C.c_int, [vp, P(C.c_int), C.c_int, P(vp)]),
    "b200_mv_global_length": (c_i64, [vp]),
    "b200_mv_local_length": (c_i64, [vp]),
    "b200_mv_num_vecs": (C.c_int, [vp]),
    "b200_mv_ctx": (vp, [vp]),
    "b200_mv_times_mat_add_mv": (C.c_int, [vp, c_dbl, vp, P(c_dbl), C.c_int, C.c_int, C.c_int, c_dbl]),
    "b200_mv_add_mv": (C.c_int, [vp, c_dbl, vp, c_dbl, vp]),
    "b200_mv_scale": (C.c_int, [vp, c_dbl]),
    "b200_mv_scale_cols": (C.c_int, [vp, P(c_dbl), C.c_int]),
    "b200_mv_trans_mv": (C.c_int, [vp, c_dbl, vp, P(c_dbl), C.c_int, C.c_int, C.c_int]),
    "b200_mv_dot": (C.c_int, [vp, vp, P(c_dbl), C.c_int]),
    "b200_mv_norm": (C.c_int, [vp, C.c_int, P(c_dbl), C.c_int]),
    "b200_mv_set_block": (C.c_int, [vp, vp, P(C.c_int), C.c_int]),
    "b200_mv_random": (C.c_int, [vp]),
    "b200_mv_init": (C.c_int, [vp, c_dbl]),
    "b200_mv_factor_explicit": (C.c_int, [vp, vp, P(c_dbl), C.c_int, C.c_int]),
    "b200_mv_reveal_rank": (C.c_int, [vp, P(c_dbl), C.c_int, c_dbl, P(C.c_int)]),
    "b200_mv_copy_to_host": (C.c_int, [vp, P(c_dbl), c_i64]),
    "b200_mv_copy_from_host": (C.c_int, [vp, P(c_dbl), c_i64]),
    "b200_mv_device_ptr": (C.c_int, [vp, C.c_int, P(vp), P(c_i64)]),
    "b200_dot_noalloc": (C.c_int, [vp, c_i64, vp, c_i64, c_i64, C.c_int, vp, vp]),
    "b200_op_create_csr": (C.c_int, [vp, c_i64, c_i64, c_i64, P(c_i32), P(c_i64), P(c_dbl), P(vp)]),
    "b200_op_destroy": (C.c_int, [vp]),
    "b200_op_apply": (C.c_int, [vp, vp, vp, C.c_int]),
    "b200_op_info": (C.c_int, [vp, P(c_i64), P(c_i64), P(c_i64), P(c_i64), P(c_i64)]),
    "b200_read_hb": (C.c_int, [C.c_char_p, c_i64, c_i64, P(c_i64), P(c_i64), P(c_i32), P(c_i64), P(c_dbl), P(c_dbl),
                               P(c_dbl), P(C.c_int)]),
    "b200_read_matrix_market": (C.c_int, [C.c_char_p, c_i64, c_i64, P(c_i64), P(c_i64), P(c_i32), P(c_i64), P(c_dbl)]),
    "b200_mv_multiply_elementwise": (C.c_int, [vp, c_dbl, vp, vp, c_dbl]),
    "b200_op_extract_diagonal": (C.c_int, [vp, P(vp)]),
    "b200_cheby_default_params": (None, [vp]),
    "b200_cheby_create": (C.c_int, [vp, vp, P(vp)]),
    "b200_cheby_apply": (C.c_int, [vp, vp, vp]),
    "b200_cheby_lambda": (C.c_int, [vp, P(c_dbl), P(c_dbl)]),
    "b200_cheby_destroy": (C.c_int, [vp]),
    "b200_linear_partition": (C.c_int, [c_i64, C.c_int, C.c_int, P(c_i64), P(c_i64)]),
    "b200_build_colmap": (c_i64, [c_i64, C.c_int, C.c_int, c_i64, P(c_i64), P(c_i32), P(c_i64), P(c_i32)]),
    "b200_halo_recv_plan": (C.c_int, [C.c_int, c_i64, P(c_i32), P(c_i64), P(c_i64)]),
    "b200_halo_send_index": (C.c_int, [c_i64, C.c_int, C.c_int, c_i64, P(c_i64), P(c_i32)]),
    "b200_spmm_window_plan": (C.c_int, [c_i64, c_i64, P(c_i32), P(c_i32), P(c_dbl), C.c_int, P(c_i64), P(c_i64),
                                        P(c_i32), P(c_i32), vp, vp]),
    "b200_galeri_generate": (C.c_int, [C.c_char_p, C.c_int, C.c_int, C.c_int, c_dbl, c_dbl, c_i64, c_i64,
                                       P(c_i64), P(c_i32), P(c_i64), P(c_dbl)]),
}

for _name, (_res, _args) in _SIGS.items():
    _f = getattr(lib, _name)  # AttributeError here == header/library mismatch: fail loudly
    _f.restype = _res
    _f.argtypes = _args

EXPORTED = tuple(_SIGS)


def check(rc):
    if rc != 0:
        raise B200Error(rc, lib.b200_last_error().decode(errors="replace"))
    return rc
