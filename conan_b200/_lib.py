"""ctypes binding of the C-ABI library (include/conan_b200.h).

There is no CPU fallback: if the library is missing the import fails loudly, and every device
entry point needs a CUDA device.  Host-only entry points (xtc index / decode / encode) work
without one.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIBPATH = os.path.join(_HERE, "lib", "libconan_b200.so")

CONAN_OK = 0

c_void_p = ctypes.c_void_p
c_i32 = ctypes.c_int32
c_i64 = ctypes.c_int64
c_size = ctypes.c_size_t
c_f32 = ctypes.c_float


class ConanError(RuntimeError):
    """A C-ABI entry point returned a negative status."""

    def __init__(self, code, message):
        super().__init__("conan_b200 error %d: %s" % (code, message))
        self.code = code


class PairStatsOut(ctypes.Structure):
    _fields_ = [(n, c_void_p) for n in (
        "d_sum_k", "d_sum_k2", "d_n_int", "d_n_enc", "d_first_f", "d_last_f", "d_status", "d_decided")]


class FrameStatsOut(ctypes.Structure):
    _fields_ = [(n, c_void_p) for n in ("d_native", "d_nonnative", "d_ssd_prev", "d_ssd_first")]


# name -> (restype, argtypes); this table is also what tests check against the header.
SIGNATURES = {
    "conan_version": (ctypes.c_int, []),
    "conan_last_error": (ctypes.c_char_p, []),
    "conan_xtc_index": (ctypes.c_int, [c_void_p, c_size, c_i64, c_void_p, c_void_p, c_void_p,
                                       c_void_p, ctypes.POINTER(c_i64), ctypes.POINTER(c_i32)]),
    "conan_xtc_decode": (ctypes.c_int, [c_void_p, c_size, c_void_p, c_void_p, c_i64, c_i32,
                                        c_void_p, c_i32, c_void_p, c_i32]),
    "conan_xtc_encode": (ctypes.c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_i64, c_i32,
                                        c_f32, c_void_p, c_size, ctypes.POINTER(c_size)]),
    "conan_xtc_decode_device": (ctypes.c_int, [c_void_p, c_i64, c_void_p, c_i64, c_i32, c_void_p, c_i32, c_void_p,
                                               c_void_p, c_void_p]),
    "conan_xtc_decode_workspace_bytes": (ctypes.c_size_t, [c_i64, c_i32]),
    "conan_xtc_decode_device_ws": (ctypes.c_int, [c_void_p, c_i64, c_void_p, c_i64, c_i32, c_void_p, c_i32, c_void_p,
                                                  c_void_p, c_void_p, ctypes.c_size_t, c_void_p]),
    "conan_xtc_walk_device": (ctypes.c_int, [c_void_p, c_i64, c_void_p, c_i64, c_i32, c_void_p, c_i32, c_void_p,
                                             c_void_p, c_void_p, ctypes.c_size_t, c_void_p]),
    "conan_xtc_groups_device": (ctypes.c_int, [c_void_p, c_i64, c_i64, c_i32, c_void_p, c_i32, c_void_p,
                                               c_void_p, ctypes.c_size_t, c_void_p]),
    "conan_write_matrices": (ctypes.c_int, [c_void_p, c_i64, c_i64, c_i64, c_i32, c_i32, c_i32, c_i32, c_i32,
                                            c_void_p, c_i32, c_i32, c_void_p, c_void_p, ctypes.c_char_p, c_i32]),
    "conan_write_pair_table": (ctypes.c_int, [ctypes.c_char_p, ctypes.c_char_p, c_i32, c_i32, c_i32, c_i32, c_void_p, c_i32,
                                              ctypes.c_char_p, c_i32]),
    "conan_write_xpm": (ctypes.c_int, [c_void_p, c_i64, c_i64, c_i32, c_i32, c_i32, c_void_p, ctypes.c_char_p,
                                       ctypes.c_char_p, ctypes.c_char_p, c_i32, c_i32, ctypes.c_char_p, c_i32, c_i32]),
    "conan_mdmat_workspace_bytes": (c_size, [c_i64, c_i32, c_i32, c_i32]),
    "conan_mdmat_frames": (ctypes.c_int, [c_void_p, c_void_p, c_i64, c_i32, c_void_p, c_i32, c_i32,
                                          c_i32, c_i32, c_i32, c_i32, c_f32, c_i32, c_i32, c_void_p,
                                          c_i64, c_void_p, c_void_p, c_size, c_void_p]),
    "conan_mdmat_frames_ex": (ctypes.c_int, [c_void_p, c_void_p, c_i64, c_i32, c_void_p, c_i32, c_i32,
                                             c_i32, c_i32, c_i32, c_i32, c_f32, c_i32, c_i32, c_void_p,
                                             c_i64, c_void_p, c_void_p, c_size, c_void_p, c_i32]),
    "conan_mdmat_frames_static": (ctypes.c_int, [c_void_p, c_void_p, c_i64, c_i32, c_void_p, c_i32, c_i32,
                                                 c_i32, c_i32, c_i32, c_i32, c_f32, c_i32, c_i32, c_void_p,
                                                 c_i64, c_void_p, c_void_p]),
    "conan_shard_summary": (ctypes.c_int, [c_void_p, c_i64, c_i64, c_i64, c_i32, c_i32, c_i32, c_i32,
                                           c_void_p, c_void_p, c_void_p]),
    "conan_pair_stats_workspace_bytes": (c_size, [c_i64, c_i64]),
    "conan_pair_stats": (ctypes.c_int, [c_void_p, c_i64, c_i64, c_i64, c_i64, c_i32, c_i32, c_i32,
                                        c_i32, c_void_p, c_void_p, c_void_p,
                                        ctypes.POINTER(PairStatsOut), ctypes.POINTER(FrameStatsOut),
                                        c_void_p, c_size, c_void_p]),
    "conan_pair_stats_ex": (ctypes.c_int, [c_void_p, c_i64, c_i64, c_i64, c_i64, c_i32, c_i32, c_i32,
                                        c_i32, c_void_p, c_void_p, c_void_p,
                                        ctypes.POINTER(PairStatsOut), ctypes.POINTER(FrameStatsOut),
                                        c_void_p, c_size, c_i32, c_void_p]),
    "conan_dimer_orient": (ctypes.c_int, [c_void_p, c_i64, c_i64, c_i32, c_i32, c_void_p, c_void_p]),
    "conan_dimer_ssd": (ctypes.c_int, [c_void_p, c_i64, c_i64, c_i32, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "conan_pearson_sums": (ctypes.c_int, [c_void_p, c_i64, c_i64, c_i64, c_void_p, c_i32, c_void_p, c_i32,
                                          c_void_p, c_void_p, c_void_p, c_void_p]),
    "conan_level_range": (ctypes.c_int, [c_void_p, c_i64, c_i64, c_i64, c_void_p, c_void_p, c_void_p, c_void_p]),
    "conan_residue_weights": (ctypes.c_int, [c_void_p, c_i64, c_i64, c_i32, c_i32, c_i32, c_i32, c_void_p, c_void_p]),
    "conan_build_operand": (ctypes.c_int, [c_void_p, c_i64, c_i64, c_void_p, c_i32, c_void_p, c_i32, c_void_p,
                                           c_f32, c_i32, c_void_p, c_void_p, c_i64, c_void_p, c_void_p]),
    "conan_gemm_3xtf32": (ctypes.c_int, [c_void_p, c_void_p, c_i32, c_void_p, c_void_p, c_i32, c_i64, c_i64,
                                         c_i32, c_f32, c_void_p, c_void_p, c_i32, c_void_p, c_i64, c_void_p]),
    "conan_sym_block_matvec": (ctypes.c_int, [c_void_p, c_i64, c_i32, c_void_p, c_i64, c_i32, c_void_p, c_i64, c_void_p]),
    "conan_block_gram": (ctypes.c_int, [c_void_p, c_i64, c_i32, c_void_p, c_i64, c_i32, c_i32, c_void_p, c_void_p]),
    "conan_block_combine": (ctypes.c_int, [c_void_p, c_i64, c_i32, c_void_p, c_i32, c_i32, c_void_p, c_i64, c_void_p]),
    "conan_measure_fp32_peak": (ctypes.c_int, [ctypes.POINTER(ctypes.c_double), c_void_p]),
}

_lib = None


def load():
    """Load the shared library once; raise if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIBPATH):
        raise ImportError(
            "conan_b200: %s is missing -- build it with `python -m conan_b200.build` "
            "(there is no CPU fallback)" % LIBPATH)
    lib = ctypes.CDLL(LIBPATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(status):
    """Turn a negative status into ConanError carrying conan_last_error()."""
    if status != CONAN_OK:
        msg = load().conan_last_error()
        raise ConanError(status, msg.decode("utf-8", "replace") if msg else "")
    return status
