"""ctypes binding of ``libtxb.so`` (declared in ``include/txb.h``).

The product path has NO CPU fallback: if the shared library is missing or a
CUDA device is unavailable the calls raise.
"""

from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_LIB_PATH = Path(__file__).resolve().parent / "lib" / "libtxb.so"

TXB_OK = 0
TXB_I64, TXB_I32, TXB_U16, TXB_U8 = 0, 1, 2, 3
MODE_COMOMENTS, MODE_X_IS_U = 0, 1
SAMPLER_PREPARED, SAMPLER_NO_BUCKETS = 1, 2

_p = C.c_void_p
_i64 = C.c_int64
_int = C.c_int
_sz = C.c_size_t

#: name -> (restype, argtypes); mirrors include/txb.h one to one
SIGNATURES = {
    "txb_version": (_int, []),
    "txb_last_error_string": (C.c_char_p, []),
    "txb_max_order": (_int, []),
    "txb_pivot_f64": (_int, [_p, _p, _i64, _i64, _i64, _i64, _p, _p]),
    "txb_reduce_workspace_bytes": (_sz, [_i64, _i64, _int]),
    "txb_reduce_sums_f64": (_int, [_p, _p, _p, _i64, _i64, _i64, _i64, _int, _p, _p, _p, _p, _sz, _p]),
    "txb_finalize_f64": (_int, [_p, _p, _p, _i64, _i64, _i64, _int, _int, _p, _p]),
    "txb_reduce_vals_f64": (_int, [_p, _p, _p, _i64, _i64, _i64, _i64, _int, _int, _p, _p, _sz, _p]),
    "txb_indices_to_freq": (_int, [_p, _i64, _i64, _i64, _p, _int, _i64, _p, _p]),
    "txb_resample_workspace_bytes": (_sz, [_i64, _i64, _int, _i64]),
    "txb_resample_sums_f64": (
        _int,
        [_p, _p, _p, _i64, _i64, _i64, _i64, _int, _p, _p, _int, _i64, _i64, _p, _p, _p, _sz, _p],
    ),
    "txb_resample_vals_f64": (
        _int,
        [_p, _p, _p, _i64, _i64, _i64, _i64, _int, _int, _p, _int, _i64, _i64, _p, _p, _sz, _p],
    ),
    "txb_resample_indices_workspace_bytes": (_sz, [_i64, _int, _i64, _i64]),
    "txb_resample_indices_sums_f64": (
        _int,
        [_p, _p, _p, _i64, _i64, _int, _p, _p, _i64, _i64, _i64, _p, _p, _p, _p, _sz, _p],
    ),
    "txb_resample_indices_f64": (
        _int,
        [_p, _p, _p, _i64, _i64, _int, _int, _p, _i64, _i64, _i64, _p, _p, _p, _sz, _p],
    ),
    "txb_reduce_vals_static_f64": (_int, [_p, _p, _p, _i64, _i64, _i64, _i64, _int, _int, _p, _p]),
    "txb_reduce_planar_workspace_bytes": (_sz, [_int, _int]),
    "txb_reduce_sums_planar_f64": (_int, [_p, _int, _p, _i64, _int, _p, _p, _p, _p, _sz, _p]),
    "txb_sampler_workspace_bytes": (_sz, [_i64, _i64]),
    "txb_sampler_layout": (_int, [_i64, _i64, _p]),
    "txb_pcg64_advance": (_int, [_p, _i64]),
    "txb_sampler_prepare": (_int, [_p, _i64, _i64, _p, _p, _p, _sz, _p]),
    "txb_sampler_nblocks": (_int, [_p, _i64, _i64]),
    "txb_sampler_count": (_int, [_p, _i64, _i64, _int, _int, _p, _sz, _p]),
    "txb_sampler_scan": (_int, [_p, _i64, _i64, _p, _p, _p, _sz, _p]),
    "txb_sampler_indices": (_int, [_p, _i64, _i64, _i64, _i64, _p, _p, _p, _int, _p, _sz, _p]),
    "txb_sampler_freq_workspace_bytes": (_sz, [_i64, _i64, _i64, _i64]),
    "txb_sampler_freq": (_int, [_p, _i64, _i64, _i64, _i64, _i64, _p, _int, _i64, _p, _p, _int, _p, _sz, _p]),
    "txb_sampler_freq_cols": (
        _int, [_p, _i64, _i64, _i64, _i64, _i64, _i64, _i64, _p, _int, _i64, _p, _i64, _p, _int, _p, _sz, _p]),
    "txb_sampler_resample_workspace_bytes": (_sz, [_i64, _int, _i64, _i64, _i64]),
    "txb_sampler_resample_fast_workspace_bytes": (_sz, [_i64, _int, _int, _i64, _i64, _i64]),
    "txb_sampler_resample_f64": (
        _int,
        [_p, _p, _p, _p, _i64, _i64, _int, _int, _i64, _i64, _i64, _i64, _p, _p, _p, _int, _p, _sz, _p],
    ),
    "txb_reduce_data_f64": (_int, [_p, _i64, _i64, _int, _p, _p]),
    "txb_resample_data_f64": (_int, [_p, _p, _i64, _i64, _i64, _int, _p, _p]),
    "txb_convert_f64": (_int, [_p, _p, _i64, _int, _int, _p]),
    "txb_eval_exprs_f64": (_int, [_p, _i64, _int, _p, _p, _p, _p, _p, _p, _p]),
    "txb_poly_predict_f64": (_int, [_p, _int, _i64, _p, _int, _int, _p, _p]),
    "txb_perturb_workspace_bytes": (_sz, [_i64, _i64, _int]),
    "txb_perturb_predict_f64": (_int, [_p, _p, _i64, _i64, _i64, _p, _int, _p, _p, _sz, _p]),
    "txb_beta_predict_f64": (_int, [_p, _i64, _int, _int, _int, _int, _int, _p, _int, _p, _p]),
    "txb_lnpi_predict_f64": (_int, [_p, _i64, _int, _int, _int, _int, _int, _p, _int, _p, _p, _i64, _p, _p]),
    "txb_combine_f64": (_int, [_p, _int, _int, _p, _i64, _int, _p, _p]),
    "txb_mean_var_f64": (_int, [_p, _i64, _i64, _i64, _int, _p, _p, _p]),
    "txb_set_resample_kernel": (_int, [_int]),
    "txb_set_scatter_threads": (_int, [_int]),
    "txb_profile_begin": (_int, [C.c_char_p]),
    "txb_profile_end": (_int, [C.POINTER(C.c_double), C.POINTER(_int)]),
    "txb_fp64_peak_launch": (_int, [_int, _int, _p, C.POINTER(C.c_double), _p]),
    "txb_fp64_tile_peak_launch": (_int, [_int, _int, _int, _p, C.POINTER(C.c_double), _p]),
    "txb_smem_atomic_peak_launch": (_int, [_int, _int, _int, _p, C.POINTER(C.c_double), _p]),
}



class Pcg64State(C.Structure):
    """``txb_pcg64``: numpy ``Generator(PCG64).bit_generator.state`` as plain integers."""

    _fields_ = [
        ("state_hi", C.c_uint64), ("state_lo", C.c_uint64), ("inc_hi", C.c_uint64), ("inc_lo", C.c_uint64),
        ("has_uint32", C.c_uint32), ("uinteger", C.c_uint32),
    ]


_lib = None


class TxbError(RuntimeError):
    """Non-zero status from libtxb."""


def lib_path() -> Path:
    return Path(os.environ.get("TXB_LIBRARY", str(_LIB_PATH)))


def load():
    """Load libtxb.so (once) and attach the signatures."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not path.exists():
        msg = (
            f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C thermoextrap_b200/csrc`.  There is no CPU fallback."
        )
        raise ImportError(msg)
    lib = C.CDLL(str(path))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(status: int, what: str) -> None:
    if status != TXB_OK:
        msg = load().txb_last_error_string().decode(errors="replace")
        exc = ValueError if status in (1, 2) else TxbError
        raise exc(f"{what} failed (status {status}): {msg}")
