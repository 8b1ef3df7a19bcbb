"""
ctypes binding of libmagicsoup_b200.so (C ABI in include/magicsoup_b200.h).

There is no CPU fallback: if the library is missing it is built with nvcc, and if
that fails or no CUDA device is present the hot-path calls raise.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import torch

_HERE = Path(__file__).resolve().parent
LIB_PATH = _HERE / "libmagicsoup_b200.so"
N_FLAGS = 32
FLAG_BAD_POSITION = 24

_vp, _i64, _i32, _sz = C.c_void_p, C.c_int64, C.c_int32, C.c_size_t


class KineticsParams(C.Structure):
    """struct ms_kinetics_params"""

    _fields_ = [(k, _vp) for k in ("N", "Nf", "Nb", "A", "Kmr", "Ke", "Kmf", "Kmb", "Vmax")] + [
        ("c", _i64), ("p", _i64), ("s", _i64)]


class KineticsCache(C.Structure):
    """struct ms_kinetics_cache"""

    _fields_ = [("records", _vp), ("rec_words", _i64), ("row_map", _vp)]


class TokenTables(C.Structure):
    """struct ms_token_tables"""

    _fields_ = [(k, _vp) for k in ("vmax_weights", "km_weights", "signs", "hill_numbers", "reaction_M", "transport_M",
                                   "effector_M")]


class KineticsParamsOut(C.Structure):
    """struct ms_kinetics_params_out"""

    _fields_ = [(k, _vp) for k in ("N", "Nf", "Nb", "A", "Kmr", "Ke", "Kmf", "Kmb", "Vmax")] + [("p", _i64), ("s", _i64)]


_PP = C.POINTER(KineticsParams)
_PC = C.POINTER(KineticsCache)
_SIGNATURES = {
    "ms_version": (C.c_int, []),
    "ms_error_string": (C.c_char_p, [C.c_int]),
    "ms_last_cuda_error": (C.c_char_p, []),
    "ms_device_count": (C.c_int, []),
    "ms_integrate_signals": (C.c_int, [_PP, _vp, _vp, C.POINTER(_i32), _vp, _vp, _sz, _vp]),
    "ms_integrate_signals_part": (C.c_int, [_PP, _vp, _vp, _vp, _i32, _vp, _vp, _vp, _sz, _vp]),
    "ms_kinetics_scratch_bytes": (_sz, [_i64, _i64, _i64]),
    "ms_kinetics_reload_switches": (None, []),
    "ms_kinetics_cache_record_words": (_i64, [_i64, _i64]),
    "ms_kinetics_cache_invalidate": (C.c_int, [_vp, _i64, _vp, _i64, _vp]),
    "ms_integrate_signals_cached": (C.c_int, [_PP, _PC, _vp, _vp, C.POINTER(_i32), _vp, _vp, _sz, _vp]),
    "ms_enzymatic_activity_cached": (C.c_int, [_PP, _PC, _vp, _vp, _vp, _i64, _i64, _i64, _vp, C.POINTER(_i32), _vp,
                                               _vp, _sz, _vp]),
    "ms_set_cell_params": (C.c_int, [C.POINTER(TokenTables), _vp, C.c_float, _vp, _vp, _vp, _vp, _i64,
                                     C.POINTER(KineticsParamsOut), _vp]),
    "ms_multiply_signals": (C.c_int, [_vp, _vp, _i64, _i64, _i64, _vp, _vp, _vp]),
    "ms_get_quotient": (C.c_int, [_vp, _vp, _vp, _i64, _i64, _i64, _vp, _vp]),
    "ms_get_velocities": (C.c_int, [_PP, _vp, _vp, _vp, _vp]),
    "ms_get_negative_adjusted_nv": (C.c_int, [_vp, _vp, _i64, _i64, _i64, _vp, _vp, _vp]),
    "ms_add_protein_rates": (C.c_int, [_vp, _vp, _i64, _i64, _i64, _vp, _vp, _vp]),
    "ms_get_equilibrium_adjusted_x": (C.c_int, [_vp] * 7 + [_i64] * 3 + [_vp, _vp, _i32, _vp, _vp]),
    "ms_enzymatic_activity": (C.c_int, [_PP, _vp, _vp, _vp, _i64, _i64, _i64, _vp, C.POINTER(_i32), _vp, _vp, _sz, _vp]),
    "ms_scatter_signals": (C.c_int, [_vp, _vp, _i64, _i64, _i64, _i64, _vp, _vp, _vp]),
    "ms_diffuse_molecules": (C.c_int, [_vp, _vp, _vp, _i64, _i64, _i64, _vp, _sz, _vp]),
    "ms_diffuse_scratch_bytes": (_sz, [_i64, _i64, _i64]),
    "ms_diffuse_stencil": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _vp, _vp, _vp]),
    "ms_diffuse_correct": (C.c_int, [_vp, _vp, _vp, _i64, _i64, _i64, _i64, _vp]),
    "ms_diffuse_molecules_static": (C.c_int, [_vp, _vp, _vp, _i64, _i64, _i64, _vp]),
    "ms_diffuse_static_max_floats": (_i64, []),
    "ms_diffuse_static_last_error": (C.c_char_p, []),
    "ms_permeate": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _i64, _vp]),
    "ms_degrade_molecules": (C.c_int, [_vp, _vp, _vp, _i64, _i64, _i64, _i64, _vp]),
    "ms_build_pixel_index": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _vp]),
    "ms_scatter_signals_by_pixel": (C.c_int, [_vp, _vp, _i64, _i64, _i64, _vp, _vp, _vp]),
    "ms_permeate_by_pixel": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _vp]),
    "ms_degrade_diffuse_stencil": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _vp, _vp, _vp]),
    "ms_degrade_permeate": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _i64, _vp]),
    "ms_early_exit_verdict": (C.c_int, [_vp, C.POINTER(C.c_int32), C.c_int32, _vp, _vp]),
    "ms_step_gate": (None, [_vp]),
    "ms_peer_buffer_bytes": (C.c_size_t, [C.c_int32, _i64, _i64]),
    "ms_peer_halo_offset": (C.c_size_t, [C.c_int32, _i64, _i64, C.c_int32, C.c_int32]),
    "ms_peer_allreduce": (C.c_int, [C.POINTER(C.c_uint64), C.c_int32, C.c_int32, C.c_uint64, _vp, C.c_int32, C.c_int32, _vp, _vp]),
    "ms_peer_halo": (C.c_int, [C.POINTER(C.c_uint64), C.c_int32, C.c_int32, C.c_uint64, _vp, _i64, _i64, _i64, _vp, _vp]),
}

_lib: C.CDLL | None = None


class NativeError(RuntimeError):
    pass


def exported_symbols() -> list[str]:
    return list(_SIGNATURES)


def lib() -> C.CDLL:
    """loads (building first if needed) the CUDA library; raises if that is impossible"""
    global _lib
    if _lib is None:
        from magicsoup_b200 import build as _build

        if _build.needs_build():
            _build.build()  # raises if nvcc is missing or the compile fails: no fallback
        handle = C.CDLL(str(LIB_PATH))
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        l = lib()
        msg = l.ms_error_string(rc).decode()
        if rc == -2:
            msg += ": " + l.ms_last_cuda_error().decode()
        raise NativeError(f"magicsoup_b200 native call failed ({rc}): {msg}")


_cuda_seen = False


def require_cuda() -> None:
    """raises without a CUDA device (checked until one was seen: it is called in every step)"""
    global _cuda_seen
    if _cuda_seen:
        return
    if not torch.cuda.is_available() or lib().ms_device_count() < 1:
        raise RuntimeError(
            "magicsoup_b200 runs its hot path as CUDA kernels on a B200 only: "
            "no CUDA device is visible and there is no CPU fallback")
    _cuda_seen = True


def ptr(t: torch.Tensor | None):
    return None if t is None else C.c_void_p(t.data_ptr())


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def stream_ptr(device: torch.device):
    """torch's current stream on `device` as a cudaStream_t (called for every launch: the short way if torch has it)"""
    if _raw_stream is not None and device.index is not None:
        return C.c_void_p(_raw_stream(device.index))
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def eq_array(eq: list[int]):
    return (_i32 * 3)(*eq)
