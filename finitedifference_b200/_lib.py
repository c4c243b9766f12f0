"""ctypes binding of the C ABI declared in include/burgers_b200.h.

The shared library is built in-tree (``finitedifference_b200/_burgers_b200.so``) by
``__graft_entry__.build()`` / ``finitedifference_b200._build``.  There is no CPU fallback:
if the library is missing, loading raises.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("BFD_LIB") or os.path.join(HERE, "_burgers_b200.so")   # BFD_LIB: profiling variants

BFD_OK = 0
ERRORS = {-1: "BFD_EINVAL", -2: "BFD_ECUDA", -3: "BFD_EWORKSPACE", -4: "BFD_ETIMEOUT",
          -5: "BFD_ENOTCONVERGED", -6: "BFD_ETOOBIG"}


class FomDesc(C.Structure):
    """``bfd_fom_desc``."""
    _fields_ = [("nx", C.c_int32), ("ny", C.c_int32), ("batch", C.c_int32),
                ("num_steps", C.c_int32), ("max_its", C.c_int32), ("chunk_steps", C.c_int32),
                ("newton_hint", C.c_int32), ("strict", C.c_int32), ("uniform_y", C.c_int32),
                ("kernel", C.c_int32), ("rel_tol", C.c_double)]


def fom_desc(nx, ny, batch, num_steps, max_its=100, rel_tol=1e-12, chunk_steps=0, newton_hint=0,
             strict=True, uniform_y=False, kernel=0):
    return FomDesc(int(nx), int(ny), int(batch), int(num_steps), int(max_its), int(chunk_steps),
                   int(newton_hint), int(bool(strict)), int(bool(uniform_y)), int(kernel), float(rel_tol))


FOM_KMAX = 6
FOM_MAX_CHUNK = 64


class FomCoef(C.Structure):
    """``bfd_fom_coef``."""
    _fields_ = [(n, C.c_void_p) for n in ("cx", "cxl", "cy", "cyd", "src", "lbc")]


class FomSlab(C.Structure):
    """``bfd_fom_slab``."""
    _fields_ = [("rank", C.c_int32), ("world", C.c_int32), ("strip_begin", C.c_int32), ("strip_end", C.c_int32),
                ("home", C.c_void_p), ("next_halo", C.c_void_p)]


IPC_HANDLE_BYTES = 64


class RomMesh(C.Structure):
    """``bfd_rom_mesh``."""
    _fields_ = [("nrows", C.c_int64), ("ncells", C.c_int64)] + \
               [(n, C.c_void_p) for n in ("p", "xl", "yd", "ja", "jb", "cx", "cxl", "cy", "cyd", "src", "lbc", "weight")]


_P = C.c_void_p
# name -> (restype, argtypes): every symbol include/burgers_b200.h declares
SIGNATURES = {
    "bfd_abi_version": (C.c_int, []),
    "bfd_last_cuda_error": (C.c_char_p, []),
    "bfd_launch_count": (C.c_uint64, []),
    "bfd_fom_workspace_bytes": (C.c_size_t, [C.POINTER(FomDesc)]),
    "bfd_fom_init": (C.c_int, [C.POINTER(FomDesc), _P, C.c_size_t, _P]),
    "bfd_fom_set_initial": (C.c_int, [C.POINTER(FomDesc), _P, _P, C.c_size_t, _P]),
    "bfd_fom_run": (C.c_int, [C.POINTER(FomDesc), _P, C.POINTER(FomCoef), _P, _P]),
    "bfd_fom_export": (C.c_int, [C.POINTER(FomDesc), _P, C.c_int, C.c_int, C.c_int, C.c_int,
                                 C.c_size_t, _P, _P]),
    "bfd_fom_trace": (C.c_int, [C.POINTER(FomDesc), _P, _P, _P]),
    "bfd_fom_diagnostics": (C.c_int, [C.POINTER(FomDesc), _P, _P, _P, _P]),
    "bfd_fom_slab_workspace_bytes": (C.c_size_t, [C.POINTER(FomDesc), C.POINTER(FomSlab)]),
    "bfd_fom_slab_halo_offset": (C.c_size_t, [C.POINTER(FomDesc), C.POINTER(FomSlab)]),
    "bfd_fom_slab_init": (C.c_int, [C.POINTER(FomDesc), C.POINTER(FomSlab), _P, C.c_size_t, _P]),
    "bfd_fom_slab_set_initial": (C.c_int, [C.POINTER(FomDesc), C.POINTER(FomSlab), _P, _P, C.c_size_t, _P]),
    "bfd_fom_slab_run": (C.c_int, [C.POINTER(FomDesc), C.POINTER(FomSlab), _P, C.POINTER(FomCoef), _P, _P]),
    "bfd_fom_slab_export": (C.c_int, [C.POINTER(FomDesc), C.POINTER(FomSlab), _P, C.c_int, C.c_int, C.c_int, C.c_int,
                                      C.c_size_t, _P, _P]),
    "bfd_fom_slab_diagnostics": (C.c_int, [C.POINTER(FomDesc), C.POINTER(FomSlab), _P, _P, _P, _P]),
    "bfd_mem_alloc": (C.c_int, [C.c_size_t, C.POINTER(C.c_void_p)]),
    "bfd_mem_free": (C.c_int, [_P]),
    "bfd_ipc_export": (C.c_int, [_P, C.c_char_p]),
    "bfd_ipc_open": (C.c_int, [C.c_char_p, C.POINTER(C.c_void_p)]),
    "bfd_ipc_close": (C.c_int, [_P]),
    "bfd_rom_residual": (C.c_int, [C.POINTER(RomMesh), _P, _P, _P, _P, _P]),
    "bfd_rom_jv": (C.c_int, [C.POINTER(RomMesh), _P, _P, C.c_int, _P, _P]),
    "bfd_rom_jv_grid": (C.c_int, [C.POINTER(RomMesh), C.c_int, C.c_int, C.c_int, C.c_int, _P, _P, C.c_int, _P, _P]),
    "bfd_rom_jv_grid_slab": (C.c_int, [C.POINTER(RomMesh), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P, _P, _P,
                                       C.c_int, _P, _P]),
    "bfd_jac_csr_values": (C.c_int, [C.POINTER(RomMesh), _P, _P, _P]),
    "bfd_ecsw_c_fused": (C.c_int, [C.POINTER(RomMesh), _P, _P, _P, C.c_int, _P, C.c_size_t, _P]),
    "bfd_ecsw_c_grid": (C.c_int, [C.POINTER(RomMesh), C.c_int, C.c_int, _P, _P, _P, C.c_int, _P, C.c_size_t, _P]),
    "bfd_gram_workspace_bytes": (C.c_size_t, [C.c_int64, C.c_int]),
    "bfd_gram": (C.c_int, [_P, _P, C.c_int64, C.c_int, _P, _P, _P, C.c_size_t, _P]),
    "bfd_gram_static": (C.c_int, [_P, _P, C.c_int64, C.c_int, _P, _P, _P]),
    "bfd_spd_solve": (C.c_int, [_P, _P, C.c_int, C.c_double, _P, _P]),
    "bfd_basis_times_y": (C.c_int, [_P, _P, C.c_int64, C.c_int, _P, _P]),
    "bfd_ecsw_c_block": (C.c_int, [C.c_int64, C.c_int, _P, _P, _P, C.c_size_t, _P]),
    "bfd_probe_dfma": (C.c_int, [_P, C.c_int, C.c_int, _P]),
    "bfd_residual": (C.c_int, [C.c_int, C.c_int, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
}


class BurgersError(RuntimeError):
    pass


def bind(path: str) -> C.CDLL:
    """Load a library exporting the ABI and attach the signatures."""
    if not os.path.exists(path):
        raise BurgersError(
            f"{path} not found: build the CUDA extension first "
            "(python -c 'import __graft_entry__ as g; g.build()'); there is no CPU fallback")
    lib = C.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is missing: loud by design
        fn.restype, fn.argtypes = res, args
    return lib


_lib = None


def get_lib() -> C.CDLL:
    global _lib
    if _lib is None:
        _lib = bind(LIB_PATH)
    return _lib


def check(rc: int, lib: C.CDLL, what: str) -> None:
    if rc != BFD_OK:
        detail = lib.bfd_last_cuda_error().decode() if rc in (-2, -6) else ""
        raise BurgersError(f"{what} failed: {ERRORS.get(rc, rc)} {detail}")
