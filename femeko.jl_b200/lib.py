"""ctypes binding of ``csrc/libfemeko_cuda.so`` (C ABI declared in ``include/femeko_cuda.h``).

This is the Python twin of the Julia ``@ccall`` shim: same symbols, same argument
order.  There is no fallback: if the shared library is missing it is built with nvcc, and if that
fails, or no CUDA device is present at ``fk_create`` time, an exception is raised.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.path.join(CSRC, "libfemeko_cuda.so")

FK_NUM_SLOTS = 4
FK_MAX_SPLINES = 8
BC = {"extrapolate": 0, "zero": 1, "error": 2, "nearest": 3}

c_double_p = C.POINTER(C.c_double)
c_int_p = C.POINTER(C.c_int)
c_ubyte_p = C.POINTER(C.c_ubyte)


class FemekoError(RuntimeError):
    pass


class fk_create_opts(C.Structure):
    _fields_ = [("tile_max_elems", C.c_int), ("tile_max_rows", C.c_int), ("natural_order", C.c_int),
                ("sell_sigma", C.c_int), ("device", C.c_int), ("n_owned", C.c_int), ("asm_cfg", C.c_int),
                ("spmv_ctas_per_sm", C.c_int), ("sell_wide_cols", C.c_int), ("nodes_per_element", C.c_int)]


class fk_pcg_opts(C.Structure):
    _fields_ = [("tol", C.c_double), ("max_iter", C.c_int), ("warm_start", C.c_int), ("check_every", C.c_int),
                ("tol_disable", C.c_int), ("no_precond", C.c_int), ("recurrence", C.c_int), ("project_nullspace", C.c_int)]


def build(verbose: bool = False) -> str:
    """Compile the CUDA library in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    r = subprocess.run(["make", "-C", CSRC, "libfemeko_cuda.so"], capture_output=True, text=True,
                       env={**os.environ, "MAKEFLAGS": ""})
    if r.returncode != 0:
        raise FemekoError("building libfemeko_cuda.so failed:\n" + r.stdout[-4000:] + r.stderr[-4000:])
    if verbose:
        print(r.stdout)
    return LIB_PATH


_H = C.c_void_p
_SIGS = {
    # name: (restype, argtypes)
    "fk_last_error": (C.c_char_p, []),
    "fk_version": (C.c_int, []),
    "fk_create": (C.c_int, [c_double_p, c_int_p, C.c_int, C.c_int, c_double_p, C.POINTER(fk_create_opts), C.POINTER(_H)]),
    "fk_destroy": (C.c_int, [_H]),
    "fk_set_stream": (C.c_int, [_H, C.c_void_p]),
    "fk_synchronize": (C.c_int, [_H]),
    "fk_tune": (C.c_int, [_H, C.c_int, C.c_int]),
    "fk_launch_count": (C.c_int, [_H, C.POINTER(C.c_longlong)]),
    "fk_geo_stats": (C.c_int, [_H] + [C.POINTER(C.c_longlong)] * 4),
    "fk_trace_read": (C.c_int, [_H, C.POINTER(C.c_ulonglong), C.c_int, C.POINTER(C.c_int)]),
    "fk_sizes": (C.c_int, [_H, c_int_p, c_int_p, C.POINTER(C.c_longlong), c_int_p]),
    "fk_csr_pattern": (C.c_int, [_H, c_int_p, c_int_p]),
    "fk_elem2nnz": (C.c_int, [_H, c_int_p]),
    "fk_volumes": (C.c_int, [_H, c_double_p]),
    "fk_tile_stats": (C.c_int, [_H, C.POINTER(C.c_longlong), c_int_p, c_int_p, c_int_p]),
    "fk_sell_stats": (C.c_int, [_H] + [C.POINTER(C.c_longlong)] * 4),
    "fk_device_ptr": (C.c_int, [_H, C.c_int, C.POINTER(C.c_void_p)]),
    "fk_symbolic_create": (C.c_int, [c_double_p, c_int_p, C.c_int, C.c_int, C.POINTER(fk_create_opts), C.POINTER(_H)]),
    "fk_symbolic_sizes": (C.c_int, [_H, C.POINTER(C.c_longlong), c_int_p, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong),
                                    C.POINTER(C.c_longlong), c_int_p, C.POINTER(C.c_longlong)]),
    "fk_symbolic_export": (C.c_int, [_H] + [C.c_void_p] * 15),
    "fk_symbolic_colz": (C.c_int, [_H, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong), C.c_void_p, C.c_void_p]),
    "fk_heat_step": (C.c_int, [_H, C.c_int, C.c_int, c_double_p, C.POINTER(fk_pcg_opts), c_int_p, c_double_p]),
    "fk_heat_step_dev": (C.c_int, [_H, C.c_int, C.c_int, C.c_void_p, C.POINTER(fk_pcg_opts), c_int_p, c_double_p]),
    "fk_multi_create": (C.c_int, [c_double_p, c_int_p, C.c_int, C.c_int, c_double_p, C.c_int, c_int_p, C.POINTER(fk_create_opts), C.POINTER(_H)]),
    "fk_multi_destroy": (C.c_int, [_H]),
    "fk_multi_info": (C.c_int, [_H, c_int_p, c_int_p, c_int_p, c_int_p, c_int_p]),
    "fk_multi_handle": (C.c_int, [_H, C.c_int, C.POINTER(_H)]),
    "fk_multi_partition": (C.c_int, [_H, C.c_int, c_int_p, c_int_p]),
    "fk_rcb": (C.c_int, [c_double_p, C.c_int, C.c_int, c_int_p]),
    "fk_partition_local": (C.c_int, [c_double_p, c_int_p, C.c_int, C.c_int, c_int_p, C.c_int] + [c_int_p] * 9),
    "fk_multi_assemble_stiffness": (C.c_int, [_H, C.c_int, c_double_p]),
    "fk_multi_assemble_mass": (C.c_int, [_H, C.c_int, c_double_p]),
    "fk_multi_matrix_axpby": (C.c_int, [_H, C.c_int, C.c_double, C.c_int, C.c_double, C.c_int]),
    "fk_multi_set_spline": (C.c_int, [_H, C.c_int, c_double_p, c_double_p, C.c_int, C.c_int, C.c_int]),
    "fk_multi_set_inside": (C.c_int, [_H, c_int_p, C.c_int]),
    "fk_multi_set_mu": (C.c_int, [_H, c_double_p]),
    "fk_multi_get_mu": (C.c_int, [_H, c_double_p]),
    "fk_multi_get_field": (C.c_int, [_H, c_double_p, c_double_p]),
    "fk_multi_reset_field": (C.c_int, [_H]),
    "fk_multi_spmv": (C.c_int, [_H, C.c_int, c_double_p, c_double_p]),
    "fk_multi_pcg": (C.c_int, [_H, C.c_int, c_double_p, c_double_p, C.POINTER(fk_pcg_opts), c_int_p, c_double_p]),
    "fk_multi_picard_step": (C.c_int, [_H, C.c_int, c_double_p, C.c_int, c_double_p]),
    "fk_multi_demag_setup": (C.c_int, [_H, C.c_int, c_int_p, C.c_int, c_double_p]),
    "fk_multi_demag_field": (C.c_int, [_H, C.c_int, c_double_p, C.POINTER(fk_pcg_opts), C.c_int, c_double_p, c_double_p, c_int_p, c_double_p]),
    "fk_host_register": (C.c_int, [C.c_void_p, C.c_longlong]),
    "fk_host_unregister": (C.c_int, [C.c_void_p]),
    "fk_assemble_convection_p2": (C.c_int, [_H, C.c_int, c_double_p, c_double_p, C.c_int, c_double_p]),
    "fk_create_edge": (C.c_int, [c_double_p, C.c_int, c_int_p, c_int_p, C.c_int, C.c_int, c_double_p, C.c_void_p, C.POINTER(_H)]),
    "fk_nedelec_local_stiffness": (C.c_int, [_H, c_double_p]),
    "fk_assemble_nedelec_stiffness": (C.c_int, [_H, C.c_int, c_double_p, c_double_p]),
    "fk_assemble_nedelec_tangential": (C.c_int, [_H, C.c_int, c_double_p, c_double_p, c_double_p, c_double_p]),
    "fk_nedelec_load_vector": (C.c_int, [_H, c_double_p, c_double_p]),
    "fk_nedelec_recover_field": (C.c_int, [_H, c_double_p, c_double_p, c_double_p]),
    "fk_residual_norm_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p, c_double_p, C.c_void_p]),
    "fk_assemble_stiffness2d": (C.c_int, [_H, C.c_int, c_double_p, c_double_p]),
    "fk_assemble_mass2d": (C.c_int, [_H, C.c_int, c_double_p, c_double_p]),
    "fk_comm_set_timeout": (C.c_int, [_H, C.c_int]),
    "fk_comm_reset": (C.c_int, [_H]),
    "fk_assemble_stiffness_p2": (C.c_int, [_H, C.c_int, c_double_p, c_double_p]),
    "fk_assemble_mass_p2": (C.c_int, [_H, C.c_int, c_double_p, c_double_p]),
    "fk_set_host_threads": (C.c_int, [C.c_int]),
    "fk_symbolic_rescale": (C.c_int, [_H, C.POINTER(C.c_longlong), c_int_p, C.POINTER(C.c_longlong), C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_void_p]),
    "fk_symbolic_destroy": (C.c_int, [_H]),
    "fk_assemble_stiffness": (C.c_int, [_H, C.c_int, c_double_p, c_double_p]),
    "fk_assemble_mass": (C.c_int, [_H, C.c_int, c_double_p, c_double_p]),
    "fk_assemble_tangential": (C.c_int, [_H, C.c_int, c_double_p, c_double_p, c_double_p]),
    "fk_local_stiffness": (C.c_int, [_H, c_double_p, c_double_p]),
    "fk_assemble_stiffness_dev": (C.c_int, [_H, C.c_int, C.c_void_p]),
    "fk_assemble_mass_dev": (C.c_int, [_H, C.c_int, C.c_void_p]),
    "fk_assemble_tangential_dev": (C.c_int, [_H, C.c_int, C.c_void_p, C.c_void_p]),
    "fk_matrix_axpby": (C.c_int, [_H, C.c_int, C.c_double, C.c_int, C.c_double, C.c_int]),
    "fk_matrix_get": (C.c_int, [_H, C.c_int, c_double_p]),
    "fk_matrix_set": (C.c_int, [_H, C.c_int, c_double_p]),
    "fk_lagrange": (C.c_int, [_H, c_double_p]),
    "fk_boundary_integral": (C.c_int, [_H, c_int_p, c_double_p, c_double_p, C.c_int, c_double_p, c_int_p, C.c_int, c_double_p]),
    "fk_spmv": (C.c_int, [_H, C.c_int, c_double_p, c_double_p]),
    "fk_spmv_dev": (C.c_int, [_H, C.c_int, C.c_void_p, C.c_void_p]),
    "fk_pcg": (C.c_int, [_H, C.c_int, c_double_p, c_double_p, C.POINTER(fk_pcg_opts), c_int_p, c_double_p]),
    "fk_pcg_dev": (C.c_int, [_H, C.c_int, C.c_void_p, C.c_void_p, C.POINTER(fk_pcg_opts), c_int_p, c_double_p]),
    "fk_recover_field": (C.c_int, [_H, c_double_p, c_double_p, C.c_int, c_double_p, c_double_p, c_double_p]),
    "fk_recover_field_dev": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "fk_spline_fit": (C.c_int, [c_double_p, c_double_p, C.c_int, c_double_p, c_double_p]),
    "fk_spline_eval": (C.c_int, [c_double_p, c_double_p, C.c_int, C.c_int, C.c_int, c_double_p, C.c_int, c_double_p]),
    "fk_set_spline": (C.c_int, [_H, C.c_int, c_double_p, c_double_p, C.c_int, C.c_int, C.c_int]),
    "fk_spline_apply": (C.c_int, [_H, C.c_int, c_double_p, C.c_int, c_double_p]),
    "fk_spline_apply_dev": (C.c_int, [_H, C.c_int, C.c_void_p, C.c_int, C.c_void_p]),
    "fk_set_material": (C.c_int, [_H, c_ubyte_p]),
    "fk_set_inside": (C.c_int, [_H, c_int_p, C.c_int]),
    "fk_set_mu": (C.c_int, [_H, c_double_p]),
    "fk_get_mu": (C.c_int, [_H, c_double_p, c_double_p]),
    "fk_get_field": (C.c_int, [_H, c_double_p, c_double_p]),
    "fk_reset_field": (C.c_int, [_H]),
    "fk_picard_step": (C.c_int, [_H, C.c_int, c_double_p, C.c_int, c_double_p]),
    "fk_picard_step_dev": (C.c_int, [_H, C.c_int, C.c_void_p, C.c_int, c_double_p]),
    "fk_residual": (C.c_int, [_H, C.c_int, c_double_p, c_double_p, c_double_p, c_double_p]),
    "fk_residual_dev": (C.c_int, [_H, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, c_double_p]),
    "fk_line_search": (C.c_int, [_H, C.c_int, c_double_p, c_double_p, c_double_p, C.c_double, c_double_p, c_double_p]),
    "fk_line_search_dev": (C.c_int, [_H, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, c_double_p, c_double_p]),
    "fk_picard_result": (C.c_int, [_H, c_double_p]),
    "fk_set_dirichlet": (C.c_int, [_H, c_int_p, C.c_int, c_double_p]),
    "fk_apply_dirichlet": (C.c_int, [_H, C.c_int, c_double_p]),
    "fk_apply_dirichlet_dev": (C.c_int, [_H, C.c_int, C.c_void_p]),
    "fk_set_nullspace": (C.c_int, [_H, c_double_p]),
    "fk_picard_pcg": (C.c_int, [_H, C.c_int, c_double_p, C.c_int, c_double_p, c_double_p, C.POINTER(fk_pcg_opts), c_double_p,
                                c_int_p, c_double_p]),
    "fk_demag_setup": (C.c_int, [_H, C.c_int, c_int_p, C.c_int, c_double_p]),
    "fk_demag_field": (C.c_int, [_H, C.c_int, c_double_p, C.POINTER(fk_pcg_opts), C.c_int, c_double_p, c_double_p, c_int_p,
                                 c_double_p]),
    "fk_demag_field_dev": (C.c_int, [_H, C.c_int, C.c_void_p, C.POINTER(fk_pcg_opts), C.c_int, C.c_void_p, C.c_void_p, c_int_p,
                                     c_double_p]),
    "fk_nccl_unique_id": (C.c_int, [c_ubyte_p]),
    "fk_comm_init": (C.c_int, [_H, c_ubyte_p, C.c_int, C.c_int]),
    "fk_set_halo": (C.c_int, [_H, C.c_int, C.c_int, c_int_p, c_int_p, c_int_p, c_int_p, c_int_p]),
    "fk_comm_p2p_export": (C.c_int, [_H, c_ubyte_p]),
    "fk_comm_p2p_attach": (C.c_int, [_H, c_ubyte_p, c_int_p, c_int_p]),
    "fk_comm_mode": (C.c_int, [_H, c_int_p]),
    "scalarPotential": (None, [c_double_p, c_double_p, c_int_p, c_int_p, c_double_p, C.c_int, C.c_int, C.c_int,
                               c_double_p, c_double_p, c_double_p, C.c_int]),
}

_lib = None


def load() -> C.CDLL:
    """Load (building first if necessary) the C-ABI library.  Raises if it cannot be had."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        build()
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in _SIGS.items():
        fn = getattr(lib, name)  # AttributeError if the ABI and the header drifted apart
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, allow=()) -> int:
    if rc != 0 and rc not in allow:
        raise FemekoError(load().fk_last_error().decode() or f"libfemeko_cuda error {rc}")
    return rc


def dptr(a):
    """Host numpy array -> double* (None -> NULL).  The array must be float64 and contiguous in
    either order (Fortran-ordered matrices are passed as Julia would pass them)."""
    if a is None:
        return None
    assert a.dtype == np.float64 and (a.flags.c_contiguous or a.flags.f_contiguous), "need contiguous float64"
    return a.ctypes.data_as(c_double_p)


def iptr(a):
    if a is None:
        return None
    assert a.dtype == np.int32 and (a.flags.c_contiguous or a.flags.f_contiguous), "need contiguous int32"
    return a.ctypes.data_as(c_int_p)


def f64(a, order_like=None):
    """Coerce to float64 without changing the memory order of an already-contiguous matrix."""
    if a is None:
        return None
    a = np.asarray(a, dtype=np.float64)
    if a.flags.c_contiguous or a.flags.f_contiguous:
        return a
    return np.asfortranarray(a)


def i32(a):
    if a is None:
        return None
    a = np.asarray(a, dtype=np.int32)
    if a.flags.c_contiguous or a.flags.f_contiguous:
        return a
    return np.asfortranarray(a)
