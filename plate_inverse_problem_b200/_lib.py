"""ctypes binding of libjaxplate_b200.so (the C-ABI in include/jax_plate_b200.h).

There is no CPU fallback: if the shared library is missing, or no CUDA device can be
opened, everything that computes raises.  Build the library with
`python -c "import __graft_entry__ as g; g.build()"` (nvcc, sm_100a, in-tree).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# JPB_LIB_PATH: development override to load an alternative build of the same library
LIB_PATH = os.environ.get("JPB_LIB_PATH") or os.path.join(_HERE, "libjaxplate_b200.so")

JPB_OK, JPB_EINVAL, JPB_ECUDA, JPB_ESINGULAR, JPB_ENOMEM, JPB_ESTATE = range(6)
LOSS_TYPES = {"MSE": 0, "RMSE": 1, "MSE_AFC": 2, "MSE_LOG_AFC": 3}
COMPLEX_LOSS_TYPES = {"MSE": 5, "RMSE": 6, "MSE_AFC": 2, "MSE_LOG_AFC": 3}

# every symbol include/jax_plate_b200.h declares
EXPORTED_SYMBOLS = (
    "jpb_create", "jpb_destroy", "jpb_last_error", "jpb_set_memory_budget", "jpb_set_refinement",
    "jpb_set_streams", "jpb_set_residual_tolerance", "jpb_set_adaptive_refinement", "jpb_set_refinement_threshold", "jpb_last_refined", "jpb_set_equilibration",
    "jpb_add_mat", "jpb_symbolic_info", "jpb_get_ordering", "jpb_solve", "jpb_matvec",
    "jpb_assemble", "jpb_register_components", "jpb_eval_forward", "jpb_eval_loss_grad",
    "jpb_eval_loss_grad_dev", "jpb_eval_candidates", "jpb_eval_candidates_dev", "jpb_eval_jacobian",
    "jpb_eval_jacobian_dev", "jpb_last_timings", "jpb_dmma_peak", "jpb_selftest", "jpb_residual_noalloc",
    "jpb_host_analyze", "jpb_host_sizes", "jpb_host_export", "jpb_host_free",
)

_lib = None


def load_library() -> C.CDLL:
    """dlopen the in-tree CUDA library; raise (never fall back) when it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: the CUDA library has not been built "
            "(run __graft_entry__.build()); plate_inverse_problem_b200 has no CPU fallback")
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64, dp = C.c_void_p, C.c_int, C.c_int64, C.c_void_p
    lib.jpb_create.argtypes = [C.POINTER(vp), i32]
    lib.jpb_destroy.argtypes = [vp]
    lib.jpb_destroy.restype = None
    lib.jpb_last_error.argtypes = [vp]
    lib.jpb_last_error.restype = C.c_char_p
    lib.jpb_set_memory_budget.argtypes = [vp, i64]
    lib.jpb_set_refinement.argtypes = [vp, i32]
    lib.jpb_set_streams.argtypes = [vp, i32]
    lib.jpb_add_mat.argtypes = [vp, i64, i64, dp, dp, dp, dp, dp, i32, C.POINTER(i32)]
    lib.jpb_symbolic_info.argtypes = [vp, i32, dp]
    lib.jpb_get_ordering.argtypes = [vp, i32, dp]
    lib.jpb_solve.argtypes = [vp, i32, i32, i32, dp, dp, i64, i64, dp, dp]
    lib.jpb_matvec.argtypes = [vp, i32, i32, i32, dp, dp, i64, i64, dp]
    lib.jpb_assemble.argtypes = [vp, i64, i32, i64, dp, dp, dp]
    lib.jpb_register_components.argtypes = [vp, i32, i32, dp, dp, i32, dp]
    lib.jpb_eval_forward.argtypes = [vp, i32, i64, dp, dp, dp, dp, dp]
    lib.jpb_eval_loss_grad.argtypes = [vp, i32, i64, dp, dp, dp, dp, i32, i64, dp, dp, dp, dp, dp]
    lib.jpb_eval_loss_grad_dev.argtypes = [vp, i32, i64, dp, dp, dp, dp, i32, i64, dp, dp, dp, dp, dp, vp]
    lib.jpb_eval_candidates.argtypes = [vp, i32, i64, dp, dp, dp, dp, i32, i64, dp, i32, dp, dp, dp, dp, dp]
    lib.jpb_eval_candidates_dev.argtypes = [vp, i32, i64, dp, dp, dp, dp, i32, i64, dp, i32, dp, dp, dp, dp, dp, vp]
    lib.jpb_eval_jacobian.argtypes = [vp, i32, i64, dp, dp, dp, dp, dp, dp, dp]
    lib.jpb_eval_jacobian_dev.argtypes = [vp, i32, i64, dp, dp, dp, dp, dp, dp, dp, vp]
    lib.jpb_set_residual_tolerance.argtypes = [vp, C.c_double, C.c_double]
    lib.jpb_set_adaptive_refinement.argtypes = [vp, i32, C.c_double]
    lib.jpb_set_refinement_threshold.argtypes = [vp, C.c_double, C.c_double]
    lib.jpb_last_refined.argtypes = [vp, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    lib.jpb_set_equilibration.argtypes = [vp, i32]
    lib.jpb_last_timings.argtypes = [vp, dp]
    lib.jpb_dmma_peak.argtypes = [vp, dp]
    lib.jpb_selftest.argtypes = [vp, dp]
    lib.jpb_residual_noalloc.argtypes = [vp, i32, i32, i64, vp, vp, vp, vp, vp, vp]
    lib.jpb_host_analyze.argtypes = [i64, i64, dp, dp, i32, C.POINTER(vp), C.c_char_p, i32]
    lib.jpb_host_sizes.argtypes = [vp, dp, dp]
    lib.jpb_host_export.argtypes = [vp] + [dp] * 14
    lib.jpb_host_free.argtypes = [vp]
    lib.jpb_host_free.restype = None
    for name in EXPORTED_SYMBOLS:
        if name not in ("jpb_destroy", "jpb_last_error", "jpb_host_free"):
            getattr(lib, name).restype = i32
    _lib = lib
    return lib


def ptr(a):
    """Raw data pointer of a C-contiguous numpy array (or None)."""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)


class State:
    """Owner of one `jpb_state` (one CUDA device)."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.jpb_create(C.byref(h), int(device))
        if rc != JPB_OK:
            raise RuntimeError(
                f"jpb_create(device={device}) failed with code {rc}: no usable CUDA device "
                "(plate_inverse_problem_b200 has no CPU fallback)")
        self.h = h
        self.device = int(device)

    def check(self, rc: int):
        if rc != JPB_OK:
            msg = self.lib.jpb_last_error(self.h).decode()
            raise RuntimeError(msg)

    def close(self):
        if getattr(self, "h", None):
            self.lib.jpb_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def timings(self) -> dict:
        t = np.zeros(10)
        self.check(self.lib.jpb_last_timings(self.h, ptr(t)))
        keys = ("factor_ms", "fwd_solve_ms", "adj_solve_ms", "spmv_ms", "grad_ms", "readout_ms",
                "h2d_ms", "d2h_ms", "launches", "wave")
        return dict(zip(keys, t.tolist()))

    def refined(self):
        """(vectors that took the residual-driven refinement step, vectors solved) of the last call."""
        a, b = C.c_int64(0), C.c_int64(0)
        self.check(self.lib.jpb_last_refined(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def symbolic_info(self, solver_num: int) -> dict:
        t = np.zeros(10)
        self.check(self.lib.jpb_symbolic_info(self.h, int(solver_num), ptr(t)))
        keys = ("N", "nnz", "nnz_lu", "flops", "front_slots", "nfronts", "nlevels", "max_front",
                "max_pivots", "vec_slots")
        return dict(zip(keys, t.tolist()))


def host_symbolic(N: int, indices: np.ndarray, indptr: np.ndarray) -> dict:
    """Run the symbolic analysis on the host (no GPU needed) and return its arrays."""
    lib = load_library()
    indices = np.ascontiguousarray(indices)
    indptr = np.ascontiguousarray(indptr, dtype=indices.dtype)
    h = C.c_void_p()
    err = C.create_string_buffer(512)
    rc = lib.jpb_host_analyze(int(N), int(indices.size), ptr(indices), ptr(indptr),
                              indices.dtype.itemsize, C.byref(h), err, 512)
    if rc != JPB_OK:
        raise RuntimeError(err.value.decode())
    try:
        z = np.zeros(8, dtype=np.int64)
        fl = np.zeros(1)
        lib.jpb_host_sizes(h, ptr(z), ptr(fl))
        nf, nrow, nrel, nnz = (int(v) for v in z[:4])
        i32, i64 = np.int32, np.int64
        out = dict(perm=np.zeros(N, i32), f_s=np.zeros(nf, i32), f_k=np.zeros(nf, i32),
                   f_ld=np.zeros(nf, i32), f_off=np.zeros(nf, i64), f_rowptr=np.zeros(nf + 1, i32),
                   rowidx=np.zeros(nrow, i32), f_parent=np.zeros(nf, i32), f_level=np.zeros(nf, i32),
                   f_reloff=np.zeros(nf, i64), relmap=np.zeros(max(nrel, 1), i32),
                   f_entptr=np.zeros(nf + 1, i64), ent_src=np.zeros(nnz, i32), ent_slot=np.zeros(nnz, i32))
        lib.jpb_host_export(h, *[ptr(out[k]) for k in ("perm", "f_s", "f_k", "f_ld", "f_off", "f_rowptr",
                                                        "rowidx", "f_parent", "f_level", "f_reloff",
                                                        "relmap", "f_entptr", "ent_src", "ent_slot")])
        out.update(nfronts=nf, nlevels=int(z[4]), front_slots=int(z[5]), nnz_lu=int(z[6]), N=int(z[7]),
                   flops=float(fl[0]), nnz=nnz)
        return out
    finally:
        lib.jpb_host_free(h)
