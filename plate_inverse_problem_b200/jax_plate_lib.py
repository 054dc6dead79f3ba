"""Drop-in for the reference's pybind11 module `jax_plate.jax_plate_lib`
(`/root/reference/source/jax_plate_lib/src/main.cpp:4-20`): class `InnerState` with
`add_mat`, `solve`, `matvec` (numpy in, numpy out, same argument order, same `mode` table,
RuntimeError on failure) -- executed by libjaxplate_b200.so on a B200.

Differences that are deliberate and documented in DESIGN.md:
* `n_cpu` is accepted and ignored (the OpenMP loop over batch items,
  `InnerState.h:251,264,277,291`, is the CUDA grid's batch dimension here);
* within one `solve` call every distinct matrix is factorised once and the factors are
  reused for all of its right-hand sides (modes 2 and 4), where the reference
  re-factorises per right-hand side (`InnerState.h:262-304`);
* real (float64) inputs are promoted to complex128 for the computation and the real part
  is returned (the reference dispatches to umfpack_di_*; only complex128 is ever used by
  jax_plate, `Problem.py:242-244,331-332`).
"""
from __future__ import annotations

import numpy as np

from . import _lib


def _as_c128(a):
    return np.ascontiguousarray(a, dtype=np.complex128)


class InnerState:
    """`InnerState` of `InnerState.h:86-471`."""

    def __init__(self, device: int = 0):
        self._st = _lib.State(device)
        self._N = []
        self._nnz = []

    # InnerState::add_mat, InnerState.h:120-162
    def add_mat(self, N, indices, indptr, indices_T, indptr_T, permutation, data=None):
        indices = np.ascontiguousarray(indices)
        if indices.dtype not in (np.int32, np.int64):
            raise TypeError("add_mat: index arrays must be int32 or int64")
        dt = indices.dtype
        arrs = [np.ascontiguousarray(a, dtype=dt) for a in (indptr, indices_T, indptr_T, permutation)]
        indptr, indices_T, indptr_T, permutation = arrs
        num = _lib.C.c_int(-1)
        st = self._st
        st.check(st.lib.jpb_add_mat(st.h, int(N), int(indices.size), _lib.ptr(indices),
                                    _lib.ptr(indptr), _lib.ptr(indices_T), _lib.ptr(indptr_T),
                                    _lib.ptr(permutation), dt.itemsize, _lib.C.byref(num)))
        self._N.append(int(N))
        self._nnz.append(int(indices.size))
        return None   # the reference returns void; solver_num is the insertion index

    def _shapes(self, data, b, solver_num, mode, what):
        if not 0 <= solver_num < len(self._N):
            raise RuntimeError(f"{what}: no pattern registered under solver_num={solver_num}")
        N, nnz = self._N[solver_num], self._nnz[solver_num]
        mode = int(mode)
        want = {0: ((nnz,), (N,)), 1: ((-1, nnz), (N,)), 2: ((nnz,), (-1, N)),
                3: ((-1, nnz), (-1, N)), 4: ((-1, nnz), (-1, -1, N))}
        if mode not in want:
            raise RuntimeError(f"Invalid `mode` argument in InnerState::{what}().")
        wd, wb = want[mode]
        if data.ndim != len(wd) or b.ndim != len(wb) or data.shape[-1] != nnz or b.shape[-1] != N:
            raise RuntimeError(f"{what}: array shapes {data.shape}, {b.shape} do not fit mode {mode}")
        nb0 = nb1 = 0
        if mode == 1:
            nb0 = data.shape[0]
            out_shape = (nb0, N)
        elif mode == 2:
            nb0 = b.shape[0]
            out_shape = b.shape
        elif mode == 3:
            if data.shape[0] != b.shape[0]:
                raise RuntimeError(f"{what}: batch sizes of data and b differ in mode 3")
            nb0 = b.shape[0]
            out_shape = b.shape
        elif mode == 4:
            if data.shape[0] != b.shape[1]:
                raise RuntimeError(f"{what}: data.shape[0] must equal b.shape[1] in mode 4")
            nb0, nb1 = b.shape[0], b.shape[1]
            out_shape = b.shape
        else:
            out_shape = b.shape
        return mode, nb0, nb1, out_shape

    # InnerState::solve, InnerState.h:164-308
    def solve(self, data, b, solver_num, transpose, n_cpu=0, mode=0):
        real = np.isrealobj(data) and np.isrealobj(b)
        data, b = _as_c128(data), _as_c128(b)
        mode, nb0, nb1, out_shape = self._shapes(data, b, int(solver_num), mode, "solve")
        out = np.empty(out_shape, dtype=np.complex128)
        st = self._st
        st.check(st.lib.jpb_solve(st.h, int(solver_num), mode, int(bool(transpose)), _lib.ptr(data),
                                  _lib.ptr(b), nb0, nb1, _lib.ptr(out), None))
        return out.real.copy() if real else out

    # InnerState::matvec, InnerState.h:310-470
    def matvec(self, mat, vec, solver_num, transpose, n_cpu=0, mode=0):
        real = np.isrealobj(mat) and np.isrealobj(vec)
        mat, vec = _as_c128(mat), _as_c128(vec)
        mode, nb0, nb1, out_shape = self._shapes(mat, vec, int(solver_num), mode, "matvec")
        out = np.empty(out_shape, dtype=np.complex128)
        st = self._st
        st.check(st.lib.jpb_matvec(st.h, int(solver_num), mode, int(bool(transpose)), _lib.ptr(mat),
                                   _lib.ptr(vec), nb0, nb1, _lib.ptr(out)))
        return out.real.copy() if real else out

    # extras (not in the reference module)
    def symbolic_info(self, solver_num):
        return self._st.symbolic_info(solver_num)

    def timings(self):
        return self._st.timings()

    def set_refinement(self, steps: int):
        self._st.check(self._st.lib.jpb_set_refinement(self._st.h, int(steps)))
