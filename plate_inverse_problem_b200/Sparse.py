"""Host-side mirror of `/root/reference/source/jax_plate/Sparse.py`: `create_symbolic`,
`find_permutation`, `spsolve`, `matvec`, the `_mode` table of the vmap batchers
(`Sparse.py:245-282`) and the derivative rules (`:162-176, 200-222`) -- with the native
calls going to libjaxplate_b200.so instead of the pybind11/UMFPACK module.

jax is not available in this image, so the two JAX primitives (`cpu_spsolve`,
`custom_matvec`) are provided as plain functions over numpy arrays with the same
signature and parameters (`solver_num, transpose, n_cpu, _mode`), plus:
* `batch_mode(data, b)` : the `_mode` the JAX batchers would have chosen for these shapes,
* `spsolve_jvp` / `spsolve_vjp` / `matvec_vjp_matrix` : the JVP / transpose rules, so that a
  host framework (JAX custom_vjp, torch.autograd.Function, ...) can register them verbatim,
* `SpSolve` : a torch.autograd.Function doing exactly that for torch.
INTEGRATION.md shows the XLA-FFI registration a jax_plate maintainer would add.
"""
from __future__ import annotations

import os

import numpy as np
import scipy.sparse as sps

from .jax_plate_lib import InnerState

FAMILIES = [np.float64, np.complex128]   # Sparse.py:87-90 (complex64/float32 are commented out there)


class SolverState:
    """`Sparse.py:19-44`."""

    def __init__(self, device: int = 0):
        self.state = InnerState(device)
        self.patterns = []   # _SOLVER_STATE.patterns[num] -> [nnz, 2] (row, col), used by the AD rules

    def add_mat(self, mat, mat_T, coo_indices, perm):
        # Sparse.py:33-34 passes mat.indptr where indptr_T is expected; kept (it is unused natively)
        self.state.add_mat(mat.shape[0], mat.indices, mat.indptr, mat_T.indices, mat.indptr, perm,
                           mat.data)
        self.patterns.append(np.asarray(coo_indices))
        return len(self.patterns) - 1

    def solve(self, data, b, solver_num, transpose, n_cpu, mode):
        return self.state.solve(data, b, solver_num, transpose, n_cpu, mode)

    def matvec(self, mat, vec, solver_num, transpose, n_cpu, mode):
        return self.state.matvec(mat, vec, solver_num, transpose, n_cpu, mode)


_SOLVER_STATE = None


def solver_state() -> SolverState:
    """The module-level singleton of `Sparse.py:44`, created on first use (needs a GPU)."""
    global _SOLVER_STATE
    if _SOLVER_STATE is None:
        _SOLVER_STATE = SolverState(int(os.environ.get("LOCAL_RANK", "0")))
    return _SOLVER_STATE


def find_permutation(arr1: np.ndarray, arr2: np.ndarray, max_val=None) -> np.ndarray:
    """`Sparse.py:46-85`, evaluated with int64 keys (the reference's int32 key
    `a[:,0] + a[:,1]*max_val` overflows for N > 46 340)."""
    assert arr1.shape == arr2.shape
    assert arr1.shape[1] == 2
    if max_val is None:
        max_val = int(np.max(arr1)) + 1
    u1 = arr1[:, 0].astype(np.int64) + arr1[:, 1].astype(np.int64) * np.int64(max_val)
    u2 = arr2[:, 0].astype(np.int64) + arr2[:, 1].astype(np.int64) * np.int64(max_val)
    is1 = np.argsort(u1)
    is2 = np.argsort(u2)
    inverse_is2 = is2[is2[is2]]
    res = np.arange(u1.size)[is1][inverse_is2]
    return np.array(res, dtype=arr1.dtype)


def symbolic_indices(N: int, indices: np.ndarray):
    """The index construction of `create_symbolic` (`Sparse.py:98-113`) without the native
    registration: returns (mat, mat_T, coo_indices, perm).  Pure integer work; bit-exact."""
    nnz = indices.shape[0]
    data = np.linspace(1.0, 42.0, nnz, dtype=np.complex128)
    mat = sps.coo_matrix((data, (indices[:, 0], indices[:, 1])), (N, N), dtype=np.complex128,
                         copy=True).tocsc()
    mat_T = sps.coo_matrix((data, (indices[:, 1], indices[:, 0])), (N, N), dtype=np.complex128,
                           copy=True).tocsc()
    m1 = mat.tocoo()
    m1_T = mat_T.tocoo()
    coo_indices = np.vstack((m1.row, m1.col)).astype(indices.dtype).T
    coo_indices_T = np.vstack((m1_T.row, m1_T.col)).astype(indices.dtype).T
    perm = find_permutation(coo_indices[:, ::-1], coo_indices_T, N)
    return mat, mat_T, coo_indices, perm


def create_symbolic(N: int, indices: np.ndarray, mat_dtype=np.complex128, state: SolverState = None):
    """`Sparse.py:92-116`: returns ((row, col), solver_num)."""
    if mat_dtype not in FAMILIES:
        raise TypeError(f"dtype {mat_dtype} is not supported")
    mat, mat_T, coo_indices, perm = symbolic_indices(N, np.asarray(indices))
    st = state or solver_state()
    num = st.add_mat(mat, mat_T, coo_indices, perm)
    return (coo_indices[:, 0], coo_indices[:, 1]), num


def _check(data, b, n_cpu):
    """`_abstract_eval`, `Sparse.py:120-140`."""
    if data.dtype != b.dtype:
        raise TypeError(f"data types do not match: type(data)={data.dtype}, type(b)={b.dtype}")
    if not isinstance(n_cpu, (int, np.integer)):
        raise TypeError(f"Argument n_cpu should have type int, got {type(n_cpu)}.")


def batch_mode(data: np.ndarray, b: np.ndarray) -> int:
    """`_mode` as produced by `_batch_template` (`Sparse.py:245-277`) for arrays whose batch
    axes have already been moved to the front."""
    assert b.ndim < 4 and data.ndim < 3
    if data.ndim == 1 and b.ndim == 1:
        return 0
    if data.ndim == 2 and b.ndim == 1:
        return 1
    if data.ndim == 1 and b.ndim == 2:
        return 2
    if data.ndim == 2 and b.ndim == 2:
        assert data.shape[0] == b.shape[0]
        return 3
    if data.ndim == 2 and b.ndim == 3:
        return 4
    raise NotImplementedError("Batching for this combination of axes is not implemented.")


def spsolve(data, b, *, solver_num: int, transpose: bool = False, n_cpu=0, _mode=None,
            state: SolverState = None):
    """`Sparse.py:231-236`: solve A x = b (or A^T x = b) for every batch item."""
    data, b = np.asarray(data), np.asarray(b)
    _check(data, b, n_cpu)
    if _mode is None:
        _mode = batch_mode(data, b)
    return (state or solver_state()).solve(data, b, solver_num, transpose, n_cpu, _mode)


def matvec(mat, vec, *, solver_num: int, transpose: bool = False, n_cpu=0, _mode=None,
           state: SolverState = None):
    """`Sparse.py:144-148`."""
    mat, vec = np.asarray(mat), np.asarray(vec)
    _check(mat, vec, n_cpu)
    if _mode is None:
        _mode = batch_mode(mat, vec)
    return (state or solver_state()).matvec(mat, vec, solver_num, transpose, n_cpu, _mode)


# ---- derivative rules (Sparse.py:162-176, 200-222); no complex conjugation anywhere ----
def spsolve_jvp(data, b, ddata, db, *, solver_num, transpose=False, state=None):
    """Tangent of x = A^-1 b: A^-1 db - A^-1 (dA x)."""
    kw = dict(solver_num=solver_num, transpose=transpose, state=state)
    x = spsolve(data, b, **kw)
    dx = np.zeros_like(x)
    if db is not None:
        dx = dx + spsolve(data, db, **kw)
    if ddata is not None:
        dx = dx - spsolve(data, matvec(ddata, x, **kw), **kw)
    return x, dx


def spsolve_vjp(data, x, ct, *, solver_num, transpose=False, state=None):
    """Transpose rule: ct_b = A^-T ct (`Sparse.py:211-222`), ct_data[e] = -ct_b[row e] x[col e]
    (`:168-176` composed with the JVP sign); for transpose=True rows and columns swap."""
    st = state or solver_state()
    ct_b = spsolve(data, ct, solver_num=solver_num, transpose=not transpose, state=st)
    row, col = st.patterns[solver_num][:, 0], st.patterns[solver_num][:, 1]
    if transpose:
        row, col = col, row
    ct_data = -ct_b[..., row] * x[..., col]
    return ct_data, ct_b


def matvec_vjp_matrix(ct, vec, *, solver_num, transpose=False, state=None, as_reference=False):
    """`_matvec_transpose`, matrix branch (`Sparse.py:172-176`): cotangent of the matrix
    entries of y = A v, `ct[row] * v[col]`; for y = A^T v rows and columns trade places.
    The reference's rule ignores `transpose` there and always returns `ct[row] * v[col]`
    (found by running it, tests/test_reference_pin.py::test_ad_rules_vs_reference); that only
    matters for second derivatives through transposed solves.  `as_reference=True`
    reproduces it."""
    st = state or solver_state()
    row, col = st.patterns[solver_num][:, 0], st.patterns[solver_num][:, 1]
    if transpose and not as_reference:
        row, col = col, row
    return ct[..., row] * vec[..., col]


def make_torch_spsolve(solver_num: int, state: SolverState = None):
    """torch.autograd.Function with the rules above.  torch's convention for complex
    cotangents is the conjugate of JAX's, hence the conj() pairs."""
    import torch

    class SpSolve(torch.autograd.Function):
        @staticmethod
        def forward(ctx, data, b):
            x = spsolve(data.detach().cpu().numpy(), b.detach().cpu().numpy(), solver_num=solver_num,
                        state=state)
            xt = torch.from_numpy(x)
            ctx.save_for_backward(data, xt)
            return xt

        @staticmethod
        def backward(ctx, g):
            data, x = ctx.saved_tensors
            ct = np.conj(g.detach().cpu().numpy())
            ct_data, ct_b = spsolve_vjp(data.detach().cpu().numpy(), x.numpy(), ct,
                                        solver_num=solver_num, state=state)
            return torch.from_numpy(np.conj(ct_data)), torch.from_numpy(np.conj(ct_b))

    return SpSolve
