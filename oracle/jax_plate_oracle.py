"""CPU oracle for the jax_plate frequency-domain hot path.  TEST INFRASTRUCTURE ONLY.

This file is a numpy/scipy restatement of the reference algorithm.  It is imported only by
`tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline legs, always as the
checker / the timed CPU baseline — never by the product package
`plate_inverse_problem_b200` (which fails loudly when its CUDA library is missing).

PARITY STATUS: pinned to the reference's own code except for the arithmetic inside UMFPACK.
`tests/ref_shim.py` executes `/root/reference/source/jax_plate/*.py` unmodified (numpy `jax`
shim) on top of the reference's pybind11 module compiled from its sources (`oracle/Makefile`),
and `tests/golden/make_reference_golden.py` stores what that run produces
(`tests/golden/ref_*.npz`: index arrays, material transforms, FRF, losses, batching modes,
AD rules, `Compressor`, optimiser trajectories, the mid-plane symmetric branch);
`tests/test_reference_pin.py` and `tests/test_symm.py` hold the functions below to those
values.  What stays unpinned ("parity unpinned" at the level of rounding): UMFPACK itself
(SuiteSparse v5.x, `/root/reference/README.md:27`; not vendored, no pin file) is absent from
this image, as are jax and FreeFem++, so the reference's golden numbers that need a FreeFem++
mesh (`examples/cpu_benchmark.py:24,35`, `examples/basics.py:30-33`) cannot be regenerated.
The linear solves use scipy's SuperLU (`splu`, complex128) as a stand-in for
`umfpack_zi_numeric/solve` with the same call structure (one numeric factorisation + one
solve per batch item, `UMFPACK_A` / `UMFPACK_Aat` = plain transpose), cross-checked by
residuals, by an 80-bit refined solution and, at small N, by dense `numpy.linalg.solve`.

Each function cites the reference lines it restates (paths relative to /root/reference).
"""
from __future__ import annotations

import numpy as np
import scipy.sparse as sps
import scipy.sparse.linalg as spla


# ----------------------------------------------------------------------------------------
# index construction  (source/jax_plate/Sparse.py:46-116, Problem.py:317-345)
# ----------------------------------------------------------------------------------------
def find_permutation(arr1: np.ndarray, arr2: np.ndarray, max_val=None) -> np.ndarray:
    """`source/jax_plate/Sparse.py:46-85`, evaluated in exact (int64) integers."""
    assert arr1.shape == arr2.shape
    assert arr1.shape[1] == 2
    if max_val is None:
        max_val = int(np.max(arr1)) + 1
    a1 = arr1.astype(np.int64)
    a2 = arr2.astype(np.int64)
    u1 = a1[:, 0] + a1[:, 1] * np.int64(max_val)
    u2 = a2[:, 0] + a2[:, 1] * np.int64(max_val)
    is1 = np.argsort(u1)
    is2 = np.argsort(u2)
    inverse_is2 = is2[is2[is2]]
    res = np.arange(u1.size)[is1][inverse_is2]
    return np.array(res, dtype=arr1.dtype)


def create_symbolic_indices(N: int, indices: np.ndarray):
    """`source/jax_plate/Sparse.py:92-116` up to (not including) the native `add_mat` call.

    Returns ((row, col), native) where `native` holds exactly the arrays handed to
    `InnerState.add_mat` by `SolverState.add_mat` (`Sparse.py:33-34`): note `indptr_T`
    is `mat.indptr` (sic)."""
    nnz = indices.shape[0]
    data = np.linspace(1.0, 42.0, nnz, dtype=np.complex128)
    mat = sps.coo_matrix((data, (indices[:, 0], indices[:, 1])), (N, N),
                         dtype=np.complex128, copy=True).tocsc()
    mat_T = sps.coo_matrix((data, (indices[:, 1], indices[:, 0])), (N, N),
                           dtype=np.complex128, copy=True).tocsc()
    m1 = mat.tocoo()
    m1_T = mat_T.tocoo()
    coo_indices = np.vstack((m1.row, m1.col)).astype(indices.dtype).T
    coo_indices_T = np.vstack((m1_T.row, m1_T.col)).astype(indices.dtype).T
    indices_T = coo_indices[:, ::-1]
    perm = find_permutation(indices_T, coo_indices_T, N)
    native = dict(N=N, indices=mat.indices.copy(), indptr=mat.indptr.copy(),
                  indices_T=mat_T.indices.copy(), indptr_T=mat.indptr.copy(),
                  permutation=perm, pattern=coo_indices)
    return (m1.row, m1.col), native


def union_pattern_and_align(mats, N: int, index_dtype=np.int64):
    """`source/jax_plate/Problem.py:317-345` with int64 keys (the
    reference's int32 key `row + N*col` overflows for N > 46 340).

    Returns nz_idx [nnz,2], (row, col) canonical CSC order, native index dict,
    mats_aligned [K, nnz] float64."""
    def unwind(m):
        return m.row.astype(np.int64) + np.int64(N) * m.col.astype(np.int64)

    mats = [m.tocoo() for m in mats]
    # the reference folds np.union1d over the 26 key arrays (`Problem.py:320-326`); the sorted
    # set of unique keys it ends with is what one sort + adjacent-difference mask gives
    # (numpy 2.3's hash-based unique degrades badly on these structured keys at N ~ 1e5)
    pattern = np.sort(np.concatenate([unwind(m) for m in mats]))
    pattern = pattern[np.concatenate(([True], pattern[1:] != pattern[:-1]))] if pattern.size else pattern
    rows = pattern % N
    cols = pattern // N
    nz_idx = np.vstack((rows, cols)).T.astype(index_dtype)
    nz_mask, native = create_symbolic_indices(N, nz_idx)
    perm = find_permutation(nz_idx, np.vstack(nz_mask).T.astype(index_dtype), N)
    out = []
    for m in mats:
        idx = unwind(m)
        idx_perm = np.argsort(idx)
        mask = np.isin(pattern, idx[idx_perm], assume_unique=True)
        data = np.zeros(rows.size, dtype=np.float64)
        data[mask] = m.data[idx_perm]
        out.append(data[perm])
    return nz_idx, nz_mask, native, np.array(out, dtype=np.float64)


def umfpack_standin_lu(A):
    """Stand-in for `umfpack_zi_symbolic` + `umfpack_zi_numeric` with default Control
    (`InnerState.h:159-160, 243-245`).  For a structurally symmetric pattern with a
    zero-free diagonal UMFPACK's automatic strategy selects its *symmetric* strategy:
    fill-reducing ordering of A+A^T and diagonal pivots accepted down to a relative
    threshold of 1e-3 (SuiteSparse user guide, `UMFPACK_SYM_PIVOT_TOLERANCE`).  SuperLU's
    equivalent is MMD on A^T+A with `SymmetricMode` and the same diagonal threshold.
    (Plain `MMD_AT_PLUS_A` with SuperLU's default threshold 1.0 pivots off the diagonal on
    these matrices and fills catastrophically: 312x nnz(A) at N = 1e4.)"""
    return spla.splu(A, permc_spec="MMD_AT_PLUS_A", diag_pivot_thresh=1e-3,
                     options=dict(SymmetricMode=True))


class OrderedLU:
    """SuperLU with a caller-supplied symmetric fill-reducing ordering: factor P A P^T with
    `permc_spec="NATURAL"`, same diagonal-pivot threshold as `umfpack_standin_lu`.  Used where
    SuperLU's own MMD ordering does not finish (N ~ 1e5: SURVEY 8d asks for a CPU arm there);
    `perm[new] = old`, e.g. the nested-dissection ordering of the symbolic phase."""

    def __init__(self, A, perm, diag_pivot_thresh=1e-5):
        # Threshold: with 1e-3 (UMFPACK's sym_pivot_tolerance, used by umfpack_standin_lu)
        # SuperLU leaves the diagonal on the 210x80 plate and the fill triples (7.1e7 instead
        # of 2.0e7 entries, 26 s instead of 2.5-3 s per factorisation on one core here, residual
        # 4e-10 instead of 2e-11); with <= 1e-5 every pivot stays on the diagonal.  The faster,
        # more accurate setting is the one timed as CPU baseline (conservative for the GPU).
        self.perm = np.asarray(perm)
        Ap = sps.csc_matrix(A)[self.perm][:, self.perm].tocsc()
        self.lu = spla.splu(Ap, permc_spec="NATURAL", diag_pivot_thresh=diag_pivot_thresh,
                            options=dict(SymmetricMode=True))

    def solve(self, b, trans="N"):
        x = np.empty_like(b)
        x[self.perm] = self.lu.solve(np.ascontiguousarray(b[self.perm]), trans=trans)
        return x

    @property
    def nnz_lu(self):
        return int(self.lu.L.nnz + self.lu.U.nnz - self.perm.size)


class ExtendedLU:
    """A factorisation whose `solve` refines in 80-bit extended precision (`solve_extended`
    with a factorisation the caller already has): the answer is the exact solution of the
    assembled system to ~1e-16, which is what BOTH the stand-in LU and the CUDA path are
    judged against where cond(A) * eps exceeds the parity tolerances (next to resonances)."""

    def __init__(self, A, lu, iters=6):
        self.lu, self.iters = lu, iters
        Ac = sps.coo_matrix(A)
        self.r, self.c, self.v = Ac.row, Ac.col, Ac.data.astype(np.clongdouble)

    def solve(self, b, trans="N"):
        r_, c_ = (self.c, self.r) if trans == "T" else (self.r, self.c)
        bl = np.asarray(b).astype(np.clongdouble)
        x = self.lu.solve(np.asarray(b, dtype=np.complex128), trans=trans).astype(np.clongdouble)
        for _ in range(self.iters):
            Ax = np.zeros(bl.size, dtype=np.clongdouble)
            np.add.at(Ax, r_, self.v * x[c_])
            x = x + self.lu.solve((bl - Ax).astype(np.complex128), trans=trans).astype(np.clongdouble)
        return x.astype(np.complex128)


def solve_extended(A, b, trans=False, iters=6):
    """High-accuracy solve used to judge BOTH the stand-in LU and the CUDA path: SuperLU as
    preconditioner, residual accumulated in 80-bit extended precision (np.clongdouble), so
    the result is correct to ~1e-16 relative even when cond(A) ~ 1e8.  (UMFPACK's own
    `umfpack_zi_solve` refines in working precision; its forward error, like SuperLU's, is
    ~cond(A)*eps.)"""
    A = sps.csc_matrix(A)
    Aop = A.T.tocoo() if trans else A.tocoo()
    lu = umfpack_standin_lu(A)
    t = "T" if trans else "N"
    r_, c_, v_ = Aop.row, Aop.col, Aop.data.astype(np.clongdouble)
    bl = np.asarray(b).astype(np.clongdouble)
    x = lu.solve(np.asarray(b, dtype=np.complex128), trans=t).astype(np.clongdouble)
    for _ in range(iters):
        Ax = np.zeros(bl.size, dtype=np.clongdouble)
        np.add.at(Ax, r_, v_ * x[c_])
        r = bl - Ax
        x = x + lu.solve(r.astype(np.complex128), trans=t).astype(np.clongdouble)
    return x.astype(np.complex128)


# ----------------------------------------------------------------------------------------
# native module semantics  (source/jax_plate_lib/include/InnerState.h, csc_matvec.h)
# ----------------------------------------------------------------------------------------
def _mode_iter(mode: int, data: np.ndarray, b: np.ndarray):
    """Shape table of `InnerState.h:187-232` / SURVEY App. B.  Yields
    (out_index, data_row, b_row) and returns the result shape via .shape attr."""
    if mode == 0:
        return b.shape, [((), data, b)]
    if mode == 1:
        return (data.shape[0], b.shape[0]), [((i,), data[i], b) for i in range(data.shape[0])]
    if mode == 2:
        return b.shape, [((i,), data, b[i]) for i in range(b.shape[0])]
    if mode == 3:
        assert data.shape[0] == b.shape[0]
        return b.shape, [((i,), data[i], b[i]) for i in range(b.shape[0])]
    if mode == 4:
        assert b.shape[1] == data.shape[0]
        return b.shape, [((j, i), data[i], b[j, i])
                         for j in range(b.shape[0]) for i in range(b.shape[1])]
    raise RuntimeError("Invalid `mode` argument in InnerState::solve().")


class OracleInnerState:
    """Restatement of `jax_plate_lib.InnerState` (`InnerState.h:86-471`)."""

    def __init__(self):
        self.mats = []

    def add_mat(self, N, indices, indptr, indices_T, indptr_T, permutation, data=None):
        self.mats.append(dict(N=int(N), indices=np.array(indices), indptr=np.array(indptr),
                              indices_T=np.array(indices_T), indptr_T=np.array(indptr_T),
                              permutation=np.array(permutation)))
        return len(self.mats) - 1

    # InnerState.h:164-308 — numeric LU + solve for every batch item
    def solve(self, data, b, solver_num, transpose, n_cpu=1, mode=0):
        p = self.mats[solver_num]
        N = p["N"]
        shape, items = _mode_iter(int(mode), np.asarray(data), np.asarray(b))
        res = np.zeros(shape, dtype=np.result_type(data, b))
        for out_idx, d, rhs in items:
            A = sps.csc_matrix((d, p["indices"], p["indptr"]), shape=(N, N))
            lu = umfpack_standin_lu(A)
            res[out_idx] = lu.solve(np.ascontiguousarray(rhs), trans="T" if transpose else "N")
        return res

    # InnerState.h:310-470 + csc_matvec.h:31-66
    def matvec(self, mat, vec, solver_num, transpose, n_cpu=1, mode=0):
        p = self.mats[solver_num]
        shape, items = _mode_iter(int(mode), np.asarray(mat), np.asarray(vec))
        res = np.zeros(shape, dtype=np.result_type(mat, vec))
        for out_idx, d, x in items:
            if transpose:
                res[out_idx] = csc_matvec_transpose(p["N"], p["indptr_T"], p["indices_T"], d, x,
                                                    p["permutation"])
            else:
                res[out_idx] = csc_matvec(p["N"], p["indptr"], p["indices"], d, x)
        return res


def csc_matvec(n_col, Ap, Ai, Ax, Xx):
    """`csc_matvec.h:31-47`: Y[Ai[ii]] += Ax[ii] * X[j] for ii in column j."""
    col_of = np.repeat(np.arange(n_col), np.diff(Ap))
    y = np.zeros(n_col, dtype=np.result_type(Ax, Xx))
    np.add.at(y, Ai, Ax * Xx[col_of])
    return y


def csc_matvec_transpose(n_col, Ap, Ai, Ax, Xx, perm):
    """`csc_matvec.h:49-66`: Ap/Ai index the transposed matrix, values fetched via perm."""
    col_of = np.repeat(np.arange(n_col), np.diff(Ap))
    y = np.zeros(n_col, dtype=np.result_type(Ax, Xx))
    np.add.at(y, Ai, Ax[perm] * Xx[col_of])
    return y


# ----------------------------------------------------------------------------------------
# material transforms  (source/jax_plate/Material.py) — numpy, plus sympy Jacobians
# ----------------------------------------------------------------------------------------
def isotropic_abd(params, h):
    """`Material.py:373-391`."""
    E, G, beta = params
    nu = E / (2.0 * G) - 1.0
    A = E * h / (1 - nu ** 2)
    D = A * h ** 2 / 12.0
    arr = np.array([1.0, nu, 0.0, 1.0, 0.0, (1 - nu) / 2]) * (1 + 1j * beta)
    return A * arr, np.zeros_like(arr), D * arr


def orthotropic_abd(params, h):
    """`Material.py:455-482`."""
    E1, E2, G12, nu12, beta = params
    E_ratio = E2 / E1
    nu21 = E_ratio * nu12
    A11 = E1 * h / (1 - nu12 * nu21)
    A12 = nu21 * A11
    A22 = E2 / E1 * A11
    A66 = G12 * h
    D11 = E1 * h ** 3 / (12 * (1 - nu12 * nu21))
    D66 = G12 * h ** 3 / 12
    D12 = nu21 * D11
    D22 = D11 / E_ratio
    As = np.array([A11, A12, 0.0, A22, 0.0, A66]) * (1 + 1j * beta)
    Ds = np.array([D11, D12, 0.0, D22, 0.0, D66]) * (1 + 1j * beta)
    return As, np.zeros_like(As), Ds


def sol_q_to_abd(angles_deg, h):
    """Numeric evaluation of the classical-lamination matrices the reference derives with
    sympy (`Material.py:660-741`): equal-thickness plies, T(theta) with n = -sin."""
    n = len(angles_deg)
    zs = -h / 2 + np.arange(n + 1) * (h / n)
    MA = np.zeros((6, 6)); MB = np.zeros((6, 6)); MD = np.zeros((6, 6))
    pairs = [(0, 0), (0, 1), (0, 2), (1, 1), (1, 2), (2, 2)]          # 11,12,16,22,26,66
    # Q as a linear function of (Q11,Q12,Q16,Q22,Q26,Q66); the reference's symbolic Q has
    # Q16 = Q26 = 0 entries (Material.py:664-668) so those columns are zero.
    basis = []
    for k, (i, j) in enumerate(pairs):
        Q = np.zeros((3, 3))
        if k not in (2, 4):
            Q[i, j] = Q[j, i] = 1.0
        basis.append(Q)
    for a, z0, z1 in zip(angles_deg, zs[:-1], zs[1:]):
        t = a * np.pi / 180
        m, nn = np.cos(t), -np.sin(t)
        mn = m * nn
        T = np.array([[m * m, nn * nn, -2 * mn], [nn * nn, m * m, 2 * mn],
                      [mn, -mn, m * m - nn * nn]])
        for k, Q in enumerate(basis):
            QT = T @ Q @ T.T
            for r, (i, j) in enumerate(pairs):
                MA[r, k] += QT[i, j] * (z1 - z0)
                MB[r, k] += QT[i, j] * (z1 ** 2 - z0 ** 2) / 2
                MD[r, k] += QT[i, j] * (z1 ** 3 - z0 ** 3) / 3
    return MA, MB, MD


def sol_abd(params, h, angles_deg):
    """`Material.py:743-764`."""
    MA, MB, MD = sol_q_to_abd(angles_deg, h)
    E1, E2, G12, nu12, beta = params
    den = 1 - E2 / E1 * nu12 ** 2
    Q = np.array([E1 / den, nu12 * E2 / den, 0, E2 / den, 0, G12])
    f = (1 + 1j * beta)
    return (MA @ Q) * f, (MB @ Q) * f, (MD @ Q) * f


def make_transform(kind: str, h: float, angles=None):
    if kind == "isotropic":
        return lambda p, omega=None: isotropic_abd(p, h)
    if kind == "orthotropic":
        return lambda p, omega=None: orthotropic_abd(p, h)
    if kind == "sol":
        return lambda p, omega=None: sol_abd(p, h, angles)
    raise ValueError(kind)


def inertia_scalars(h, rho, acc_mass, acc_radius, acc_height):
    """`Problem.py:356-374`."""
    rho_corr = acc_mass / (np.pi * acc_radius ** 2) / acc_height
    I0 = h * rho
    I0Corr = acc_height * rho_corr
    I2 = rho * h ** 3 / 12
    I2Corr = rho_corr / 3 * ((h / 2 + acc_height) ** 3 - h ** 3 / 8)
    return I0, I0Corr, I2, I2Corr


def coefficients(freqs, params, transform, inertia):
    """SURVEY App. C restating `Problem.py:437-449`: coef [F,26] and rhs scale s [F]."""
    I0, I0Corr, I2, I2Corr = inertia
    freqs = np.atleast_1d(np.asarray(freqs, dtype=np.float64))
    coef = np.zeros((freqs.size, 26), dtype=np.complex128)
    s = np.zeros(freqs.size, dtype=np.complex128)
    for i, f in enumerate(freqs):
        omega = 2 * np.pi * f
        A, B, D = transform(params, omega)
        coef[i, 0:6], coef[i, 6:12], coef[i, 12:18] = A, B, D
        coef[i, [18, 20, 22]] = -omega ** 2 * I0
        coef[i, [19, 21, 23]] = -omega ** 2 * I0Corr
        coef[i, 24] = -omega ** 2 * I2
        coef[i, 25] = -omega ** 2 * I2Corr
        s[i] = (D[0] + 2 * D[1] + 4 * D[2] + 4 * D[4] + 4 * D[5] + D[3]
                - omega ** 2 * (I0 + I0Corr + I2 + I2Corr))
    return coef, s


def coefficient_jacobian(freqs, params, transform, inertia, rel_step=1e-3):
    """d coef / d p and d s / d p.  The coefficient functions are holomorphic (rational)
    in the real parameters, so a 4th-order central difference with a 1e-3 relative
    step is accurate to ~1e-12 relative (truncation 1e-12, round-off 1e-13)."""
    params = np.asarray(params, dtype=np.float64)
    P = params.size
    c0, s0 = coefficients(freqs, params, transform, inertia)
    dc = np.zeros(c0.shape + (P,), dtype=np.complex128)
    ds = np.zeros(s0.shape + (P,), dtype=np.complex128)
    for j in range(P):
        hstep = rel_step * max(abs(params[j]), 1e-300)
        acc_c, acc_s = 0, 0
        for k, w in ((-2, 1 / 12), (-1, -8 / 12), (1, 8 / 12), (2, -1 / 12)):
            p = params.copy()
            p[j] += k * hstep
            c, s = coefficients(freqs, p, transform, inertia)
            acc_c = acc_c + w * c
            acc_s = acc_s + w * s
        dc[..., j] = acc_c / hstep
        ds[..., j] = acc_s / hstep
    # s = D0 + 2 D1 + 4 D2 + D3 + 4 D4 + 4 D5 - omega^2 (...) is linear in coef[12:18]; the
    # direct difference quotient of s cancels the (parameter-independent, ~1e7 times larger)
    # inertia term only to round-off, so use the exact linear relation instead.
    comb = np.zeros(26)
    comb[12:18] = [1.0, 2.0, 4.0, 1.0, 4.0, 4.0]
    ds = np.einsum("k,fkp->fp", comb, dc)
    return dc, ds


# ----------------------------------------------------------------------------------------
# assembly, read-out, losses, adjoint gradient (Problem.py:437-477, 948-975; Sparse.py AD)
# ----------------------------------------------------------------------------------------
def assemble(coef_row, mats):
    """`Problem.py:440-445` for one frequency: same grouping of operations."""
    m = mats
    c = coef_row
    mat = np.zeros(m.shape[1], dtype=np.complex128)
    # -omega^2 (I0 (m18+m20+m22) + I0Corr (m19+m21+m23) + I2 m24 + I2Corr m25)
    mat += (c[18] * (m[18] + m[20] + m[22]) + c[19] * (m[19] + m[21] + m[23])
            + c[24] * m[24] + c[25] * m[25])
    for i in range(6):
        mat += c[i] * m[i] + c[i + 6] * m[i + 6] + c[i + 12] * m[i + 12]
    return mat


def readout_vectors(plate, acc_h, acc_h_eff):
    """Row-mean reduction of `Problem.py:454-462`: u_bar = q_u . sol etc.; Q [3, N] real."""
    Lh, Mh = plate.Lh_size, plate.Mh_size
    N = 2 * Lh + Mh
    mL = np.asarray(plate.interp_mat_Lh).mean(axis=0)
    mW = np.asarray(plate.interp_mat).mean(axis=0)
    mWx = np.asarray(plate.interp_mat_Wx).mean(axis=0)
    mWy = np.asarray(plate.interp_mat_Wy).mean(axis=0)
    Q = np.zeros((3, N))
    Q[0, 0:Lh] = mL
    Q[0, 2 * Lh:] = -acc_h_eff * acc_h * mWx
    Q[1, Lh:2 * Lh] = mL
    Q[1, 2 * Lh:] = -acc_h_eff * acc_h * mWy
    Q[2, 2 * Lh:] = mW
    return Q


def readout(sol, plate, acc_h, acc_h_eff, acc_ts):
    """`Problem.py:454-477` literally (dense interp matrices, means), one solution vector."""
    Lh = plate.Lh_size
    u_sol = plate.interp_mat_Lh @ sol[:Lh]
    v_sol = plate.interp_mat_Lh @ sol[Lh:2 * Lh]
    w_sol = plate.interp_mat @ sol[2 * Lh:]
    wx_sol = plate.interp_mat_Wx @ sol[2 * Lh:]
    wy_sol = plate.interp_mat_Wy @ sol[2 * Lh:]
    u = np.mean(u_sol - acc_h_eff * acc_h * wx_sol)
    v = np.mean(v_sol - acc_h_eff * acc_h * wy_sol)
    w = np.mean(w_sol)
    return np.sqrt((np.abs(u) * acc_ts) ** 2 + (np.abs(v) * acc_ts) ** 2 + np.abs(w) ** 2)


def loss_value(fr, ref, func_type):
    """`Problem.py:948-975`."""
    if func_type == "MSE":
        return np.mean(np.abs(fr - ref) ** 2)
    if func_type == "RMSE":
        return np.mean(np.abs((fr - ref) / ref) ** 2)
    if func_type == "MSE_AFC":
        return np.mean((np.abs(fr) - np.abs(ref)) ** 2)
    if func_type == "MSE_LOG_AFC":
        return np.mean((np.log(np.abs(fr)) - np.log(np.abs(ref))) ** 2)
    raise ValueError(f'Function type "{func_type}" is not supported!')


def loss_dfr(fr, ref, func_type):
    """dL/dfr for real fr (unsymm path read-out is real)."""
    F = fr.size
    if func_type == "MSE":
        return 2 * (fr - ref) / F
    if func_type == "RMSE":
        return 2 * (fr - ref) / ref ** 2 / F
    if func_type == "MSE_AFC":
        return 2 * (np.abs(fr) - np.abs(ref)) * np.sign(fr) / F
    if func_type == "MSE_LOG_AFC":
        return 2 * (np.log(np.abs(fr)) - np.log(np.abs(ref))) / fr / F
    raise ValueError(func_type)


class OracleProblem:
    """Forward / adjoint / gradient evaluation exactly as the reference composes it
    (SURVEY §3.2, §3.3, App. D): per frequency one numeric LU for the forward solve and a
    second one for the transposed (adjoint) solve (`InnerState.h:276-288` called twice)."""

    def __init__(self, plate, transform, inertia, acc_h, acc_h_eff, acc_ts, ordering=None):
        self.ordering = ordering      # None: SuperLU's MMD; else perm[new] = old for OrderedLU
        self.exact = False            # True: every solve refined in extended precision (ExtendedLU)
        self.plate = plate
        self.N = plate.N
        self.nz_idx, (self.row, self.col), self.native, self.mats = \
            union_pattern_and_align(plate.mats, plate.N)
        self.transform = transform
        self.inertia = inertia
        self.acc = (acc_h, acc_h_eff, acc_ts)
        self.Q = readout_vectors(plate, acc_h, acc_h_eff)
        self.state = OracleInnerState()
        n = self.native
        self.solver_num = self.state.add_mat(n["N"], n["indices"], n["indptr"], n["indices_T"],
                                             n["indptr_T"], n["permutation"])

    def factor(self, A):
        lu = umfpack_standin_lu(A) if self.ordering is None else OrderedLU(A, self.ordering)
        return ExtendedLU(A, lu) if self.exact else lu

    def matrix(self, coef_row):
        n = self.native
        return sps.csc_matrix((assemble(coef_row, self.mats), n["indices"], n["indptr"]),
                              shape=(self.N, self.N))

    def forward(self, freqs, params, return_u=False, reuse_factor=False):
        coef, s = coefficients(freqs, params, self.transform, self.inertia)
        acc_h, acc_h_eff, ts = self.acc
        U = np.zeros((coef.shape[0], self.N), dtype=np.complex128)
        fr = np.zeros(coef.shape[0])
        self._lus = []
        for i in range(coef.shape[0]):
            lu = self.factor(self.matrix(coef[i]))
            U[i] = lu.solve(s[i] * self.plate.rhs_vec.astype(np.complex128))
            y = self.Q @ U[i]
            fr[i] = np.sqrt(ts ** 2 * abs(y[0]) ** 2 + ts ** 2 * abs(y[1]) ** 2 + abs(y[2]) ** 2)
            if reuse_factor:
                self._lus.append(lu)
        return (fr, U) if return_u else fr

    def fr_jacobian(self, freqs, params):
        """J[f, p] = d fr_f / d p_p, the matrix `jax.jacobian` of `getFRFunction` would give
        (the Gauss-Newton model of `Optimizers.py:139-144` is built from it): the adjoint
        formula of `loss_and_grad` with the cotangent of fr itself (dL/dfr = 1 per row)."""
        freqs = np.atleast_1d(np.asarray(freqs, dtype=np.float64))
        J = np.zeros((freqs.size, len(params)))
        fr = np.zeros(freqs.size)
        for i, f in enumerate(freqs):
            # MSE against ref = 0 with F = 1: L = fr^2, dL/dp = 2 fr dfr/dp
            L, g, aux = self.loss_and_grad(np.array([f]), params, np.zeros(1), "MSE", reuse_factor=True)
            fr[i] = aux["fr"][0]
            J[i] = g / (2.0 * fr[i])
        return fr, J

    def loss_and_grad(self, freqs, params, ref, func_type, reuse_factor=False):
        """JAX convention (App. D): u_bar = dL/dRe(u) - i dL/dIm(u), lambda = A^-T u_bar,
        c_bar_k = -lambda^T M_k u, s_bar = lambda^T rhs_vec,
        dL/dp = Re(sum c_bar_k dc_k/dp + s_bar ds/dp)."""
        coef, s = coefficients(freqs, params, self.transform, self.inertia)
        dc, ds = coefficient_jacobian(freqs, params, self.transform, self.inertia)
        fr, U = self.forward(freqs, params, return_u=True, reuse_factor=reuse_factor)
        acc_h, acc_h_eff, ts = self.acc
        L = loss_value(fr, ref, func_type)
        dfr = loss_dfr(fr, ref, func_type)
        wts = np.array([ts ** 2, ts ** 2, 1.0])
        grad = np.zeros(len(params))
        n = self.native
        ct_coef = np.zeros_like(coef)
        ct_s = np.zeros_like(s)
        for i in range(coef.shape[0]):
            y = self.Q @ U[i]
            # fr = sqrt(sum w_r |y_r|^2): y_bar_r = dfr * w_r * conj(y_r) / fr
            ybar = dfr[i] * wts * np.conj(y) / fr[i]
            ubar = ybar @ self.Q                       # [N] complex
            if reuse_factor:
                lam = self._lus[i].solve(ubar, trans="T")
            else:
                lu = self.factor(self.matrix(coef[i]))
                lam = lu.solve(ubar, trans="T")
            ctA = -lam[self.row] * U[i][self.col]      # Sparse.py:173-176 (with the JVP sign)
            ct_coef[i] = self.mats @ ctA
            ct_s[i] = lam @ self.plate.rhs_vec
            grad += np.real(ct_coef[i] @ dc[i] + ct_s[i] * ds[i])
        return L, grad, dict(fr=fr, U=U, ct_coef=ct_coef, ct_s=ct_s, coef=coef, s=s)


# ----------------------------------------------------------------------------------------
# mid-plane symmetric (bending-only) path  (Problem.py:228-308 set-up, :391-417 body)
# ----------------------------------------------------------------------------------------
def isotropic_d(params, h):
    """`Material.py:394-404`."""
    E, G, beta = params
    nu = E / (2.0 * G) - 1.0
    D = E * h ** 3 / (12.0 * (1.0 - nu ** 2))
    return np.array([D, nu * D, 0.0, D, 0.0, D * (1 - nu) / 2]) * (1 + 1j * beta)


def make_d_transform(kind: str, h: float, angles=None):
    """`get_D_transform` of the reference's materials (`Material.py:394-406, 485-504, 766-784`)."""
    if kind == "isotropic":
        return lambda p, omega=None: isotropic_d(p, h)
    if kind == "orthotropic":
        return lambda p, omega=None: orthotropic_abd(p, h)[2]
    if kind == "sol":
        return lambda p, omega=None: sol_abd(p, h, angles)[2]
    raise ValueError(kind)


class OracleSymmProblem:
    """The symmetric branch as the reference composes it: on the free DOFs
    `A = -w^2 MInertia + sum_k D_k Ks_k / 2e`, `b = -w^2 fInertia + sum_k D_k fKs_k / 2e`,
    `fr = c0 + c . u` (complex), the four losses on it, and the gradient by the adjoint
    (JAX convention, as `OracleProblem.loss_and_grad`): with u_bar = dL/dRe(fr) - i dL/dIm(fr) times c,
    lambda = A^-T u_bar, dL/dD_k = lambda^T (fKs_k - Ks_k u) / 2e."""

    def __init__(self, ff: dict, d_transform, rho: float, h: float):
        self.e = e = h / 2.0
        self.Ks = [sps.csc_matrix(K) / 2.0 / e for K in ff["Ks"]]                 # Problem.py:419
        self.fKs = np.asarray(ff["fKs"], dtype=np.float64) / 2.0 / e               # :420
        self.MInertia = (rho * (sps.csc_matrix(ff["M"]) + 1.0 / 3.0 * e ** 2 * sps.csc_matrix(ff["L"]))).tocsc()
        self.fInertia = rho * (np.asarray(ff["fM"]) + 1.0 / 3.0 * e ** 2 * np.asarray(ff["fL"]))
        self.c = np.asarray(ff["interpolation_vector"], dtype=np.float64)
        self.c0 = float(ff["interpolation_value_from_bc"])
        self.transform = d_transform
        self.n = self.MInertia.shape[0]
        self.exact = False

    def system(self, f, params):
        omega = 2.0 * np.pi * f
        D = self.transform(np.asarray(params, dtype=np.float64), omega)
        A = -(omega ** 2) * self.MInertia.astype(np.complex128)
        b = -(omega ** 2) * self.fInertia.astype(np.complex128)
        for k in range(6):
            A = A + D[k] * self.Ks[k]
            b = b + D[k] * self.fKs[k]
        return sps.csc_matrix(A), b, D

    def factor(self, A):
        lu = umfpack_standin_lu(A)
        return ExtendedLU(A, lu) if self.exact else lu

    def forward(self, freqs, params, return_u=False):
        freqs = np.atleast_1d(np.asarray(freqs, dtype=np.float64))
        fr = np.zeros(freqs.size, dtype=np.complex128)
        U = np.zeros((freqs.size, self.n), dtype=np.complex128)
        for i, f in enumerate(freqs):
            A, b, _ = self.system(f, params)
            U[i] = self.factor(A).solve(b)
            fr[i] = self.c0 + self.c @ U[i]
        return (fr, U) if return_u else fr

    def d_jacobian(self, params, rel_step=1e-3):
        """d D_k / d p_j by the 4th-order central difference of `coefficient_jacobian`."""
        params = np.asarray(params, dtype=np.float64)
        dD = np.zeros((6, params.size), dtype=np.complex128)
        for j in range(params.size):
            hstep = rel_step * max(abs(params[j]), 1e-300)
            for k, w in ((-2, 1 / 12), (-1, -8 / 12), (1, 8 / 12), (2, -1 / 12)):
                p = params.copy()
                p[j] += k * hstep
                dD[:, j] += w * self.transform(p, None) / hstep
        return dD

    def loss_and_grad(self, freqs, params, ref, func_type):
        freqs = np.atleast_1d(np.asarray(freqs, dtype=np.float64))
        ref = np.asarray(ref)
        F = freqs.size
        dD = self.d_jacobian(params)
        grad = np.zeros(len(params))
        fr = np.zeros(F, dtype=np.complex128)
        for i, f in enumerate(freqs):
            A, b, D = self.system(f, params)
            lu = self.factor(A)
            u = lu.solve(b)
            fr[i] = self.c0 + self.c @ u
            if func_type == "MSE":
                ybar = 2 * np.conj(fr[i] - ref[i]) / F
            elif func_type == "RMSE":
                ybar = 2 * np.conj(fr[i] - ref[i]) / abs(ref[i]) ** 2 / F
            else:
                a, ar = abs(fr[i]), abs(ref[i])
                dfr = loss_dfr(np.array([a]), np.array([ar]), func_type)[0] * 1 / F
                ybar = dfr * np.conj(fr[i]) / a
            lam = lu.solve(ybar * self.c.astype(np.complex128), trans="T")
            ctD = np.array([lam @ (self.fKs[k] - self.Ks[k] @ u) for k in range(6)])
            grad += np.real(ctD @ dD)
        return loss_value(fr, ref, func_type), grad, dict(fr=fr)
