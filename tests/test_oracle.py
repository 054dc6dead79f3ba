"""CPU tests pinning the oracle: against the committed golden vectors, dense LAPACK,
finite differences, and the one piece of the reference that compiles here (csc_matvec.h,
built into oracle/_ref by oracle/Makefile)."""
import ctypes
import os

import numpy as np
import pytest
import scipy.sparse as sps

from conftest import PARAMS, ROOT, relerr
from oracle import jax_plate_oracle as orc

GOLD = np.load(os.path.join(ROOT, "tests", "golden", "tiny_sol.npz"))


def test_golden_indices_bit_exact(tiny):
    n = tiny.native
    for k in ("indices", "indptr", "indices_T", "permutation"):
        assert np.array_equal(GOLD[k], n[k]), k
    assert np.array_equal(GOLD["row"], tiny.op.row) and np.array_equal(GOLD["col"], tiny.op.col)
    assert np.allclose(GOLD["mats_checksum"], [tiny.op.mats.sum(), np.abs(tiny.op.mats).sum()], rtol=1e-13)


def test_golden_forward_and_gradients(tiny):
    freqs = GOLD["freqs"]
    fr, U = tiny.op.forward(freqs, PARAMS, return_u=True)
    assert relerr(fr, GOLD["fr"]) < 1e-10
    assert max(relerr(U[i], GOLD["U"][i]) for i in range(freqs.size)) < 1e-9   # SuperLU vs extended
    for lt in ("MSE", "RMSE", "MSE_AFC", "MSE_LOG_AFC"):
        L, g, _ = tiny.op.loss_and_grad(freqs, PARAMS, GOLD["ref"], lt)
        assert abs(L - GOLD[f"loss_{lt}"]) <= 1e-9 * abs(GOLD[f"loss_{lt}"])
        assert relerr(g, GOLD[f"grad_{lt}"]) < 1e-8


def test_oracle_solve_against_dense_lapack(tiny):
    n = tiny.native
    st = orc.OracleInnerState()
    num = st.add_mat(n["N"], n["indices"], n["indptr"], n["indices_T"], n["indptr_T"], n["permutation"])
    data, _, _ = tiny.data(np.array([40.0, 320.0, 600.0]))
    rng = np.random.default_rng(0)
    b = rng.standard_normal((3, tiny.N)) + 1j * rng.standard_normal((3, tiny.N))
    for tr in (False, True):
        x = st.solve(data, b, num, tr, 1, 3)
        for i in range(3):
            A = sps.csc_matrix((data[i], n["indices"], n["indptr"]), shape=(tiny.N,) * 2).toarray()
            xd = np.linalg.solve(A.T if tr else A, b[i])
            xe = orc.solve_extended(A, b[i], trans=tr)
            assert relerr(x[i], xe) < 1e-9 and relerr(xd, xe) < 1e-9


def test_oracle_mode_table_shapes(tiny):
    n = tiny.native
    st = orc.OracleInnerState()
    num = st.add_mat(n["N"], n["indices"], n["indptr"], n["indices_T"], n["indptr_T"], n["permutation"])
    data, _, _ = tiny.data(np.array([100.0, 200.0]))
    N = tiny.N
    b = np.ones((3, 2, N), dtype=np.complex128)
    assert st.solve(data[0], b[0, 0], num, False, 1, 0).shape == (N,)
    assert st.solve(data, b[0, 0], num, False, 1, 1).shape == (2, N)
    assert st.solve(data[0], b[:, 0], num, False, 1, 2).shape == (3, N)
    assert st.solve(data, b[0], num, False, 1, 3).shape == (2, N)
    assert st.solve(data, b, num, False, 1, 4).shape == (3, 2, N)
    with pytest.raises(RuntimeError):
        st.solve(data, b, num, False, 1, 7)


def _ref_lib():
    path = os.path.join(ROOT, "oracle", "_ref", "libref_csc_matvec.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref not built (needs /root/reference; run `make -C oracle`)")
    return ctypes.CDLL(path)


def test_csc_matvec_restatement_against_reference_build(small):
    """oracle.csc_matvec{,_transpose} vs the reference's own templates compiled from
    /root/reference/source/jax_plate_lib/include/csc_matvec.h."""
    lib = _ref_lib()
    n = small.native
    N = small.N
    data, _, _ = small.data(np.array([123.0]))
    d = np.ascontiguousarray(data[0])
    rng = np.random.default_rng(3)
    x = rng.standard_normal(N) + 1j * rng.standard_normal(N)
    i32 = lambda a: np.ascontiguousarray(a, dtype=np.int32)
    vp = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    Ap, Ai, AiT, perm = i32(n["indptr"]), i32(n["indices"]), i32(n["indices_T"]), i32(n["permutation"])
    y = np.zeros(N, dtype=np.complex128)
    lib.ref_csc_matvec_z32(ctypes.c_int32(N), vp(Ap), vp(Ai), vp(d), vp(x), vp(y))
    assert relerr(orc.csc_matvec(N, Ap, Ai, d, x), y) < 1e-14
    A = sps.csc_matrix((d, Ai, Ap), shape=(N, N))
    assert relerr(A @ x, y) < 1e-14
    yt = np.zeros(N, dtype=np.complex128)
    lib.ref_csc_matvec_transpose_z32(ctypes.c_int32(N), vp(Ap), vp(AiT), vp(d), vp(x), vp(perm), vp(yt))
    assert relerr(orc.csc_matvec_transpose(N, Ap, AiT, d, x, perm), yt) < 1e-14
    assert relerr(A.T @ x, yt) < 1e-14    # structurally symmetric pattern (App. E)


def test_gradient_against_finite_differences(tiny):
    freqs = np.linspace(60.0, 500.0, 4)
    ref = tiny.op.forward(freqs, PARAMS * np.array([1.1, 1.1, 1.2, 1.0, 1.5]))
    for lt in ("MSE", "MSE_LOG_AFC"):
        L, g, _ = tiny.op.loss_and_grad(freqs, PARAMS, ref, lt)
        for j in range(PARAMS.size):
            hstep = 2e-3 * PARAMS[j]
            fd = 0.0
            for k, w in ((-2, 1 / 12), (-1, -8 / 12), (1, 8 / 12), (2, -1 / 12)):   # 4th-order stencil
                p = PARAMS.copy()
                p[j] += k * hstep
                fd += w * orc.loss_value(tiny.op.forward(freqs, p), ref, lt)
            fd /= hstep
            assert abs(fd - g[j]) <= 1e-5 * abs(g[j]) + 1e-9 * np.abs(g).max(), (lt, j, fd, g[j])


def test_reuse_factor_variant_equals_reference_variant(tiny):
    freqs = GOLD["freqs"]
    a = tiny.op.loss_and_grad(freqs, PARAMS, GOLD["ref"], "MSE")
    b = tiny.op.loss_and_grad(freqs, PARAMS, GOLD["ref"], "MSE", reuse_factor=True)
    assert a[0] == b[0] and relerr(a[1], b[1]) < 1e-12
