"""`jpb_residual_noalloc` (include/jax_plate_b200.h): the residual / backward-error kernel with its
norm accumulators in a file-scope device array, for callers that cannot allocate.  No counterpart
in the reference.  Checked through the C-ABI against the existing entry point `jpb_matvec`
(r = b - op(A) x) and the backward error formed from it, on two calls in a row (the second with
different inputs and fewer systems: the shared scratch keeps the first call's contents).
The file name sorts last on purpose."""
import ctypes as C

import numpy as np
import pytest

from conftest import get_case

pytestmark = pytest.mark.gpu
TOL = 1e-10          # the tolerance of the parity tests on vectors (test_gpu_parity.TOL_U)


def _expected(st, num, transpose, mat, x, b):
    """r and omega from the EXISTING entry point jpb_matvec (mode 3: one matrix and one vector per system)."""
    from plate_inverse_problem_b200 import _lib
    y = np.empty_like(x)
    st.check(st.lib.jpb_matvec(st.h, num, 3, transpose, _lib.ptr(mat), _lib.ptr(x), x.shape[0], 0, _lib.ptr(y)))
    r = b - y
    return r, y


def _omega(case, transpose, mat, x, b, r):
    import scipy.sparse as sps
    n = case.native
    m1 = lambda v: np.abs(v.real) + np.abs(v.imag)
    out = np.empty(x.shape[0])
    for i in range(x.shape[0]):
        A = sps.csc_matrix((m1(mat[i]), n["indices"], n["indptr"]), shape=(case.N, case.N))
        t = (A.T if transpose else A) @ m1(x[i]) + m1(b[i])
        out[i] = np.linalg.norm(r[i]) / np.linalg.norm(t)
    return out


def test_residual_noalloc_matches_matvec_on_two_calls_in_a_row():
    import torch
    from plate_inverse_problem_b200 import _lib
    from plate_inverse_problem_b200.jax_plate_lib import InnerState
    case = get_case("small")
    state = InnerState()
    n = case.native
    state.add_mat(case.N, n["indices"], n["indptr"], n["indices_T"], n["indptr_T"], n["permutation"])
    num, st, vp = 0, state._st, C.c_void_p
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(7)
    # call 1: random x (residual of the size of |A||x|), A x; call 2: fewer systems, A^T x, x = the solutions
    # of the systems (tiny residuals, the regime the check is for) with the last one perturbed
    calls = [(np.array([60.0, 99.0, 101.0, 250.0, 480.0]), 0, False), (np.array([75.0, 333.0, 590.0]), 1, True)]
    for freqs, transpose, solved in calls:
        mat, _, _ = case.data(freqs)
        mat = np.ascontiguousarray(mat, dtype=np.complex128)
        F = freqs.size
        b = rng.standard_normal((F, case.N)) + 1j * rng.standard_normal((F, case.N))
        if solved:
            x = np.ascontiguousarray(state.solve(mat, b, num, bool(transpose), 0, 3))
            x[-1] *= 1.0 + 1e-6
        else:
            x = rng.standard_normal((F, case.N)) + 1j * rng.standard_normal((F, case.N))
        r_ref, _ = _expected(st, num, transpose, mat, x, b)
        om_ref = _omega(case, transpose, mat, x, b, r_ref)
        d = [torch.from_numpy(a).to(dev) for a in (mat, x, b)]
        d_r = torch.full((F, case.N), complex(np.nan, np.nan), dtype=torch.complex128, device=dev)
        d_om = torch.full((F,), np.nan, dtype=torch.float64, device=dev)
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        st.check(st.lib.jpb_residual_noalloc(st.h, num, transpose, F, vp(d[0].data_ptr()), vp(d[1].data_ptr()),
                                             vp(d[2].data_ptr()), vp(d_r.data_ptr()), vp(d_om.data_ptr()), None))
        assert torch.cuda.memory_allocated() == before
        r, om = d_r.cpu().numpy(), d_om.cpu().numpy()
        for i in range(F):      # per system, relative to the residual itself
            assert np.max(np.abs(r[i] - r_ref[i])) <= TOL * np.max(np.abs(r_ref[i])), (i, solved)
        assert np.all(np.isfinite(om)) and np.allclose(om, om_ref, rtol=TOL, atol=0.0), (om, om_ref)
        if solved:
            assert om[:-1].max() < 1e-14 and om[-1] > 1e3 * om[:-1].max()      # the perturbed system stands out
    # sizes beyond the static scratch are refused, with a message that names the limit
    limit = 4096
    rc = st.lib.jpb_residual_noalloc(st.h, num, 0, limit + 1, vp(d[0].data_ptr()), vp(d[1].data_ptr()),
                                     vp(d[2].data_ptr()), vp(d_r.data_ptr()), vp(d_om.data_ptr()), None)
    assert rc == _lib.JPB_EINVAL
    assert b"at most 4096" in st.lib.jpb_last_error(st.h)
