"""GPU parity of the mid-plane symmetric (bending-only) path: `Problem(geometry, material)` without
an accelerometer -> ProblemSymm.SymmetricProblem -> the fused CUDA path on the bordered system,
against values produced by the reference's own branch (tests/golden/ref_symm.npz) and against the
oracle restatement (tolerances of BASELINE.json: 1e-10 on the FRF, 1e-8 on loss / gradient)."""
import numpy as np
import pytest

from test_symm import CASES, H, REF, oracle_case, product_material

pytestmark = pytest.mark.gpu
TOL_FR, TOL_LG = 1e-10, 1e-8
LOSSES = ("MSE", "RMSE", "MSE_AFC", "MSE_LOG_AFC")


def product(name, ff):
    from plate_inverse_problem_b200.Problem import Geometry, Problem
    from plate_inverse_problem_b200.ProblemSymm import SymmetricProblem
    prob = Problem(Geometry(H, ff), product_material(name))            # no accelerometer + is_mps -> symmetric path (Problem.py:228)
    assert isinstance(prob, SymmetricProblem) and prob.mat_size == ff["M"].shape[0] + 1
    return prob


@pytest.mark.parametrize("name", list(CASES))
def test_symmetric_path_against_reference_values_and_oracle(name):
    ff, op = oracle_case(name)
    prob = product(name, ff)
    p, freqs, ref = REF[f"{name}_params"], REF[f"{name}_freqs"], REF[f"{name}_ref"]
    fr = prob.solveForward(freqs)
    assert fr.dtype == np.complex128
    op.exact = True                                   # yardstick: the stand-in refined in extended precision
    fr_x = op.forward(freqs, p)
    op.exact = False
    # cond(A) ~ 1e10 next to the resonances: every LU carries a forward error ~1e-10 there.  The CUDA path is
    # held to the tolerance against the exact answer or to the distance of the two CPU LUs from it (SuperLU
    # stand-in; the dense-LU stand-in inside the reference's own run), as in test_gpu_parity
    err = lambda a: float(np.max(np.abs(a - fr_x) / np.abs(fr_x)))
    yard = max(err(op.forward(freqs, p)), err(REF[f"{name}_fr"]))
    assert err(fr) <= max(TOL_FR, 2 * yard), (err(fr), yard)
    assert np.max(np.abs(fr - REF[f"{name}_fr"]) / np.abs(fr_x)) <= 1e-9      # the reference's own numbers
    U = prob.solveFree(freqs[:2])
    _, U_o = op.forward(freqs[:2], p, return_u=True)
    assert np.max(np.abs(U - U_o)) <= 1e-8 * np.max(np.abs(U_o))      # the DOF mapping; accuracy is judged on fr above
    for lt in LOSSES:
        loss = prob.getLossFunction(freqs, ref, lt)
        L, g = loss.value_and_grad(p)
        L_o, g_o, _ = op.loss_and_grad(freqs, p, ref, lt)
        assert abs(L - L_o) <= TOL_LG * abs(L_o), lt
        assert np.max(np.abs(g - g_o)) <= TOL_LG * np.max(np.abs(g_o)), lt
        assert abs(L - float(REF[f"{name}_loss_{lt}"])) <= TOL_LG * abs(L_o), lt
        assert abs(loss(p) - L) <= 1e-12 * abs(L)
        fd = REF[f"{name}_fd_grad_{lt}"]              # central differences of the REFERENCE's loss
        assert np.max(np.abs(g - fd)) <= 2e-5 * np.max(np.abs(fd)), lt
    # scaled parameters (Problem.py:945-946)
    sc = p.copy()
    Ls, gs = prob.getLossFunction(freqs, ref, "MSE", scaling_params=sc).value_and_grad(np.ones(p.size))
    L1, g1 = prob.getLossFunction(freqs, ref, "MSE").value_and_grad(p)
    assert abs(Ls - L1) <= 1e-12 * abs(L1) and np.max(np.abs(gs - g1 * sc)) <= 1e-10 * np.max(np.abs(g1 * sc))


def test_symmetric_path_candidates_jacobian_and_a_descent_step():
    name = "iso_tiny"
    ff, op = oracle_case(name)
    prob = product(name, ff)
    p, freqs, ref = REF[f"{name}_params"], REF[f"{name}_freqs"], REF[f"{name}_ref"]
    cands = np.stack([p, p * np.array([1.03, 0.98, 1.4])])
    for lt in ("MSE", "MSE_LOG_AFC"):
        loss = prob.getLossFunction(freqs, ref, lt)
        Lc, gc = loss.value_and_grad(cands)
        for c in range(2):
            L1, g1 = loss.value_and_grad(cands[c])
            assert abs(Lc[c] - L1) <= 1e-12 * abs(L1) and np.max(np.abs(gc[c] - g1)) <= 1e-10 * np.max(np.abs(g1))
    # Jacobian of the amplitude |fr| against a 4th-order difference of the product's own forward sweep
    a, J = prob.getJacobianFunction(freqs)(p)
    assert np.max(np.abs(a - np.abs(prob.solveForward(freqs)))) <= 1e-12 * np.max(a)
    for j in (0, 2):
        h = 1e-3 * p[j]
        fd = sum(w * np.abs(prob.solveForward(freqs, p + k * h * np.eye(p.size)[j]))
                 for k, w in ((-2, 1 / 12), (-1, -8 / 12), (1, 8 / 12), (2, -1 / 12))) / h
        assert np.max(np.abs(J[:, j] - fd)) <= 1e-6 * np.max(np.abs(fd))
    model = prob.getGaussNewtonModel(freqs, ref, "MSE_LOG_AFC")
    L, g, Hm = model(p)
    L1, g1 = prob.getLossFunction(freqs, ref, "MSE_LOG_AFC").value_and_grad(p)
    assert abs(L - L1) <= 1e-10 * abs(L1) and np.max(np.abs(g - g1)) <= 1e-8 * np.max(np.abs(g1))
    assert np.all(np.linalg.eigvalsh(Hm) >= -1e-12 * np.abs(Hm).max())
    # the reference's optimisers run on it (Optimizers.optimize_gd, relative step): the loss goes down
    from plate_inverse_problem_b200 import Optimizers
    start = p * np.array([1.05, 1.02, 1.5])
    lf = prob.getLossFunction(freqs, ref, "MSE_LOG_AFC", scaling_params=start)
    res = Optimizers.optimize_gd(lf, np.ones(3), N_steps=5, h=0.05)
    assert res.f_history[-1] < res.f_history[0]


def test_complex_losses_need_a_single_readout_row():
    """C-ABI loss types 5 / 6 (complex read-out) are refused on a pattern with three read-out rows."""
    from conftest import PARAMS, get_case
    from plate_inverse_problem_b200 import _lib
    prob = get_case("tiny").problem()
    freqs = np.array([100.0, 200.0])
    coef, s, _, _ = prob.coefficients(freqs, PARAMS)
    coef, s = np.ascontiguousarray(coef), np.ascontiguousarray(s)
    ref = np.ones(2, dtype=np.complex128).view(np.float64)
    fr, loss = np.zeros(2), np.zeros(1)
    ct, cts = np.zeros((2, coef.shape[1]), dtype=np.complex128), np.zeros(2, dtype=np.complex128)
    st = prob._st
    for lt in (5, 6, 7, -1):
        rc = st.lib.jpb_eval_loss_grad(st.h, prob.solver_num, 2, _lib.ptr(coef), _lib.ptr(s), _lib.ptr(ref),
                                       _lib.ptr(prob.weights), lt, 2, _lib.ptr(fr), _lib.ptr(loss), _lib.ptr(ct),
                                       _lib.ptr(cts), None)
        assert rc == _lib.JPB_EINVAL
        msg = st.lib.jpb_last_error(st.h)
        assert (b"single read-out row" in msg) if lt in (5, 6) else (b"unknown loss type" in msg)
