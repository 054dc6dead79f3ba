"""Mid-plane symmetric (bending-only) path on the CPU: the oracle restatement against values
produced by the reference's own branch (`tests/golden/ref_symm.npz`, made by
`tests/golden/make_reference_golden.py` through `ref_shim.reference_symm_problem`), and the host
logic of `ProblemSymm` (bordered components, coefficient functions) against the oracle."""
import os

import numpy as np
import pytest
import scipy.sparse as sps

from conftest import ROOT
from oracle import jax_plate_oracle as orc

REF = np.load(os.path.join(ROOT, "tests", "golden", "ref_symm.npz"))
H = 2e-3
CASES = {"iso_tiny": (6, 2, "isotropic"), "ortho_small": (30, 4, "orthotropic"), "sol_sym_tiny": (8, 2, "sol")}
SOL_ANGLES = (0, 90, 90, 0)       # a mid-plane symmetric lay-up of the benchmark laminate


def oracle_case(name):
    from plate_inverse_problem_b200.synthetic import make_bending_plate
    nx, ny, kind = CASES[name]
    ff = make_bending_plate(nx, ny)
    return ff, orc.OracleSymmProblem(ff, orc.make_d_transform(kind, H, SOL_ANGLES), float(REF[f"{name}_rho"]), H)


def product_material(name):
    from plate_inverse_problem_b200 import Material
    p, rho = REF[f"{name}_params"], float(REF[f"{name}_rho"])
    if name.startswith("iso"):
        return Material.Isotropic(rho, *p)
    if name.startswith("ortho"):
        return Material.Orthotropic(rho, *p)
    return Material.SOL(rho, SOL_ANGLES, *p)


@pytest.mark.parametrize("name", list(CASES))
def test_oracle_symm_matches_the_reference_branch(name):
    ff, op = oracle_case(name)
    p, freqs, ref = REF[f"{name}_params"], REF[f"{name}_freqs"], REF[f"{name}_ref"]
    assert np.max(np.abs(op.transform(p, None) - REF[f"{name}_D"]) / np.abs(REF[f"{name}_D"]).max()) <= 1e-15
    fr = op.forward(freqs, p)
    assert np.max(np.abs(fr - REF[f"{name}_fr"]) / np.abs(REF[f"{name}_fr"])) <= 1e-9
    for lt in ("MSE", "RMSE", "MSE_AFC", "MSE_LOG_AFC"):
        L, g, _ = op.loss_and_grad(freqs, p, ref, lt)
        assert abs(L - float(REF[f"{name}_loss_{lt}"])) <= 1e-8 * abs(float(REF[f"{name}_loss_{lt}"]))
        fd = REF[f"{name}_fd_grad_{lt}"]              # central differences of the REFERENCE's loss
        assert np.max(np.abs(g - fd)) <= 2e-5 * np.max(np.abs(fd))


def test_dense_and_sparse_producer_agree_and_eliminate_the_clamp():
    from plate_inverse_problem_b200.synthetic import make_bending_plate
    a, b = make_bending_plate(6, 2), make_bending_plate(6, 2, dense=True)
    for k in range(6):
        assert np.array_equal(np.asarray(a["Ks"][k].todense()), b["Ks"][k])
    assert np.array_equal(a["fKs"], b["fKs"]) and np.array_equal(np.asarray(a["M"].todense()), b["M"])
    n = a["M"].shape[0]
    assert n == int((~a["constrained_idx"]).sum()) and a["interpolation_vector"].shape == (n,)
    # a rigid translation w = 1 has no curvature: K (1 on the free DOFs) = f (minus the clamped part)
    nv = a["mesh"].points.shape[0]
    w_all = np.concatenate([np.ones(nv), np.zeros(a["constrained_idx"].size - nv)])     # values 1, normal slopes 0
    free = ~a["constrained_idx"]
    for k in range(6):
        assert np.max(np.abs(a["Ks"][k] @ w_all[free] - a["fKs"][k])) <= 1e-6 * np.abs(a["Ks"][k]).max()


@pytest.mark.parametrize("name", list(CASES))
def test_bordered_system_reproduces_the_reduced_one(name):
    """Host logic of ProblemSymm: sum_k coef_k M_k on N = n + 1 unknowns with rhs e_t has the
    solution (u, 1), u the solution of the reference's reduced system; Q (u, 1) = c0 + c.u."""
    from plate_inverse_problem_b200.ProblemSymm import bordered_components, d_coefficients_and_jacobian
    ff, op = oracle_case(name)
    p, rho = REF[f"{name}_params"], float(REF[f"{name}_rho"])
    mat = product_material(name)
    assert mat.is_mps
    mats = bordered_components(ff, rho, H / 2)
    coef, s, dc, ds = d_coefficients_and_jacobian(mat.get_D_transform(H), p, [100.0, 250.0], with_jacobian=True)
    assert coef.shape == (2, 8) and np.all(s == 1.0) and np.all(ds == 0.0)
    assert np.max(np.abs(coef[0, :6] - op.transform(p, None)) / np.abs(coef[0, :6]).max()) <= 1e-15
    assert np.max(np.abs(dc[0, :6] - op.d_jacobian(p))) <= 1e-9 * np.max(np.abs(dc[0, :6]))
    assert np.all(dc[:, 6:] == 0.0) and coef[1, 6] == -(2 * np.pi * 250.0) ** 2 and coef[1, 7] == 1.0
    n = op.n
    A = sum(c * sps.csr_matrix(m) for c, m in zip(coef[1], mats)).tocsc()
    rhs = np.zeros(n + 1, dtype=np.complex128)
    rhs[n] = 1.0
    x = orc.solve_extended(A, rhs)
    fr, U = op.forward([250.0], p, return_u=True)
    assert abs(x[n] - 1.0) <= 1e-14
    assert np.max(np.abs(x[:n] - U[0])) <= 1e-9 * np.max(np.abs(U[0]))
    Q = np.concatenate([ff["interpolation_vector"], [ff["interpolation_value_from_bc"]]])
    assert abs(Q @ x - fr[0]) <= 1e-9 * abs(fr[0])
