"""CPU tests of the host logic: C-ABI surface, symbolic analysis (replayed in numpy),
index construction, material transforms, no-CPU-fallback behaviour."""
import ctypes
import os
import re

import numpy as np
import pytest
import scipy.sparse as sps

from conftest import PARAMS, ROOT, H, relerr
from oracle import jax_plate_oracle as orc
from plate_inverse_problem_b200 import _lib
from mf_emulator import Emulator


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "jax_plate_b200.h")).read()
    declared = set(re.findall(r"\b(jpb_[a-z_0-9]+)\s*\(", hdr))
    assert declared == set(_lib.EXPORTED_SYMBOLS)
    lib = _lib.load_library()
    for name in declared:
        assert hasattr(lib, name), name


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _lib.State(0)
    from plate_inverse_problem_b200.jax_plate_lib import InnerState
    with pytest.raises(RuntimeError):
        InnerState()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "plate_inverse_problem_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, re.M), f


@pytest.mark.parametrize("name", ["tiny", "small", "cfg0"])
def test_symbolic_replay_matches_direct_solve(name):
    from conftest import get_case
    case = get_case(name)
    n = case.native
    S = _lib.host_symbolic(case.N, n["indices"], n["indptr"])
    assert sorted(S["perm"].tolist()) == list(range(case.N))
    assert S["f_k"].sum() == case.N
    # every entry owns a distinct slot inside its front
    for f in range(S["nfronts"]):
        sl = S["ent_slot"][S["f_entptr"][f]:S["f_entptr"][f + 1]]
        assert len(set(sl.tolist())) == sl.size
    assert sorted(S["ent_src"].tolist()) == list(range(S["nnz"]))
    data, _, _ = case.data(np.array([40.0, 311.0]))
    rng = np.random.default_rng(1)
    b = rng.standard_normal(case.N) + 1j * rng.standard_normal(case.N)
    em = Emulator(S, case.N)
    for d in data:
        em.factor(d)
        A = sps.csc_matrix((d, n["indices"], n["indptr"]), shape=(case.N, case.N))
        lu = orc.umfpack_standin_lu(A)
        assert relerr(em.solve(b), lu.solve(b)) < 1e-9
        assert relerr(em.solve(b, trans=True), lu.solve(b, trans="T")) < 1e-9


def test_symbolic_fill_not_worse_than_superlu():
    """SURVEY 8d guard: nnz(L+U) of our ordering vs SuperLU MMD_AT_PLUS_A on the same matrix."""
    from conftest import get_case
    case = get_case("cfg0")
    n = case.native
    S = _lib.host_symbolic(case.N, n["indices"], n["indptr"])
    data, _, _ = case.data(np.array([100.0]))
    lu = orc.umfpack_standin_lu(sps.csc_matrix((data[0], n["indices"], n["indptr"]), shape=(case.N,) * 2))
    assert S["nnz_lu"] <= 1.5 * (lu.L.nnz + lu.U.nnz - case.N)   # measured 1.38 here, 1.19 at 2e4, 0.92 at 1e5 DOF


def test_host_analyze_rejects_bad_input():
    with pytest.raises(RuntimeError):
        _lib.host_symbolic(4, np.array([0, 1, 2, 9], dtype=np.int32), np.array([0, 1, 2, 3, 4], dtype=np.int32))
    with pytest.raises(RuntimeError):
        _lib.host_symbolic(4, np.array([0, 1, 2, 3], dtype=np.int32), np.array([0, 1, 2, 3, 5], dtype=np.int32))


def test_index_construction_bit_exact(small):
    """App. A: our create_symbolic index arrays == the oracle's == scipy's own machinery."""
    from plate_inverse_problem_b200.Problem import union_pattern_and_align
    from plate_inverse_problem_b200.Sparse import find_permutation
    plate = small.plate
    nz_idx, (mat, mat_T, coo, perm), mats = union_pattern_and_align(plate.mats, plate.N)
    n = small.native
    assert np.array_equal(nz_idx, small.op.nz_idx)
    assert np.array_equal(mat.indices, n["indices"]) and np.array_equal(mat.indptr, n["indptr"])
    assert np.array_equal(mat_T.indices, n["indices_T"])
    assert np.array_equal(perm, n["permutation"])
    assert np.array_equal(mats, small.op.mats)
    # canonical order is CSC (col, then row), perm = CSR->CSC map
    key = coo[:, 1].astype(np.int64) * plate.N + coo[:, 0]
    assert np.all(np.diff(key) > 0)
    assert np.array_equal(perm, np.argsort(coo[:, 0].astype(np.int64) * plate.N + coo[:, 1]))
    # int64 keys: no overflow for N > 46 340 (reference quirk, App. E)
    big = np.array([[50000, 60000], [1, 2]], dtype=np.int32)
    assert np.array_equal(find_permutation(big, big[::-1].copy(), 70000), [1, 0])


def test_batch_mode_table():
    from plate_inverse_problem_b200.Sparse import batch_mode
    z = np.zeros
    assert batch_mode(z(5), z(3)) == 0
    assert batch_mode(z((2, 5)), z(3)) == 1
    assert batch_mode(z(5), z((2, 3))) == 2
    assert batch_mode(z((2, 5)), z((2, 3))) == 3
    assert batch_mode(z((2, 5)), z((4, 2, 3))) == 4
    with pytest.raises(AssertionError):
        batch_mode(z((2, 5)), z((3, 3)))
    with pytest.raises(NotImplementedError):
        batch_mode(z(5), z((4, 2, 3)))


def test_material_transforms_match_oracle():
    from plate_inverse_problem_b200 import Material as M
    fr = np.linspace(40, 600, 7)
    inertia = orc.inertia_scalars(H, 1600.0, 0.0017, 0.0038, 0.0087)
    cases = [(M.SOL(1600, (0, 45, -45, 90)), orc.make_transform("sol", H, (0, 45, -45, 90)), PARAMS),
             (M.Orthotropic(1600), orc.make_transform("orthotropic", H), PARAMS),
             (M.Isotropic(7800), orc.make_transform("isotropic", H), np.array([2e11, 8e10, 0.003]))]
    for mat, tr, p in cases:
        c, s, dc, ds = M.coefficients_and_jacobian(mat.get_ABD_transform(H), p, fr, inertia)
        co, so = orc.coefficients(fr, p, tr, inertia)
        dco, dso = orc.coefficient_jacobian(fr, p, tr, inertia)
        assert relerr(c, co) < 1e-14 and relerr(s, so) < 1e-14
        assert relerr(dc, dco) < 1e-9 and relerr(ds, dso) < 1e-9
