"""Development check run on the GPU box: every layer against the numpy/scipy oracle."""
import sys, time, os
import numpy as np
import scipy.sparse as sps
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from plate_inverse_problem_b200 import _lib
from plate_inverse_problem_b200.jax_plate_lib import InnerState
from plate_inverse_problem_b200.synthetic import make_plate, NAMED_SIZES
from oracle import jax_plate_oracle as orc

ACC = dict(mass=0.0017, radius=0.0038, height=0.0087, eff=0.5, ts=0.03)
H, RHO = 2e-3, 1600.0
PARAMS = np.array([130e9, 9e9, 5e9, 0.3, 0.01])
ANGLES = (0, 45, -45, 90)


def setup(name):
    nx, ny = NAMED_SIZES[name]
    plate = make_plate(nx, ny)
    inertia = orc.inertia_scalars(H, RHO, ACC["mass"], ACC["radius"], ACC["height"])
    tr = orc.make_transform("sol", H, ANGLES)
    op = orc.OracleProblem(plate, tr, inertia, ACC["height"], ACC["eff"], ACC["ts"])
    return plate, op


def relerr(a, b):
    return float(np.max(np.abs(a - b)) / np.max(np.abs(b)))


def check_solve(name, B=3, refine=1):
    plate, op = setup(name)
    n = op.native
    N = op.N
    freqs = np.linspace(40, 600, B)
    coef, s = orc.coefficients(freqs, PARAMS, op.transform, op.inertia)
    data = np.stack([orc.assemble(coef[i], op.mats) for i in range(B)])
    rng = np.random.default_rng(0)
    b = rng.standard_normal((B, N)) + 1j * rng.standard_normal((B, N))
    st = InnerState()
    t0 = time.time()
    st.add_mat(N, n["indices"], n["indptr"], n["indices_T"], n["indptr_T"], n["permutation"])
    t_sym = time.time() - t0
    st.set_refinement(refine)
    print(f"[{name}] N={N} nnz={n['indices'].size} symbolic {t_sym:.2f}s info={st.symbolic_info(0)}")
    for tr in (False, True):
        x = st.solve(data, b, 0, tr, 0, 3)
        tm = st.timings()
        worst_res, worst_err = 0, 0
        for i in range(B):
            A = sps.csc_matrix((data[i], n["indices"], n["indptr"]), shape=(N, N))
            Aop = A.T if tr else A
            r = Aop @ x[i] - b[i]
            worst_res = max(worst_res, np.linalg.norm(r) / np.linalg.norm(b[i]))
            if i < 2:
                xo = orc.umfpack_standin_lu(A).solve(b[i], trans="T" if tr else "N")
                worst_err = max(worst_err, relerr(x[i], xo))
        print(f"[{name}] solve trans={tr} refine={refine}: max resid {worst_res:.2e}  max rel err vs SuperLU {worst_err:.2e}  timings {tm}")
    # matvec
    for tr in (False, True):
        y = st.matvec(data, b, 0, tr, 0, 3)
        yo = np.stack([(sps.csc_matrix((data[i], n["indices"], n["indptr"]), shape=(N, N)).T if tr else
                        sps.csc_matrix((data[i], n["indices"], n["indptr"]), shape=(N, N))) @ b[i] for i in range(B)])
        print(f"[{name}] matvec trans={tr}: rel err {relerr(y, yo):.2e}")
    # modes
    x2 = st.solve(data[0], b, 0, False, 0, 2)
    A0 = sps.csc_matrix((data[0], n["indices"], n["indptr"]), shape=(N, N))
    print(f"[{name}] mode2 resid {max(np.linalg.norm(A0 @ x2[i] - b[i]) / np.linalg.norm(b[i]) for i in range(B)):.2e}")
    b4 = rng.standard_normal((2, B, N)) + 0j
    x4 = st.solve(data, b4, 0, True, 0, 4)
    res4 = 0
    for j in range(2):
        for i in range(B):
            A = sps.csc_matrix((data[i], n["indices"], n["indptr"]), shape=(N, N))
            res4 = max(res4, np.linalg.norm(A.T @ x4[j, i] - b4[j, i]) / np.linalg.norm(b4[j, i]))
    print(f"[{name}] mode4(T) resid {res4:.2e}")
    x1 = st.solve(data, b[0], 0, False, 0, 1)
    print(f"[{name}] mode1 resid {max(np.linalg.norm(sps.csc_matrix((data[i], n['indices'], n['indptr']), shape=(N, N)) @ x1[i] - b[0]) / np.linalg.norm(b[0]) for i in range(B)):.2e}")
    x0 = st.solve(data[0], b[0], 0, False, 0, 0)
    print(f"[{name}] mode0 resid {np.linalg.norm(A0 @ x0 - b[0]) / np.linalg.norm(b[0]):.2e}")


def check_fused(name, F=4, refine=1, loss="MSE_LOG_AFC"):
    from plate_inverse_problem_b200 import _lib as L
    plate, op = setup(name)
    n = op.native
    N = op.N
    freqs = np.linspace(40, 600, F)
    coef, s = orc.coefficients(freqs, PARAMS, op.transform, op.inertia)
    st = L.State(0)
    num = L.C.c_int(-1)
    st.check(st.lib.jpb_add_mat(st.h, N, n["indices"].size, L.ptr(n["indices"]), L.ptr(n["indptr"]),
                                None, None, None, n["indices"].dtype.itemsize, L.C.byref(num)))
    st.check(st.lib.jpb_set_refinement(st.h, refine))
    mats = np.ascontiguousarray(op.mats)
    rhs_vec = np.ascontiguousarray(plate.rhs_vec, dtype=np.float64)
    Q = np.ascontiguousarray(op.Q)
    st.check(st.lib.jpb_register_components(st.h, 0, 26, L.ptr(mats), L.ptr(rhs_vec), 3, L.ptr(Q)))
    u = np.zeros((F, N), dtype=np.complex128)
    y = np.zeros((F, 3), dtype=np.complex128)
    coef = np.ascontiguousarray(coef)
    s = np.ascontiguousarray(s)
    st.check(st.lib.jpb_eval_forward(st.h, 0, F, L.ptr(coef), L.ptr(s), L.ptr(u), L.ptr(y), None))
    fr_o, U_o = op.forward(freqs, PARAMS, return_u=True)
    print(f"[{name}] fused forward: u rel err {max(relerr(u[i], U_o[i]) for i in range(F)):.2e}; timings {st.timings()}")
    ref = fr_o * (1 + 0.1 * np.cos(np.arange(F)))
    L_o, g_o, aux = op.loss_and_grad(freqs, PARAMS, ref, loss)
    w = np.array([ACC["ts"] ** 2, ACC["ts"] ** 2, 1.0])
    fr = np.zeros(F); loss_sum = np.zeros(1)
    ct = np.zeros((F, 26), dtype=np.complex128); cts = np.zeros(F, dtype=np.complex128)
    st.check(st.lib.jpb_eval_loss_grad(st.h, 0, F, L.ptr(coef), L.ptr(s), L.ptr(ref), L.ptr(w),
                                       L.LOSS_TYPES[loss], F, L.ptr(fr), L.ptr(loss_sum), L.ptr(ct),
                                       L.ptr(cts), None))
    print(f"[{name}] fused loss/grad: fr {relerr(fr, aux['fr']):.2e} loss {abs(loss_sum[0] / F - L_o) / abs(L_o):.2e} "
          f"ct_coef {relerr(ct, aux['ct_coef']):.2e} ct_s {relerr(cts, aux['ct_s']):.2e}; timings {st.timings()}")
    dc, ds = orc.coefficient_jacobian(freqs, PARAMS, op.transform, op.inertia)
    g = np.real(np.einsum("fk,fkp->p", ct, dc) + np.einsum("f,fp->p", cts, ds))
    print(f"[{name}] grad rel err {relerr(g, g_o):.2e}  g={g} g_o={g_o}")
    st.close()


def bench_fused(name, F, refine=1):
    from plate_inverse_problem_b200 import _lib as L
    plate, op = setup(name)
    n = op.native
    N = op.N
    freqs = np.linspace(40, 600, F)
    coef, s = orc.coefficients(freqs, PARAMS, op.transform, op.inertia)
    st = L.State(0)
    num = L.C.c_int(-1)
    t0 = time.time()
    st.check(st.lib.jpb_add_mat(st.h, N, n["indices"].size, L.ptr(n["indices"]), L.ptr(n["indptr"]),
                                None, None, None, n["indices"].dtype.itemsize, L.C.byref(num)))
    print(f"[{name}] symbolic {time.time() - t0:.2f}s {st.symbolic_info(0)}")
    st.check(st.lib.jpb_set_refinement(st.h, refine))
    mats = np.ascontiguousarray(op.mats)
    rhs_vec = np.ascontiguousarray(plate.rhs_vec, dtype=np.float64)
    Q = np.ascontiguousarray(op.Q)
    st.check(st.lib.jpb_register_components(st.h, 0, 26, L.ptr(mats), L.ptr(rhs_vec), 3, L.ptr(Q)))
    coef = np.ascontiguousarray(coef); s = np.ascontiguousarray(s)
    ref = np.ones(F); w = np.array([ACC["ts"] ** 2, ACC["ts"] ** 2, 1.0])
    fr = np.zeros(F); loss_sum = np.zeros(1)
    ct = np.zeros((F, 26), dtype=np.complex128); cts = np.zeros(F, dtype=np.complex128)
    for it in range(3):
        t0 = time.time()
        st.check(st.lib.jpb_eval_loss_grad(st.h, 0, F, L.ptr(coef), L.ptr(s), L.ptr(ref), L.ptr(w), 3, F,
                                           L.ptr(fr), L.ptr(loss_sum), L.ptr(ct), L.ptr(cts), None))
        dt = time.time() - t0
        print(f"[{name}] F={F} refine={refine} eval_loss_grad {dt * 1e3:.1f} ms -> {F / dt:.1f} evals/s; {st.timings()}")
    # residual check of forward on 2 systems
    u = np.zeros((F, N), dtype=np.complex128); y = np.zeros((F, 3), dtype=np.complex128)
    st.check(st.lib.jpb_eval_forward(st.h, 0, F, L.ptr(coef), L.ptr(s), L.ptr(u), L.ptr(y), None))
    for i in (0, F // 2, F - 1):
        A = op.matrix(coef[i]); bb = s[i] * rhs_vec
        print(f"   resid sys {i}: {np.linalg.norm(A @ u[i] - bb) / np.linalg.norm(bb):.2e}")
    st.close()


if __name__ == "__main__":
    st = _lib.State(0)
    e = np.zeros(1); st.check(st.lib.jpb_selftest(st.h, _lib.ptr(e))); print("selftest max err", e[0])
    p = np.zeros(1); st.check(st.lib.jpb_dmma_peak(st.h, _lib.ptr(p))); print("DMMA peak TFLOP/s", p[0])
    st.close()
    what = sys.argv[1:] or ["tiny", "small"]
    for name in what:
        if name.startswith("bench:"):
            _, nm, F, rf = name.split(":")
            bench_fused(nm, int(F), int(rf))
        else:
            for rf in (0, 1):
                check_solve(name, refine=rf)
            check_fused(name, refine=1)
            check_fused(name, refine=0, loss="MSE")
