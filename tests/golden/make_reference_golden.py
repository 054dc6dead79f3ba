"""Generates tests/golden/ref_*.npz by EXECUTING THE REFERENCE'S OWN CODE in this container.

    make -C oracle && python tests/golden/make_reference_golden.py

What runs (see tests/ref_shim.py): `/root/reference/source/jax_plate/{Sparse,Problem,Material,
Accelerometer}.py` unmodified, on top of the reference's pybind11 module compiled from its
own sources (`InnerState.h`, `umfpack_interface.h`, `csc_matvec.h`, `main.cpp`) against the
dense-LU stand-in for the UMFPACK library functions, with `jax` replaced by a numpy shim
and FreeFem++'s output replaced by the synthetic plate of
`plate_inverse_problem_b200.synthetic` (same tuple format, `pyFFInterface.py:503-509`).

What the fixtures therefore pin against the reference itself: the index construction
(`create_symbolic`, `find_permutation`, pattern union, data alignment), the inertia
scalars, every material transform, assembly + right-hand side + read-out (`_solve`),
the four losses, the vmap batching `_mode` table, the JVP / transpose rules, and
`InnerState::solve / matvec` in all five modes.  What stays a stand-in: the arithmetic of
the UMFPACK library (dense LU with partial pivoting here): "parity unpinned" only for
UMFPACK's pivot order / refinement, i.e. at the level of rounding.

The gradient has no reference value without JAX's AD; it is pinned by central finite
differences of the REFERENCE's loss function (`fd_grad_*`, good to ~1e-6 relative).
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ref_shim  # noqa: E402

H, RHO = 2e-3, 1600.0
PARAMS = np.array([130e9, 9e9, 5e9, 0.3, 0.01])
ANGLES = (0, 45, -45, 90)
OUT = os.path.join(ROOT, "tests", "golden")


def sha(a):
    a = np.ascontiguousarray(a)
    return np.frombuffer(hashlib.sha256(a.tobytes()).digest(), dtype=np.uint8).copy()


class Recorder:
    """Wraps the InnerState the reference's SolverState holds and records what reaches the
    native boundary (add_mat arguments, the `_mode` of every solve / matvec call)."""

    def __init__(self, inner):
        self.inner, self.add_mat_args, self.calls = inner, [], []

    def add_mat(self, *a):
        self.add_mat_args.append([np.array(x) for x in a])
        return self.inner.add_mat(*a)

    def solve(self, data, b, solver_num, transpose, n_cpu, mode):
        self.calls.append(("solve", int(mode), bool(transpose), np.shape(data), np.shape(b)))
        return self.inner.solve(data, b, solver_num, transpose, n_cpu, mode)

    def matvec(self, mat, vec, solver_num, transpose, n_cpu, mode):
        self.calls.append(("matvec", int(mode), bool(transpose), np.shape(mat), np.shape(vec)))
        return self.inner.matvec(mat, vec, solver_num, transpose, n_cpu, mode)


def materials(pkg):
    M = pkg.Material
    return {
        "isotropic": (M.Isotropic(7920.0, 2.2e11, 8.25e10, 0.003), np.array([2.2e11, 8.25e10, 0.003])),
        "orthotropic": (M.Orthotropic(RHO, *PARAMS), PARAMS),
        "orthotropic_d4": (M.OrthotropicD4(RHO, 130e9, 9e9, 5e9, 0.3, 0.01, 0.012, 0.02, 0.004),
                           np.array([130e9, 9e9, 5e9, 0.3, 0.01, 0.012, 0.02, 0.004])),
        "sol": (M.SOL(RHO, ANGLES, *PARAMS), PARAMS),
        "sol_sym": (M.SOL(RHO, (0, 90, 90, 0), *PARAMS), PARAMS),
        "symmetrical_sol": (M.SymmetricalSOL(RHO, (0, 45, -45), 70e9, 5e9, 0.3, 0.01),
                            np.array([70e9, 5e9, 0.3, 0.01])),
    }


def transform_fixture(pkg):
    """(A, B, D) of every material class (`Material.py:357-885`), ABD and D transforms."""
    out = {}
    for name, (mat, p) in materials(pkg).items():
        A, B, D = mat.get_ABD_transform(H)(p, 2 * np.pi * 100.0)
        out[f"abd_{name}"] = np.stack([np.asarray(A, dtype=complex), np.asarray(B, dtype=complex),
                                       np.asarray(D, dtype=complex)])
        out[f"params_{name}"] = p
        if mat.is_mps:
            out[f"d_{name}"] = np.asarray(mat.get_D_transform(H)(p, 2 * np.pi * 100.0), dtype=complex)
    Ms = pkg.Material.SOL(RHO, ANGLES, *PARAMS)._Q_to_ABD_matrices
    out["clt_sol"] = np.stack([np.array(m.evalf(subs={"h": H}), dtype=np.float64) for m in Ms])
    return out


def problem_fixture(pkg, name, nfreq, full):
    from plate_inverse_problem_b200.synthetic import NAMED_SIZES, make_plate
    S = pkg.Sparse
    rec = Recorder(S._SOLVER_STATE.state)
    S._SOLVER_STATE.state = rec
    plate = make_plate(*NAMED_SIZES[name])
    mat = pkg.Material.SOL(RHO, ANGLES, *PARAMS)
    acc = pkg.Accelerometer.Accelerometer("AP1030")
    prob = ref_shim.reference_problem(pkg, plate, mat, acc, H)
    N, idx, iptr, idxT, iptrT, perm, _ = rec.add_mat_args[-1]
    out = dict(N=np.array(prob.mat_size), solver_num=np.array(prob.solver_num),
               inertia=np.array([prob.I0, prob.I0Corr, prob.I2, prob.I2Corr]),
               params=PARAMS, nnz=np.array(prob.mats.shape[1]),
               mats_sha=sha(np.asarray(prob.mats, dtype=np.float64)),
               pattern_sha=sha(np.asarray(S._SOLVER_STATE.patterns[prob.solver_num], dtype=np.int64)))
    for k, v in (("indices", idx), ("indptr", iptr), ("indices_T", idxT), ("indptr_T", iptrT),
                 ("permutation", perm)):
        out[k + "_sha"] = sha(np.asarray(v, dtype=np.int64))
        out[k + "_dtype"] = np.array(str(np.asarray(v).dtype))
        if full:
            out[k] = np.asarray(v)
    if full:
        out["mats"] = np.asarray(prob.mats)
        out["pattern"] = np.asarray(S._SOLVER_STATE.patterns[prob.solver_num])
    # FRF sweep through the reference's getFRFunction (vmap over frequencies -> one native call
    # per frequency under the shim), with a point next to the first resonances
    freqs = np.sort(np.concatenate([np.linspace(40.0, 600.0, nfreq - 1), [100.35]]))
    fr = np.asarray(prob.solveForward(freqs))
    out["freqs"], out["fr"] = freqs, fr
    pert = PARAMS * np.array([1.1, 1.1, 1.2, 1.0, 1.5])
    ref = np.asarray(prob.solveForward(freqs, pert))
    out["ref"] = ref
    for lt in ("MSE", "RMSE", "MSE_AFC", "MSE_LOG_AFC"):
        lf = prob.getLossFunction(freqs, ref, lt)
        out[f"loss_{lt}"] = np.array(float(lf(PARAMS)))
        if full or lt == "MSE_LOG_AFC":
            g = np.zeros(PARAMS.size)
            for j in range(PARAMS.size):
                h = 1e-4 * PARAMS[j]
                acc_ = 0.0
                for k, w in ((-2, 1 / 12), (-1, -8 / 12), (1, 8 / 12), (2, -1 / 12)):
                    p = PARAMS.copy()
                    p[j] += k * h
                    acc_ += w * float(lf(p))
                g[j] = acc_ / h
            out[f"fd_grad_{lt}"] = g
    sc = PARAMS.copy()
    out["loss_scaled_MSE"] = np.array(float(prob.getLossFunction(freqs, ref, "MSE", scaling_params=sc)(np.ones(5))))
    S._SOLVER_STATE.state = rec.inner
    return out, prob, rec


def rules_fixture(pkg, prob, rec):
    """The reference's own batching rule, JVP and transpose rule functions on seeded inputs
    (`Sparse.py:162-176, 200-222, 245-282`), and `InnerState.solve / matvec` in every mode."""
    S = pkg.Sparse
    S._SOLVER_STATE.state = rec
    num, N, nnz = prob.solver_num, prob.mat_size, prob.mats.shape[1]
    rng = np.random.default_rng(77)
    c = lambda *s: rng.standard_normal(s) + 1j * rng.standard_normal(s)
    # well-conditioned data: two assembled plate matrices of the sweep
    m = prob.mats
    tr = prob.material.get_ABD_transform(H)
    data = []
    for f in (80.0, 333.0, 590.0):
        om = 2 * np.pi * f
        A, B, D = tr(PARAMS, om)
        mat = -om ** 2 * (prob.I0 * (m[18] + m[20] + m[22]) + prob.I0Corr * (m[19] + m[21] + m[23])
                          + prob.I2 * m[24] + prob.I2Corr * m[25]) + 0j
        for i in range(6):
            mat = mat + A[i] * m[i] + B[i] * m[i + 6] + D[i] * m[i + 12]
        data.append(mat)
    data = np.array(data)
    b = c(2, 3, N)
    ddata = c(3, nnz) * np.abs(data).mean()
    out = dict(r_data=data, r_b=b, r_ddata=ddata)
    kw = dict(solver_num=int(num), transpose=False, n_cpu=1, _mode=0)
    # batching rule: (data axis, b axis) -> chosen _mode, result
    spb = pkg.Sparse.batching.primitive_batchers[S._spsolve_p]
    mvb = pkg.Sparse.batching.primitive_batchers[S._matvec_p]
    cases = {"m1": ((data, b[0, 0]), (0, None)), "m2": ((data[0], b[:, 0]), (None, 0)),
             "m3": ((data, b[0]), (0, 0)), "m4": ((data, b), (0, 0)),
             "m3_moved": ((np.ascontiguousarray(data.T), np.ascontiguousarray(b[0].T)), (1, 1))}
    for trn in (False, True):
        kw["transpose"] = trn
        for cname, (vals, axes) in cases.items():
            for pname, fn in (("solve", spb), ("matvec", mvb)):
                rec.calls.clear()
                res, ax = fn(vals, axes, **{k: v for k, v in kw.items() if k != "_mode"})
                assert ax == 0
                out[f"batch_{pname}_{cname}_T{int(trn)}"] = np.asarray(res)
                out[f"batch_{pname}_{cname}_T{int(trn)}_mode"] = np.array(rec.calls[-1][1])
        x0 = S.spsolve(data[0], b[0, 0], **kw)
        out[f"solve_m0_T{int(trn)}"] = np.asarray(x0)
        out[f"matvec_m0_T{int(trn)}"] = np.asarray(S.matvec(data[0], b[0, 0], **kw))
    kw["transpose"] = False
    # JVP rules (registered with ad.defjvp): tangent wrt matrix entries, wrt rhs
    jl, jr = S._spsolve_p.jvps
    out["jvp_lhs"] = np.asarray(jl(ddata[0], data[0], b[0, 0], **kw))
    out["jvp_rhs"] = np.asarray(jr(b[1, 1], data[0], b[0, 0], **kw))
    ml, mr = S._matvec_p.jvps
    out["mv_jvp_mat"] = np.asarray(ml(ddata[0], data[0], b[0, 0], **kw))
    out["mv_jvp_vec"] = np.asarray(mr(b[1, 1], data[0], b[0, 0], **kw))
    # transpose rules
    ad = pkg.Sparse.ad
    und = ad.UndefinedPrimal(ref_shim.ShapedArray((N,), np.complex128))
    for trn in (False, True):
        kw["transpose"] = trn
        _, ctb = ad.primitive_transposes[S._spsolve_p](b[0, 1], data[0], und, **dict(kw))
        out[f"tr_solve_ct_b_T{int(trn)}"] = np.asarray(ctb)
        _, ctv = ad.primitive_transposes[S._matvec_p](b[0, 1], data[0], und, **dict(kw))
        out[f"tr_matvec_ct_v_T{int(trn)}"] = np.asarray(ctv)
        ctm, _ = ad.primitive_transposes[S._matvec_p](b[0, 1], ad.UndefinedPrimal(None), b[1, 2], **dict(kw))
        out[f"tr_matvec_ct_mat_T{int(trn)}"] = np.asarray(ctm)
    S._SOLVER_STATE.state = rec.inner
    return out


def input_fixture(pkg):
    """`Input.Compressor` (`Input.py:5-158`) of the reference on a synthetic FRF with several
    resonances (the oracle's forward sweep of the tiny plate, 1500 points)."""
    from conftest import get_case
    case = get_case("tiny")
    freqs = np.linspace(40.0, 1200.0, 1500)
    fr = case.op.forward(freqs, PARAMS).astype(np.complex128) * np.exp(0.3j)
    out = dict(in_freqs=freqs, in_fr=fr)
    for alg in (0, 1):
        for want in (30, 200, 317, 1000):
            c = pkg.Input.Compressor(freqs, fr, 1200, alg)
            f_sel, _ = c(want)
            out[f"mask_alg{alg}_{want}"] = np.isin(freqs, f_sel)
    return out


class Rosenbrock:
    """Analytic objective carrying its own derivatives, so that the reference's optimiser
    loops (`Optimizers.py`) run under the shim: f, value_and_grad, grad_fn(.jacobian = Hessian)."""

    def __init__(self, a=10.0):
        self.a = a
        outer = self

        class _Grad:
            def __call__(self, x):
                return outer.value_and_grad(x)[1]

            def jacobian(self, x):
                return outer.hessian(x)
        self.grad_fn = _Grad()

    def __call__(self, x):
        x = np.asarray(x, dtype=np.float64)
        return float(np.sum(self.a * (x[1:] - x[:-1] ** 2) ** 2 + (1 - x[:-1]) ** 2))

    def value_and_grad(self, x):
        x = np.asarray(x, dtype=np.float64)
        g = np.zeros_like(x)
        d = x[1:] - x[:-1] ** 2
        g[:-1] += -4 * self.a * d * x[:-1] - 2 * (1 - x[:-1])
        g[1:] += 2 * self.a * d
        return self(x), g

    def hessian(self, x):
        x = np.asarray(x, dtype=np.float64)
        n = x.size
        H = np.zeros((n, n))
        for i in range(n - 1):
            H[i, i] += -4 * self.a * (x[i + 1] - 3 * x[i] ** 2) + 2
            H[i, i + 1] += -4 * self.a * x[i]
            H[i + 1, i] += -4 * self.a * x[i]
            H[i + 1, i + 1] += 2 * self.a
        return H


def optimizer_fixture(pkg):
    """The reference's optimiser loops (`Optimizers.py:153-286`) and trust-region sub-problem
    solver (`:70-122`) on an analytic objective.  (`optimize_cd_mem2` differentiates a local
    closure with jax and cannot run under the shim.)"""
    O = pkg.Optimizers
    f = Rosenbrock()
    x0 = np.array([-0.8, 0.6, 0.3])
    out = dict(opt_x0=x0)
    runs = dict(gd=lambda: O.optimize_gd(f, x0.copy(), N_steps=40, h=0.01, f_min=1e-8),
                cd=lambda: O.optimize_cd(f, x0.copy(), N_steps=15, h=0.01, f_min=1e-8),
                tr=lambda: O.optimize_trust_region(f, x0.copy(), N_steps=25, delta_max=1.0))
    for name, run in runs.items():
        r = run()
        out[f"opt_{name}_x"] = np.asarray(r.x, dtype=np.float64)
        out[f"opt_{name}_f"] = np.array(float(r.f))
        out[f"opt_{name}_fhist"] = np.array([float(v) for v in r.f_history])
        out[f"opt_{name}_ghist"] = np.array([np.asarray(g, dtype=np.float64) for g in r.grad_history])
        out[f"opt_{name}_niter"] = np.array(int(r.niter))
        out[f"opt_{name}_status"] = np.array(str(r.status))
    rng = np.random.default_rng(5)
    for i in range(6):
        M = rng.standard_normal((4, 4))
        B = M + M.T + (3.0 if i % 2 == 0 else -0.5) * np.eye(4)      # definite and indefinite models
        g = rng.standard_normal(4)
        delta = [0.05, 0.5, 5.0][i % 3]
        sd, lam, pred = O.solve_trust_region_model(B, g, delta)
        out[f"trm_{i}_B"], out[f"trm_{i}_g"], out[f"trm_{i}_delta"] = B, g, np.array(delta)
        out[f"trm_{i}_sd"], out[f"trm_{i}_lam"], out[f"trm_{i}_pred"] = np.asarray(sd), np.array(float(lam)), np.array(float(pred))
    fp = O.FixedParameterFunction(f, 3, 1, 0.6)
    out["fixed_val"] = np.array(float(fp(np.array([-0.8, 0.3]))))
    return out


SYMM_CASES = {      # name -> (nx, ny, material kind, constructor args after density, perturbation of the parameters)
    "iso_tiny": (6, 2, "Isotropic", (200e9, 77e9, 0.003), (1.1, 1.05, 2.0)),
    "ortho_small": (30, 4, "Orthotropic", (130e9, 9e9, 5e9, 0.3, 0.01), (1.1, 1.1, 1.2, 1.0, 1.5)),
    "sol_sym_tiny": (8, 2, "SOL", ((0, 90, 90, 0), 130e9, 9e9, 5e9, 0.3, 0.01), (1.1, 1.1, 1.2, 1.0, 1.5)),
}


def symm_fixture(pkg):
    """The mid-plane symmetric branch of the reference (complex FRF, four losses, FD gradient of the
    reference's own loss) on bending-only synthetic plates; see ref_shim.reference_symm_problem."""
    from plate_inverse_problem_b200.synthetic import make_bending_plate
    out = {}
    for name, (nx, ny, kind, args, pert) in SYMM_CASES.items():
        rho = 7800.0 if kind == "Isotropic" else RHO
        mat = getattr(pkg.Material, kind)(rho, *args)
        prob = ref_shim.reference_symm_problem(pkg, make_bending_plate(nx, ny, dense=True), mat, H)
        p0 = np.asarray(mat.get_parameters(), dtype=np.float64)
        freqs = np.sort(np.concatenate([np.linspace(40.0, 600.0, 6), [83.0]]))
        fr = np.asarray(prob.solveForward(freqs))
        ref = np.asarray(prob.solveForward(freqs, p0 * np.array(pert)))
        out[f"{name}_rho"], out[f"{name}_params"], out[f"{name}_freqs"] = np.array(rho), p0, freqs
        out[f"{name}_fr"], out[f"{name}_ref"] = fr, ref
        out[f"{name}_nnz"] = np.array(prob.M.shape[0])
        out[f"{name}_D"] = np.asarray(mat.get_D_transform(H)(p0, 2 * np.pi * 100.0))
        for lt in ("MSE", "RMSE", "MSE_AFC", "MSE_LOG_AFC"):
            lf = prob.getLossFunction(freqs, ref, lt)
            out[f"{name}_loss_{lt}"] = np.array(float(lf(p0)))
            g = np.zeros(p0.size)
            for j in range(p0.size):
                h = 1e-4 * p0[j]
                for k, w in ((-2, 1 / 12), (-1, -8 / 12), (1, 8 / 12), (2, -1 / 12)):
                    p = p0.copy()
                    p[j] += k * h
                    g[j] += w * float(lf(p)) / h
            out[f"{name}_fd_grad_{lt}"] = g
    return out


def check(pkg):
    """--check: recompute the quick parts and compare with the committed fixtures."""
    bad = []

    def cmp(tag, new, old, tol):
        for k, v in new.items():
            o = old[k]
            v = np.asarray(v)
            if v.dtype.kind in "biuUS" or k.endswith("_sha"):
                ok = np.array_equal(v, o)
            else:
                ok = v.shape == o.shape and np.max(np.abs(v - o)) <= tol * max(np.max(np.abs(o)), 1e-300)
            if not ok:
                bad.append(f"{tag}:{k}")
    cmp("transforms", transform_fixture(pkg), np.load(os.path.join(OUT, "ref_transforms.npz")), 1e-14)
    cmp("input", input_fixture(pkg), np.load(os.path.join(OUT, "ref_input.npz")), 1e-11)
    cmp("optim", optimizer_fixture(pkg), np.load(os.path.join(OUT, "ref_optim.npz")), 1e-11)
    tiny, prob, rec = problem_fixture(pkg, "tiny", 9, full=False)
    tiny.update(rules_fixture(pkg, prob, rec))
    cmp("tiny", tiny, np.load(os.path.join(OUT, "ref_tiny.npz")), 1e-11)
    cmp("symm", symm_fixture(pkg), np.load(os.path.join(OUT, "ref_symm.npz")), 1e-9)
    if bad:
        print("MISMATCH", bad)
        sys.exit(1)
    print("fixtures match the reference")


def main():
    pkg = ref_shim.load("compiled")
    if "--check" in sys.argv:
        return check(pkg)
    if "--only-input" not in sys.argv:
        np.savez_compressed(os.path.join(OUT, "ref_symm.npz"), **symm_fixture(pkg))
    if "--only-symm" in sys.argv:
        return
    np.savez_compressed(os.path.join(OUT, "ref_input.npz"), **input_fixture(pkg))
    np.savez_compressed(os.path.join(OUT, "ref_optim.npz"), **optimizer_fixture(pkg))
    if "--only-input" in sys.argv:
        return
    np.savez_compressed(os.path.join(OUT, "ref_transforms.npz"), **transform_fixture(pkg))
    tiny, prob, rec = problem_fixture(pkg, "tiny", 9, full=True)
    tiny.update(rules_fixture(pkg, prob, rec))
    np.savez_compressed(os.path.join(OUT, "ref_tiny.npz"), **tiny)
    small, _, _ = problem_fixture(pkg, "small", 7, full=False)
    np.savez_compressed(os.path.join(OUT, "ref_small.npz"), **small)
    for f in ("ref_transforms.npz", "ref_tiny.npz", "ref_small.npz"):
        print(f, os.path.getsize(os.path.join(OUT, f)), "bytes")


if __name__ == "__main__":
    main()
