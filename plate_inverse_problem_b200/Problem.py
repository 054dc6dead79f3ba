"""`Problem`: the reference's problem object (`/root/reference/source/jax_plate/Problem.py`)
restricted to the live (unsymmetric, accelerometer) path, with the per-frequency body of
`_solve` (`Problem.py:432-477`), the losses (`:948-975`) and their reverse-mode derivative
executed by the fused CUDA path (`jpb_eval_forward`, `jpb_eval_loss_grad`).

What is kept from the reference interface: `Problem(geometry, material, accelerometer,
ref_fr)`, `.getFRFunction()` -> `f(freqs, params) -> fr[F]`, `.solveForward(freqs, params)`,
`.getLossFunction(frequencies, reference_fr, func_type, scaling_params)`.  What differs:
* FreeFem++ is not available, so `geometry` carries the output tuple of
  `load_matrices_unsymm` (`pyFFInterface.py:503-509`) -- e.g. from `synthetic.make_plate` --
  instead of an .edp file name;
* the callable returned by `getLossFunction` also has `.value_and_grad(params)` (the
  reference obtains it with `jax.value_and_grad`, `Optimizers.py:232,258`); with
  `params[C, P]` it evaluates C parameter candidates in one batched call (BASELINE
  configs[4]) and returns `(L[C], g[C, P])`;
* `getJacobianFunction` / `getGaussNewtonModel` give the residual Jacobian and the
  Gauss-Newton model `(r^T J / n, J^T J / n)` the reference only sketches
  (`Optimizers.py:139-144`), one adjoint solve per frequency;
* with `torch.distributed` initialised the frequency list is sharded over the ranks and the
  loss / gradient partials are summed with ONE all-reduce (NCCL over NVLink).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable

import numpy as np

from . import _lib, parallel
from .Material import Material, coefficients_and_jacobian
from .Sparse import find_permutation, symbolic_indices


@dataclass
class AccelerometerParams:
    """`Accelerometer.py:8-33`."""
    mass: float
    radius: float
    height: float
    effective_height: float
    transverse_sensitivity: float


AP1030 = AccelerometerParams(0.0017, 0.0038, 0.0087, 0.5, 0.03)   # accelerometers/AP1030.json


@dataclass
class Geometry:
    """Stand-in for `Geometry.py`: plate thickness + the FreeFEM output tuple."""
    height: float
    matrices: tuple   # (mats[26] COO, rhs_vec, interp_mat, interp_mat_Lh, Lh, Mh, mesh, interp_Wx, interp_Wy)


def union_pattern_and_align(mats, N: int):
    """`Problem.py:317-345` with int64 keys: union of the component patterns in canonical
    (CSC) order and every component's data aligned to it.
    Returns nz_idx [nnz,2], (mat, mat_T, coo_indices, perm) of `symbolic_indices`, mats [K,nnz]."""
    def unwind(m):
        return m.row.astype(np.int64) + np.int64(N) * m.col.astype(np.int64)

    mats = [m.tocoo() for m in mats]
    # the reference folds np.union1d over the 26 key arrays (`Problem.py:320-326`); the sorted
    # set of unique keys it ends with is one sort + adjacent-difference mask (numpy 2.3's
    # hash-based unique takes minutes on these structured keys at N ~ 1e5)
    pattern = np.sort(np.concatenate([unwind(m) for m in mats]))
    if pattern.size:
        pattern = pattern[np.concatenate(([True], pattern[1:] != pattern[:-1]))]
    rows, cols = pattern % N, pattern // N
    nz_idx = np.vstack((rows, cols)).T
    sym = symbolic_indices(N, nz_idx)
    coo_indices = sym[2]
    perm = find_permutation(nz_idx, coo_indices, N)
    out = np.zeros((len(mats), rows.size), dtype=np.float64)
    for i, m in enumerate(mats):
        idx = unwind(m)
        order = np.argsort(idx)
        mask = np.isin(pattern, idx[order], assume_unique=True)
        data = np.zeros(rows.size, dtype=np.float64)
        data[mask] = m.data[order]
        out[i] = data[perm]
    return nz_idx, sym, out


class Problem:
    def __new__(cls, geometry=None, material=None, accelerometer=None, *args, **kw):
        # `Problem.py:228`: a mid-plane symmetric material without an accelerometer takes the
        # bending-only path (ProblemSymm.SymmetricProblem), everything else the unsymmetric one
        if cls is Problem and accelerometer is None and material is not None and material.is_mps:
            from .ProblemSymm import SymmetricProblem
            return object.__new__(SymmetricProblem)
        return object.__new__(cls)

    def __init__(self, geometry: Geometry, material: Material, accelerometer: AccelerometerParams,
                 ref_fr=None, *, cpu: int = 0, device: int = None, refinement: int = None,
                 refinement_threshold=None):
        if accelerometer is None:
            raise ValueError("the unsymmetric path needs an accelerometer (Problem.py:228, 310); without one the "
                             "material must be mid-plane symmetric")
        self.geometry, self.material, self.accelerometer = geometry, material, accelerometer
        self.reference_fr = ref_fr
        self.n_cpu = cpu    # accepted for interface parity; the batch loop runs on the GPU
        self.rho = material.density
        self.parameters = material.get_parameters()
        ff = geometry.matrices
        ff = ff.as_tuple() if hasattr(ff, "as_tuple") else ff
        mats = list(ff[0])
        N = mats[0].shape[0]
        self.interp_mat, self.interp_mat_Lh = np.asarray(ff[2]), np.asarray(ff[3])
        self.Lh_size, self.Mh_size, self.mesh = ff[4], ff[5], ff[6]
        self.interp_mat_Wx, self.interp_mat_Wy = np.asarray(ff[7]), np.asarray(ff[8])
        acc, h = accelerometer, geometry.height
        rho_corr = acc.mass / (np.pi * acc.radius ** 2) / acc.height        # Problem.py:356-374
        self.I0 = h * self.rho
        self.I0Corr = acc.height * rho_corr
        self.I2 = self.rho * h ** 3 / 12
        self.I2Corr = rho_corr / 3 * ((h / 2 + acc.height) ** 3 - h ** 3 / 8)
        self.transform = material.get_ABD_transform(h)
        # read-out rows (row means of the dense interpolation matrices, Problem.py:454-462)
        Lh = self.Lh_size
        Q = np.zeros((3, N))
        mL = self.interp_mat_Lh.mean(axis=0)
        hh = acc.effective_height * acc.height
        Q[0, :Lh] = mL
        Q[0, 2 * Lh:] = -hh * self.interp_mat_Wx.mean(axis=0)
        Q[1, Lh:2 * Lh] = mL
        Q[1, 2 * Lh:] = -hh * self.interp_mat_Wy.mean(axis=0)
        Q[2, 2 * Lh:] = self.interp_mat.mean(axis=0)
        ts = acc.transverse_sensitivity
        self._setup_native(mats, ff[1], Q, np.array([ts ** 2, ts ** 2, 1.0]), device=device,
                           refinement=refinement, refinement_threshold=refinement_threshold)

    def _setup_native(self, mats, rhs_vec, Q, weights, *, device=None, refinement=None, refinement_threshold=None):
        """Pattern union + alignment (`Problem.py:317-345`), registration of the pattern, the
        component matrices, the right-hand-side vector and the read-out rows with the CUDA library."""
        import os
        self.mat_size = N = mats[0].shape[0]
        nz_idx, (mat, mat_T, coo_indices, perm), self.mats = union_pattern_and_align(mats, N)
        self.sparsity = nz_idx.shape[0] / N ** 2
        self.nz_mask = (coo_indices[:, 0], coo_indices[:, 1])
        self.vec = np.ascontiguousarray(rhs_vec, dtype=np.float64)
        self.Q = Q = np.ascontiguousarray(Q, dtype=np.float64)
        self.weights = np.ascontiguousarray(weights, dtype=np.float64)
        if device is None:
            device = int(os.environ.get("LOCAL_RANK", "0"))
        self._st = st = _lib.State(device)
        idx = np.ascontiguousarray(mat.indices)
        iptr = np.ascontiguousarray(mat.indptr.astype(idx.dtype))
        idxT = np.ascontiguousarray(mat_T.indices.astype(idx.dtype))
        permc = np.ascontiguousarray(perm.astype(idx.dtype))
        num = _lib.C.c_int(-1)
        st.check(st.lib.jpb_add_mat(st.h, N, idx.size, _lib.ptr(idx), _lib.ptr(iptr), _lib.ptr(idxT),
                                    _lib.ptr(iptr), _lib.ptr(permc), idx.dtype.itemsize, _lib.C.byref(num)))
        self.solver_num = num.value
        # refinement policy (include/jax_plate_b200.h): by default one unconditional step;
        # refinement=0 is the fast mode (residual pass, a step only for systems that need it)
        if refinement is not None:
            st.check(st.lib.jpb_set_refinement(st.h, int(refinement)))
        if refinement_threshold is not None:       # (omega, relative residual) or omega alone
            th = np.atleast_1d(np.asarray(refinement_threshold, dtype=np.float64))
            st.check(st.lib.jpb_set_refinement_threshold(st.h, float(th[0]), float(th[1]) if th.size > 1 else 0.0))
        self._refinement = 1 if refinement is None else int(refinement)
        self._residual_pass = self._refinement == 0 and (
            refinement_threshold is None or float(np.atleast_1d(refinement_threshold)[0]) > 0)
        self.mats = np.ascontiguousarray(self.mats)
        st.check(st.lib.jpb_register_components(st.h, self.solver_num, self.mats.shape[0],
                                                _lib.ptr(self.mats), _lib.ptr(self.vec), Q.shape[0],
                                                _lib.ptr(Q)))

    # ---- helpers -------------------------------------------------------------------------
    @property
    def inertia(self):
        return (self.I0, self.I0Corr, self.I2, self.I2Corr)

    def coefficients(self, freqs, params, with_jacobian=False):
        return coefficients_and_jacobian(self.transform, params, freqs, self.inertia, with_jacobian)

    def timings(self):
        return self._st.timings()

    def symbolic_info(self):
        return self._st.symbolic_info(self.solver_num)

    # ---- forward -------------------------------------------------------------------------
    def _forward_raw(self, freqs, params, return_u=False):
        freqs = np.atleast_1d(np.asarray(freqs, dtype=np.float64))
        F = freqs.size
        if F == 0:      # an empty sweep (the reference's vmap over zero frequencies returns an empty array)
            R = self.Q.shape[0]
            empty_u = np.zeros((0, self.mat_size), dtype=np.complex128)
            return (np.zeros(0), np.zeros((0, R), dtype=np.complex128)) + ((empty_u,) if return_u else ())
        coef, s, _, _ = self.coefficients(freqs, params)
        coef, s = np.ascontiguousarray(coef), np.ascontiguousarray(s)
        y = np.zeros((F, self.Q.shape[0]), dtype=np.complex128)
        u = np.zeros((F, self.mat_size), dtype=np.complex128) if return_u else None
        st = self._st
        st.check(st.lib.jpb_eval_forward(st.h, self.solver_num, F, _lib.ptr(coef), _lib.ptr(s),
                                         _lib.ptr(u), _lib.ptr(y), None))
        fr = np.sqrt((self.weights[None, :] * np.abs(y) ** 2).sum(axis=1))
        return (fr, y, u) if return_u else (fr, y)

    def getFRFunction(self) -> Callable:
        """`Problem.py:377-518`: f(freqs[F], params[P]) -> fr[F] (real amplitudes)."""
        def fr_function(freqs, params):
            freqs = np.atleast_1d(np.asarray(freqs, dtype=np.float64))
            return parallel.sharded_gather(lambda lo, hi: self._forward_raw(freqs[lo:hi], params)[0],
                                           freqs.size, self._st.device)
        return fr_function

    def solveForward(self, freqs, params=None):
        """`Problem.py:611-639`."""
        if params is None:
            params = self.parameters
        return self.getFRFunction()(freqs, np.asarray(params, dtype=np.float64))

    # ---- loss + gradient -----------------------------------------------------------------
    def getLossFunction(self, frequencies, reference_fr, func_type: str, scaling_params=None):
        """`Problem.py:933-980`.  Returns `loss(params) -> float` with the extra method
        `loss.value_and_grad(params) -> (float, grad[P])`."""
        frequencies = np.atleast_1d(np.asarray(frequencies, dtype=np.float64))
        if func_type not in _lib.LOSS_TYPES:
            raise ValueError(f'Function type "{func_type}" is not supported!')
        assert frequencies.shape[0] == np.asarray(reference_fr).shape[0]
        reference_in = reference_fr
        # the reference as the device takes it: `unit` doubles per frequency
        reference_fr, ltype, unit = self._loss_reference(reference_fr, func_type)
        scaling = 1.0 if scaling_params is None else np.array(scaling_params, dtype=np.float64)
        F = frequencies.size

        def value_and_grad(params):
            p = np.asarray(params, dtype=np.float64) * scaling
            if p.ndim == 2:
                L, g = self._candidates_value_and_grad(frequencies, reference_fr, ltype, p, unit)
                return L, g * scaling

            def local(lo, hi):
                nloc = hi - lo
                coef, s, _, _ = self.coefficients(frequencies[lo:hi], p, with_jacobian=False)
                coef, s = np.ascontiguousarray(coef), np.ascontiguousarray(s)
                ref = np.ascontiguousarray(reference_fr[unit * lo:unit * hi])
                fr, loss_sum = np.zeros(nloc), np.zeros(1)
                ct = np.zeros((nloc, coef.shape[1]), dtype=np.complex128)
                cts = np.zeros(nloc, dtype=np.complex128)
                st = self._st
                # the native call (blocking, GIL released by ctypes) runs on a helper thread while
                # this one differentiates the material transform: the Jacobians are only needed for
                # the chain rule after the device has finished
                fut = _native_pool().submit(
                    st.lib.jpb_eval_loss_grad, st.h, self.solver_num, nloc, _lib.ptr(coef), _lib.ptr(s),
                    _lib.ptr(ref), _lib.ptr(self.weights), ltype, F, _lib.ptr(fr), _lib.ptr(loss_sum),
                    _lib.ptr(ct), _lib.ptr(cts), None)
                _, _, dc, ds = self.coefficients(frequencies[lo:hi], p, with_jacobian=True)
                st.check(fut.result())
                value_and_grad.last_fr = fr
                # chain rule through the coefficient functions
                g = np.real(np.einsum("fk,fkp->p", ct, dc) + np.einsum("f,fp->p", cts, ds))
                return loss_sum[0] / F, g

            val, grad = parallel.sharded_value_and_grad(local, F, p.size, self._st.device)   # ONE all-reduce
            return val, grad * scaling

        fr_function = self.getFRFunction()

        def loss(params):
            fr = fr_function(frequencies, np.asarray(params, dtype=np.float64) * scaling)
            return self._loss_value(fr, reference_in, func_type)

        loss.value_and_grad = value_and_grad
        loss.grad = lambda params: value_and_grad(params)[1]
        # the Gauss-Newton model of the same loss, for Optimizers.optimize_lm / method="gn"
        loss.gauss_newton = self.getGaussNewtonModel(frequencies, reference_in, func_type, scaling_params)
        loss.__name__ = func_type + "Loss"
        return loss

    def _loss_reference(self, reference_fr, func_type):
        """(reference as the device takes it, C-ABI loss type, doubles per frequency)."""
        return _real_reference(reference_fr), _lib.LOSS_TYPES[func_type], 1

    def _loss_value(self, fr, reference_fr, func_type):
        return float(np.mean(_residuals(fr, _real_reference(reference_fr), func_type)[0] ** 2))

    # ---- batched parameter candidates (BASELINE configs[4]) --------------------------------
    def _candidates_value_and_grad(self, frequencies, reference_fr, ltype, cands, unit=1):
        """C candidates x F frequencies as ONE flattened batch of C*F systems, sharded
        contiguously over the ranks; per-candidate loss sums are segmented on the device
        (`jpb_eval_candidates`), the only exchange is one all-reduce of [C, 1+P]."""
        C_, P = cands.shape
        F = frequencies.size
        lo, hi = parallel.shard_range(C_ * F)
        acc = np.zeros((C_, 1 + P))
        if hi > lo:
            c0, c1 = lo // F, (hi - 1) // F + 1
            coefs, ss, dcs, dss = [], [], [], []
            for c in range(c0, c1):
                coef, s_, dc, ds = self.coefficients(frequencies, cands[c], with_jacobian=True)
                coefs.append(coef); ss.append(s_); dcs.append(dc); dss.append(ds)
            sl = slice(lo - c0 * F, hi - c0 * F)
            coef = np.ascontiguousarray(np.concatenate(coefs)[sl])
            s_ = np.ascontiguousarray(np.concatenate(ss)[sl])
            dc, ds = np.concatenate(dcs)[sl], np.concatenate(dss)[sl]
            n = hi - lo
            cand = np.ascontiguousarray((np.arange(lo, hi) // F).astype(np.int32))
            ref = np.ascontiguousarray(np.tile(reference_fr, c1 - c0)[unit * sl.start:unit * sl.stop])
            fr, loss_sum = np.zeros(n), np.zeros(C_)
            ct = np.zeros((n, coef.shape[1]), dtype=np.complex128)
            cts = np.zeros(n, dtype=np.complex128)
            st = self._st
            st.check(st.lib.jpb_eval_candidates(st.h, self.solver_num, n, _lib.ptr(coef), _lib.ptr(s_),
                                                _lib.ptr(ref), _lib.ptr(self.weights), ltype, F,
                                                _lib.ptr(cand), C_, _lib.ptr(fr), _lib.ptr(loss_sum),
                                                _lib.ptr(ct), _lib.ptr(cts), None))
            g_sys = np.real(np.einsum("fk,fkp->fp", ct, dc) + cts[:, None] * ds)      # [n, P]
            acc[:, 0] = loss_sum / F
            np.add.at(acc[:, 1:], cand, g_sys)
        acc = parallel.allreduce_sum(acc.ravel(), self._st.device).reshape(C_, 1 + P)   # ONE all-reduce
        return acc[:, 0].copy(), acc[:, 1:].copy()

    # ---- Gauss-Newton / Levenberg-Marquardt model (BASELINE configs[2]) --------------------
    def getJacobianFunction(self, frequencies):
        """`jac(params) -> (fr[F], J[F, P])` with `J[f, p] = d fr_f / d p_p`: one numeric
        factorisation, one forward and ONE adjoint solve per frequency (`jpb_eval_jacobian`);
        the reference would reach the same matrix through `jax.jacobian` and P-fold batched
        solves (`_mode=4`, `Sparse.py:267-270`).  Frequencies are sharded over the ranks; every
        rank receives the full (fr, J)."""
        frequencies = np.atleast_1d(np.asarray(frequencies, dtype=np.float64))
        F = frequencies.size

        def jac(params):
            p = np.asarray(params, dtype=np.float64)
            lo, hi = parallel.shard_range(F)
            out = np.zeros((F, 1 + p.size))
            if hi > lo:
                n = hi - lo
                coef, s_, dc, ds = self.coefficients(frequencies[lo:hi], p, with_jacobian=True)
                coef, s_ = np.ascontiguousarray(coef), np.ascontiguousarray(s_)
                fr = np.zeros(n)
                ct = np.zeros((n, coef.shape[1]), dtype=np.complex128)
                cts = np.zeros(n, dtype=np.complex128)
                st = self._st
                st.check(st.lib.jpb_eval_jacobian(st.h, self.solver_num, n, _lib.ptr(coef), _lib.ptr(s_),
                                                  _lib.ptr(self.weights), _lib.ptr(fr), _lib.ptr(ct),
                                                  _lib.ptr(cts), None))
                out[lo:hi, 0] = fr
                out[lo:hi, 1:] = np.real(np.einsum("fk,fkp->fp", ct, dc) + cts[:, None] * ds)
            out = parallel.allreduce_sum(out.ravel(), self._st.device).reshape(F, -1)
            return out[:, 0].copy(), out[:, 1:].copy()
        return jac

    def getGaussNewtonModel(self, frequencies, reference_fr, func_type: str, scaling_params=None):
        """The model of `Optimizers.py:139-144` (`get_model_lvmq`): for the loss
        `mean_f r_f(p)^2` returns `model(x) -> (loss, g = 2 r^T J / n, H = 2 J^T J / n)` with
        `J = d r / d x` the residual Jacobian in the (scaled) optimisation variables."""
        frequencies = np.atleast_1d(np.asarray(frequencies, dtype=np.float64))
        reference_fr = _real_reference(reference_fr)
        if func_type not in _lib.LOSS_TYPES:
            raise ValueError(f'Function type "{func_type}" is not supported!')
        scaling = 1.0 if scaling_params is None else np.array(scaling_params, dtype=np.float64)
        jac = self.getJacobianFunction(frequencies)
        n = frequencies.size

        def model(x):
            fr, J = jac(np.asarray(x, dtype=np.float64) * scaling)
            r, dr = _residuals(fr, reference_fr, func_type)
            Jr = (dr[:, None] * J) * scaling
            return float(np.mean(r ** 2)), 2.0 * (r @ Jr) / n, 2.0 * (Jr.T @ Jr) / n
        return model


_POOL = None


def _native_pool():
    """One helper thread for blocking native calls that overlap host-side differentiation."""
    global _POOL
    if _POOL is None:
        from concurrent.futures import ThreadPoolExecutor
        _POOL = ThreadPoolExecutor(max_workers=1, thread_name_prefix="jpb-native")
    return _POOL


def _real_reference(reference_fr):
    """The reference's losses take `jnp.abs(fr - reference_fr)` and so accept a complex
    reference FRF (`Problem.py:948-975`); the unsymmetric read-out here is real and the C-ABI
    takes a real reference, so a genuinely complex one is refused instead of silently losing
    its imaginary part."""
    ref = np.asarray(reference_fr)
    if np.iscomplexobj(ref):
        if np.any(ref.imag != 0.0):
            raise ValueError("reference_fr has a non-zero imaginary part: the fused CUDA losses take a real "
                             "reference FRF (amplitudes); pass np.abs(reference_fr) or the real part")
        ref = ref.real
    return np.ascontiguousarray(ref, dtype=np.float64)


def _residuals(fr, ref, func_type):
    """r_f and d r_f / d fr_f of the four losses (`Problem.py:948-975`), loss = mean r^2."""
    if func_type == "MSE":
        return fr - ref, np.ones_like(fr)
    if func_type == "RMSE":
        return (fr - ref) / ref, 1.0 / ref
    if func_type == "MSE_AFC":
        return np.abs(fr) - np.abs(ref), np.sign(fr)
    if func_type == "MSE_LOG_AFC":
        return np.log(np.abs(fr)) - np.log(np.abs(ref)), 1.0 / fr
    raise ValueError(f'Function type "{func_type}" is not supported!')
