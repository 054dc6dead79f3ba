"""Mid-plane symmetric (bending-only) plates: the branch `material.is_mps and accelerometer is
None` of the reference's `Problem` (`/root/reference/source/jax_plate/Problem.py:228-308` set-up,
`:391-417` per-frequency body, `:948-975` losses on the COMPLEX read-out), executed by the same
fused CUDA path as the unsymmetric one.

Per frequency the reference solves, on the free DOFs of one Morley field w,

    A u = b,   A = sum_k D_k(p, w) Ks_k / 2e - w^2 MInertia,   b = sum_k D_k fKs_k / 2e - w^2 fInertia,
    fr = c0 + c . u                                               (complex)

i.e. the right-hand side is the SAME linear combination of seven vectors as the matrix is of seven
matrices.  The fused path takes `A = sum_k coef_k M_k`, `rhs = s * rhs_vec` and `y = Q u`, so the
system is bordered with one unknown t = 1:

    [ A  -b ] [u]   [0]          M_k = [ Ks_k/2e  -fKs_k/2e ]  (k < 6),  M_6 = [ MInertia  -fInertia ],
    [ 0   1 ] [t] = [1]                [   0          0     ]                  [    0          0     ]
                                 M_7 = e_t e_t^T (coefficient 1),  rhs_vec = e_t, s = 1,  Q = [c, c0].

With that, forward sweeps, losses, gradients, candidate batches, Jacobians and the multi-GPU
sharding of `Problem` apply unchanged (K = 8 components, one read-out row, N = n_free + 1).  The
reference densifies its matrices (`pyFFInterface.py:99`) and so stops at a few thousand DOFs;
here they stay sparse.  `MSE` / `RMSE` compare complex numbers (loss types 5 / 6 of the C-ABI),
`MSE_AFC` / `MSE_LOG_AFC` amplitudes (types 2 / 3 with |reference|).
"""
from __future__ import annotations

import math

import numpy as np
import scipy.sparse as sps
import torch

from . import _lib, parallel
from .Problem import Problem, _residuals


def bordered_components(ff: dict, rho: float, e: float):
    """The eight (n+1) x (n+1) component matrices, in the order of the coefficient vector
    `[D11, D12, D16, D22, D26, D66, -w^2, 1]`."""
    n = ff["M"].shape[0]
    MInertia = rho * (sps.csr_matrix(ff["M"]) + 1.0 / 3.0 * e ** 2 * sps.csr_matrix(ff["L"]))     # Problem.py:267
    fInertia = rho * (np.asarray(ff["fM"]) + 1.0 / 3.0 * e ** 2 * np.asarray(ff["fL"]))           # :269

    def border(K, f):
        K = sps.coo_matrix(K)
        nzf = np.nonzero(f)[0]
        rows = np.concatenate([K.row, nzf])
        cols = np.concatenate([K.col, np.full(nzf.size, n)])
        data = np.concatenate([K.data, -f[nzf]])
        return sps.coo_matrix((data, (rows, cols)), shape=(n + 1, n + 1))

    mats = [border(sps.csr_matrix(ff["Ks"][k]) / 2.0 / e, np.asarray(ff["fKs"][k]) / 2.0 / e) for k in range(6)]
    mats.append(border(MInertia, fInertia))
    mats.append(sps.coo_matrix(([1.0], ([n], [n])), shape=(n + 1, n + 1)))
    return mats


def d_coefficients_and_jacobian(transform, params, freqs, with_jacobian=True):
    """coef[F, 8] = [D_k(p, w) (6), -w^2, 1], the rhs scale s = 1 and their Jacobians by torch
    autograd (the reference differentiates the same `get_D_transform` with JAX)."""
    p0 = torch.as_tensor(np.asarray(params, dtype=np.float64))
    om = 2 * math.pi * np.atleast_1d(np.asarray(freqs, dtype=np.float64))
    F, P = om.size, p0.numel()
    coef = np.zeros((F, 8), dtype=np.complex128)
    dc = np.zeros((F, 8, P), dtype=np.complex128) if with_jacobian else None

    def one(w):
        D = transform(p0, torch.as_tensor(w))
        J = None
        if with_jacobian:
            Jr = torch.autograd.functional.jacobian(lambda p: torch.view_as_real(transform(p, torch.as_tensor(w))), p0)
            J = torch.complex(Jr[:, 0, :], Jr[:, 1, :]).numpy()
        return D.detach().numpy(), J

    if getattr(transform, "omega_independent", True):
        D, J = one(om[0])
        coef[:, :6] = D[None, :]
        if with_jacobian:
            dc[:, :6, :] = J[None]
    else:
        for i in range(F):
            D, J = one(om[i])
            coef[i, :6] = D
            if with_jacobian:
                dc[i, :6, :] = J
    coef[:, 6] = -om ** 2
    coef[:, 7] = 1.0
    s = np.ones(F, dtype=np.complex128)
    ds = np.zeros((F, P), dtype=np.complex128) if with_jacobian else None
    return coef, s, dc, ds


class SymmetricProblem(Problem):
    """`Problem(geometry, material)` without an accelerometer and with a mid-plane symmetric
    material; `geometry.matrices` is the dict of `load_matrices_symm` (`pyFFInterface.py:149-166`,
    e.g. `synthetic.make_bending_plate`; sparse or dense matrices)."""

    def __init__(self, geometry, material, accelerometer=None, ref_fr=None, **kw):
        if accelerometer is not None or not material.is_mps:
            raise ValueError("the symmetric path needs a mid-plane symmetric material and no accelerometer "
                             "(Problem.py:228)")
        ff = geometry.matrices
        self.geometry, self.material, self.accelerometer = geometry, material, None
        self.reference_fr = ref_fr
        self.n_cpu = kw.pop("cpu", 0)
        self.rho = material.density
        self.parameters = material.get_parameters()
        self.e = geometry.height / 2.0
        self.n_free = n = ff["M"].shape[0]
        self.interpolation_vector = np.asarray(ff["interpolation_vector"], dtype=np.float64)
        self.interpolation_value_from_bc = float(ff["interpolation_value_from_bc"])
        self.test_point = ff["test_point_coord"]
        self.constrained_idx = ff["constrained_idx"]
        self.mesh = ff["mesh"]
        self.boundary_value = ff["boundary_value"]
        self.transform = material.get_D_transform(geometry.height)
        rhs_vec = np.zeros(n + 1)
        rhs_vec[n] = 1.0
        Q = np.concatenate([self.interpolation_vector, [self.interpolation_value_from_bc]])[None, :]
        self._setup_native(bordered_components(ff, self.rho, self.e), rhs_vec, Q, np.array([1.0]), **kw)

    def coefficients(self, freqs, params, with_jacobian=False):
        return d_coefficients_and_jacobian(self.transform, params, freqs, with_jacobian)

    # ---- forward: the complex amplitude in the test point (Problem.py:415) -----------------------
    def getFRFunction(self):
        def fr_function(freqs, params):
            freqs = np.atleast_1d(np.asarray(freqs, dtype=np.float64))
            lo, hi = parallel.shard_range(freqs.size)
            full = np.zeros(2 * freqs.size)
            if hi > lo:
                y = self._forward_raw(freqs[lo:hi], params)[1][:, 0]
                full[2 * lo:2 * hi] = np.ascontiguousarray(y).view(np.float64)
            return parallel.allreduce_sum(full, self._st.device).view(np.complex128)
        return fr_function

    def solveFree(self, freqs, params=None):
        """u on the free DOFs, [F, n_free] (what `getModePicture`'s `_solve` returns, Problem.py:551-563)."""
        params = self.parameters if params is None else params
        return self._forward_raw(np.atleast_1d(freqs), np.asarray(params, dtype=np.float64), return_u=True)[2][:, :-1]

    # ---- losses ------------------------------------------------------------------------------------
    def _loss_reference(self, reference_fr, func_type):
        ref = np.asarray(reference_fr)
        if func_type in ("MSE", "RMSE"):
            return np.ascontiguousarray(ref, dtype=np.complex128).view(np.float64), _lib.COMPLEX_LOSS_TYPES[func_type], 2
        return np.ascontiguousarray(np.abs(ref), dtype=np.float64), _lib.COMPLEX_LOSS_TYPES[func_type], 1

    def _loss_value(self, fr, reference_fr, func_type):
        ref = np.asarray(reference_fr)
        if func_type == "MSE":
            return float(np.mean(np.abs(fr - ref) ** 2))
        if func_type == "RMSE":
            return float(np.mean(np.abs((fr - ref) / ref) ** 2))
        return float(np.mean(_residuals(np.abs(fr), np.abs(ref), func_type)[0] ** 2))

    def getGaussNewtonModel(self, frequencies, reference_fr, func_type, scaling_params=None):
        if func_type in ("MSE", "RMSE"):
            def model(x):
                raise NotImplementedError("the Gauss-Newton model of the complex MSE / RMSE losses is not built: "
                                          "use MSE_AFC / MSE_LOG_AFC or a gradient method")
            return model
        return super().getGaussNewtonModel(frequencies, np.abs(np.asarray(reference_fr)), func_type, scaling_params)
