"""Material models: real parameters -> complex laminate coefficients A_ij, B_ij, D_ij
(6 each, order 11, 12, 16, 22, 26, 66), mirroring `/root/reference/source/jax_plate/Material.py`
(`Isotropic` :357-433, `Orthotropic` :436-542, `OrthotropicD4` :545-636, `SOL` :639-788,
`SymmetricalSOL` :791-885).  `tests/test_reference_pin.py` checks every class against values
produced by the reference's own classes (`tests/golden/ref_transforms.npz`).

The reference writes the transforms in jax.numpy and lets JAX differentiate them; jax is not
part of this image, so the same formulas are written with torch (complex128) and
differentiated with torch.autograd.  These are a few dozen scalar flops per evaluation and
stay on the host: only their values `coef[F, K]` and Jacobians enter the CUDA path.
Constraints / JSON I/O of the reference (`Material.py:149-355, 888-994`) are out of scope.
"""
from __future__ import annotations

import math
from typing import Callable

import numpy as np
import torch

_C = torch.complex128


def _stack6(a11, a12, a22, a66):
    z = torch.zeros((), dtype=a11.dtype)
    return torch.stack([a11, a12, z, a22, z, a66])


class Material:
    """Base class (`Material.py:35-147`): holds density, parameters and the flag `is_mps`
    (mid-plane symmetric => B = 0)."""
    density: float
    is_mps: bool = True
    param_names: tuple = ()

    def get_parameters(self):
        vals = [getattr(self, n) for n in self.param_names]
        if any(v is None for v in vals):
            return None
        return np.array(vals, dtype=np.float64)

    def has_params(self) -> bool:
        return self.get_parameters() is not None

    def get_ABD_transform(self, h: float) -> Callable:
        raise NotImplementedError

    def get_D_transform(self, h: float) -> Callable:
        abd = self.get_ABD_transform(h)
        return lambda params, *args: abd(params, *args)[2]


class Isotropic(Material):
    """`Material.py:357-433`: params (E, G, beta)."""
    param_names = ("E", "G", "beta")

    def __init__(self, density, E=None, G=None, beta=None):
        self.density, self.E, self.G, self.beta = density, E, G, beta

    def get_ABD_transform(self, h):
        def _transform(params, *args):
            E, G, beta = params[0], params[1], params[2]
            nu = E / (2.0 * G) - 1.0
            A = E * h / (1 - nu ** 2)
            D = A * h ** 2 / 12.0
            f = (1 + 1j * beta.to(_C))
            arr = torch.stack([torch.ones_like(nu), nu, torch.zeros_like(nu), torch.ones_like(nu),
                               torch.zeros_like(nu), (1 - nu) / 2]).to(_C) * f
            return A * arr, torch.zeros_like(arr), D * arr
        return _transform


class Orthotropic(Material):
    """`Material.py:436-542`: params (E1, E2, G12, nu12, beta).  `D22 = D11 / E_ratio` is the
    reference's formula (`:475`) and is kept as is."""
    param_names = ("E1", "E2", "G12", "nu12", "beta")

    def __init__(self, density, E1=None, E2=None, G12=None, nu12=None, beta=None):
        self.density = density
        self.E1, self.E2, self.G12, self.nu12, self.beta = E1, E2, G12, nu12, beta

    def get_ABD_transform(self, h):
        def _transform(params, *args):
            E1, E2, G12, nu12, beta = (params[i] for i in range(5))
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
            f = (1 + 1j * beta.to(_C))
            As = _stack6(A11, A12, A22, A66).to(_C) * f
            Ds = _stack6(D11, D12, D22, D66).to(_C) * f
            return As, torch.zeros_like(As), Ds
        return _transform


class OrthotropicD4(Material):
    """`Material.py:545-636`: params (E1, E2, G12, nu12, b1..b4), complex moduli."""
    param_names = ("E1", "E2", "G12", "nu12", "b1", "b2", "b3", "b4")

    def __init__(self, density, E1=None, E2=None, G12=None, nu12=None, b1=None, b2=None,
                 b3=None, b4=None):
        self.density = density
        self.E1, self.E2, self.G12, self.nu12 = E1, E2, G12, nu12
        self.b1, self.b2, self.b3, self.b4 = b1, b2, b3, b4

    def get_ABD_transform(self, h):
        def _transform(params, *args):
            p = params.to(_C)
            E1 = p[0] * (1 + 1j * p[4])
            E2 = p[1] * (1 + 1j * p[5])
            G12 = p[2] * (1 + 1j * p[6])
            nu12 = p[3] * (1 + 1j * p[7])
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
            As = _stack6(A11, A12, A22, A66)
            Ds = _stack6(D11, D12, D22, D66)
            return As, torch.zeros_like(As), Ds
        return _transform


def _clt_matrices(angles_deg, h):
    """Numeric evaluation of the classical-lamination matrices the reference derives with
    sympy (`Material.py:660-741`): Q -> (A, B, D) as 6x6 real matrices acting on
    (Q11, Q12, Q16, Q22, Q26, Q66); equal-thickness plies, T(theta) with n = -sin(theta).
    The reference's symbolic Q has no Q16/Q26 entries (`:664-668`), so those columns vanish."""
    n = len(angles_deg)
    zs = -h / 2 + np.arange(n + 1) * (h / n)
    MA, MB, MD = np.zeros((6, 6)), np.zeros((6, 6)), np.zeros((6, 6))
    pairs = [(0, 0), (0, 1), (0, 2), (1, 1), (1, 2), (2, 2)]
    for k, (i, j) in enumerate(pairs):
        if k in (2, 4):
            continue
        Q = np.zeros((3, 3))
        Q[i, j] = Q[j, i] = 1.0
        for a, z0, z1 in zip(angles_deg, zs[:-1], zs[1:]):
            t = a * math.pi / 180
            m, nn = math.cos(t), -math.sin(t)
            mn = m * nn
            T = np.array([[m * m, nn * nn, -2 * mn], [nn * nn, m * m, 2 * mn],
                          [mn, -mn, m * m - nn * nn]])
            QT = T @ Q @ T.T
            for r, (ri, rj) in enumerate(pairs):
                MA[r, k] += QT[ri, rj] * (z1 - z0)
                MB[r, k] += QT[ri, rj] * (z1 ** 2 - z0 ** 2) / 2
                MD[r, k] += QT[ri, rj] * (z1 ** 3 - z0 ** 3) / 3
    return MA, MB, MD


class SOL(Orthotropic):
    """Simple Orthotropic Laminate, `Material.py:639-788`."""

    def __init__(self, density, angles, E1=None, E2=None, G12=None, nu12=None, beta=None):
        super().__init__(density, E1, E2, G12, nu12, beta)
        self.angles = np.array(angles, dtype=np.float64)
        self.is_mps = not (np.sum(np.abs(self.angles - self.angles[::-1])) > 1e-6)

    def get_ABD_transform(self, h):
        MA, MB, MD = (torch.from_numpy(m).to(_C) for m in _clt_matrices(self.angles, h))

        def _transform(params, *args):
            E1, E2, G12, nu12, beta = (params[i] for i in range(5))
            den = 1 - E2 / E1 * nu12 ** 2
            z = torch.zeros_like(E1)
            Q = torch.stack([E1 / den, nu12 * E2 / den, z, E2 / den, z, G12]).to(_C)
            f = (1 + 1j * beta.to(_C))
            return (MA @ Q) * f, (MB @ Q) * f, (MD @ Q) * f
        return _transform


class SymmetricalSOL(SOL):
    """SOL whose plies have E1 = E2, `Material.py:791-885`: params (E1, G12, nu12, beta)."""
    param_names = ("E1", "G12", "nu12", "beta")

    def __init__(self, density, angles, E1=None, G12=None, nu12=None, beta=None):
        super().__init__(density, angles, E1, E1, G12, nu12, beta)

    def get_ABD_transform(self, h):
        sol = super().get_ABD_transform(h)

        def _transform(params, *args):
            return sol(torch.stack([params[0], params[0], params[1], params[2], params[3]]), *args)
        return _transform


def _abd_and_jacobian(transform, p0, omega, with_jacobian):
    """(A,B,D) stacked [18] complex and d/dp [18, P] complex."""
    def flat(p):
        A, B, D = transform(p, omega)
        return torch.cat([A, B, D])

    abd = flat(p0)
    if not with_jacobian:
        return abd, None
    fn = lambda p: torch.view_as_real(flat(p))
    try:        # one vmapped backward pass instead of 36 (same numbers, ~7x less host time)
        J = torch.autograd.functional.jacobian(fn, p0, vectorize=True)   # [18, 2, P]
    except RuntimeError:     # a user transform with ops vmap cannot batch
        J = torch.autograd.functional.jacobian(fn, p0)
    return abd, torch.complex(J[:, 0, :], J[:, 1, :])


def coefficients_and_jacobian(transform: Callable, params, freqs, inertia, with_jacobian=True):
    """coef[F, 26], rhs scale s[F] of `Problem.py:437-449` and their
    Jacobians d coef / d p [F, 26, P], d s / d p [F, P] by torch autograd (the reference
    gets them from JAX's transpose of the same expressions).  The reference's material
    transforms ignore omega (`*args`); a transform that does depend on it must carry the
    attribute `omega_independent = False` and is then evaluated per frequency."""
    I0, I0Corr, I2, I2Corr = inertia
    p0 = torch.as_tensor(np.asarray(params, dtype=np.float64))
    freqs = np.atleast_1d(np.asarray(freqs, dtype=np.float64))
    om = 2 * math.pi * freqs
    F, P = om.size, p0.numel()
    coef = np.zeros((F, 26), dtype=np.complex128)
    dc = np.zeros((F, 26, P), dtype=np.complex128) if with_jacobian else None
    if getattr(transform, "omega_independent", True):
        abd, J = _abd_and_jacobian(transform, p0, torch.as_tensor(om[0]), with_jacobian)
        coef[:, :18] = abd.numpy()[None, :]
        if with_jacobian:
            dc[:, :18, :] = J.numpy()[None, :, :]
    else:
        for i in range(F):
            abd, J = _abd_and_jacobian(transform, p0, torch.as_tensor(om[i]), with_jacobian)
            coef[i, :18] = abd.numpy()
            if with_jacobian:
                dc[i, :18, :] = J.numpy()
    w2 = om ** 2
    coef[:, [18, 20, 22]] = (-w2 * I0)[:, None]
    coef[:, [19, 21, 23]] = (-w2 * I0Corr)[:, None]
    coef[:, 24] = -w2 * I2
    coef[:, 25] = -w2 * I2Corr
    comb = np.zeros(26)
    comb[12:18] = [1.0, 2.0, 4.0, 1.0, 4.0, 4.0]      # D0 + 2 D1 + 4 D2 + D3 + 4 D4 + 4 D5
    s = coef @ comb - w2 * (I0 + I0Corr + I2 + I2Corr)
    ds = np.einsum("k,fkp->fp", comb, dc) if with_jacobian else None
    return coef, s, dc, ds
