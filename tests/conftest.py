import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


H, RHO = 2e-3, 1600.0
PARAMS = np.array([130e9, 9e9, 5e9, 0.3, 0.01])
ANGLES = (0, 45, -45, 90)


class Case:
    """A synthetic plate + the oracle problem on it (SURVEY 8d inputs)."""

    def __init__(self, nx, ny, nd_ordering=False):
        from oracle import jax_plate_oracle as orc
        from plate_inverse_problem_b200.Problem import AP1030
        from plate_inverse_problem_b200.synthetic import make_plate
        self.plate = make_plate(nx, ny)
        self.acc = AP1030
        self.inertia = orc.inertia_scalars(H, RHO, AP1030.mass, AP1030.radius, AP1030.height)
        self.transform = orc.make_transform("sol", H, ANGLES)
        self.op = orc.OracleProblem(self.plate, self.transform, self.inertia, AP1030.height,
                                    AP1030.effective_height, AP1030.transverse_sensitivity)
        self.N = self.plate.N
        self.native = self.op.native
        if nd_ordering:
            # SuperLU's own MMD ordering does not finish at N ~ 1e5; the oracle then factors
            # P A P^T with the nested-dissection ordering of the (host-only) symbolic phase
            from plate_inverse_problem_b200 import _lib
            self.op.ordering = _lib.host_symbolic(self.N, self.native["indices"], self.native["indptr"])["perm"]

    def problem(self, **kw):
        from plate_inverse_problem_b200.Material import SOL
        from plate_inverse_problem_b200.Problem import Geometry, Problem
        return Problem(Geometry(H, self.plate), SOL(RHO, ANGLES, *PARAMS), self.acc, **kw)

    def data(self, freqs):
        from oracle import jax_plate_oracle as orc
        coef, s = orc.coefficients(freqs, PARAMS, self.transform, self.inertia)
        return np.stack([orc.assemble(c, self.op.mats) for c in coef]), coef, s


_cases = {}


def get_case(name):
    from plate_inverse_problem_b200.synthetic import NAMED_SIZES
    if name not in _cases:
        _cases[name] = Case(*NAMED_SIZES[name], nd_ordering=name in ("40k", "cfg4", "cfg5"))
    return _cases[name]


@pytest.fixture(scope="session")
def tiny():
    return get_case("tiny")


@pytest.fixture(scope="session")
def small():
    return get_case("small")


def relerr(a, b):
    return float(np.max(np.abs(np.asarray(a) - np.asarray(b))) / np.max(np.abs(b)))
