"""Generates tests/golden/*.npz.  The reference itself cannot be imported in this
environment (jax, FreeFem++, UMFPACK, matplotlib absent -- SURVEY 0.1), so the golden
vectors come from the CPU oracle (numpy + SuperLU stand-in, refined in extended precision)
on the seeded synthetic plate "tiny".  PARITY UNPINNED at the UMFPACK boundary; the index
arrays in the fixture ARE pinned (scipy's coo->csc machinery is what the reference calls).

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np
import scipy.sparse as sps

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import PARAMS, get_case  # noqa: E402
from oracle import jax_plate_oracle as orc  # noqa: E402


def main():
    case = get_case("tiny")
    op, n = case.op, case.native
    freqs = np.linspace(40.0, 600.0, 5)
    data, coef, s = case.data(freqs)
    U = np.stack([orc.solve_extended(sps.csc_matrix((d, n["indices"], n["indptr"]), shape=(case.N,) * 2),
                                     sc * case.plate.rhs_vec) for d, sc in zip(data, s)])
    fr = op.forward(freqs, PARAMS)
    ref = fr * (1.0 + 0.1 * np.cos(np.arange(freqs.size)))
    out = dict(freqs=freqs, params=PARAMS, indices=n["indices"], indptr=n["indptr"],
               indices_T=n["indices_T"], permutation=n["permutation"], row=op.row, col=op.col,
               mats_checksum=np.array([op.mats.sum(), np.abs(op.mats).sum()]), coef=coef, s=s,
               data0=data[0], U=U, fr=fr, ref=ref)
    for lt in ("MSE", "RMSE", "MSE_AFC", "MSE_LOG_AFC"):
        L, g, aux = op.loss_and_grad(freqs, PARAMS, ref, lt)
        out[f"loss_{lt}"] = np.array(L)
        out[f"grad_{lt}"] = g
        if lt == "MSE":
            out["ct_coef_MSE"], out["ct_s_MSE"] = aux["ct_coef"], aux["ct_s"]
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "tiny_sol.npz"), **out)
    print("wrote tiny_sol.npz", {k: np.asarray(v).shape for k, v in out.items()})


if __name__ == "__main__":
    main()
