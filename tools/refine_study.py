"""Development study (GPU box): what does the refinement step of the triangular solves buy?
Forward responses, loss and gradient of the cfg1 sweep at 0, 1 and 2 refinement steps."""
import os, sys, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from conftest import get_case, PARAMS

name = sys.argv[1] if len(sys.argv) > 1 else "cfg1"
F = int(sys.argv[2]) if len(sys.argv) > 2 else 400
case = get_case(name)
freqs = np.linspace(40.0, 600.0, F)
pert = PARAMS * (1 + np.array([0.1, 0.1, 0.2, 0.1, 0.1]))
out = {}
# 2 / 1: unconditional steps (1 = default); 0: fast mode (residual pass; one step for the systems
# whose backward error is above 1e-15); -1: no residual pass at all
for r in (2, 1, 0, -1):
    prob = case.problem(refinement=max(r, 0), refinement_threshold=0.0 if r == -1 else None)
    ref = prob.solveForward(freqs, pert)
    fr, y, u = prob._forward_raw(freqs, PARAMS, return_u=True)
    loss = prob.getLossFunction(freqs, ref, "MSE_LOG_AFC")
    L, g = loss.value_and_grad(PARAMS)
    print(f"policy {r}: vectors refined / solved in the loss+gradient call: {prob._st.refined()}, "
          f"phases {dict((k, round(v, 2)) for k, v in prob.timings().items() if k.endswith('_ms'))}")
    out[r] = (fr, u, L, g)
    del prob
fr2, u2, L2, g2 = out[2]
for r in (1, 0, -1):
    fr, u, L, g = out[r]
    eu = np.max(np.abs(u - u2), axis=1) / np.max(np.abs(u2), axis=1)
    print(f"refine={r}: max rel err u {eu.max():.2e} (median {np.median(eu):.2e}, #>1e-10: {(eu > 1e-10).sum()}/{F}), "
          f"fr {np.max(np.abs(fr - fr2) / np.abs(fr2)):.2e}, loss {abs(L - L2) / abs(L2):.2e}, "
          f"grad {np.max(np.abs(g - g2)) / np.max(np.abs(g2)):.2e}")

# the worst systems against an extended-precision solve (the oracle's own yardstick)
from oracle import jax_plate_oracle as orc
coef, s = orc.coefficients(freqs, PARAMS, case.transform, case.inertia)
eu1 = np.max(np.abs(out[1][1] - u2), axis=1) / np.max(np.abs(u2), axis=1)
worst = np.argsort(eu1)[::-1][:3].tolist() + [int(np.argsort(eu1)[F // 2])]
for i in worst:
    A = case.op.matrix(coef[i])
    b = s[i] * case.plate.rhs_vec.astype(np.complex128)
    ue = orc.solve_extended(A, b)
    xo = orc.umfpack_standin_lu(A).solve(b)
    rel = lambda x: float(np.max(np.abs(x - ue)) / np.max(np.abs(ue)))
    res = lambda x: float(np.linalg.norm(A @ x - b) / np.linalg.norm(b))
    print(f"f[{i}]={freqs[i]:.2f} Hz: err vs extended  none {rel(out[-1][1][i]):.2e} default {rel(out[0][1][i]):.2e}  "
          f"r1 {rel(out[1][1][i]):.2e}  r2 {rel(u2[i]):.2e}  SuperLU {rel(xo):.2e} | resid none {res(out[-1][1][i]):.2e} "
          f"default {res(out[0][1][i]):.2e} r1 {res(out[1][1][i]):.2e} r2 {res(u2[i]):.2e} SuperLU {res(xo):.2e}")
