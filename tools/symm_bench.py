"""End-to-end timing of the mid-plane symmetric (bending-only) path through the host API
(`loss.value_and_grad`, host buffers) next to the oracle's per-frequency CPU cost.
Usage: python tools/symm_bench.py [nx ny nfreq]   (default 200 16 400)"""
import json, sys, time
import numpy as np
sys.path.insert(0, ".")
from plate_inverse_problem_b200 import Material, synthetic
from plate_inverse_problem_b200.Problem import Geometry, Problem
from oracle import jax_plate_oracle as orc

nx, ny, F = (int(a) for a in (sys.argv[1:4] + ["200", "16", "400"][len(sys.argv) - 1:]))
H, RHO = 2e-3, 1600.0
p = np.array([130e9, 9e9, 5e9, 0.3, 0.01])
ff = synthetic.make_bending_plate(nx, ny)
prob = Problem(Geometry(H, ff), Material.Orthotropic(RHO, *p))
freqs = np.linspace(40.0, 600.0, F)
ref = prob.solveForward(freqs, p * np.array([1.1, 1.1, 1.2, 1.0, 1.5]))
loss = prob.getLossFunction(freqs, ref, "MSE")
for _ in range(3):
    L, g = loss.value_and_grad(p)
t0 = time.perf_counter()
K = 5
for _ in range(K):
    L, g = loss.value_and_grad(p)
dt = (time.perf_counter() - t0) / K
op = orc.OracleSymmProblem(ff, orc.make_d_transform("orthotropic", H), RHO, H)
sub = freqs[:: max(1, F // 8)][:8]
t0 = time.perf_counter()
Lo, go, _ = op.loss_and_grad(sub, p, ref[:: max(1, F // 8)][:8], "MSE")
cpu = (time.perf_counter() - t0) / sub.size
print(json.dumps({"workload": f"bending {nx}x{ny}", "N": prob.mat_size, "freqs": F, "ms_per_sweep": dt * 1e3,
                  "evals_per_s_e2e": F / dt, "cpu_oracle_s_per_eval_1core": cpu, "info": prob.symbolic_info(),
                  "timings": prob.timings()}))
