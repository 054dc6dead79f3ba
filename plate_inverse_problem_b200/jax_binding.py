"""JAX side of the drop-in: the reference's two primitives and the fused sweep as XLA FFI
custom calls (handlers in `csrc/xla_ffi.cc`), with the reference's own derivative rules.

jax / jaxlib are not part of this image, so this module cannot be imported or tested here;
importing it without jax raises ImportError (nothing else in the package imports it).  It is
what `source/jax_plate/Sparse.py` becomes in a deployment that has jax on a CUDA platform:

* `spsolve` / `matvec` keep the signatures of `Sparse.py:144-148, 231-236` and the batching
  `_mode` table of `:245-277` (through `vmap_method="sequential"` replaced by explicit modes),
  with `jax.custom_jvp` rules equal to `Sparse.py:162-166, 200-209`; JAX derives the transposes
  (`:168-176, 211-222`) from the linear JVPs exactly as it does for the reference's primitives;
* `sweep_loss` is the whole per-sweep body of `Problem._solve` + loss with a `custom_vjp`: one
  factorisation per frequency serves value and gradient.
"""
from __future__ import annotations

import ctypes
import os

import jax
import jax.numpy as jnp
import numpy as np

from . import _lib

_HERE = os.path.dirname(os.path.abspath(__file__))
_FFI = ctypes.CDLL(os.path.join(_HERE, "libjaxplate_b200_ffi.so"))
for _name, _sym in (("jpb_solve", "JpbSolve"), ("jpb_matvec", "JpbMatvec"),
                    ("jpb_eval_loss_grad", "JpbEvalLossGrad"), ("jpb_eval_jacobian", "JpbEvalJacobian")):
    jax.ffi.register_ffi_target(_name, jax.ffi.pycapsule(getattr(_FFI, _sym)), platform="CUDA")


def _out_shape(data, b, mode):
    if mode == 1:
        return (data.shape[0], b.shape[0])
    return b.shape


def _call(name, data, b, state: _lib.State, solver_num, transpose, mode):
    out = jax.ShapeDtypeStruct(_out_shape(data, b, mode), b.dtype)
    return jax.ffi.ffi_call(name, out)(data, b, state=np.int64(state.h.value), solver_num=np.int64(solver_num),
                                       transpose=bool(transpose), mode=np.int64(mode))


def make_primitives(state: _lib.State, solver_num: int, pattern):
    """`spsolve(data, b, transpose=False, _mode=0)` and `matvec(...)` bound to one registered
    pattern; `pattern` is `[nnz, 2]` (row, col), `_SOLVER_STATE.patterns[num]` of the reference."""
    row, col = jnp.asarray(pattern[:, 0]), jnp.asarray(pattern[:, 1])

    @jax.custom_jvp
    def matvec(mat, vec, transpose=False, _mode=0):
        return _call("jpb_matvec", mat, vec, state, solver_num, transpose, _mode)

    @matvec.defjvp
    def _matvec_jvp(primals, tangents):
        mat, vec, transpose, mode = primals
        dmat, dvec = tangents[0], tangents[1]
        out = matvec(mat, vec, transpose, mode)
        return out, matvec(dmat, vec, transpose, mode) + matvec(mat, dvec, transpose, mode)

    @jax.custom_jvp
    def spsolve(data, b, transpose=False, _mode=0):
        return _call("jpb_solve", data, b, state, solver_num, transpose, _mode)

    @spsolve.defjvp
    def _spsolve_jvp(primals, tangents):
        data, b, transpose, mode = primals
        ddata, db = tangents[0], tangents[1]
        x = spsolve(data, b, transpose, mode)
        # Sparse.py:200-209:  dx = A^-1 db - A^-1 (dA x)
        return x, spsolve(data, db - matvec(ddata, x, transpose, mode), transpose, mode)

    return spsolve, matvec, (row, col)


def make_sweep_loss(state: _lib.State, solver_num: int, ref, weights, loss_type: int, F: int):
    """`loss(coef[F, K], rhs_scale[F]) -> scalar` with a custom_vjp: the cotangents come out of
    the same native call as the value (JAX convention, no conjugation)."""
    ref, weights = jnp.asarray(ref), jnp.asarray(weights)

    def _run(coef, scale):
        B, K = coef.shape
        outs = (jax.ShapeDtypeStruct((B,), jnp.float64), jax.ShapeDtypeStruct((1,), jnp.float64),
                jax.ShapeDtypeStruct((B, K), coef.dtype), jax.ShapeDtypeStruct((B,), coef.dtype))
        return jax.ffi.ffi_call("jpb_eval_loss_grad", outs)(
            coef, scale, ref, weights, state=np.int64(state.h.value), solver_num=np.int64(solver_num),
            loss_type=np.int64(loss_type), f_total=np.int64(F))

    @jax.custom_vjp
    def sweep_loss(coef, scale):
        return _run(coef, scale)[1][0] / F

    def fwd(coef, scale):
        fr, loss_sum, ct_coef, ct_s = _run(coef, scale)
        return loss_sum[0] / F, (ct_coef, ct_s)

    def bwd(res, g):
        ct_coef, ct_s = res
        return g * ct_coef, g * ct_s

    sweep_loss.defvjp(fwd, bwd)
    return sweep_loss
