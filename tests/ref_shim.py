"""Run the reference's OWN Python code (`/root/reference/source/jax_plate/*.py`) in this
container, where jax, matplotlib, FreeFem++ and UMFPACK are absent.  TEST INFRASTRUCTURE ONLY.

What is the reference's and what is a stand-in when the shim is active:

* `jax_plate.Sparse`, `jax_plate.Problem`, `jax_plate.Material`, `jax_plate.Accelerometer`,
  `jax_plate.Optimizers`, `jax_plate.Input` ... are imported from `/root/reference/source`
  UNMODIFIED (nothing is copied into this repository);
* `jax_plate.jax_plate_lib` is the reference's pybind11 module compiled from its own sources by
  `oracle/Makefile` (`oracle/_ref/jax_plate_lib*.so`): `InnerState.h`, `umfpack_interface.h`,
  `csc_matvec.h`, `main.cpp` as they lie, linked against the dense-LU stand-in for the UMFPACK
  library functions (`oracle/umfpack_standin/`) -- or, for sizes where a dense LU is too slow,
  `oracle.jax_plate_oracle.OracleInnerState` (scipy SuperLU);
* `jax` is a small numpy-backed shim: `jax.numpy` = numpy, `jit` = identity, `vmap` = a Python
  loop over the mapped axis, `core.Primitive.bind` = abstract-eval checks + the primitive's
  registered CPU lowering, whose `mlir.emit_python_callback` simply calls the callback.  The
  reference's JVP / transpose / batching rule FUNCTIONS are collected exactly as registered
  (`ad.defjvp`, `ad.primitive_transposes`, `batching.primitive_batchers`) so tests can call
  them; there is no tracing and no automatic differentiation (`jax.grad` of a function works
  only if the function carries its own `.value_and_grad`, as this repository's losses do);
* `matplotlib`, `pylab`, `pyFreeFem` are inert stubs (plots / FreeFem++ are out of scope);
  `load_matrices_unsymm` is replaced by a function returning a synthetic plate in the same
  tuple format (`pyFFInterface.py:503-509`).

Used by `tests/golden/make_reference_golden.py` (fixtures that travel to the GPU box) and by
`tests/test_reference_pin.py` (live comparison when `/root/reference` is present).
"""
from __future__ import annotations

import functools
import glob
import importlib
import importlib.util
import os
import sys
import types

import numpy as np

REF_SRC = "/root/reference/source"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def available() -> bool:
    return os.path.isdir(os.path.join(REF_SRC, "jax_plate"))


# ------------------------------------------------------------------------------ jax shim
class _Inert(types.ModuleType):
    """Module whose every attribute is an inert callable / namespace."""

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        obj = _InertObj(f"{self.__name__}.{name}")
        setattr(self, name, obj)
        return obj


class _InertObj:
    def __init__(self, name):
        self._name = name

    def __call__(self, *a, **k):
        return _InertObj(self._name + "()")

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _InertObj(f"{self._name}.{name}")

    def __iter__(self):
        return iter(())


class ShapedArray:
    def __init__(self, shape, dtype):
        self.shape, self.dtype = tuple(shape), np.dtype(dtype)


class UndefinedPrimal:
    """`jax.interpreters.ad.UndefinedPrimal`: marks the linear argument of a transpose rule."""

    def __init__(self, aval):
        self.aval = aval


class Zero:
    def __init__(self, aval=None):
        self.aval = aval


class _Ctx:
    def __init__(self, avals_in, avals_out):
        self.avals_in, self.avals_out = avals_in, avals_out


class Primitive:
    """`jax.core.Primitive` reduced to what `Sparse.py` uses at import time, with `bind`
    evaluating eagerly: abstract-eval (the reference's own shape / type checks), then the
    reference's own CPU lowering, whose host callback does the work."""
    multiple_results = False

    def __init__(self, name):
        self.name = name
        self.abstract_eval = None
        self.lowering = None
        self.jvps = None

    def def_impl(self, impl):
        self.impl = impl

    def def_abstract_eval(self, fn):
        self.abstract_eval = fn

    def bind(self, *args, **params):
        args = [np.asarray(a) for a in args]
        out_aval = self.abstract_eval(*args, **params)
        ctx = _Ctx([ShapedArray(a.shape, a.dtype) for a in args], [out_aval])
        res = self.lowering(ctx, *args, **params)
        out = np.asarray(res[0])
        assert tuple(out.shape) == tuple(out_aval.shape), (self.name, out.shape, out_aval.shape)
        return out


class _At:
    def __init__(self, arr, idx=None):
        self.arr, self.idx = arr, idx

    def __getitem__(self, idx):
        return _At(self.arr, idx)

    def set(self, value):
        out = np.array(self.arr)
        out[self.idx] = value
        return out.view(_JArr)


class _JArr(np.ndarray):
    """ndarray with jax's functional update `x.at[idx].set(v)` (`Optimizers.py:54-55`)."""

    @property
    def at(self):
        return _At(self)


def _jnp_array(*a, **k):
    return np.array(*a, **k).view(_JArr)


def _vmap(fn, in_axes=0, out_axes=0):
    def mapped(*args):
        axes = in_axes if isinstance(in_axes, (tuple, list)) else (in_axes,) * len(args)
        n = next(np.asarray(a).shape[ax] for a, ax in zip(args, axes) if ax is not None)
        outs = [fn(*[np.take(np.asarray(a), i, axis=ax) if ax is not None else a
                     for a, ax in zip(args, axes)]) for i in range(n)]
        return np.stack(outs, axis=out_axes)
    return mapped


def _jit(fn=None, **kw):
    if fn is None:
        return lambda f: f
    return fn


def _value_and_grad(f, *a, **k):
    if hasattr(f, "value_and_grad"):
        return f.value_and_grad
    raise NotImplementedError("ref_shim: no automatic differentiation (the function needs .value_and_grad)")


def _grad(f, *a, **k):
    if hasattr(f, "grad_fn"):          # an object that may in turn carry .jacobian (Hessian)
        return f.grad_fn
    vg = _value_and_grad(f)
    return lambda x: vg(x)[1]


def _no_ad(*a, **k):
    raise NotImplementedError("ref_shim: no automatic differentiation")


def _jacobian(g, *a, **k):
    if hasattr(g, "jacobian"):
        return g.jacobian
    return _no_ad()


def _install_stubs(inner_state_cls):
    jnp = types.ModuleType("jax.numpy")
    jnp.__getattr__ = lambda name: getattr(np, name)
    jnp.array = _jnp_array
    jax = types.ModuleType("jax")
    jax.numpy = jnp
    jax.Array = type("Array", (), {})       # annotations only; must NOT be np.ndarray (scipy's array-api probes)
    jax.jit = _jit
    jax.vmap = _vmap
    jax.value_and_grad = _value_and_grad
    jax.grad = _grad
    jax.jacobian = _jacobian
    jax.hessian = jax.jacfwd = jax.jacrev = _no_ad
    jax.config = types.SimpleNamespace(update=lambda *a, **k: None)
    tree_util = types.ModuleType("jax.tree_util")
    tree_util.Partial = functools.partial
    jax.tree_util = tree_util
    core = types.ModuleType("jax.core")
    core.Primitive = Primitive
    core.ShapedArray = ShapedArray
    core.Tracer = type("Tracer", (), {})
    jax.core = core
    interp = types.ModuleType("jax.interpreters")
    ad = types.ModuleType("jax.interpreters.ad")
    ad.primitive_transposes = {}
    ad.UndefinedPrimal = UndefinedPrimal
    ad.Zero = Zero
    ad.is_undefined_primal = lambda x: isinstance(x, UndefinedPrimal)

    def defjvp(prim, *rules):
        prim.jvps = rules
    ad.defjvp = defjvp
    batching = types.ModuleType("jax.interpreters.batching")
    batching.primitive_batchers = {}
    mlir = types.ModuleType("jax.interpreters.mlir")

    def register_lowering(prim, fn, platform=None):
        prim.lowering = fn

    def emit_python_callback(ctx, callback, token, args, avals_in, avals_out, has_side_effect=False, **kw):
        return callback(*args), None, None
    mlir.register_lowering = register_lowering
    mlir.emit_python_callback = emit_python_callback
    xla = types.ModuleType("jax.interpreters.xla")
    xla.apply_primitive = lambda prim, *a, **k: prim.bind(*a, **k)
    interp.ad, interp.batching, interp.mlir, interp.xla = ad, batching, mlir, xla
    jax.interpreters = interp
    import scipy.linalg
    jscipy = types.ModuleType("jax.scipy")
    jscipy.linalg = scipy.linalg
    jax.scipy = jscipy
    mods = {"jax": jax, "jax.numpy": jnp, "jax.tree_util": tree_util, "jax.core": core,
            "jax.interpreters": interp, "jax.interpreters.ad": ad, "jax.interpreters.batching": batching,
            "jax.interpreters.mlir": mlir, "jax.interpreters.xla": xla, "jax.scipy": jscipy,
            "jax.scipy.linalg": scipy.linalg}
    for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.tri", "pylab", "pyFreeFem"):
        mods[name] = _Inert(name)
    mods["matplotlib"].pyplot = mods["matplotlib.pyplot"]
    mods["matplotlib"].tri = mods["matplotlib.tri"]
    lib = types.ModuleType("jax_plate.jax_plate_lib")
    lib.InnerState = inner_state_cls
    mods["jax_plate.jax_plate_lib"] = lib
    sys.modules.update(mods)
    return mods


def compiled_reference_lib():
    """The reference's pybind11 module built from its sources by oracle/Makefile (dense-LU
    stand-in for the UMFPACK functions), or None when it has not been built."""
    cands = glob.glob(os.path.join(ROOT, "oracle", "_ref", "jax_plate_lib*.so"))
    if not cands:
        return None
    spec = importlib.util.spec_from_file_location("jax_plate_lib", cands[0])
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


class _OracleBackedInnerState:
    """`InnerState` with scipy SuperLU behind it (for sizes the dense stand-in cannot do)."""

    def __init__(self):
        sys.path.insert(0, ROOT)
        from oracle.jax_plate_oracle import OracleInnerState
        self._o = OracleInnerState()

    def add_mat(self, N, indices, indptr, indices_T, indptr_T, permutation, data):
        self._o.add_mat(N, indices, indptr, indices_T, indptr_T, permutation)

    def solve(self, data, b, solver_num, transpose, n_cpu, mode):
        return self._o.solve(data, b, solver_num, transpose, n_cpu, mode)

    def matvec(self, mat, vec, solver_num, transpose, n_cpu, mode):
        return self._o.matvec(mat, vec, solver_num, transpose, n_cpu, mode)


_LOADED = {}


def load(backend: str = "compiled"):
    """Import the reference package under the shim.  backend: "compiled" (the reference's own
    native module, dense-LU stand-in for UMFPACK) or "superlu" (oracle-backed InnerState).
    Returns the `jax_plate` package (attributes Sparse, Problem, Material, ...).  The shim
    replaces `jax` & co. in `sys.modules` for the rest of the process."""
    if not available():
        raise RuntimeError("/root/reference is not present")
    if _LOADED:
        if _LOADED["backend"] != backend:
            raise RuntimeError("ref_shim: one backend per process")
        return _LOADED["pkg"]
    if backend == "compiled":
        lib = compiled_reference_lib()
        if lib is None:
            raise RuntimeError("oracle/_ref/jax_plate_lib*.so has not been built (make -C oracle)")
        cls = lib.InnerState
    else:
        cls = _OracleBackedInnerState
    for name in [m for m in sys.modules if m == "jax" or m.startswith("jax.") or m.startswith("jax_plate")]:
        del sys.modules[name]
    _install_stubs(cls)
    if REF_SRC not in sys.path:
        sys.path.insert(0, REF_SRC)
    pkg = importlib.import_module("jax_plate")
    _LOADED.update(pkg=pkg, backend=backend)
    return pkg


class FakeGeometry:
    """What `Problem.__init__` reads of a `Geometry` on the unsymmetric path: `.height` and
    `.current_file` (handed to `load_matrices_unsymm`, which the shim replaces)."""

    def __init__(self, height, plate):
        self.height = height
        self.current_file = plate


def reference_problem(pkg, plate, material, accelerometer, height):
    """`jax_plate.Problem.Problem(geometry, material, accel)` executed by the reference's own
    constructor (`Problem.py:43-374`) on a synthetic plate: only `load_matrices_unsymm` is
    substituted (FreeFem++ is absent), returning the tuple of `pyFFInterface.py:503-509`."""
    P = pkg.Problem

    def fake_load(geom_file):
        t = geom_file.as_tuple() if hasattr(geom_file, "as_tuple") else geom_file
        mats = [m.copy() for m in t[0]]
        return (mats,) + tuple(t[1:])
    P.load_matrices_unsymm = fake_load
    return P.Problem(FakeGeometry(height, plate), material, accelerometer)


def reference_symm_problem(pkg, ff_dense, material, height):
    """The reference's mid-plane symmetric branch (`Problem.py:228-308, 391-417`) executed as
    written, on the dense dict of `synthetic.make_bending_plate(..., dense=True)` in place of
    `load_matrices_symm`'s.  The branch needs `self.accelerometer is None`, which the reference's
    own constructor refuses (`Problem.py:91-94` and `:201-204` raise when an accelerometer is
    missing) -- it is dead code there.  To run it unmodified, a subclass pins the attribute
    `accelerometer` to None and the constructor gets a placeholder object."""
    P = pkg.Problem
    P.load_matrices_symm = lambda geom_file: geom_file

    class SymmetricBranch(P.Problem):
        accelerometer = property(lambda self: None, lambda self, value: None)

    return SymmetricBranch(FakeGeometry(height, ff_dense), material, object())
