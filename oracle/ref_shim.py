"""
Import the reference's OWN observable / task / bridge modules from
/root/reference without jax, flax, optax, pint or h5py installed.

TEST INFRASTRUCTURE ONLY.  Used in the build container by
oracle/make_golden.py to generate tests/golden/*.npz and by the
`reference`-marked CPU tests; /root/reference does not exist on the GPU box, so
nothing that runs there imports this module.

How: `import swarmrl` would execute swarmrl/__init__.py (which pulls in every
sub-package and their heavy dependencies, /root/reference/swarmrl/__init__.py:7-22),
so bare parent packages whose __path__ points into the reference tree are
pre-registered, together with a ~60-line NumPy-backed `jax` stand-in
(jit = identity, vmap = a Python loop, jax.numpy = numpy).  The reference source
files themselves are executed unmodified.  Caveat: the stand-in computes in
float64 where real JAX computes in float32.
"""
from __future__ import annotations

import importlib
import pathlib
import sys
import types

import numpy as np

REFERENCE_ROOT = pathlib.Path("/root/reference")


def available() -> bool:
    return (REFERENCE_ROOT / "swarmrl" / "observables").is_dir()


def _make_jax():
    jax = types.ModuleType("jax")
    jnp = types.ModuleType("jax.numpy")
    for name in dir(np):
        if not name.startswith("__"):
            setattr(jnp, name, getattr(np, name))

    def nonzero(a, size=None, fill_value=0):
        idx = np.nonzero(a)
        if size is None:
            return idx
        out = []
        for ix in idx:
            buf = np.full(size, fill_value, dtype=ix.dtype)
            m = min(size, len(ix))
            buf[:m] = ix[:m]
            out.append(buf)
        return tuple(out)

    jnp.nonzero = nonzero
    jnp.linalg = np.linalg
    jnp.ndarray = np.ndarray

    def jit(fn=None, **kwargs):
        if fn is None:
            return lambda f: f
        return fn

    def vmap(fn, in_axes=0, out_axes=0):
        def mapped(*args):
            axes = in_axes if isinstance(in_axes, (tuple, list)) else (in_axes,) * len(args)
            n = None
            for a, ax in zip(args, axes):
                if ax is not None:
                    n = np.shape(a)[ax]
                    break
            res = []
            for i in range(n):
                call = [a if ax is None else np.take(a, i, axis=ax) for a, ax in zip(args, axes)]
                res.append(fn(*call))
            if res and isinstance(res[0], tuple):   # tuple-valued functions map to a tuple of stacked arrays, as in JAX
                return tuple(np.stack([r[k] for r in res], axis=out_axes) for k in range(len(res[0])))
            return np.stack(res, axis=out_axes)

        return mapped

    jax.jit = jit
    jax.vmap = vmap
    jax.numpy = jnp
    tree_util = types.ModuleType("jax.tree_util")
    tree_util.register_pytree_node_class = lambda cls: cls
    jax.tree_util = tree_util
    return jax, jnp, tree_util


def _bare_package(name: str, path: pathlib.Path):
    mod = types.ModuleType(name)
    mod.__path__ = [str(path)]
    mod.__package__ = name
    sys.modules[name] = mod
    return mod


_installed = False


def install():
    """Register the stand-ins. Idempotent. Raises if the reference tree is absent."""
    global _installed
    if _installed:
        return
    if not available():
        raise RuntimeError("reference tree not present (expected only in the build container)")
    if "jax" not in sys.modules:
        jax, jnp, tree_util = _make_jax()
        sys.modules["jax"] = jax
        sys.modules["jax.numpy"] = jnp
        sys.modules["jax.tree_util"] = tree_util
    if "pint" not in sys.modules:
        pint = types.ModuleType("pint")
        pint.UnitRegistry = type("UnitRegistry", (), {})
        pint.Quantity = type("Quantity", (), {})
        sys.modules["pint"] = pint
    root = REFERENCE_ROOT / "swarmrl"
    top = _bare_package("swarmrl", root)
    for sub in ("utils", "agents", "observables", "tasks", "components", "actions", "force_functions"):
        _bare_package(f"swarmrl.{sub}", root / sub)
    _bare_package("swarmrl.tasks.searching", root / "tasks" / "searching")
    # the few names the real __init__ files would have re-exported
    comp = sys.modules["swarmrl.components"]
    comp.Colloid = importlib.import_module("swarmrl.components.colloid").Colloid
    act = sys.modules["swarmrl.actions"]
    act.Action = importlib.import_module("swarmrl.actions.actions").Action
    ff = sys.modules["swarmrl.force_functions"]
    ff.ForceFunction = importlib.import_module("swarmrl.force_functions.force_fn").ForceFunction
    top.__version__ = "reference-under-shim"
    _installed = True


def ref(module: str):
    """Import `swarmrl.<module>` from the reference tree under the shim."""
    install()
    return importlib.import_module(f"swarmrl.{module}")
