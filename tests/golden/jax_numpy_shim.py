"""NumPy stand-ins for the third-party modules the reference's Kolmogorov path imports (TEST INFRASTRUCTURE).

The reference (`/root/reference/hydrogym/jax`) is pure Python over `jax`, `flax`, `gymnax`, `chex`, `tree_math`
(pyproject.toml:76-81, all unpinned and none installable here: no network).  `install()` registers minimal
stand-ins in `sys.modules` so that the reference's OWN, UNMODIFIED source files
(`envs/kolmogorov.py`, `solvers/base.py`, `utils/utils.py`, `env_core.py`, `flow.py`, `equation.py`, `core.py`)
import and run on NumPy fp64 — the same arithmetic as JAX on CPU with `jax_enable_x64`:

* `jax.numpy`  -> numpy (same call semantics for everything the path uses: `fft.rfftn/irfftn/rfft2/fftfreq/rfftfreq`,
  `linspace`, `meshgrid(indexing="ij")`, `clip`, `mean`, `gradient`, ...), arrays wrapped in an ndarray subclass that
  carries JAX's functional `.at[idx].set(v)` (returns a NEW array, the operand is untouched — which is what makes the
  reference's `dealiasing` a no-op, utils.py:83-84);
* `jax.lax.scan`, `jax.vmap`, `jax.jit`, `jax.grad` (complex-step derivative: exact to rounding for the analytic
  `cos`/`sin` stream-function derivatives of kolmogorov.py:114-124), `jax.random.PRNGKey/split`, `jax.tree_map`;
* `flax.struct.dataclass` -> frozen dataclass with `.replace`;
* `gymnax.environments.environment.{Environment, EnvState, EnvParams}` and `spaces.Box`: the published gymnax
  semantics: `reset` = `reset_env`, `step` = `step_env` + `reset_env` + leaf-wise select on
  `done`, `discount` = 1 - is_terminal;
* `tree_math.wrap/unwrap` -> identity (they only lift pytrees to vectors; the state here is one array);
* `chex` type names; inert placeholders for modules that are imported but never touched on this path
  (`gymnasium`, `navix`, `omegaconf`, `matplotlib`).

Only `tests/golden/make_reference_golden.py` (fixture generation, in the build container where `/root/reference`
exists) and `tests/test_reference_shim.py` use this file.  Nothing in the product imports it.
"""
from __future__ import annotations

import dataclasses
import importlib
import sys
import types
import typing

import numpy as np


# ----------------------------------------------------------------------------------------------- arrays
class _At:
    def __init__(self, arr):
        self._arr = arr

    def __getitem__(self, idx):
        return _AtIdx(self._arr, idx)


class _AtIdx:
    def __init__(self, arr, idx):
        self._arr, self._idx = arr, idx

    def set(self, value):
        out = np.array(self._arr, copy=True).view(JArray)
        out[self._idx] = value
        return out

    def add(self, value):
        out = np.array(self._arr, copy=True).view(JArray)
        out[self._idx] += value
        return out


class JArray(np.ndarray):
    """ndarray with JAX's functional `.at[...]` updates."""

    @property
    def at(self):
        return _At(self)

    def block_until_ready(self):
        return self


# x64 = False mimics JAX's DEFAULT mode (jax_enable_x64 off): every float64 / complex128 result is narrowed to
# float32 / complex64 as it is produced, so the reference's arithmetic runs in single precision end to end
# (NumPy >= 2 keeps float32 through np.fft and treats Python scalars as weak).  Set through install(x64=...).
_X64 = True


def _wrap(x):
    if isinstance(x, np.ndarray):
        if not _X64 and x.dtype == np.float64:
            x = x.astype(np.float32)
        elif not _X64 and x.dtype == np.complex128:
            x = x.astype(np.complex64)
        return x if isinstance(x, JArray) else x.view(JArray)
    if not _X64 and isinstance(x, np.float64):
        return np.float32(x)
    if not _X64 and isinstance(x, np.complex128):
        return np.complex64(x)
    if isinstance(x, tuple):
        return tuple(_wrap(v) for v in x)
    if isinstance(x, list):
        return [_wrap(v) for v in x]
    return x


class _NumpyProxy(types.ModuleType):
    def __init__(self, name, target):
        super().__init__(name)
        self.__dict__["_target"] = target

    def __getattr__(self, name):
        obj = getattr(self._target, name)
        if isinstance(obj, types.ModuleType):
            return _NumpyProxy(self.__name__ + "." + name, obj)
        if callable(obj) and not isinstance(obj, type):
            def fn(*a, **k):
                return _wrap(obj(*a, **k))
            fn.__name__ = name
            return fn
        return obj


# ----------------------------------------------------------------------------------------------- pytrees
def tree_map(f, tree, *rest):
    if tree is None:
        return None
    if dataclasses.is_dataclass(tree) and not isinstance(tree, type):
        kw = {fl.name: tree_map(f, getattr(tree, fl.name), *[getattr(r, fl.name) for r in rest])
              for fl in dataclasses.fields(tree)}
        return type(tree)(**kw)
    if isinstance(tree, dict):
        return {k: tree_map(f, v, *[r[k] for r in rest]) for k, v in tree.items()}
    if isinstance(tree, (tuple, list)):
        out = [tree_map(f, v, *[r[i] for r in rest]) for i, v in enumerate(tree)]
        return type(tree)(out) if not hasattr(tree, "_fields") else type(tree)(*out)
    return f(tree, *rest)


def _tree_stack(trees):
    return tree_map(lambda *xs: _wrap(np.stack([np.asarray(x) for x in xs])), trees[0], *trees[1:])


def _tree_index(tree, i):
    return tree_map(lambda x: _wrap(np.asarray(x)[i]), tree)


# ----------------------------------------------------------------------------------------------- jax
def _scan(f, init, xs=None, length=None):
    if xs is not None:
        leaves = []
        tree_map(lambda x: leaves.append(len(x)), xs)
        length = leaves[0]
    carry, ys = init, []
    for i in range(int(length)):
        carry, y = f(carry, None if xs is None else _tree_index(xs, i))
        ys.append(y)
    if not ys or ys[0] is None:
        return carry, None
    return carry, _tree_stack(ys)


def _vmap(f, in_axes=0, out_axes=0):
    def batched(*args):
        axes = in_axes if isinstance(in_axes, (tuple, list)) else (in_axes,) * len(args)
        n = None
        for a, ax in zip(args, axes):
            if ax is not None:
                sizes = []
                tree_map(lambda x: sizes.append(np.shape(x)[0]), a)
                n = sizes[0]
                break
        outs = []
        for i in range(n):
            outs.append(f(*[a if ax is None else _tree_index(a, i) for a, ax in zip(args, axes)]))
        return _tree_stack(outs)
    return batched


def _grad(f, argnums=0):
    """Complex-step derivative Im f(x + ih)/h, h = 1e-30: exact to rounding for real-analytic f."""
    def df(*args):
        a = list(args)
        a[argnums] = complex(float(a[argnums]), 1e-30)
        return float(np.imag(f(*a)) / 1e-30)
    return df


def _jit(f=None, **kw):
    if f is None:
        return lambda g: g
    return f


def _select(pred, on_true, on_false):
    return _wrap(np.where(pred, on_true, on_false))


def _make_jax():
    jax = types.ModuleType("jax")
    jnp = _NumpyProxy("jax.numpy", np)
    lax = types.ModuleType("jax.lax")
    lax.scan, lax.select = _scan, _select
    lax.cond = lambda p, t, f, *ops: t(*ops) if p else f(*ops)
    lax.stop_gradient = lambda x: x
    random = types.ModuleType("jax.random")
    random.PRNGKey = lambda seed: _wrap(np.array([0, seed], dtype=np.uint32))
    random.key = random.PRNGKey

    def split(key, num=2):
        base = int(np.asarray(key).ravel()[-1])
        return _wrap(np.array([[base + 1, base * 7919 + i] for i in range(num)], dtype=np.uint32))
    random.split = split
    tree_util = types.ModuleType("jax.tree_util")
    tree_util.tree_map = tree_map
    jax.numpy, jax.lax, jax.random, jax.tree_util = jnp, lax, random, tree_util
    jax.tree_map = tree_map
    jax.vmap, jax.grad, jax.jit = _vmap, _grad, _jit
    jax.value_and_grad = lambda f, **k: (lambda *a: (f(*a), None))
    jax.Array = np.ndarray
    jax.config = types.SimpleNamespace(update=lambda *a, **k: None)
    return {"jax": jax, "jax.numpy": jnp, "jax.lax": lax, "jax.random": random, "jax.tree_util": tree_util}


# ----------------------------------------------------------------------------------------------- flax / gymnax / chex / tree_math
def _struct_dataclass(cls):
    cls = dataclasses.dataclass(frozen=True)(cls)
    cls.replace = lambda self, **kw: dataclasses.replace(self, **kw)
    return cls


def _make_flax():
    flax = types.ModuleType("flax")
    struct = types.ModuleType("flax.struct")
    struct.dataclass = _struct_dataclass
    struct.field = lambda pytree_node=True, **kw: dataclasses.field(**kw)
    flax.struct = struct
    return {"flax": flax, "flax.struct": struct}


def _make_gymnax(jax):
    gymnax = types.ModuleType("gymnax")
    envs = types.ModuleType("gymnax.environments")
    environment = types.ModuleType("gymnax.environments.environment")
    spaces = types.ModuleType("gymnax.environments.spaces")

    @_struct_dataclass
    class EnvState:
        time: int

    @_struct_dataclass
    class EnvParams:
        max_steps_in_episode: int = 1

    TS, TP = typing.TypeVar("TS"), typing.TypeVar("TP")

    class Environment(typing.Generic[TS, TP]):
        """gymnax.environments.environment.Environment (published semantics)."""

        @property
        def default_params(self):
            return EnvParams()

        def step(self, key, state, action, params=None):
            if params is None:
                params = self.default_params
            key, key_reset = jax.random.split(key)
            obs_st, state_st, reward, done, info = self.step_env(key, state, action, params)
            obs_re, state_re = self.reset_env(key_reset, params)
            state = tree_map(lambda x, y: _select(done, x, y), state_re, state_st)
            obs = _select(done, obs_re, obs_st)
            return obs, state, reward, done, info

        def reset(self, key, params=None):
            if params is None:
                params = self.default_params
            return self.reset_env(key, params)

        def discount(self, state, params):
            return _select(self.is_terminal(state, params), 0.0, 1.0)

    class Box:
        def __init__(self, low, high, shape, dtype=np.float32):
            self.low, self.high, self.shape, self.dtype = low, high, shape, dtype

    environment.Environment, environment.EnvState, environment.EnvParams = Environment, EnvState, EnvParams
    spaces.Box = Box
    envs.environment, envs.spaces = environment, spaces
    gymnax.environments = envs
    return {"gymnax": gymnax, "gymnax.environments": envs, "gymnax.environments.environment": environment,
            "gymnax.environments.spaces": spaces}


class _Inert(types.ModuleType):
    """Module that is imported by the reference but never used on the Kolmogorov path."""

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        obj = type(name, (), {"__init__": lambda self, *a, **k: None})
        setattr(self, name, obj)
        return obj


def _make_misc():
    chex = types.ModuleType("chex")
    chex.Array = chex.PRNGKey = chex.ArrayTree = chex.Scalar = chex.Numeric = np.ndarray
    chex.dataclass = _struct_dataclass
    tm = types.ModuleType("tree_math")
    tm.wrap = lambda f, **k: f
    tm.unwrap = lambda f, **k: f
    mods = {"chex": chex, "tree_math": tm}
    for name in ("gymnasium", "gymnasium.spaces", "navix", "omegaconf", "matplotlib", "matplotlib.animation",
                 "matplotlib.pyplot"):
        mods[name] = _Inert(name)
    mods["gymnasium"].spaces = mods["gymnasium.spaces"]
    mods["matplotlib"].animation = mods["matplotlib.animation"]
    mods["matplotlib"].pyplot = mods["matplotlib.pyplot"]
    return mods


def install(reference_root="/root/reference", x64=True):
    """Register the stand-ins (only for modules that are really absent) and put the reference on sys.path.
    ``x64=False``: single precision throughout, as JAX runs without ``jax_enable_x64``."""
    global _X64
    _X64 = bool(x64)
    mods = {}
    mods.update(_make_jax())
    mods.update(_make_flax())
    mods.update(_make_gymnax(mods["jax"]))
    mods.update(_make_misc())
    tops = sorted({name.split(".")[0] for name in mods})
    for top in tops:
        try:
            importlib.import_module(top)  # a real install wins
            continue
        except ImportError:
            pass
        for name, mod in mods.items():
            if name.split(".")[0] == top:
                sys.modules[name] = mod
    if reference_root not in sys.path:
        sys.path.insert(0, reference_root)
    return mods
