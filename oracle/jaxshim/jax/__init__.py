"""A NumPy stand-in for the handful of `jax` entry points the reference's sampling controller uses.

TEST INFRASTRUCTURE ONLY (lives under oracle/, never imported by the product).  `jax`/`jaxlib` are not installed in
this image and cannot be (no network), so the reference's own Python source
(/root/reference/quadruped_pympc/controllers/sampling/*.py, helpers/periodic_gait_generator_jax.py) cannot be executed
as shipped.  This package lets tools/make_golden.py execute THAT SOURCE, UNMODIFIED, line by line, with NumPy float32
arithmetic in place of XLA:

  * `jax.numpy` arrays are `JArr` (an ndarray subclass) with JAX's x64-disabled type rules: every float is float32,
    every int is int32, Python scalars are weakly typed, int/int true division gives float32 (NumPy would give
    float64), results never decay to NumPy scalars, `.at[idx].set(v)` is a functional update;
  * `jax.jit` is the identity apart from the f64->f32 cast JAX applies at the jit boundary; `jax.vmap` is a Python loop
    over the mapped axis (one un-batched call of the reference function per sample); `jax.lax.fori_loop` is a Python
    loop with an int32 counter;
  * `jax.random.*` does NOT reproduce threefry bit streams: draws come from `numpy.random.RandomState` seeded from the
    key and are logged (`jax.random.DRAW_LOG`) so a test can hand the very same raw draws to the checker.

What this does and does not pin: it pins the restatement in oracle/sampling_oracle.py against the reference's own
formulas, indexing, operation order and control flow as written in its source.  It does not exercise XLA's fusion /
re-association choices (those are covered by the float32 tolerance the north_star states).
"""
from . import numpy, lax, random, config, debug  # noqa: F401
from .core import jit, vmap, devices, JArr  # noqa: F401

__version__ = '0.0-numpy-shim'
