"""jax.random subset.  NOT threefry: numpy RandomState streams keyed by the key words; every draw is logged."""
import numpy as _np

from .core import base, canon

DRAW_LOG = []     # [(kind, key words, seed, raw draws)]


def _seed_of(key):
    k = _np.asarray(base(key)).astype(_np.uint64).ravel()
    return int((k[0] * _np.uint64(0x9E3779B1) + k[1] * _np.uint64(0x85EBCA77) + _np.uint64(12345)) % _np.uint64(2**32 - 1))


def PRNGKey(seed):
    return canon(_np.array([0, int(seed) & 0xFFFFFFFF], dtype=_np.uint32))


def split(key, num=2):
    rs = _np.random.RandomState(_seed_of(key) ^ 0x5BD1E995)
    return canon(rs.randint(0, 2**32 - 1, size=(num, 2), dtype=_np.uint64).astype(_np.uint32))


def normal(key, shape=(), dtype=None):
    seed = _seed_of(key)
    raw = _np.random.RandomState(seed).standard_normal(shape).astype(_np.float32)
    DRAW_LOG.append(('normal', _np.asarray(base(key)).copy(), seed, raw))
    return canon(raw)


def uniform(key, shape=(), dtype=None, minval=0.0, maxval=1.0):
    seed = _seed_of(key)
    u = _np.random.RandomState(seed ^ 0x2545F491).random_sample(shape).astype(_np.float32)
    u = _np.minimum(u, _np.nextafter(_np.float32(1), _np.float32(0)))
    DRAW_LOG.append(('uniform01', _np.asarray(base(key)).copy(), seed, u))
    lo, hi = _np.float32(minval), _np.float32(maxval)
    return canon(_np.maximum(lo, u * (hi - lo) + lo))           # jax.random.uniform's own formula


def choice(key, a, shape=()):
    seed = _seed_of(key)
    a = _np.asarray(base(a))
    idx = _np.random.RandomState(seed ^ 0x1B873593).randint(0, a.shape[0], size=shape)
    DRAW_LOG.append(('choice', _np.asarray(base(key)).copy(), seed, idx.astype(_np.int32)))
    return canon(a[idx])


def randint(key, shape, minval, maxval):
    seed = _seed_of(key)
    return canon(_np.random.RandomState(seed ^ 0x68E31DA4).randint(minval, maxval, size=shape).astype(_np.int32))
