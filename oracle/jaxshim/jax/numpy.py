"""jax.numpy subset (see package docstring)."""
import numpy as _np

from .core import JArr, base, canon, _canon_dtype

ndarray = JArr
newaxis = None
pi = _np.pi
inf = _np.inf
nan = _np.nan


def _dt(dtype):
    return None if dtype is None else _canon_dtype(dtype)


def _w(f):
    def g(*a, **k):
        return canon(f(*[base(x) for x in a], **{kk: base(v) for kk, v in k.items()}))
    g.__name__ = f.__name__
    return g


class _Scalar:
    """jnp.float32 / jnp.int32 / jnp.int16: usable as a constructor and as a dtype."""

    def __init__(self, np_type):
        self.np_type = np_type
        self.dtype = _np.dtype(np_type)

    def __call__(self, x):
        return canon(_np.asarray(_np.asarray(base(x)).astype(self.np_type)))


float32, int32, int16, uint32, bool_ = (_Scalar(_np.float32), _Scalar(_np.int32), _Scalar(_np.int16),
                                        _Scalar(_np.uint32), _Scalar(_np.bool_))


def _resolve(dtype):
    if isinstance(dtype, _Scalar):
        return dtype.dtype
    return _dt(dtype)


def array(obj, dtype=None, **_kw):
    return canon(_np.array(base(obj), dtype=_resolve(dtype)))


asarray = array


def zeros(shape, dtype=None):
    return canon(_np.zeros(shape, dtype=_resolve(dtype) or _np.float32))


def ones(shape, dtype=None):
    return canon(_np.ones(shape, dtype=_resolve(dtype) or _np.float32))


def identity(n, dtype=None):
    return canon(_np.identity(n, dtype=_resolve(dtype) or _np.float32))


def eye(n, dtype=None):
    return identity(n, dtype)


def arange(*a, dtype=None):
    return canon(_np.arange(*a, dtype=_resolve(dtype)))


def linspace(start, stop, num=50, dtype=None):
    # jnp.linspace computes in the (float32) result dtype
    return canon(_np.linspace(start, stop, num).astype(_resolve(dtype) or _np.float32))


def where(cond, x=None, y=None):
    if x is None:
        return canon(_np.where(base(cond)))
    xb, yb = base(x), base(y)
    dt = _canon_dtype(_np.result_type(xb, yb))
    # weak typing: a Python scalar adopts the other operand's dtype
    if isinstance(xb, (int, float)) and not isinstance(yb, (int, float)):
        dt = _canon_dtype(_np.asarray(yb).dtype) if not (isinstance(xb, float) and _np.asarray(yb).dtype.kind in 'iub') \
            else _np.dtype(_np.float32)
    elif isinstance(yb, (int, float)) and not isinstance(xb, (int, float)):
        dt = _canon_dtype(_np.asarray(xb).dtype) if not (isinstance(yb, float) and _np.asarray(xb).dtype.kind in 'iub') \
            else _np.dtype(_np.float32)
    return canon(_np.where(base(cond), _np.asarray(xb, dtype=dt), _np.asarray(yb, dtype=dt)))


def dot(a, b):
    a, b = _np.asarray(base(a)), _np.asarray(base(b))
    if a.dtype.kind == 'f' or b.dtype.kind == 'f':
        a, b = a.astype(_np.float32), b.astype(_np.float32)
    if a.ndim == 1 and b.ndim == 1:
        # sequential float32 accumulation in index order (no BLAS pairwise / FMA surprises)
        acc = _np.float32(0)
        for x, y in zip(a, b):
            acc = _np.float32(acc + _np.float32(x * y))
        return canon(_np.asarray(acc))
    return canon(_np.dot(a, b))


def sum(a, axis=None, dtype=None, keepdims=False):
    return canon(_np.sum(base(a), axis=axis, dtype=_resolve(dtype), keepdims=keepdims))


def max(a, axis=None):
    return canon(_np.max(base(a), axis=axis))


def min(a, axis=None):
    return canon(_np.min(base(a), axis=axis))


def argsort(a, axis=-1, **_kw):
    return canon(_np.argsort(base(a), axis=axis, kind='stable'))       # jnp.argsort is stable by default


def nanargmin(a, axis=None):
    return canon(_np.asarray(_np.nanargmin(base(a), axis=axis)))


def argmin(a, axis=None):
    return canon(_np.asarray(_np.argmin(base(a), axis=axis)))


def split(a, n, axis=0):
    return [canon(x) for x in _np.split(base(a), n, axis=axis)]


def tile(a, reps):
    return canon(_np.tile(_np.asarray(base(a), dtype=_np.float32 if isinstance(a, float) else None), base(reps)))


def concatenate(arrs, axis=0):
    return canon(_np.concatenate([_np.atleast_1d(base(x)) for x in arrs], axis=axis))


def diag(a, k=0):
    return canon(_np.diag(base(a), k))


def cov(m, rowvar=True, ddof=1):
    """jnp.cov's own algorithm, in float32 (np.cov would promote to float64)."""
    X = _np.array(base(m), ndmin=2, dtype=_np.float32)
    if not rowvar and X.shape[0] != 1:
        X = X.T
    avg = X.mean(axis=1, dtype=_np.float32)
    norm = _np.float32(X.shape[1] - ddof)
    X = X - avg[:, None]
    return canon((_np.dot(X, X.T) / norm).squeeze())


sin, cos, exp, sqrt, isnan, isinf, abs, log = (_w(_np.sin), _w(_np.cos), _w(_np.exp), _w(_np.sqrt), _w(_np.isnan),
                                               _w(_np.isinf), _w(_np.abs), _w(_np.log))
maximum, minimum, clip, mean, stack, cross = (_w(_np.maximum), _w(_np.minimum), _w(_np.clip), _w(_np.mean),
                                              _w(_np.stack), _w(_np.cross))
