import numpy as _np


def _canon_dtype(dt):
    dt = _np.dtype(dt)
    if dt.kind == 'f' and dt.itemsize != 4:
        return _np.dtype(_np.float32)
    if dt.kind == 'i' and dt.itemsize == 8:
        return _np.dtype(_np.int32)
    if dt.kind == 'u' and dt.itemsize == 8:
        return _np.dtype(_np.uint32)
    return dt


def base(x):
    """Strip the JArr subclass (so NumPy functions do not re-enter __array_ufunc__)."""
    if isinstance(x, JArr):
        return x.view(_np.ndarray)
    if isinstance(x, (list, tuple)):
        return type(x)(base(v) for v in x)
    return x


def canon(x):
    """Result of any jnp operation: canonical dtype (x64 disabled), JArr, never a NumPy scalar."""
    if isinstance(x, (list, tuple)):
        return type(x)(canon(v) for v in x)
    if isinstance(x, (_np.ndarray, _np.generic)):
        a = _np.asarray(x)
        dt = _canon_dtype(a.dtype)
        if dt != a.dtype:
            a = a.astype(dt)
        return a.view(JArr)
    return x


def _is_float(x):
    if isinstance(x, (bool, _np.bool_)):
        return False
    if isinstance(x, float):
        return True
    if isinstance(x, (_np.ndarray, _np.generic)):
        return _np.asarray(x).dtype.kind == 'f'
    if isinstance(x, (list, tuple)):
        return any(_is_float(v) for v in x)
    return False


_FLOAT_ONLY = {_np.true_divide, _np.sin, _np.cos, _np.tan, _np.exp, _np.log, _np.sqrt, _np.arctan2, _np.tanh}


class _At:
    def __init__(self, arr):
        self.arr = arr

    def __getitem__(self, idx):
        return _AtIdx(self.arr, idx)


class _AtIdx:
    def __init__(self, arr, idx):
        self.arr, self.idx = arr, idx

    def set(self, value):
        out = _np.array(self.arr.view(_np.ndarray), copy=True)
        idx = base(self.idx)
        out[idx] = _np.asarray(base(value)).astype(out.dtype)     # jnp casts the update to the operand's dtype
        return canon(out)

    def add(self, value):
        out = _np.array(self.arr.view(_np.ndarray), copy=True)
        out[base(self.idx)] += _np.asarray(base(value)).astype(out.dtype)
        return canon(out)


class JArr(_np.ndarray):
    __array_priority__ = 100.0

    @property
    def at(self):
        return _At(self)

    def __array_ufunc__(self, ufunc, method, *inputs, out=None, **kwargs):
        ins = [base(x) for x in inputs]
        if method == '__call__':
            if any(_is_float(x) for x in ins) or ufunc in _FLOAT_ONLY:
                ins = [_np.asarray(x, dtype=_np.float32) if not isinstance(x, str) else x for x in ins]
        res = getattr(ufunc, method)(*ins, **kwargs)
        return canon(res)

    def __getitem__(self, idx):
        if isinstance(idx, tuple):
            idx = tuple(base(i) for i in idx)
        else:
            idx = base(idx)
        if isinstance(idx, _np.ndarray) and idx.ndim == 0 and idx.dtype.kind in 'iu':
            idx = int(idx)
        r = self.view(_np.ndarray)[idx]
        return canon(r)

    def __iter__(self):
        for i in range(self.shape[0]):
            yield self[i]

    def __index__(self):
        return int(self.view(_np.ndarray))

    def __bool__(self):
        return bool(self.view(_np.ndarray))

    def __float__(self):
        return float(self.view(_np.ndarray))

    def __int__(self):
        return int(self.view(_np.ndarray))

    def take(self, indices, axis=None, **kw):
        return canon(self.view(_np.ndarray).take(base(indices), axis=axis))

    @property
    def T(self):
        return canon(self.view(_np.ndarray).T)

    def block_until_ready(self):
        return self


def _to_device_arg(x):
    """What the jit boundary does to an argument: arrays -> canonical-dtype device arrays; Python scalars stay weak."""
    if isinstance(x, (_np.ndarray, _np.generic)):
        return canon(x)
    if isinstance(x, (list, tuple)):
        return type(x)(_to_device_arg(v) for v in x)
    return x


def jit(fun, device=None, **_kw):
    def wrapped(*args, **kwargs):
        return fun(*[_to_device_arg(a) for a in args], **{k: _to_device_arg(v) for k, v in kwargs.items()})
    wrapped.__wrapped__ = fun
    return wrapped


def vmap(fun, in_axes=0, out_axes=0):
    def wrapped(*args):
        axes = in_axes if isinstance(in_axes, (tuple, list)) else (in_axes,) * len(args)
        args = [_to_device_arg(a) for a in args]
        n = None
        for a, ax in zip(args, axes):
            if ax is not None:
                assert ax == 0, 'shim supports in_axes 0 / None'
                n = a.shape[0] if n is None else n
                assert a.shape[0] == n
        outs = []
        for i in range(n):
            outs.append(fun(*[a if ax is None else a[i] for a, ax in zip(args, axes)]))
        if isinstance(outs[0], (tuple, list)):
            return tuple(canon(_np.stack([base(o[j]) for o in outs])) for j in range(len(outs[0])))
        return canon(_np.stack([_np.asarray(base(o)) for o in outs]))
    return wrapped


class _Device:
    platform = 'cpu'

    def __repr__(self):
        return 'NumpyShimDevice(cpu)'


def devices(kind=None):
    if kind == 'gpu':
        raise RuntimeError('no gpu backend in the numpy shim')
    return [_Device()]
