import numpy as _np

from .core import canon


def fori_loop(lower, upper, body_fun, init_val):
    val = init_val
    for i in range(int(lower), int(upper)):
        val = body_fun(canon(_np.asarray(_np.int32(i))), val)
    return val
