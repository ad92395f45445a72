/* qpfast.c -- CPython binding of the ONE call on the real-time path: qpmpc_tick / qpmpc_step of a single controller.
 *
 * Everything else in the package talks to libqpmpc_b200.so through ctypes (_lib.py).  For the tick itself ctypes costs
 * ~10 us of Python per call (numpy staging copies, struct marshalling, result unpacking) on top of a ~26 us C call; this
 * module reads the caller's numpy arrays in place (float64 or float32, any strides), fills the qpmpc_tick_in struct on the
 * stack, calls the same C-ABI entry point and builds the result arrays directly: ~1.5 us.  No arithmetic of the path lives
 * here -- it is marshalling only; the entry points are those of include/qpmpc.h.
 *
 *   tick(handle, mode, state, reference, contact_sequence, best, key0, key1, sigma, timing, nominal, optimize_swing,
 *        iterations, sigma0, P, H, elite_k)
 *     mode 0: qpmpc_tick | 1: qpmpc_tick_sharded | 2: qpmpc_tick_iterations | 3: qpmpc_step
 *     -> (grfs f32[12], predicted_state f32[24], best f32[P], best_cost, best_freq, best_index, sigma_new f32[P] | None,
 *         elite_indices i32[num_elite] | None)        or an int error code (the message is in qpmpc_last_error())
 */
#define PY_SSIZE_T_CLEAN
#define NPY_NO_DEPRECATED_API NPY_1_7_API_VERSION
#include <Python.h>
#include <numpy/arrayobject.h>
#include <string.h>
#include "../../include/qpmpc.h"

#define QPF_MAX_P 1024
#define QPF_ERR_ARG (-1)

/* copy `n` elements of a 1-D-viewable numpy array (float64 / float32 / anything castable) into dst as float32 */
static int to_f32(PyObject *obj, float *dst, npy_intp n) {
    if (PyArray_Check(obj)) {
        PyArrayObject *a = (PyArrayObject *)obj;
        if (PyArray_SIZE(a) == n && PyArray_ISCARRAY_RO(a)) {
            const int t = PyArray_TYPE(a);
            if (t == NPY_FLOAT64) { const double *s = (const double *)PyArray_DATA(a); for (npy_intp i = 0; i < n; ++i) dst[i] = (float)s[i]; return 0; }
            if (t == NPY_FLOAT32) { memcpy(dst, PyArray_DATA(a), (size_t)n * sizeof(float)); return 0; }
        }
    }
    PyArrayObject *c = (PyArrayObject *)PyArray_FROM_OTF(obj, NPY_FLOAT32, NPY_ARRAY_IN_ARRAY | NPY_ARRAY_FORCECAST);
    if (!c) return -1;
    if (PyArray_SIZE(c) != n) { Py_DECREF(c); PyErr_Format(PyExc_ValueError, "expected %zd values, got %zd", (Py_ssize_t)n, (Py_ssize_t)PyArray_SIZE(c)); return -1; }
    memcpy(dst, PyArray_DATA(c), (size_t)n * sizeof(float));
    Py_DECREF(c);
    return 0;
}

/* contact_sequence (4, >= H) -> dst[4][H]: only [:, :H] is read (FULL_STANCE tables are (4, 2H)) */
static int contact_to_f32(PyObject *obj, float *dst, int H) {
    PyArrayObject *a = NULL;
    int own = 0;
    if (PyArray_Check(obj) && (PyArray_TYPE((PyArrayObject *)obj) == NPY_FLOAT64 || PyArray_TYPE((PyArrayObject *)obj) == NPY_FLOAT32) &&
        PyArray_ISALIGNED((PyArrayObject *)obj) && PyArray_ISNOTSWAPPED((PyArrayObject *)obj)) {
        a = (PyArrayObject *)obj;
    } else {
        a = (PyArrayObject *)PyArray_FROM_OTF(obj, NPY_FLOAT64, NPY_ARRAY_IN_ARRAY | NPY_ARRAY_FORCECAST);
        if (!a) return -1;
        own = 1;
    }
    if (PyArray_NDIM(a) != 2 || PyArray_DIM(a, 0) != 4 || PyArray_DIM(a, 1) < H) {
        PyErr_Format(PyExc_ValueError, "contact_sequence must be (4, >= %d)", H);
        if (own) Py_DECREF(a);
        return -1;
    }
    const int f64 = PyArray_TYPE(a) == NPY_FLOAT64;
    for (int l = 0; l < 4; ++l)
        for (int n = 0; n < H; ++n) {
            const void *p = PyArray_GETPTR2(a, l, n);
            dst[l * H + n] = f64 ? (float)*(const double *)p : *(const float *)p;
        }
    if (own) Py_DECREF(a);
    return 0;
}

static PyObject *f32_array(const float *src, npy_intp n) {
    PyObject *a = PyArray_SimpleNew(1, &n, NPY_FLOAT32);
    if (a) memcpy(PyArray_DATA((PyArrayObject *)a), src, (size_t)n * sizeof(float));
    return a;
}
