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

static PyObject *qpfast_tick(PyObject *self, PyObject *const *args, Py_ssize_t nargs) {
    (void)self;
    if (nargs != 17) { PyErr_SetString(PyExc_TypeError, "tick() takes 17 positional arguments"); return NULL; }
    qpmpc_handle *h = (qpmpc_handle *)PyLong_AsVoidPtr(args[0]);
    const long mode = PyLong_AsLong(args[1]);
    const long P = PyLong_AsLong(args[14]), H = PyLong_AsLong(args[15]), elite_k = PyLong_AsLong(args[16]);
    if (PyErr_Occurred()) return NULL;
    if (!h || P < 1 || P > QPF_MAX_P || H < 1 || H > QPMPC_MAX_HORIZON || elite_k < 0 || elite_k > QPMPC_MAX_ELITE) {
        PyErr_SetString(PyExc_ValueError, "bad handle / sizes");
        return NULL;
    }
    float state[24], reference[24], contact[4 * QPMPC_MAX_HORIZON], best[QPF_MAX_P], sigma[QPF_MAX_P], timing[4];
    float best_out[QPF_MAX_P], sigma_out[QPF_MAX_P];
    int32_t elite[QPMPC_MAX_ELITE];
    qpmpc_tick_in in;
    qpmpc_tick_out out;
    memset(&in, 0, sizeof(in));
    memset(&out, 0, sizeof(out));
    if (to_f32(args[2], state, 24) || to_f32(args[3], reference, 24) || contact_to_f32(args[4], contact, (int)H)) return NULL;
    in.state = state; in.reference = reference; in.contact_sequence = contact; in.contact_ld = (int32_t)H;
    if (args[5] != Py_None) { if (to_f32(args[5], best, P)) return NULL; in.best = best; }
    in.key[0] = (uint32_t)PyLong_AsUnsignedLongMask(args[6]);
    in.key[1] = (uint32_t)PyLong_AsUnsignedLongMask(args[7]);
    const int have_sigma = args[8] != Py_None;
    if (have_sigma) {
        if (PyFloat_Check(args[8]) || PyLong_Check(args[8])) {           /* with_newsigma(cfg scalar), I:128-129 */
            const float sv = (float)PyFloat_AsDouble(args[8]);
            for (long p = 0; p < P; ++p) sigma[p] = sv;
        } else if (PyArray_Check(args[8]) && PyArray_SIZE((PyArrayObject *)args[8]) == 1) {
            PyObject *f = PyNumber_Float(args[8]);
            if (!f) return NULL;
            const float sv = (float)PyFloat_AsDouble(f);
            Py_DECREF(f);
            for (long p = 0; p < P; ++p) sigma[p] = sv;
        } else if (to_f32(args[8], sigma, P)) return NULL;
        in.sigma = sigma;
    }
    if (args[9] != Py_None) { if (to_f32(args[9], timing, 4)) return NULL; in.phase_signal = timing; }
    in.nominal_step_freq = args[10] == Py_None ? 0.0f : (float)PyFloat_AsDouble(args[10]);
    in.optimize_swing = args[11] == Py_None ? 0 : (int32_t)PyLong_AsLong(args[11]);
    const long iterations = PyLong_AsLong(args[12]);
    const float sigma0 = (float)PyFloat_AsDouble(args[13]);
    if (PyErr_Occurred()) return NULL;
    out.best = best_out; out.sigma_new = sigma_out; out.elite_indices = elite;
    int rc;
    Py_BEGIN_ALLOW_THREADS
    switch (mode) {
        case 0: rc = qpmpc_tick(h, &in, &out); break;
        case 1: rc = qpmpc_tick_sharded(h, &in, &out); break;
        case 2: rc = qpmpc_tick_iterations(h, &in, (int32_t)iterations, &out); break;
        case 3: rc = qpmpc_step(h, &in, (int32_t)iterations, sigma0, &out); break;
        default: rc = QPF_ERR_ARG;
    }
    Py_END_ALLOW_THREADS
    if (rc != 0) return PyLong_FromLong(rc);
    PyObject *g = f32_array(out.grfs, 12), *pr = f32_array(out.predicted_state, 24), *b = f32_array(best_out, P);
    PyObject *sg = NULL, *el = NULL;
    if (elite_k > 0) {
        sg = f32_array(sigma_out, P);
        npy_intp n = out.num_elite < 0 ? 0 : out.num_elite;
        el = PyArray_SimpleNew(1, &n, NPY_INT32);
        if (el) memcpy(PyArray_DATA((PyArrayObject *)el), elite, (size_t)n * sizeof(int32_t));
    } else {
        sg = Py_None; Py_INCREF(sg);
        el = Py_None; Py_INCREF(el);
    }
    if (!g || !pr || !b || !sg || !el) { Py_XDECREF(g); Py_XDECREF(pr); Py_XDECREF(b); Py_XDECREF(sg); Py_XDECREF(el); return NULL; }
    PyObject *res = PyTuple_New(8);
    if (!res) return NULL;
    PyTuple_SET_ITEM(res, 0, g); PyTuple_SET_ITEM(res, 1, pr); PyTuple_SET_ITEM(res, 2, b);
    PyTuple_SET_ITEM(res, 3, PyFloat_FromDouble((double)out.best_cost));
    PyTuple_SET_ITEM(res, 4, PyFloat_FromDouble((double)out.best_freq));
    PyTuple_SET_ITEM(res, 5, PyLong_FromLong((long)out.best_index));
    PyTuple_SET_ITEM(res, 6, sg); PyTuple_SET_ITEM(res, 7, el);
    return res;
}

static PyMethodDef qpfast_methods[] = {
    {"tick", (PyCFunction)(void (*)(void))qpfast_tick, METH_FASTCALL, "one MPC tick through the C ABI (see module docstring)"},
    {NULL, NULL, 0, NULL}};
static struct PyModuleDef qpfast_module = {PyModuleDef_HEAD_INIT, "_qpfast",
                                           "CPython binding of qpmpc_tick / qpmpc_step (marshalling only)", -1, qpfast_methods};
PyMODINIT_FUNC PyInit__qpfast(void) {
    import_array();
    return PyModule_Create(&qpfast_module);
}
