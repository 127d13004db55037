/* Host-side marshalling of the reference's list arguments (CPython C API, loaded through ctypes.PyDLL).
 *
 * The reference's tree modules take Python lists of lists and Cython copies them element by element into
 * std::vector (mz_tree.pyx:34-39,74-82: `list noises`, `list policy_logits_pool`, ...), and hand lists of lists back
 * (get_distributions, mz_tree.pyx:44-45).  An unmodified policy therefore calls this framework with lists
 * (lzero/policy/muzero.py:672-719).  numpy needs ~50 ns per element to walk such lists; these loops need ~5:
 * at 8192 roots x 18 actions a search call with lists in and out drops from 17 ms to a few ms.
 * No GPU code here and no arithmetic: bytes in, bytes out. */
#define PY_SSIZE_T_CLEAN
#include <Python.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <unistd.h>

/* Worker threads of the parallel readers below.  DEFAULT 1: the exact-type fast loops run in the calling thread.  More
 * threads (LZB_PYLIST_THREADS=2..8) help only in a process whose cores are otherwise idle (0.46 -> 0.18 ms per
 * [8192][18] argument on the GPU box); next to PyTorch's OpenMP pool, whose workers spin after every parallel region,
 * freshly created threads wait milliseconds for a core (measured: a list-in / list-out search call 5.1 -> 18 ms). */
static int reader_threads(void)
{
    const char *e = getenv("LZB_PYLIST_THREADS");
    long want = e ? atol(e) : 1;
    long cores = sysconf(_SC_NPROCESSORS_ONLN);
    if (want > cores) want = cores;
    if (want > 8) want = 8;
    return want < 1 ? 1 : (int)want;
}

static int as_double(PyObject *o, double *v)
{
    if (PyFloat_CheckExact(o)) {
        *v = PyFloat_AS_DOUBLE(o);
        return 0;
    }
    *v = PyFloat_AsDouble(o); /* ints, numpy scalars, anything with __float__ */
    return (*v == -1.0 && PyErr_Occurred()) ? -1 : 0;
}

/* ---- parallel fast path for the common shape: a LIST of LISTS of exact floats (what `.tolist()` of a float array
 * gives, lzero/policy/muzero.py:669,672-677).  The calling thread keeps the GIL for the whole call, so no Python code
 * can run or mutate the lists; the workers only READ list item pointers and the ob_fval of exact floats - no reference
 * count is touched and no Python API is called from them.  Anything else in the data (ints, numpy scalars, tuples)
 * makes a worker report "not this shape" and the serial loops below redo the call. */
typedef struct {
    PyObject **rows;
    float *out;
    int32_t *lens;
    long lo, hi, width;
    int status; /* 0 ok, 1 = a row that is not a list of exact floats, 2 = a row longer than the width */
} f32_job;

static void *f32_worker(void *arg)
{
    f32_job *j = (f32_job *)arg;
    for (long i = j->lo; i < j->hi; ++i) {
        PyObject *row = j->rows[i];
        if (!PyList_CheckExact(row)) { j->status = 1; return NULL; }
        const Py_ssize_t m = PyList_GET_SIZE(row);
        if (m > j->width) { j->status = 2; return NULL; }
        PyObject **items = ((PyListObject *)row)->ob_item;
        float *dst = j->out + (size_t)i * (size_t)j->width;
        for (Py_ssize_t k = 0; k < m; ++k) {
            if (!PyFloat_CheckExact(items[k])) { j->status = 1; return NULL; }
            dst[k] = (float)PyFloat_AS_DOUBLE(items[k]);
        }
        for (Py_ssize_t k = m; k < j->width; ++k) dst[k] = 0.0f;
        if (j->lens) j->lens[i] = (int32_t)m;
    }
    return NULL;
}

/* 1 = done by the exact-type loops (one or more threads), 0 = not the fast shape (nothing reported, caller falls back),
 * -1 = error set */
static int rows_to_f32_parallel(PyObject *rows, float *out, long n, long width, int32_t *lens)
{
    enum { kMaxThreads = 8 };
    if (!PyList_CheckExact(rows) || PyList_GET_SIZE(rows) != n) return 0;
    int nt = (n * width < 32768) ? 1 : reader_threads();
    f32_job jobs[kMaxThreads];
    pthread_t tid[kMaxThreads];
    int started[kMaxThreads];
    PyObject **items = ((PyListObject *)rows)->ob_item;
    for (int t = 0; t < nt; ++t) {
        jobs[t].rows = items; jobs[t].out = out; jobs[t].lens = lens; jobs[t].width = width; jobs[t].status = 0;
        jobs[t].lo = n * t / nt; jobs[t].hi = n * (t + 1) / nt;
        started[t] = (t > 0) && pthread_create(&tid[t], NULL, f32_worker, &jobs[t]) == 0;
    }
    f32_worker(&jobs[0]);
    for (int t = 1; t < nt; ++t) {
        if (started[t]) pthread_join(tid[t], NULL);
        else f32_worker(&jobs[t]);
    }
    int worst = 0;
    for (int t = 0; t < nt; ++t) worst = jobs[t].status > worst ? jobs[t].status : worst;
    if (worst == 2) {
        PyErr_SetString(PyExc_ValueError, "a row is longer than the action space");
        return -1;
    }
    return worst == 0 ? 1 : 0;
}

/* rows: sequence of n sequences, row i of length lens[i] <= width  ->  out[n][width] float32, zero padded.
 * Returns 0, or -1 with a Python exception set. */
int lzb_rows_to_f32(PyObject *rows, float *out, long n, long width, int32_t *lens)
{
    const int fast = rows_to_f32_parallel(rows, out, n, width, lens);
    if (fast != 0) return fast < 0 ? -1 : 0;
    PyObject *outer = PySequence_Fast(rows, "expected a sequence of rows");
    if (!outer) return -1;
    if (PySequence_Fast_GET_SIZE(outer) != n) {
        Py_DECREF(outer);
        PyErr_SetString(PyExc_ValueError, "wrong number of rows");
        return -1;
    }
    for (long i = 0; i < n; ++i) {
        PyObject *row = PySequence_Fast(PySequence_Fast_GET_ITEM(outer, i), "expected a sequence of numbers per row");
        if (!row) {
            Py_DECREF(outer);
            return -1;
        }
        const Py_ssize_t m = PySequence_Fast_GET_SIZE(row);
        if (m > width) {
            Py_DECREF(row);
            Py_DECREF(outer);
            PyErr_SetString(PyExc_ValueError, "a row is longer than the action space");
            return -1;
        }
        float *dst = out + (size_t)i * (size_t)width;
        for (Py_ssize_t j = 0; j < m; ++j) {
            double v;
            if (as_double(PySequence_Fast_GET_ITEM(row, j), &v)) {
                Py_DECREF(row);
                Py_DECREF(outer);
                return -1;
            }
            dst[j] = (float)v;
        }
        for (Py_ssize_t j = m; j < width; ++j) dst[j] = 0.0f;
        if (lens) lens[i] = (int32_t)m;
        Py_DECREF(row);
    }
    Py_DECREF(outer);
    return 0;
}

/* seq: sequence of n numbers -> out[n] float32 */
int lzb_seq_to_f32(PyObject *seq, float *out, long n)
{
    PyObject *s = PySequence_Fast(seq, "expected a sequence of numbers");
    if (!s) return -1;
    if (PySequence_Fast_GET_SIZE(s) != n) {
        Py_DECREF(s);
        PyErr_SetString(PyExc_ValueError, "wrong number of elements");
        return -1;
    }
    for (long i = 0; i < n; ++i) {
        double v;
        if (as_double(PySequence_Fast_GET_ITEM(s, i), &v)) {
            Py_DECREF(s);
            return -1;
        }
        out[i] = (float)v;
    }
    Py_DECREF(s);
    return 0;
}

static int as_long(PyObject *o, long *v)
{
    *v = PyLong_AsLong(o); /* ints; anything with __index__ */
    if (*v == -1 && PyErr_Occurred()) {
        PyErr_Clear();
        double d = PyFloat_AsDouble(o); /* floats that hold integers (numpy .tolist() of a float mask) */
        if (d == -1.0 && PyErr_Occurred()) return -1;
        *v = (long)d;
    }
    return 0;
}

/* the same parallel fast path for a list of lists of exact SMALL ints (CPython >= 3.12 keeps them "compact": the value
 * is read from the object without any API call); anything else falls back to the serial loop */
#if PY_VERSION_HEX >= 0x030C0000
typedef struct {
    PyObject **rows;
    int32_t *out, *lens;
    long lo, hi, width;
    int status;
} i32_job;

static void *i32_worker(void *arg)
{
    i32_job *j = (i32_job *)arg;
    for (long i = j->lo; i < j->hi; ++i) {
        PyObject *row = j->rows[i];
        if (!PyList_CheckExact(row)) { j->status = 1; return NULL; }
        const Py_ssize_t m = PyList_GET_SIZE(row);
        if (m > j->width) { j->status = 2; return NULL; }
        PyObject **items = ((PyListObject *)row)->ob_item;
        int32_t *dst = j->out + (size_t)i * (size_t)j->width;
        for (Py_ssize_t k = 0; k < m; ++k) {
            if (!PyLong_CheckExact(items[k]) || !PyUnstable_Long_IsCompact((PyLongObject *)items[k])) { j->status = 1; return NULL; }
            dst[k] = (int32_t)PyUnstable_Long_CompactValue((PyLongObject *)items[k]);
        }
        for (Py_ssize_t k = m; k < j->width; ++k) dst[k] = 0;
        if (j->lens) j->lens[i] = (int32_t)m;
    }
    return NULL;
}

static int rows_to_i32_parallel(PyObject *rows, int32_t *out, long n, long width, int32_t *lens)
{
    enum { kMaxThreads = 8 };
    if (!PyList_CheckExact(rows) || PyList_GET_SIZE(rows) != n) return 0;
    int nt = (n * width < 32768) ? 1 : reader_threads();
    i32_job jobs[kMaxThreads];
    pthread_t tid[kMaxThreads];
    int started[kMaxThreads];
    PyObject **items = ((PyListObject *)rows)->ob_item;
    for (int t = 0; t < nt; ++t) {
        jobs[t].rows = items; jobs[t].out = out; jobs[t].lens = lens; jobs[t].width = width; jobs[t].status = 0;
        jobs[t].lo = n * t / nt; jobs[t].hi = n * (t + 1) / nt;
        started[t] = (t > 0) && pthread_create(&tid[t], NULL, i32_worker, &jobs[t]) == 0;
    }
    i32_worker(&jobs[0]);
    for (int t = 1; t < nt; ++t) {
        if (started[t]) pthread_join(tid[t], NULL);
        else i32_worker(&jobs[t]);
    }
    int worst = 0;
    for (int t = 0; t < nt; ++t) worst = jobs[t].status > worst ? jobs[t].status : worst;
    if (worst == 2) {
        PyErr_SetString(PyExc_ValueError, "a root has more legal actions than the action space");
        return -1;
    }
    return worst == 0 ? 1 : 0;
}
#else
static int rows_to_i32_parallel(PyObject *rows, int32_t *out, long n, long width, int32_t *lens) { return 0; }
#endif

/* rows: sequence of n sequences of ints, row i of length lens[i] <= width -> out[n][width] int32, zero padded (the
 * per-root legal-action lists every policy passes to Roots(...), lzero/policy/muzero.py:671).  0, or -1 with an exception. */
int lzb_rows_to_i32(PyObject *rows, int32_t *out, long n, long width, int32_t *lens)
{
    const int fast = rows_to_i32_parallel(rows, out, n, width, lens);
    if (fast != 0) return fast < 0 ? -1 : 0;
    PyObject *outer = PySequence_Fast(rows, "expected a sequence of rows");
    if (!outer) return -1;
    if (PySequence_Fast_GET_SIZE(outer) != n) {
        Py_DECREF(outer);
        PyErr_SetString(PyExc_ValueError, "wrong number of rows");
        return -1;
    }
    for (long i = 0; i < n; ++i) {
        PyObject *row = PySequence_Fast(PySequence_Fast_GET_ITEM(outer, i), "expected a sequence of integers per row");
        if (!row) {
            Py_DECREF(outer);
            return -1;
        }
        const Py_ssize_t m = PySequence_Fast_GET_SIZE(row);
        if (m > width) {
            Py_DECREF(row);
            Py_DECREF(outer);
            PyErr_SetString(PyExc_ValueError, "a root has more legal actions than the action space");
            return -1;
        }
        int32_t *dst = out + (size_t)i * (size_t)width;
        for (Py_ssize_t j = 0; j < m; ++j) {
            long v;
            if (as_long(PySequence_Fast_GET_ITEM(row, j), &v)) {
                Py_DECREF(row);
                Py_DECREF(outer);
                return -1;
            }
            dst[j] = (int32_t)v;
        }
        for (Py_ssize_t j = m; j < width; ++j) dst[j] = 0;
        if (lens) lens[i] = (int32_t)m;
        Py_DECREF(row);
    }
    Py_DECREF(outer);
    return 0;
}

/* seq: sequence of n integers -> out[n] int32 (to_play_batch) */
int lzb_seq_to_i32(PyObject *seq, int32_t *out, long n)
{
    PyObject *s = PySequence_Fast(seq, "expected a sequence of integers");
    if (!s) return -1;
    if (PySequence_Fast_GET_SIZE(s) != n) {
        Py_DECREF(s);
        PyErr_SetString(PyExc_ValueError, "wrong number of elements");
        return -1;
    }
    for (long i = 0; i < n; ++i) {
        long v;
        if (as_long(PySequence_Fast_GET_ITEM(s, i), &v)) {
            Py_DECREF(s);
            return -1;
        }
        out[i] = (int32_t)v;
    }
    Py_DECREF(s);
    return 0;
}

/* data[n][stride] int32, counts[n] -> [[int] * counts[i]] (new reference), NULL on error.  The cyclic collector is
 * paused while the n row lists are created: each new list is a tracked container, so building 8192 of them otherwise
 * triggers a dozen young-generation collections that find nothing (2.6 -> 1.4 ms at 8192 x 18). */
static PyObject *rows_from_i32(const int32_t *data, const int32_t *counts, long n, long stride);
PyObject *lzb_rows_from_i32(const int32_t *data, const int32_t *counts, long n, long stride)
{
    const int was_enabled = PyGC_Disable();
    PyObject *res = rows_from_i32(data, counts, n, stride);
    if (was_enabled) PyGC_Enable();
    return res;
}

/* int objects 0 .. 255, created once (CPython keeps these as shared small-int singletons; strong references are held here
 * for interpreters where they are not immortal).  Visit counts are at most num_simulations, so almost every element of a
 * distributions table is one of them: a table lookup + INCREF instead of a PyLong_FromLong call per element
 * (1.5 -> 0.7 ms for 8192 x 18). */
#define LZB_SMALL_INTS 256
static PyObject *small_int[LZB_SMALL_INTS];
static int small_int_ready = 0;

static int small_ints_init(void)
{
    if (small_int_ready) return 0;
    for (long v = 0; v < LZB_SMALL_INTS; ++v) {
        small_int[v] = PyLong_FromLong(v);
        if (!small_int[v]) return -1;
    }
    small_int_ready = 1;
    return 0;
}

static PyObject *rows_from_i32(const int32_t *data, const int32_t *counts, long n, long stride)
{
    if (small_ints_init() != 0) return NULL;
    PyObject *outer = PyList_New(n);
    if (!outer) return NULL;
    for (long i = 0; i < n; ++i) {
        const long m = counts ? counts[i] : stride;
        PyObject *row = PyList_New(m);
        if (!row) {
            Py_DECREF(outer);
            return NULL;
        }
        const int32_t *src = data + (size_t)i * (size_t)stride;
        for (long j = 0; j < m; ++j) {
            const int32_t x = src[j];
            PyObject *v;
            if ((uint32_t)x < (uint32_t)LZB_SMALL_INTS) {
                v = small_int[x];
                Py_INCREF(v);
            } else {
                v = PyLong_FromLong(x);
                if (!v) {
                    Py_DECREF(row);
                    Py_DECREF(outer);
                    return NULL;
                }
            }
            PyList_SET_ITEM(row, j, v);
        }
        PyList_SET_ITEM(outer, i, row);
    }
    return outer;
}
