/* CPython extension: reference-style record lists -> wire columns, in C.
 *
 * The adapter's one-decision entry point receives what the reference's method receives: a Python
 * list of NodeMetrics objects (scheduler.py:72-84).  Turning that list into columns with a Python
 * loop costs ~1.3 us per node -- three times the reference's whole decision loop (~0.4 us per
 * node) -- so the walk is done here with the C API: attribute reads by interned name, the same
 * conversions and checks as tables.py (cores_to_milli, gib_to_bytes, is_ready), writing straight
 * into caller-provided numpy buffers.  Results are identical to the Python loop, which stays in
 * tables.py as the specification and as the fallback when this module has not been built.
 */
#define PY_SSIZE_T_CLEAN
#include <Python.h>
#include <math.h>
#include <stdint.h>
#include <string.h>

static PyObject *s_available_cpu, *s_available_memory, *s_pod_count, *s_max_pods, *s_conditions,
    *s_cpu_usage, *s_mem_usage, *s_cpu_request, *s_memory_request, *s_type, *s_status, *s_get;

#define WIRE_LIMIT 9007199254740992.0 /* 2^53 */

static int get_double(PyObject *obj, PyObject *name, double *out)
{
    PyObject *v = PyObject_GetAttr(obj, name);
    if (!v) return -1;
    const double d = PyFloat_AsDouble(v);
    Py_DECREF(v);
    if (d == -1.0 && PyErr_Occurred()) return -1;
    *out = d;
    return 0;
}

static int get_int64(PyObject *obj, PyObject *name, int64_t *out)
{
    PyObject *v = PyObject_GetAttr(obj, name);
    if (!v) return -1;
    const long long x = PyLong_AsLongLong(v);
    Py_DECREF(v);
    if (x == -1 && PyErr_Occurred()) return -1;
    *out = (int64_t)x;
    return 0;
}

static void value_error(const char *fmt, double x)
{
    PyObject *f = PyFloat_FromDouble(x);
    if (f) {
        PyErr_Format(PyExc_ValueError, fmt, f);
        Py_DECREF(f);
    }
}

/* tables.cores_to_milli: int(round(cores * 1000)), exact -> must reproduce cores as m / 1000.0 */
static int cores_to_milli(double cores, int exact, int64_t *out)
{
    const double scaled = cores * 1000.0;
    if (!(fabs(scaled) < WIRE_LIMIT)) {
        value_error("%R cores is outside the wire format's range (|value| < 2^53)", cores);
        return -1;
    }
    const double r = nearbyint(scaled); /* round half to even, like Python's round() */
    const int64_t m = (int64_t)r;
    if (exact && (double)m / 1000.0 != cores) {
        value_error("%R cores is not an integer number of milli-cores in fp64", cores);
        return -1;
    }
    *out = m;
    return 0;
}

/* tables.gib_to_bytes */
static int gib_to_bytes(double gib, int exact, int64_t *out)
{
    const double scaled = gib * 1073741824.0;
    if (!(fabs(scaled) < WIRE_LIMIT)) {
        value_error("%R GiB is outside the wire format's range (|value| < 2^53)", gib);
        return -1;
    }
    const double r = nearbyint(scaled);
    const int64_t b = (int64_t)r;
    if (exact && (double)b / 1024 / 1024 / 1024 != gib) {
        value_error("%R GiB is not an integer number of bytes in fp64", gib);
        return -1;
    }
    *out = b;
    return 0;
}

/* c.get(key) == literal, for dicts and for anything else with a .get method */
static int cond_field_is(PyObject *cond, PyObject *key, const char *literal)
{
    PyObject *v;
    int owned = 0;
    if (PyDict_CheckExact(cond)) {
        v = PyDict_GetItemWithError(cond, key); /* borrowed */
        if (!v) return PyErr_Occurred() ? -1 : 0;
    } else {
        v = PyObject_CallMethodObjArgs(cond, s_get, key, NULL);
        if (!v) return -1;
        owned = 1;
    }
    int eq = 0;
    if (PyUnicode_Check(v)) eq = PyUnicode_CompareWithASCIIString(v, literal) == 0;
    if (owned) Py_DECREF(v);
    return eq;
}

/* scheduler.py:532-533: any(c.get('type') == 'Ready' and c.get('status') == 'True' for c in conditions) */
static int is_ready(PyObject *conditions)
{
    PyObject *it = PyObject_GetIter(conditions);
    if (!it) return -1;
    int ready = 0;
    PyObject *c;
    while (!ready && (c = PyIter_Next(it))) {
        int a = cond_field_is(c, s_type, "Ready");
        int b = a > 0 ? cond_field_is(c, s_status, "True") : 0;
        Py_DECREF(c);
        if (a < 0 || b < 0) { Py_DECREF(it); return -1; }
        ready = a > 0 && b > 0;
    }
    Py_DECREF(it);
    if (PyErr_Occurred()) return -1;
    return ready;
}

static int writable_at_least(PyObject *obj, Py_buffer *view, Py_ssize_t itemsize, Py_ssize_t n, const char *what, int larger_ok)
{
    if (PyObject_GetBuffer(obj, view, PyBUF_WRITABLE | PyBUF_C_CONTIGUOUS) < 0) return -1;
    if (larger_ok ? view->len < itemsize * n : view->len != itemsize * n) {
        PyBuffer_Release(view);
        PyErr_Format(PyExc_ValueError, "%s: buffer of %zd bytes, expected %zd", what, view->len, itemsize * n);
        return -1;
    }
    return 0;
}
static int writable(PyObject *obj, Py_buffer *view, Py_ssize_t itemsize, Py_ssize_t n, const char *what)
{
    return writable_at_least(obj, view, itemsize, n, what, 0);
}

/* node_metrics_to_columns(nodes, cpu_milli, mem_bytes, pod_count, max_pods, ready, cpu_usage, mem_usage, mode)
 * mode 0: allocatable rounded to the nearest wire integer, 1: must be exact, 2: allocatable columns not read at all
 * (zeros) -- the reference's own decision (scheduler.py:531-549) never looks at available_cpu / available_memory,
 * so under the Ready-only filter whatever they hold (inf, nan, 1e300, ...) must not matter.
 * mode + 4: the buffers may be LARGER than the list (column buffers kept across calls; rows [0, n) are written). */
static PyObject *node_metrics_to_columns(PyObject *self, PyObject *args)
{
    PyObject *nodes, *o[7];
    int mode;
    if (!PyArg_ParseTuple(args, "OOOOOOOOi", &nodes, &o[0], &o[1], &o[2], &o[3], &o[4], &o[5], &o[6], &mode)) return NULL;
    const int larger_ok = (mode & 4) != 0;
    mode &= 3;
    const int exact = mode == 1, skip_alloc = mode == 2;
    PyObject *seq = PySequence_Fast(nodes, "node_metrics must be a sequence");
    if (!seq) return NULL;
    const Py_ssize_t n = PySequence_Fast_GET_SIZE(seq);
    Py_buffer v[7];
    const Py_ssize_t size[7] = {8, 8, 8, 8, 1, 8, 8};
    static const char *names[7] = {"alloc_cpu_milli", "alloc_mem_bytes", "pod_count", "max_pods", "ready", "cpu_usage_pct", "mem_usage_pct"};
    int got = 0, ok = 1;
    for (; got < 7; ++got)
        if (writable_at_least(o[got], &v[got], size[got], n, names[got], larger_ok) < 0) { ok = 0; break; }
    if (ok) {
        int64_t *milli = v[0].buf, *bytes = v[1].buf, *pc = v[2].buf, *mp = v[3].buf;
        uint8_t *ready = v[4].buf;
        double *cu = v[5].buf, *mu = v[6].buf;
        PyObject **items = PySequence_Fast_ITEMS(seq);
        for (Py_ssize_t i = 0; i < n && ok; ++i) {
            PyObject *nd = items[i];
            double cores, gib;
            if (skip_alloc) milli[i] = bytes[i] = 0;
            else if (get_double(nd, s_available_cpu, &cores) < 0 || get_double(nd, s_available_memory, &gib) < 0 ||
                     cores_to_milli(cores, exact, &milli[i]) < 0 || gib_to_bytes(gib, exact, &bytes[i]) < 0) { ok = 0; break; }
            if (get_int64(nd, s_pod_count, &pc[i]) < 0 || get_int64(nd, s_max_pods, &mp[i]) < 0 ||
                get_double(nd, s_cpu_usage, &cu[i]) < 0 || get_double(nd, s_mem_usage, &mu[i]) < 0) { ok = 0; break; }
            PyObject *conds = PyObject_GetAttr(nd, s_conditions);
            if (!conds) { ok = 0; break; }
            const int r = is_ready(conds);
            Py_DECREF(conds);
            if (r < 0) { ok = 0; break; }
            ready[i] = (uint8_t)r;
        }
    }
    for (int k = 0; k < got; ++k) PyBuffer_Release(&v[k]);
    Py_DECREF(seq);
    if (!ok) return NULL;
    Py_RETURN_NONE;
}

/* pod_specs_to_columns(pods, cpu_milli, mem_bytes, exact) */
static PyObject *pod_specs_to_columns(PyObject *self, PyObject *args)
{
    PyObject *pods, *o[2];
    int exact;
    if (!PyArg_ParseTuple(args, "OOOp", &pods, &o[0], &o[1], &exact)) return NULL;
    PyObject *seq = PySequence_Fast(pods, "pod_specs must be a sequence");
    if (!seq) return NULL;
    const Py_ssize_t n = PySequence_Fast_GET_SIZE(seq);
    Py_buffer v[2];
    int got = 0, ok = 1;
    for (; got < 2; ++got)
        if (writable(o[got], &v[got], 8, n, got ? "req_mem_bytes" : "req_cpu_milli") < 0) { ok = 0; break; }
    if (ok) {
        int64_t *milli = v[0].buf, *bytes = v[1].buf;
        PyObject **items = PySequence_Fast_ITEMS(seq);
        for (Py_ssize_t i = 0; i < n; ++i) {
            double cores, gib;
            if (get_double(items[i], s_cpu_request, &cores) < 0 || get_double(items[i], s_memory_request, &gib) < 0 ||
                cores_to_milli(cores, exact, &milli[i]) < 0 || gib_to_bytes(gib, exact, &bytes[i]) < 0) { ok = 0; break; }
        }
    }
    for (int k = 0; k < got; ++k) PyBuffer_Release(&v[k]);
    Py_DECREF(seq);
    if (!ok) return NULL;
    Py_RETURN_NONE;
}

/* ---- pack_pod_requests(pods) -> (begin, cpu_text, cpu_off, mem_text, mem_off, odd) -------------------------
 * ingest.flatten_pod_requests + pack_strings in one pass: per pod a sequence of container `requests`
 * mappings (None / empty = container without requests, dropped like scheduler.py:738); a missing key
 * reads as "0" (:739-740); a pod holding a non-string quantity is listed in `odd` (the host mirror
 * redoes it) and contributes "0" placeholders.  Buffers come back as bytearrays (int64 little-endian
 * for the offset arrays). */
typedef struct { char *p; Py_ssize_t len, cap; } buf_t;

static int buf_put(buf_t *b, const void *src, Py_ssize_t n)
{
    if (b->len + n > b->cap) {
        Py_ssize_t cap = b->cap ? b->cap * 2 : 1 << 16;
        while (cap < b->len + n) cap *= 2;
        char *q = PyMem_Realloc(b->p, (size_t)cap);
        if (!q) { PyErr_NoMemory(); return -1; }
        b->p = q;
        b->cap = cap;
    }
    memcpy(b->p + b->len, src, (size_t)n);
    b->len += n;
    return 0;
}

static int put_quantity(PyObject *req, PyObject *key, buf_t *text, buf_t *off, int *odd)
{
    PyObject *v = PyDict_CheckExact(req) ? PyDict_GetItemWithError(req, key) : NULL;   /* borrowed */
    PyObject *owned = NULL;
    if (!v) {
        if (PyErr_Occurred()) return -1;
        if (!PyDict_CheckExact(req)) {                  /* any mapping with .get */
            owned = PyObject_CallMethodObjArgs(req, s_get, key, NULL);
            if (!owned) return -1;
            v = owned == Py_None ? NULL : owned;
        }
    }
    const char *str = "0";
    Py_ssize_t n = 1;
    if (v) {
        if (PyUnicode_Check(v)) {
            str = PyUnicode_AsUTF8AndSize(v, &n);
            if (!str) { Py_XDECREF(owned); return -1; }
        } else {
            *odd = 1;
        }
    }
    int rc = buf_put(text, str, n);
    Py_XDECREF(owned);
    if (rc < 0) return -1;
    const int64_t end = (int64_t)text->len;
    return buf_put(off, &end, 8);
}

static PyObject *pack_pod_requests(PyObject *self, PyObject *arg)
{
    static PyObject *s_cpu = NULL, *s_memory = NULL;
    if (!s_cpu) { s_cpu = PyUnicode_InternFromString("cpu"); s_memory = PyUnicode_InternFromString("memory"); }
    PyObject *seq = PySequence_Fast(arg, "pods must be a sequence");
    if (!seq) return NULL;
    const Py_ssize_t n = PySequence_Fast_GET_SIZE(seq);
    buf_t begin = {0}, ct = {0}, co = {0}, mt = {0}, mo = {0};
    PyObject *odd = PyList_New(0), *result = NULL;
    const int64_t zero = 0;
    int ok = odd && buf_put(&begin, &zero, 8) == 0 && buf_put(&co, &zero, 8) == 0 && buf_put(&mo, &zero, 8) == 0;
    int64_t rows = 0;
    for (Py_ssize_t j = 0; ok && j < n; ++j) {
        PyObject *containers = PySequence_Fast(PySequence_Fast_GET_ITEM(seq, j), "a pod's containers must be a sequence");
        if (!containers) { ok = 0; break; }
        int is_odd = 0;
        const Py_ssize_t k = PySequence_Fast_GET_SIZE(containers);
        for (Py_ssize_t c = 0; ok && c < k; ++c) {
            PyObject *req = PySequence_Fast_GET_ITEM(containers, c);
            const int truth = PyObject_IsTrue(req);
            if (truth < 0) { ok = 0; break; }
            if (!truth) continue;
            int o2 = 0;
            if (put_quantity(req, s_cpu, &ct, &co, &o2) < 0 || put_quantity(req, s_memory, &mt, &mo, &o2) < 0) { ok = 0; break; }
            is_odd |= o2;
            ++rows;
        }
        Py_DECREF(containers);
        if (!ok) break;
        if (is_odd) {
            PyObject *idx = PyLong_FromSsize_t(j);
            if (!idx || PyList_Append(odd, idx) < 0) { Py_XDECREF(idx); ok = 0; break; }
            Py_DECREF(idx);
        }
        if (buf_put(&begin, &rows, 8) < 0) ok = 0;
    }
    if (ok) {
        result = Py_BuildValue("(NNNNNO)", PyByteArray_FromStringAndSize(begin.p, begin.len),
                               PyByteArray_FromStringAndSize(ct.p ? ct.p : "", ct.len), PyByteArray_FromStringAndSize(co.p, co.len),
                               PyByteArray_FromStringAndSize(mt.p ? mt.p : "", mt.len), PyByteArray_FromStringAndSize(mo.p, mo.len), odd);
    }
    PyMem_Free(begin.p); PyMem_Free(ct.p); PyMem_Free(co.p); PyMem_Free(mt.p); PyMem_Free(mo.p);
    Py_XDECREF(odd);
    Py_DECREF(seq);
    return result;
}

static PyMethodDef methods[] = {
    {"node_metrics_to_columns", node_metrics_to_columns, METH_VARARGS, "NodeMetrics list -> node table columns"},
    {"pod_specs_to_columns", pod_specs_to_columns, METH_VARARGS, "PodSpec list -> pod table columns"},
    {"pack_pod_requests", pack_pod_requests, METH_O, "per-pod container requests -> CSR + packed quantity strings"},
    {NULL, NULL, 0, NULL}};

static struct PyModuleDef module = {PyModuleDef_HEAD_INIT, "_kp_pyobj", "record lists -> wire columns", -1, methods};

PyMODINIT_FUNC PyInit__kp_pyobj(void)
{
    s_available_cpu = PyUnicode_InternFromString("available_cpu");
    s_available_memory = PyUnicode_InternFromString("available_memory");
    s_pod_count = PyUnicode_InternFromString("pod_count");
    s_max_pods = PyUnicode_InternFromString("max_pods");
    s_conditions = PyUnicode_InternFromString("conditions");
    s_cpu_usage = PyUnicode_InternFromString("cpu_usage_percent");
    s_mem_usage = PyUnicode_InternFromString("memory_usage_percent");
    s_cpu_request = PyUnicode_InternFromString("cpu_request");
    s_memory_request = PyUnicode_InternFromString("memory_request");
    s_type = PyUnicode_InternFromString("type");
    s_status = PyUnicode_InternFromString("status");
    s_get = PyUnicode_InternFromString("get");
    return PyModule_Create(&module);
}
