/* task_encode.c -- CPython extension `pilottai_b200._taskenc`: the reference's task dicts ->
 * the bit-set columns pilott_route_score takes, at C speed.
 *
 * The reference routes `task: Dict` objects one at a time (TaskRouter.route_task,
 * pilott/core/router.py:34; LoadBalancer._find_best_agent, pilott/orchestration/load_balancer.py:268-273)
 * and reads three keys of each: "required_capabilities" and "type"
 * (BaseAgent.evaluate_task_suitability, pilott/core/agent.py:549-563) and "tags"
 * (TaskRouter._calculate_specialization_score, pilott/core/router.py:116-126).  A batch of a
 * million such dicts has to be turned into arrays before one kernel launch can score it; done in
 * Python (routing.encode_tasks) that costs ~10 us per task, four orders of magnitude more than the
 * kernel.  This module walks the list with the C API: three dict probes per task with pre-built
 * key objects, one vocabulary probe per string (str hashes are cached), bits set in place in the
 * caller's numpy buffers.  Same results as routing.encode_tasks, which stays as the definition
 * (tests/test_routing_host.py compares the two) and as the fallback for inputs this fast path
 * declines (it returns -1 at the first task that is not a plain dict of str / list / tuple / set).
 *
 *   encode(tasks, cap_vocab, spec_vocab, words, req, tags, typ, flags, urgent, cx, ndeps) -> int
 * tasks: list of dicts; *_vocab: dict str -> bit; req/tags: writable uint64 [n, words] (zeroed);
 * typ int32 [n]; flags uint8 [n] (bit0 has 'required_capabilities', bit1 has 'type', bit2 a
 * required capability outside the vocabulary); urgent uint8 [n]; cx int32 [n] (complexity class:
 * 9 if complexity > 8, 6 if > 5, else 0 -- the only comparisons get_task_priority makes,
 * router.py:135-145); ndeps int32 [n].  Returns n on success, -1 - i when task i needs the slow path.
 */
#define PY_SSIZE_T_CLEAN
#include <Python.h>
#include <stdint.h>

static PyObject *k_req, *k_type, *k_tags, *k_urgent, *k_cx, *k_deps, *k_empty;

/* bit of `s` in `vocab`, -1 if absent, -2 on error / non-str */
static inline long vocab_bit(PyObject* vocab, PyObject* s) {
    if (!PyUnicode_CheckExact(s)) return -2;
    PyObject* v = PyDict_GetItemWithError(vocab, s);       /* borrowed */
    if (!v) return PyErr_Occurred() ? -2 : -1;
    return PyLong_AsLong(v);
}

/* ORs the vocabulary bits of every string of `seq` into row; returns 0 ok, 1 some string outside the
 * vocabulary (only reported, never an error), -1 slow path needed. */
static int set_bits(PyObject* seq, PyObject* vocab, uint64_t* row, int words) {
    int missing = 0;
    if (PyList_CheckExact(seq) || PyTuple_CheckExact(seq)) {
        const Py_ssize_t m = PySequence_Fast_GET_SIZE(seq);
        PyObject** items = PySequence_Fast_ITEMS(seq);
        for (Py_ssize_t j = 0; j < m; ++j) {
            const long b = vocab_bit(vocab, items[j]);
            if (b == -2) return -1;
            if (b < 0) { missing = 1; continue; }
            if ((b >> 6) >= words) return -1;
            row[b >> 6] |= (uint64_t)1 << (b & 63);
        }
        return missing;
    }
    if (PyAnySet_CheckExact(seq)) {
        Py_ssize_t pos = 0;
        PyObject* key;
        Py_hash_t h;
        while (_PySet_NextEntry(seq, &pos, &key, &h)) {
            const long b = vocab_bit(vocab, key);
            if (b == -2) return -1;
            if (b < 0) { missing = 1; continue; }
            if ((b >> 6) >= words) return -1;
            row[b >> 6] |= (uint64_t)1 << (b & 63);
        }
        return missing;
    }
    return -1;
}

static PyObject* encode(PyObject* self, PyObject* args) {
    PyObject *tasks, *cap_vocab, *spec_vocab;
    int words;
    Py_buffer req, tags, typ, flags, urgent, cx, ndeps;
    if (!PyArg_ParseTuple(args, "O!O!O!iw*w*w*w*w*w*w*", &PyList_Type, &tasks, &PyDict_Type, &cap_vocab, &PyDict_Type,
                          &spec_vocab, &words, &req, &tags, &typ, &flags, &urgent, &cx, &ndeps))
        return NULL;
    const Py_ssize_t n = PyList_GET_SIZE(tasks);
    Py_ssize_t result = n;
    if (words < 1 || req.len < n * words * 8 || tags.len < n * words * 8 || typ.len < n * 4 || flags.len < n ||
        urgent.len < n || cx.len < n * 4 || ndeps.len < n * 4) {
        PyErr_SetString(PyExc_ValueError, "output buffers are too small for the task list");
        result = -2;
        goto done;
    }
    {
        uint64_t* r = (uint64_t*)req.buf;
        uint64_t* tg = (uint64_t*)tags.buf;
        int32_t* ty = (int32_t*)typ.buf;
        uint8_t* fl = (uint8_t*)flags.buf;
        uint8_t* ur = (uint8_t*)urgent.buf;
        int32_t* cxx = (int32_t*)cx.buf;
        int32_t* nd = (int32_t*)ndeps.buf;
        for (Py_ssize_t i = 0; i < n; ++i) {
            PyObject* t = PyList_GET_ITEM(tasks, i);
            if (!PyDict_CheckExact(t)) { result = -1 - i; break; }
            uint8_t f = 0;
            PyObject* v = PyDict_GetItemWithError(t, k_req);
            if (v) {
                f |= 1;
                const int rc = set_bits(v, cap_vocab, r + i * words, words);
                if (rc < 0) { result = -1 - i; break; }
                if (rc > 0) f |= 4;                 /* a capability nobody has: every agent misses it */
            } else if (PyErr_Occurred()) { result = -2; break; }
            v = PyDict_GetItemWithError(t, k_type);
            long tb = -1;
            if (v) {
                f |= 2;
                if (PyUnicode_CheckExact(v)) {
                    tb = vocab_bit(spec_vocab, v);
                    if (tb == -2) { result = -1 - i; break; }
                } else { result = -1 - i; break; }
            } else if (PyErr_Occurred()) { result = -2; break; }
            else {
                tb = vocab_bit(spec_vocab, k_empty);  /* task.get('type', '') */
                if (tb == -2) { result = -2; break; }
            }
            ty[i] = (int32_t)tb;
            v = PyDict_GetItemWithError(t, k_tags);
            if (v) {
                if (set_bits(v, spec_vocab, tg + i * words, words) < 0) { result = -1 - i; break; }
            } else if (PyErr_Occurred()) { result = -2; break; }
            fl[i] = f;
            /* get_task_priority inputs (router.py:135-145) */
            v = PyDict_GetItemWithError(t, k_urgent);
            int u = 0;
            if (v) { u = PyObject_IsTrue(v); if (u < 0) { result = -2; break; } }
            ur[i] = (uint8_t)u;
            v = PyDict_GetItemWithError(t, k_cx);
            int32_t cclass = 0;
            if (v) {
                const double c = PyFloat_AsDouble(v);
                if (c == -1.0 && PyErr_Occurred()) { PyErr_Clear(); result = -1 - i; break; }
                cclass = c > 8 ? 9 : (c > 5 ? 6 : 0);
            }
            cxx[i] = cclass;
            v = PyDict_GetItemWithError(t, k_deps);
            Py_ssize_t d = 0;
            if (v) { d = PyObject_Length(v); if (d < 0) { PyErr_Clear(); result = -1 - i; break; } }
            nd[i] = (int32_t)(d > 1000000 ? 1000000 : d);
        }
    }
done:
    PyBuffer_Release(&req); PyBuffer_Release(&tags); PyBuffer_Release(&typ); PyBuffer_Release(&flags);
    PyBuffer_Release(&urgent); PyBuffer_Release(&cx); PyBuffer_Release(&ndeps);
    if (result == -2) {
        if (!PyErr_Occurred()) PyErr_SetString(PyExc_RuntimeError, "task encoding failed");
        return NULL;
    }
    return PyLong_FromSsize_t(result);
}

static PyMethodDef methods[] = {
    {"encode", encode, METH_VARARGS, "encode(tasks, cap_vocab, spec_vocab, words, req, tags, typ, flags, urgent, cx, ndeps) -> n or -1-i"},
    {NULL, NULL, 0, NULL}};

static struct PyModuleDef moddef = {PyModuleDef_HEAD_INIT, "_taskenc", "task dicts -> routing kernel columns", -1, methods};

PyMODINIT_FUNC PyInit__taskenc(void) {
    k_req = PyUnicode_InternFromString("required_capabilities");
    k_type = PyUnicode_InternFromString("type");
    k_tags = PyUnicode_InternFromString("tags");
    k_urgent = PyUnicode_InternFromString("urgent");
    k_cx = PyUnicode_InternFromString("complexity");
    k_deps = PyUnicode_InternFromString("dependencies");
    k_empty = PyUnicode_InternFromString("");
    if (!k_req || !k_type || !k_tags || !k_urgent || !k_cx || !k_deps || !k_empty) return NULL;
    return PyModule_Create(&moddef);
}
