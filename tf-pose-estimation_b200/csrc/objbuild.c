/* objbuild.c -- CPython extension `_objbuild`: builds the list[list[Human]] of a decoded batch from the
 * tfpd_human tables in one call.
 *
 * What it replaces: the object loop of PoseEstimator.estimate_paf (tf_pose/estimator.py:276-299), which the
 * reference runs once per frame over 73 swig getter calls per human.  The objects are instances of THIS
 * package's Human / BodyPart classes (tf-pose-estimation_b200/humans.py; same __slots__, uidx and value types as
 * the reference's, estimator.py:43-265) -- the classes are passed in, nothing is re-declared here.  Equivalent,
 * attribute for attribute, to decoder.humans_from_table (tests/test_host.py checks that).
 *
 * Why native: building ~3,500 BodyPart objects per 64-frame batch through BodyPart.__init__ costs ~1.5 us each
 * (5 ms per batch: 12 k frames/s on one core), an order of magnitude below what the GPU path delivers to the
 * host.  Here an instance is allocated with the type's own allocator and its slots are stored directly
 * (offsets taken from the classes' member descriptors), uidx strings come from a cache.
 *
 * Not part of the C-ABI (include/tfpose_decode.h): libtfpose_decode.so has no Python dependency; this module
 * sits beside the ctypes binding on the Python side of the boundary.
 */
#define PY_SSIZE_T_CLEAN
#include <Python.h>
#include <descrobject.h>
#include <stdint.h>
#include <structmember.h>

#include "../../include/tfpose_decode.h"

#define NPARTS TFPD_NUM_PARTS
#define UIDX_HUMANS 512 /* uidx strings '<human>-<part>' cached for human ids below this */

static PyObject *g_uidx[UIDX_HUMANS][NPARTS];
static PyObject *g_part_int[NPARTS];
static PyObject *g_empty_tuple;

/* byte offset of slot `name` inside instances of `type` (a __slots__ member descriptor), or -1 */
static Py_ssize_t slot_offset(PyObject *type, const char *name) {
    PyObject *d = PyObject_GetAttrString(type, name);
    Py_ssize_t off = -1;
    if (!d) {
        PyErr_Clear();
        PyErr_Format(PyExc_TypeError, "%R has no __slots__ member '%s'", type, name);
        return -1;
    }
    if (Py_TYPE(d) == &PyMemberDescr_Type) {
        PyMemberDef *m = ((PyMemberDescrObject *)d)->d_member;
        if (m && m->type == T_OBJECT_EX) off = m->offset;
    }
    Py_DECREF(d);
    if (off < 0) PyErr_Format(PyExc_TypeError, "%s is not a __slots__ member of %R", name, type);
    return off;
}

static inline void set_slot(PyObject *obj, Py_ssize_t off, PyObject *v /* stolen */) {
    PyObject **p = (PyObject **)((char *)obj + off);
    PyObject *old = *p;
    *p = v;
    Py_XDECREF(old);
}

static PyObject *uidx_of(int human, int part) {
    if (human < UIDX_HUMANS) {
        PyObject *s = g_uidx[human][part];
        if (!s) {
            s = PyUnicode_FromFormat("%d-%d", human, part); /* '%d-%d' % (human_id, part_idx), estimator.py:288 */
            if (!s) return NULL;
            g_uidx[human][part] = s;
        }
        Py_INCREF(s);
        return s;
    }
    return PyUnicode_FromFormat("%d-%d", human, part);
}

/* build_humans(BodyPart, Human, table, counts, max_humans, H, W) -> list[list[Human]]
 *   table  : buffer of N * max_humans tfpd_human records (frame i at i * max_humans)
 *   counts : buffer of N int32 = humans per frame (tfpd_frame_info.n_humans) */
static PyObject *build_humans(PyObject *self, PyObject *args) {
    PyObject *bp_type, *hm_type, *result = NULL;
    Py_buffer tab = {0}, cnt = {0};
    int max_humans, H, W;
    if (!PyArg_ParseTuple(args, "OOy*y*iii", &bp_type, &hm_type, &tab, &cnt, &max_humans, &H, &W)) return NULL;
    if (!PyType_Check(bp_type) || !PyType_Check(hm_type) || max_humans < 1 || H < 1 || W < 1) {
        PyErr_SetString(PyExc_ValueError, "build_humans(BodyPart, Human, table, counts, max_humans, H, W)");
        goto done;
    }
    const Py_ssize_t N = cnt.len / (Py_ssize_t)sizeof(int32_t);
    if (tab.len < N * (Py_ssize_t)max_humans * (Py_ssize_t)sizeof(tfpd_human)) {
        PyErr_SetString(PyExc_ValueError, "table buffer is smaller than len(counts) * max_humans records");
        goto done;
    }
    PyTypeObject *bpt = (PyTypeObject *)bp_type, *hmt = (PyTypeObject *)hm_type;
    /* one lookup at a time, stopping at the first failure: no Python API call may run with an exception pending */
    static const char *const bp_slots[5] = {"uidx", "part_idx", "x", "y", "score"};
    static const char *const hm_slots[4] = {"body_parts", "pairs", "uidx_list", "score"};
    Py_ssize_t bo[5], ho[4];
    for (int i = 0; i < 5; ++i)
        if ((bo[i] = slot_offset(bp_type, bp_slots[i])) < 0) goto done;
    for (int i = 0; i < 4; ++i)
        if ((ho[i] = slot_offset(hm_type, hm_slots[i])) < 0) goto done;
    const Py_ssize_t o_uidx = bo[0], o_pidx = bo[1], o_x = bo[2], o_y = bo[3], o_score = bo[4];
    const Py_ssize_t o_parts = ho[0], o_pairs = ho[1], o_ulist = ho[2], o_hscore = ho[3];

    const tfpd_human *table = (const tfpd_human *)tab.buf;
    const int32_t *counts = (const int32_t *)cnt.buf;
    const double dW = (double)W, dH = (double)H;
    result = PyList_New(N);
    if (!result) goto done;
    /* thousands of GC-tracked instances are allocated below, none of which can be part of a cycle yet: keep the
     * young-generation collections they would trigger (one per 700 allocations) out of the loop */
    const int gc_was_enabled = PyGC_Disable();
    for (Py_ssize_t f = 0; f < N; ++f) {
        int n = counts[f];
        if (n < 0) n = 0;
        if (n > max_humans) n = max_humans;
        PyObject *frame = PyList_New(0);
        if (!frame) goto fail;
        PyList_SET_ITEM(result, f, frame);
        for (int hid = 0; hid < n; ++hid) {
            const tfpd_human *r = table + f * max_humans + hid;
            PyObject *parts = PyDict_New();
            if (!parts) goto fail;
            for (int k = 0; k < NPARTS; ++k) {
                if (r->cid[k] < 0) continue; /* estimator.py:283-284 */
                PyObject *bp = bpt->tp_alloc(bpt, 0);
                if (!bp) {
                    Py_DECREF(parts);
                    goto fail;
                }
                PyObject *u = uidx_of(hid, k);
                /* float(get_part_x(c_idx)) / W, float(get_part_y(c_idx)) / H : Python float division (:289-290) */
                PyObject *x = PyFloat_FromDouble((double)r->x[k] / dW);
                PyObject *y = PyFloat_FromDouble((double)r->y[k] / dH);
                PyObject *s = PyFloat_FromDouble((double)r->part_score[k]);
                if (!u || !x || !y || !s) {
                    Py_XDECREF(u), Py_XDECREF(x), Py_XDECREF(y), Py_XDECREF(s);
                    Py_DECREF(bp);
                    Py_DECREF(parts);
                    goto fail;
                }
                set_slot(bp, o_uidx, u);
                Py_INCREF(g_part_int[k]);
                set_slot(bp, o_pidx, g_part_int[k]);
                set_slot(bp, o_x, x);
                set_slot(bp, o_y, y);
                set_slot(bp, o_score, s);
                const int rc = PyDict_SetItem(parts, g_part_int[k], bp); /* ascending part order (:281-292) */
                Py_DECREF(bp);
                if (rc < 0) {
                    Py_DECREF(parts);
                    goto fail;
                }
            }
            if (PyDict_GET_SIZE(parts) == 0) { /* is_added stays False: the human is dropped (:294) */
                Py_DECREF(parts);
                continue;
            }
            PyObject *hm = hmt->tp_alloc(hmt, 0);
            PyObject *pairs = PyList_New(0), *ulist = PySet_New(NULL); /* Human([]): pairs=[], uidx_list=set() (:49-55) */
            PyObject *score = PyFloat_FromDouble((double)r->score);
            if (!hm || !pairs || !ulist || !score) {
                Py_XDECREF(hm), Py_XDECREF(pairs), Py_XDECREF(ulist), Py_XDECREF(score);
                Py_DECREF(parts);
                goto fail;
            }
            set_slot(hm, o_parts, parts);
            set_slot(hm, o_pairs, pairs);
            set_slot(hm, o_ulist, ulist);
            set_slot(hm, o_hscore, score);
            const int rc = PyList_Append(frame, hm);
            Py_DECREF(hm);
            if (rc < 0) goto fail;
        }
    }
    if (gc_was_enabled) PyGC_Enable();
    goto done;
fail:
    if (gc_was_enabled) PyGC_Enable();
    Py_CLEAR(result);
done:
    PyBuffer_Release(&tab);
    PyBuffer_Release(&cnt);
    return result;
}

static PyMethodDef methods[] = {
    {"build_humans", build_humans, METH_VARARGS,
     "build_humans(BodyPart, Human, table, counts, max_humans, H, W) -> list[list[Human]]"},
    {NULL, NULL, 0, NULL}};

static struct PyModuleDef moddef = {PyModuleDef_HEAD_INIT, "_objbuild",
                                    "Human/BodyPart construction from tfpd_human tables (estimator.py:276-299)", -1,
                                    methods};

PyMODINIT_FUNC PyInit__objbuild(void) {
    PyObject *m = PyModule_Create(&moddef);
    if (!m) return NULL;
    g_empty_tuple = PyTuple_New(0);
    for (int k = 0; k < NPARTS; ++k) {
        g_part_int[k] = PyLong_FromLong(k);
        if (!g_part_int[k]) return NULL;
    }
    PyModule_AddIntConstant(m, "HUMAN_RECORD_BYTES", (long)sizeof(tfpd_human));
    return m;
}
