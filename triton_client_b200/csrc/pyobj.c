/* _pyobj: the two loops over Python objects that the BYTES drop-ins cannot avoid, in C.
 *
 * The reference's serialize_byte_tensor / deserialize_bytes_tensor take and return np.object_
 * arrays of bytes objects (/root/reference/src/python/library/tritonclient/utils/__init__.py:
 * 193-276). The GPU kernels work on (payload, lengths / offsets); what stays on the host is
 *   - going in:  visit every element, str()-encode what is not exactly bytes (:229-237), note its
 *                length and copy its bytes into one staging buffer;
 *   - coming out: make one bytes object per element from (payload, offsets).
 * In Python these loops cost 170 ms and 160 ms per million elements - far more than the copies and
 * kernels together. Here: one pass with the GIL for the object work, the payload copy without it
 * on a few threads.
 *
 * Plain CPython C API, no numpy headers: arrays are passed as addresses (ndarray.ctypes.data) of
 * C-contiguous buffers whose lifetime the Python caller guarantees.
 */
#define PY_SSIZE_T_CLEAN
#include <Python.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
  const char* ptr;
  int64_t len;
} Piece;

typedef struct {
  Piece* pieces;
  int64_t n;
  int64_t total;
  PyObject* keep; /* list of the bytes objects made for elements that were not bytes, or NULL */
} Gather;

static void gather_free(PyObject* capsule) {
  Gather* g = (Gather*)PyCapsule_GetPointer(capsule, "tcb.gather");
  if (!g) return;
  free(g->pieces);
  Py_XDECREF(g->keep);
  free(g);
}

/* prepare(items_addr, n, lens_addr) -> (total_bytes, handle)
 * items_addr: address of n PyObject* (data of a C-contiguous np.object_ array, kept alive and
 * unchanged by the caller until copy() has returned); lens_addr: address of n int64 to fill. */
static PyObject* py_prepare(PyObject* self, PyObject* args) {
  unsigned long long items_addr, lens_addr;
  long long n;
  if (!PyArg_ParseTuple(args, "KLK", &items_addr, &n, &lens_addr)) return NULL;
  if (n < 0) {
    PyErr_SetString(PyExc_ValueError, "negative element count");
    return NULL;
  }
  PyObject** items = (PyObject**)(uintptr_t)items_addr;
  int64_t* lens = (int64_t*)(uintptr_t)lens_addr;
  Gather* g = (Gather*)calloc(1, sizeof(Gather));
  if (!g) return PyErr_NoMemory();
  g->pieces = (Piece*)malloc(sizeof(Piece) * (size_t)(n > 0 ? n : 1));
  if (!g->pieces) {
    free(g);
    return PyErr_NoMemory();
  }
  g->n = n;
  PyObject* capsule = PyCapsule_New(g, "tcb.gather", gather_free);
  if (!capsule) {
    free(g->pieces);
    free(g);
    return NULL;
  }
  int64_t total = 0;
  for (long long i = 0; i < n; i++) {
    PyObject* o = items[i];
    PyObject* b = o;
    if (!o) {
      Py_DECREF(capsule);
      PyErr_SetString(PyExc_ValueError, "object array holds a NULL element");
      return NULL;
    }
    if (!PyBytes_CheckExact(o)) { /* the reference: type(obj) != bytes -> str(obj).encode("utf-8") */
      PyObject* s = PyObject_Str(o);
      if (!s) {
        Py_DECREF(capsule);
        return NULL;
      }
      b = PyUnicode_AsUTF8String(s);
      Py_DECREF(s);
      if (!b) {
        Py_DECREF(capsule);
        return NULL;
      }
      if (!g->keep) g->keep = PyList_New(0);
      if (!g->keep || PyList_Append(g->keep, b) < 0) {
        Py_DECREF(b);
        Py_DECREF(capsule);
        return NULL;
      }
      Py_DECREF(b); /* the list holds it */
    }
    g->pieces[i].ptr = PyBytes_AS_STRING(b);
    g->pieces[i].len = (int64_t)PyBytes_GET_SIZE(b);
    lens[i] = g->pieces[i].len;
    total += g->pieces[i].len;
  }
  g->total = total;
  return Py_BuildValue("LN", (long long)total, capsule);
}

/* copy(handle, dst_addr, dst_capacity, threads) -> bytes copied. The payloads back to back, in
 * element order. Runs without the GIL; `threads` > 1 splits the elements into equal byte shares. */
static PyObject* py_copy(PyObject* self, PyObject* args) {
  PyObject* capsule;
  unsigned long long dst_addr, dst_cap;
  int threads;
  if (!PyArg_ParseTuple(args, "OKKi", &capsule, &dst_addr, &dst_cap, &threads)) return NULL;
  Gather* g = (Gather*)PyCapsule_GetPointer(capsule, "tcb.gather");
  if (!g) return NULL;
  if ((unsigned long long)g->total > dst_cap) {
    PyErr_SetString(PyExc_ValueError, "destination too small for the gathered payload");
    return NULL;
  }
  char* dst = (char*)(uintptr_t)dst_addr;
  const Piece* p = g->pieces;
  const int64_t n = g->n;
  if (threads < 1) threads = 1;
  if (threads > 16) threads = 16;
  if (g->total < (4 << 20)) threads = 1;
  Py_BEGIN_ALLOW_THREADS
  if (threads == 1) {
    int64_t off = 0;
    for (int64_t i = 0; i < n; i++) {
      memcpy(dst + off, p[i].ptr, (size_t)p[i].len);
      off += p[i].len;
    }
  } else {
    /* element ranges of roughly equal bytes: first[t] = first element of share t, at byte at[t] */
    int64_t first[17], at[17];
    int64_t off = 0;
    int t = 0;
    const int64_t share = g->total / threads + 1;
    first[0] = 0;
    at[0] = 0;
    for (int64_t i = 0; i < n && t + 1 < threads; i++) {
      if (off >= (int64_t)(t + 1) * share) {
        t++;
        first[t] = i;
        at[t] = off;
      }
      off += p[i].len;
    }
    const int used = t + 1;
    first[used] = n;
#pragma omp parallel for num_threads(used) schedule(static, 1)
    for (int k = 0; k < used; k++) {
      int64_t o = at[k];
      for (int64_t i = first[k]; i < first[k + 1]; i++) {
        memcpy(dst + o, p[i].ptr, (size_t)p[i].len);
        o += p[i].len;
      }
    }
  }
  Py_END_ALLOW_THREADS
  return PyLong_FromLongLong((long long)g->total);
}

/* scatter(payload_addr, offsets_addr, n, items_addr) -> None
 * items[i] = bytes(payload[offsets[i] : offsets[i + 1]]) for the n slots of a C-contiguous
 * np.object_ array (whatever they held is released). */
static PyObject* py_scatter(PyObject* self, PyObject* args) {
  unsigned long long payload_addr, offsets_addr, items_addr;
  long long n;
  if (!PyArg_ParseTuple(args, "KKLK", &payload_addr, &offsets_addr, &n, &items_addr)) return NULL;
  const char* payload = (const char*)(uintptr_t)payload_addr;
  const int64_t* off = (const int64_t*)(uintptr_t)offsets_addr;
  PyObject** items = (PyObject**)(uintptr_t)items_addr;
  for (long long i = 0; i < n; i++) {
    const int64_t a = off[i], b = off[i + 1];
    if (b < a) {
      PyErr_SetString(PyExc_ValueError, "offsets are not monotonic");
      return NULL;
    }
    PyObject* v = PyBytes_FromStringAndSize(payload + a, (Py_ssize_t)(b - a));
    if (!v) return NULL;
    PyObject* old = items[i];
    items[i] = v;
    Py_XDECREF(old);
  }
  Py_RETURN_NONE;
}

static PyMethodDef methods[] = {
    {"prepare", py_prepare, METH_VARARGS, "prepare(items_addr, n, lens_addr) -> (total, handle)"},
    {"copy", py_copy, METH_VARARGS, "copy(handle, dst_addr, dst_capacity, threads) -> bytes copied"},
    {"scatter", py_scatter, METH_VARARGS, "scatter(payload_addr, offsets_addr, n, items_addr)"},
    {NULL, NULL, 0, NULL}};

static struct PyModuleDef module = {PyModuleDef_HEAD_INIT, "_pyobj", "object-array loops of the BYTES drop-ins", -1, methods};

PyMODINIT_FUNC PyInit__pyobj(void) { return PyModule_Create(&module); }
