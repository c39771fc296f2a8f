/* fastcall.c — CPython shim for the two SyncBN entry points that are called once per layer and
 * direction.  ctypes spends ~10 us converting the 19-21 arguments of b200dp_bn_forward /
 * b200dp_bn_backward; a METH_FASTCALL function does it in well under 1 us, which matters for
 * the small layers of CIFAR / 64x64 models where the host, not the GPU, sets the step time.
 * The shim holds no state but the two function addresses handed to bind(); it calls the same
 * C ABI (include/b200dp.h) and returns its int code.  Pointers travel as Python ints (0 = NULL). */
#define PY_SSIZE_T_CLEAN
#include <Python.h>
#include <stdint.h>

typedef int (*fwd_fn)(const void*, const void*, void*, int, int64_t, int, void*, float*, const float*,
                      const float*, float, float, float*, float*, float*, int, void*, uint32_t, void*,
                      uint32_t*, int64_t*);
typedef int (*bwd_fn)(const void*, const void*, const void*, void*, void*, int, int64_t, int, const float*,
                      const float*, const float*, int, void*, float*, float*, float*, void*, uint32_t,
                      void*);
static fwd_fn g_fwd = NULL;
static bwd_fn g_bwd = NULL;

static inline void* P(PyObject* o) { return (void*)(uintptr_t)PyLong_AsUnsignedLongLong(o); }
static inline long long I(PyObject* o) { return PyLong_AsLongLong(o); }

static PyObject* bind(PyObject* self, PyObject* const* a, Py_ssize_t n) {
  if (n != 2) { PyErr_SetString(PyExc_TypeError, "bind(addr_forward, addr_backward)"); return NULL; }
  g_fwd = (fwd_fn)P(a[0]);
  g_bwd = (bwd_fn)P(a[1]);
  if (PyErr_Occurred()) return NULL;
  Py_RETURN_NONE;
}

static PyObject* bn_forward(PyObject* self, PyObject* const* a, Py_ssize_t n) {
  if (n != 21 || !g_fwd) { PyErr_SetString(PyExc_TypeError, "bn_forward takes 21 arguments after bind()"); return NULL; }
  const float eps = (float)PyFloat_AsDouble(a[10]), mom = (float)PyFloat_AsDouble(a[11]);
  void* p[21];
  for (int k = 0; k < 21; ++k) p[k] = NULL;
  const int ptr_idx[] = {0, 1, 2, 6, 7, 8, 9, 12, 13, 14, 16, 18, 19, 20};
  for (unsigned k = 0; k < sizeof(ptr_idx) / sizeof(int); ++k) p[ptr_idx[k]] = P(a[ptr_idx[k]]);
  const int dt = (int)I(a[3]), C = (int)I(a[5]), relu = (int)I(a[15]);
  const int64_t M = (int64_t)I(a[4]);
  const uint32_t seq = (uint32_t)I(a[17]);
  if (PyErr_Occurred()) return NULL;
  const int rc = g_fwd(p[0], p[1], p[2], dt, M, C, p[6], (float*)p[7], (const float*)p[8], (const float*)p[9], eps,
                       mom, (float*)p[12], (float*)p[13], (float*)p[14], relu, p[16], seq, p[18], (uint32_t*)p[19],
                       (int64_t*)p[20]);
  return PyLong_FromLong(rc);
}

static PyObject* bn_backward(PyObject* self, PyObject* const* a, Py_ssize_t n) {
  if (n != 19 || !g_bwd) { PyErr_SetString(PyExc_TypeError, "bn_backward takes 19 arguments after bind()"); return NULL; }
  void* p[19];
  for (int k = 0; k < 19; ++k) p[k] = NULL;
  const int ptr_idx[] = {0, 1, 2, 3, 4, 8, 9, 10, 12, 13, 14, 15, 16, 18};
  for (unsigned k = 0; k < sizeof(ptr_idx) / sizeof(int); ++k) p[ptr_idx[k]] = P(a[ptr_idx[k]]);
  const int dt = (int)I(a[5]), C = (int)I(a[7]), relu = (int)I(a[11]);
  const int64_t M = (int64_t)I(a[6]);
  const uint32_t seq = (uint32_t)I(a[17]);
  if (PyErr_Occurred()) return NULL;
  const int rc = g_bwd(p[0], p[1], p[2], p[3], p[4], dt, M, C, (const float*)p[8], (const float*)p[9],
                       (const float*)p[10], relu, p[12], (float*)p[13], (float*)p[14], (float*)p[15], p[16], seq,
                       p[18]);
  return PyLong_FromLong(rc);
}

static PyMethodDef methods[] = {
    {"bind", (PyCFunction)(void (*)(void))bind, METH_FASTCALL, "bind(addr_forward, addr_backward)"},
    {"bn_forward", (PyCFunction)(void (*)(void))bn_forward, METH_FASTCALL, "b200dp_bn_forward with int arguments"},
    {"bn_backward", (PyCFunction)(void (*)(void))bn_backward, METH_FASTCALL, "b200dp_bn_backward with int arguments"},
    {NULL, NULL, 0, NULL}};
static struct PyModuleDef moddef = {PyModuleDef_HEAD_INIT, "_b200dp_fast", "fast-call shim for libb200dp", -1, methods};
PyMODINIT_FUNC PyInit__b200dp_fast(void) { return PyModule_Create(&moddef); }
