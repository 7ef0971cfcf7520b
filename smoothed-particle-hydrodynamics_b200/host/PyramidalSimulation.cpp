// PyramidalSimulation.cpp — muscle-signal hook on the Python 3 C-API.
// Mirrors /root/reference/src/PyramidalSimulation.cpp:21-93: import `main_sim`, instantiate
// `muscle_simulation`, call `.run()` each step, copy the list into a fixed vector of 96 floats
// (:57 — shorter lists leave zeros, longer ones are truncated instead of overflowing).
// Works both stand-alone (owns the interpreter) and inside an already running Python (ctypes host).
#include "ow/PyramidalSimulation.h"

#include <Python.h>

#include <iostream>

PyramidalSimulation::PyramidalSimulation()
    : pModule(0), pClass(0), pInstance(0), callback(0), callbackUser(0), ownsInterpreter(false) {}

PyramidalSimulation::~PyramidalSimulation() {
  if (pInstance && Py_IsInitialized()) {
    PyGILState_STATE g = PyGILState_Ensure();
    Py_XDECREF((PyObject *)pInstance);
    Py_XDECREF((PyObject *)pClass);
    Py_XDECREF((PyObject *)pModule);
    PyGILState_Release(g);
  }
}

void PyramidalSimulation::setCallback(ow_muscle_callback cb, void *user) { callback = cb; callbackUser = user; }

int PyramidalSimulation::setup() {
  if (callback) return 0;
  if (!Py_IsInitialized()) {
    Py_Initialize();
    ownsInterpreter = true;
  }
  PyGILState_STATE g = PyGILState_Ensure();
  int rc = -1;
  PyObject *name = PyUnicode_FromString("main_sim");
  PyObject *mod = PyImport_Import(name);
  Py_XDECREF(name);
  if (mod) {
    PyObject *cls = PyObject_GetAttrString(mod, "muscle_simulation");
    if (cls && PyCallable_Check(cls)) {
      PyObject *inst = PyObject_CallObject(cls, NULL);
      if (inst) {
        pModule = mod; pClass = cls; pInstance = inst;
        std::cout << "Pyramidal simulation class loaded!" << std::endl;
        rc = 0;
      }
    } else {
      std::cout << "Pyramidal simulation class not callable!" << std::endl;
    }
  } else {
    std::cout << "Module not loaded, have you set PYTHONPATH?" << std::endl;
  }
  if (rc != 0) PyErr_Clear();
  PyGILState_Release(g);
  if (ownsInterpreter) PyEval_SaveThread();  // release the GIL between steps
  return rc;
}

std::vector<float> PyramidalSimulation::run() {
  std::vector<float> out(96, 0.f);
  if (callback) { callback(callbackUser, out.data(), 96); return out; }
  if (!pInstance) return out;
  PyGILState_STATE g = PyGILState_Ensure();
  PyObject *v = PyObject_CallMethod((PyObject *)pInstance, "run", NULL);
  if (v && PyList_Check(v)) {
    Py_ssize_t n = PyList_Size(v);
    for (Py_ssize_t i = 0; i < n && i < 96; i++) out[i] = (float)PyFloat_AsDouble(PyList_GetItem(v, i));
  } else if (v) {
    out[0] = (float)PyFloat_AsDouble(v);
  }
  if (PyErr_Occurred()) PyErr_Clear();
  Py_XDECREF(v);
  PyGILState_Release(g);
  return out;
}
