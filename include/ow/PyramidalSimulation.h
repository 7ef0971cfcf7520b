// PyramidalSimulation.h — the muscle-signal hook (/root/reference/src/PyramidalSimulation.h:12-22):
// setup() imports module `main_sim` from PYTHONPATH and instantiates `muscle_simulation()`;
// run() calls its .run() and converts the returned list to floats (96 in the reference).
// Re-built on the Python 3 C-API (the reference uses Python 2.7's PyString_FromString).
// A C callback can be installed instead of the interpreter (setCallback).
#ifndef PYRAMIDALSIMULATION_H
#define PYRAMIDALSIMULATION_H
#include <vector>

typedef void (*ow_muscle_callback)(void *user, float *out96, int n);

class PyramidalSimulation {
 public:
  PyramidalSimulation();
  ~PyramidalSimulation();
  int setup();                 // 0 on success, -1 if the module/class could not be loaded
  std::vector<float> run();    // 96 activations (zeros if the hook is not available)
  void setCallback(ow_muscle_callback cb, void *user);
  bool available() const { return pInstance != 0 || callback != 0; }

 private:
  void *pModule, *pClass, *pInstance;  // PyObject*
  ow_muscle_callback callback;
  void *callbackUser;
  bool ownsInterpreter;
};
#endif
