// C ABI shim: the reference's 22 wrapper functions (wrapper/cxxnet_wrapper.cpp:236-352), forwarded to
// cxxnet_b200.wrapper.capi inside an embedded (or already running) CPython interpreter.
// Errors in the Python layer are printed and abort the process, matching the reference's
// utils::Error -> exit(-1) policy (src/utils/utils.h:67-78).
#include <Python.h>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <string>

#include "cxxnet_wrapper.h"

namespace {

PyObject* g_mod = nullptr;
std::once_flag g_once;

void init_python() {
  std::call_once(g_once, [] {
    bool we_started = false;
    if (!Py_IsInitialized()) {
      Py_InitializeEx(0);
      we_started = true;
    }
    PyGILState_STATE st = PyGILState_Ensure();
    const char* root = getenv("CXXNET_B200_ROOT");
    if (root != nullptr) {
      std::string code = std::string("import sys\nsys.path.insert(0, r'") + root + "')\n";
      PyRun_SimpleString(code.c_str());
    }
    g_mod = PyImport_ImportModule("cxxnet_b200.wrapper.capi");
    if (g_mod == nullptr) {
      PyErr_Print();
      fprintf(stderr, "cxxnet wrapper: cannot import cxxnet_b200.wrapper.capi (set PYTHONPATH or CXXNET_B200_ROOT)\n");
      exit(-1);
    }
    PyGILState_Release(st);
    if (we_started) PyEval_SaveThread();  // release the GIL held since Py_Initialize
  });
}

// call capi.<fn>(*args) -> unsigned long long.  fmt uses Py_BuildValue codes: K = handle/pointer, s/z = string, i = int
template <typename... A>
unsigned long long call(const char* fn, const char* fmt, A... args) {
  init_python();
  PyGILState_STATE st = PyGILState_Ensure();
  PyObject* f = PyObject_GetAttrString(g_mod, fn);
  PyObject* a = Py_BuildValue(fmt, args...);
  PyObject* r = (f && a) ? PyObject_CallObject(f, a) : nullptr;
  if (r == nullptr) {
    PyErr_Print();
    fprintf(stderr, "cxxnet wrapper: %s failed\n", fn);
    exit(-1);
  }
  unsigned long long v = PyLong_AsUnsignedLongLong(r);
  Py_XDECREF(r);
  Py_XDECREF(a);
  Py_XDECREF(f);
  PyGILState_Release(st);
  return v;
}
inline unsigned long long H(const void* p) { return (unsigned long long)(uintptr_t)p; }

}  // namespace

extern "C" {

void* CXNIOCreateFromConfig(const char* cfg) { return (void*)(uintptr_t)call("io_create", "(s)", cfg); }
int CXNIONext(void* h) { return (int)call("io_next", "(K)", H(h)); }
void CXNIOBeforeFirst(void* h) { call("io_before_first", "(K)", H(h)); }
const cxx_real_t* CXNIOGetData(void* h, cxx_uint oshape[4], cxx_uint* ostride) {
  return (const cxx_real_t*)(uintptr_t)call("io_get_data", "(KKK)", H(h), H(oshape), H(ostride));
}
const cxx_real_t* CXNIOGetLabel(void* h, cxx_uint oshape[2], cxx_uint* ostride) {
  return (const cxx_real_t*)(uintptr_t)call("io_get_label", "(KKK)", H(h), H(oshape), H(ostride));
}
void CXNIOFree(void* h) { call("free", "(K)", H(h)); }

void* CXNNetCreate(const char* device, const char* cfg) {
  return (void*)(uintptr_t)call("net_create", "(zz)", device, cfg);
}
void CXNNetFree(void* h) { call("free", "(K)", H(h)); }
void CXNNetSetParam(void* h, const char* name, const char* val) { call("net_set_param", "(Kss)", H(h), name, val); }
void CXNNetInitModel(void* h) { call("net_init_model", "(K)", H(h)); }
void CXNNetSaveModel(void* h, const char* fname) { call("net_save_model", "(Ks)", H(h), fname); }
void CXNNetLoadModel(void* h, const char* fname) { call("net_load_model", "(Ks)", H(h), fname); }
void CXNNetStartRound(void* h, int round) { call("net_start_round", "(Ki)", H(h), round); }
void CXNNetSetWeight(void* h, cxx_real_t* p_weight, cxx_uint size_weight, const char* layer_name, const char* wtag) {
  call("net_set_weight", "(KKIss)", H(h), H(p_weight), size_weight, layer_name, wtag);
}
const cxx_real_t* CXNNetGetWeight(void* h, const char* layer_name, const char* wtag, cxx_uint wshape[4],
                                  cxx_uint* out_dim) {
  return (const cxx_real_t*)(uintptr_t)call("net_get_weight", "(KssKK)", H(h), layer_name, wtag, H(wshape), H(out_dim));
}
void CXNNetUpdateIter(void* h, void* dh) { call("net_update_iter", "(KK)", H(h), H(dh)); }
void CXNNetUpdateBatch(void* h, cxx_real_t* p_data, const cxx_uint dshape[4], cxx_real_t* p_label,
                       const cxx_uint lshape[2]) {
  call("net_update_batch", "(KKKKK)", H(h), H(p_data), H(dshape), H(p_label), H(lshape));
}
const cxx_real_t* CXNNetPredictBatch(void* h, cxx_real_t* p_data, const cxx_uint dshape[4], cxx_uint* out_size) {
  return (const cxx_real_t*)(uintptr_t)call("net_predict_batch", "(KKKK)", H(h), H(p_data), H(dshape), H(out_size));
}
const cxx_real_t* CXNNetPredictIter(void* h, void* dh, cxx_uint* out_size) {
  return (const cxx_real_t*)(uintptr_t)call("net_predict_iter", "(KKK)", H(h), H(dh), H(out_size));
}
const cxx_real_t* CXNNetExtractBatch(void* h, cxx_real_t* p_data, const cxx_uint dshape[4], const char* node_name,
                                     cxx_uint oshape[4]) {
  return (const cxx_real_t*)(uintptr_t)call("net_extract_batch", "(KKKsK)", H(h), H(p_data), H(dshape), node_name,
                                            H(oshape));
}
const cxx_real_t* CXNNetExtractIter(void* h, void* dh, const char* node_name, cxx_uint oshape[4]) {
  return (const cxx_real_t*)(uintptr_t)call("net_extract_iter", "(KKsK)", H(h), H(dh), node_name, H(oshape));
}
const char* CXNNetEvaluate(void* h, void* dh, const char* data_name) {
  return (const char*)(uintptr_t)call("net_evaluate", "(KKs)", H(h), H(dh), data_name);
}

}  // extern "C"
