/*
 * C ABI of cxxnet_b200: the 22 functions of the reference's wrapper/cxxnet_wrapper.h:29-225 with
 * the same names, argument order and ownership rules (returned pointers are owned by the
 * library and stay valid until the next call on the same handle).  Link with
 * cxxnet_b200/lib/libcxxnetwrapper.so; the library hosts a CPython interpreter that runs the
 * cxxnet_b200 package (set PYTHONPATH to the repository root, or CXXNET_B200_ROOT).
 */
#ifndef CXXNET_WRAPPER_H_
#define CXXNET_WRAPPER_H_

#ifdef __cplusplus
extern "C" {
#endif

typedef unsigned int cxx_uint;
typedef float cxx_real_t;

void *CXNIOCreateFromConfig(const char *cfg);
int CXNIONext(void *handle);
void CXNIOBeforeFirst(void *handle);
const cxx_real_t *CXNIOGetData(void *handle, cxx_uint oshape[4], cxx_uint *ostride);
const cxx_real_t *CXNIOGetLabel(void *handle, cxx_uint oshape[2], cxx_uint *ostride);
void CXNIOFree(void *handle);

void *CXNNetCreate(const char *device, const char *cfg);
void CXNNetFree(void *handle);
void CXNNetSetParam(void *handle, const char *name, const char *val);
void CXNNetInitModel(void *handle);
void CXNNetSaveModel(void *handle, const char *fname);
void CXNNetLoadModel(void *handle, const char *fname);
void CXNNetStartRound(void *handle, int round);
void CXNNetSetWeight(void *handle, cxx_real_t *p_weight, cxx_uint size_weight, const char *layer_name,
                     const char *wtag);
const cxx_real_t *CXNNetGetWeight(void *handle, const char *layer_name, const char *wtag, cxx_uint wshape[4],
                                  cxx_uint *out_dim);
void CXNNetUpdateIter(void *handle, void *data_handle);
void CXNNetUpdateBatch(void *handle, cxx_real_t *p_data, const cxx_uint dshape[4], cxx_real_t *p_label,
                       const cxx_uint lshape[2]);
const cxx_real_t *CXNNetPredictBatch(void *handle, cxx_real_t *p_data, const cxx_uint dshape[4], cxx_uint *out_size);
const cxx_real_t *CXNNetPredictIter(void *handle, void *data_handle, cxx_uint *out_size);
const cxx_real_t *CXNNetExtractBatch(void *handle, cxx_real_t *p_data, const cxx_uint dshape[4],
                                     const char *node_name, cxx_uint oshape[4]);
const cxx_real_t *CXNNetExtractIter(void *handle, void *data_handle, const char *node_name, cxx_uint oshape[4]);
const char *CXNNetEvaluate(void *handle, void *data_handle, const char *data_name);

#ifdef __cplusplus
}
#endif
#endif /* CXXNET_WRAPPER_H_ */
