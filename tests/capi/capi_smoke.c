/* Exercises the 22-function C ABI from plain C: iterator, net creation, training through the
 * iterator and through raw arrays, predict / extract / evaluate, weight get / set, save / load.
 * Mirrors what the reference's example/MNIST/mnist.py does through ctypes. */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "cxxnet_wrapper.h"

#define CHECK(c) do { if (!(c)) { fprintf(stderr, "CHECK failed line %d: %s\n", __LINE__, #c); return 1; } } while (0)

static const char *ITER_CFG =
    "iter = mnist\n path_img = \"none\"\n path_label = \"none\"\n num_synthetic = 2000\n batch_size = 100\n silent = 1\niter = end\n";
static const char *NET_CFG =
    "netconfig=start\nlayer[+1:fc1] = fullc:fc1\n  nhidden = 16\n  init_sigma = 0.01\nlayer[+1:sg1] = sigmoid:se1\n"
    "layer[sg1->fc2] = fullc:fc2\n  nhidden = 10\n  init_sigma = 0.01\nlayer[+0] = softmax\nnetconfig=end\n"
    "input_shape = 1,1,784\nbatch_size = 100\neta = 0.1\nmomentum = 0.9\nmetric[label] = error\nsilent = 1\n";

int main(int argc, char **argv) {
  const char *model = argc > 1 ? argv[1] : "/tmp/capi.model";
  void *it = CXNIOCreateFromConfig(ITER_CFG);
  CHECK(it != NULL);
  void *net = CXNNetCreate("cpu", NET_CFG);
  CHECK(net != NULL);
  CXNNetSetParam(net, "seed", "1");
  CXNNetInitModel(net);
  for (int r = 0; r < 5; ++r) {
    CXNNetStartRound(net, r);
    CXNIOBeforeFirst(it);
    while (CXNIONext(it)) CXNNetUpdateIter(net, it);
  }
  const char *ev = CXNNetEvaluate(net, it, "eval");
  CHECK(strstr(ev, "eval-error:") != NULL);
  printf("evaluate: %s\n", ev);
  /* iterator data / label access */
  CXNIOBeforeFirst(it);
  CHECK(CXNIONext(it));
  cxx_uint ds[4], ls[2], stride;
  const cxx_real_t *d = CXNIOGetData(it, ds, &stride);
  CHECK(ds[0] == 100 && ds[1] == 1 && ds[2] == 1 && ds[3] == 784 && stride == 784);
  float *data = (float *)malloc(sizeof(float) * 100 * 784);
  memcpy(data, d, sizeof(float) * 100 * 784);
  const cxx_real_t *l = CXNIOGetLabel(it, ls, &stride);
  CHECK(ls[0] == 100 && ls[1] == 1);
  float *label = (float *)malloc(sizeof(float) * 100);
  memcpy(label, l, sizeof(float) * 100);
  /* predict: iterator path and array path agree */
  cxx_uint n1, n2;
  const cxx_real_t *p1 = CXNNetPredictIter(net, it, &n1);
  float pred1[100];
  CHECK(n1 == 100);
  memcpy(pred1, p1, sizeof(pred1));
  const cxx_real_t *p2 = CXNNetPredictBatch(net, data, ds, &n2);
  CHECK(n2 == 100);
  int same = 0, hit = 0;
  for (int i = 0; i < 100; ++i) { same += pred1[i] == p2[i]; hit += p2[i] == label[i]; }
  CHECK(same == 100);
  printf("train accuracy on first batch: %d/100\n", hit);
  CHECK(hit > 60);
  /* extract */
  cxx_uint os[4];
  const cxx_real_t *f = CXNNetExtractBatch(net, data, ds, "sg1", os);
  CHECK(os[0] == 100 && os[3] == 16 && f[0] >= 0.0f && f[0] <= 1.0f);
  f = CXNNetExtractIter(net, it, "top[-1]", os);
  CHECK(os[3] == 10);
  float s = 0;
  for (int i = 0; i < 10; ++i) s += f[i];
  CHECK(fabsf(s - 1.0f) < 1e-4f);
  /* weight get / set round trip */
  cxx_uint ws[4], wd;
  const cxx_real_t *w = CXNNetGetWeight(net, "fc2", "wmat", ws, &wd);
  CHECK(w != NULL && wd == 2 && ws[0] == 10 && ws[1] == 16);
  float wcopy[160];
  memcpy(wcopy, w, sizeof(wcopy));
  float zeros[160] = {0};
  CXNNetSetWeight(net, zeros, 160, "fc2", "wmat");
  w = CXNNetGetWeight(net, "fc2", "wmat", ws, &wd);
  CHECK(w[3] == 0.0f);
  CXNNetSetWeight(net, wcopy, 160, "fc2", "wmat");
  /* manual update loop with raw arrays */
  cxx_uint lshape[2] = {100, 1};
  CXNNetUpdateBatch(net, data, ds, label, lshape);
  /* save / load */
  CXNNetSaveModel(net, model);
  void *net2 = CXNNetCreate("cpu", NET_CFG);
  CXNNetLoadModel(net2, model);
  const cxx_real_t *q = CXNNetPredictBatch(net2, data, ds, &n2);
  float predq[100];
  memcpy(predq, q, sizeof(predq));
  const cxx_real_t *q1 = CXNNetPredictBatch(net, data, ds, &n1);
  for (int i = 0; i < 100; ++i) CHECK(predq[i] == q1[i]);
  CXNNetFree(net2);
  CXNNetFree(net);
  CXNIOFree(it);
  free(data);
  free(label);
  printf("C ABI OK\n");
  return 0;
}
