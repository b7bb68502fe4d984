// Independent writer of a cxxnet .model file, straight from the layout in SURVEY.md section 2.11
// (reference src/cxxnet_main.cpp:173-182, src/nnet/nnet_config.h:126-145, src/nnet/nnet_impl-inl.hpp:82-87,
// src/layer/fullc_layer-inl.hpp:46-50, src/layer/param.h:15-75).  Used by tests/test_model_io.py to prove that
// the Python loader reads files laid out by C structs, not just its own output.
//   net: in -> fullc:fc1 (3 -> 2) -> softmax ; weights w[i][j] = 10*i + j + 0.5, bias = {-1, -2}; epoch 42
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

struct NetParam {
  int num_nodes, num_layers;
  unsigned input_shape[3];
  int init_end, extra_data_num;
  int reserved[31];
};
struct LayerParam {
  int num_hidden; float init_sigma; int init_sparse; float init_uniform; float init_bias;
  int num_channel, random_type, num_group, kernel_height, kernel_width, stride, pad_y, pad_x, no_bias, temp_col_max,
      silent, num_input_channel, num_input_node;
  int reserved[64];
};
static_assert(sizeof(NetParam) == 152, "NetParam");
static_assert(sizeof(LayerParam) == 328, "LayerParam");
static_assert(sizeof(long) == 8, "LP64");

static void wstr(std::string& o, const std::string& s) { uint64_t n = s.size(); o.append((char*)&n, 8); o.append(s); }
static void wvec(std::string& o, const std::vector<int>& v) {
  uint64_t n = v.size(); o.append((char*)&n, 8); if (n) o.append((const char*)v.data(), n * 4);
}

int main(int argc, char** argv) {
  std::string f;
  int net_type = 0;
  f.append((char*)&net_type, 4);
  NetParam np; memset(&np, 0, sizeof(np));
  np.num_nodes = 2; np.num_layers = 2; np.input_shape[0] = 1; np.input_shape[1] = 1; np.input_shape[2] = 3; np.init_end = 1;
  f.append((char*)&np, sizeof(np));
  wstr(f, "in"); wstr(f, "fc1");
  int t = 1, prim = -1;  // fullc
  f.append((char*)&t, 4); f.append((char*)&prim, 4); wstr(f, "fc1"); wvec(f, {0}); wvec(f, {1});
  t = 2;                 // softmax (self loop on node 1)
  f.append((char*)&t, 4); f.append((char*)&prim, 4); wstr(f, ""); wvec(f, {1}); wvec(f, {1});
  long epoch = 42;
  f.append((char*)&epoch, 8);
  std::string blob;
  LayerParam lp; memset(&lp, 0, sizeof(lp));
  lp.num_hidden = 2; lp.init_sigma = 0.01f; lp.init_sparse = 10; lp.init_uniform = -1.0f; lp.num_group = 1; lp.stride = 1;
  lp.temp_col_max = 64 << 18; lp.num_input_node = 3;
  blob.append((char*)&lp, sizeof(lp));
  unsigned shp2[2] = {2, 3};
  blob.append((char*)shp2, 8);
  for (int i = 0; i < 2; ++i) for (int j = 0; j < 3; ++j) { float v = 10.0f * i + j + 0.5f; blob.append((char*)&v, 4); }
  unsigned shp1[1] = {2};
  blob.append((char*)shp1, 4);
  float b[2] = {-1.0f, -2.0f};
  blob.append((char*)b, 8);
  wstr(f, blob);
  FILE* fo = fopen(argc > 1 ? argv[1] : "golden.model", "wb");
  fwrite(f.data(), 1, f.size(), fo);
  fclose(fo);
  return 0;
}
