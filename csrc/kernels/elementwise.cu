// Element-wise / layout kernels (K7, K12, K14 of SURVEY.md section 2.9).
// Device activations are NHWC bf16; the public layout is NCHW fp32
// (reference: src/layer/layer.h:33-54), so the two converters below sit at the
// API boundary (input copy, extract/predict copy-out, flatten).
#include "common.cuh"

namespace cxn {

// ---------------------------------------------------------------- layout
// src fp32 [N][C][HW] -> dst bf16 [N][HW][ldc] (channels innermost, ldc >= C)
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ src, bf16* __restrict__ dst, int N, int C, int HW,
                                    int ldc) {
  __shared__ float tile[32][33];
  const int n = blockIdx.z;
  const int hw0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const float* s = src + (long long)n * C * HW;
  bf16* d = dst + (long long)n * HW * ldc;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = c0 + i, hw = hw0 + threadIdx.x;
    tile[i][threadIdx.x] = (c < C && hw < HW) ? s[(long long)c * HW + hw] : 0.0f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int hw = hw0 + i, c = c0 + threadIdx.x;
    if (hw < HW && c < C) d[(long long)hw * ldc + c] = __float2bfloat16(tile[threadIdx.x][i]);
  }
}
// small-C variant (C <= 4, ldc == 4): one pixel per thread, channel planes read coalesced,
// one 8-byte store per pixel (pad channels written as zero)
__global__ void nchw_to_nhwc4_kernel(const float* __restrict__ src, bf16* __restrict__ dst, long long total, int C,
                                     int HW) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long n = i / HW;
    const int hw = (int)(i - n * HW);
    const float* s = src + n * C * HW + hw;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int c = 0; c < 4; ++c)
      if (c < C) v[c] = s[(long long)c * HW];
    __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
    uint2 o;
    o.x = *reinterpret_cast<uint32_t*>(&a);
    o.y = *reinterpret_cast<uint32_t*>(&b);
    *reinterpret_cast<uint2*>(dst + i * 4) = o;
  }
}
// C <= 4 with an even row pitch Wp >= W (pixel pitch 4): one thread converts a pixel pair and
// stores it as 16 bytes; pixels at x >= W are written as zero.  This is the layout the first
// convolution reads as [H][Wp/2] super-pixels of 8 channels (see ConvolutionLayer pair mode).
__device__ __forceinline__ float ld_as_float(const float* p) { return *p; }
__device__ __forceinline__ float ld_as_float(const bf16* p) { return __bfloat162float(*p); }
template <typename TS>
__global__ void nchw_to_nhwc4_pitched_kernel(const TS* __restrict__ src, bf16* __restrict__ dst, long long total,
                                             int C, int H, int W, int Wp) {
  const int wp2 = Wp >> 1;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / wp2;             // n * H + y
    const int j = (int)(i - row * wp2);
    const long long n = row / H;
    const int y = (int)(row - n * H);
    const TS* s = src + (n * C * H + y) * (long long)W + 2 * j;
    float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int c = 0; c < 4; ++c)
      if (c < C) {
        if (2 * j < W) v[c] = ld_as_float(s + (long long)c * H * W);
        if (2 * j + 1 < W) v[4 + c] = ld_as_float(s + (long long)c * H * W + 1);
      }
    uint4 o;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
    for (int k = 0; k < 4; ++k) h[k] = __floats2bfloat162_rn(v[2 * k], v[2 * k + 1]);
    *reinterpret_cast<uint4*>(dst + i * 8) = o;
  }
}
// The same from raw 8-bit pixels (a data pipeline that ships a quarter of the fp32 bytes over PCIe): the
// normalisation the host would have done, (pixel - mean[c]) * scale in fp32, happens here before the bf16
// rounding, so the result is bit-identical to a bf16 batch prepared on the host.  Pad pixels stay zero.
__global__ void nchw_u8_to_nhwc4_pitched_kernel(const uint8_t* __restrict__ src, bf16* __restrict__ dst, long long total,
                                                int C, int H, int W, int Wp, float m0, float m1, float m2, float m3,
                                                float scale) {
  const int wp2 = Wp >> 1;
  const float mean[4] = {m0, m1, m2, m3};
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / wp2;             // n * H + y
    const int j = (int)(i - row * wp2);
    const long long n = row / H;
    const int y = (int)(row - n * H);
    const uint8_t* s = src + (n * C * H + y) * (long long)W + 2 * j;
    float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int c = 0; c < 4; ++c)
      if (c < C) {
        if (2 * j < W) v[c] = ((float)s[(long long)c * H * W] - mean[c]) * scale;
        if (2 * j + 1 < W) v[4 + c] = ((float)s[(long long)c * H * W + 1] - mean[c]) * scale;
      }
    uint4 o;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
    for (int k = 0; k < 4; ++k) h[k] = __floats2bfloat162_rn(v[2 * k], v[2 * k + 1]);
    *reinterpret_cast<uint4*>(dst + i * 8) = o;
  }
}
// C <= 4 into the space-to-depth layout the first convolution of a stride-4 net reads as a stride-1 convolution
// over 64 channels: dst[n][Y][X][(dy*4 + dx)*4 + c] = src[n][c][4Y + dy][4X + dx] (zero for c >= C and past the
// image).  One 4 x 4 block of pixels = 64 channels = one aligned 128-byte row of the patch operand.  One thread
// converts the four pixels of one (block, dy): it reads 4 consecutive source pixels per channel and stores 32 bytes.
// MODE 0: fp32 source, 1: bf16 source, 2: 8-bit source with (pixel - mean[c]) * scale.
template <typename TS, bool NORM>
__global__ void nchw_to_s2d4_kernel(const TS* __restrict__ src, bf16* __restrict__ dst, long long total, int C, int H,
                                    int W, int Hs, int Ws, float m0, float m1, float m2, float m3, float scale) {
  const float mean[4] = {m0, m1, m2, m3};
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    // i = ((n * Hs + Y) * 4 + dy) * Ws + X : neighbouring threads read neighbouring pixels of one source row
    const int X = (int)(i % Ws);
    long long r = i / Ws;
    const int dy = (int)(r & 3);
    r >>= 2;
    const int Y = (int)(r % Hs);
    const long long n = r / Hs;
    const int y = 4 * Y + dy;
    float v[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) v[k] = 0.f;
    if (y < H) {
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        if (c < C) {
          const TS* s = src + ((n * C + c) * H + y) * (long long)W + 4 * X;
#pragma unroll
          for (int dx = 0; dx < 4; ++dx)
            if (4 * X + dx < W) {
              float f;
              if constexpr (NORM) f = ((float)s[dx] - mean[c]) * scale;
              else f = ld_as_float(s + dx);
              v[dx * 4 + c] = f;
            }
        }
      }
    }
    uint4 o0, o1;
    __nv_bfloat162* h0 = reinterpret_cast<__nv_bfloat162*>(&o0);
    __nv_bfloat162* h1 = reinterpret_cast<__nv_bfloat162*>(&o1);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      h0[k] = __floats2bfloat162_rn(v[2 * k], v[2 * k + 1]);
      h1[k] = __floats2bfloat162_rn(v[8 + 2 * k], v[8 + 2 * k + 1]);
    }
    bf16* d = dst + (((n * Hs + Y) * (long long)Ws + X) * 64 + dy * 16);
    *reinterpret_cast<uint4*>(d) = o0;
    *reinterpret_cast<uint4*>(d + 8) = o1;
  }
}
__device__ __forceinline__ float ld_as_float(const uint8_t* p) { return (float)*p; }

// g[r][c] *= mask[c]: structural zeros of an internal filter layout must not collect a gradient
__global__ void mask_cols_kernel(float* __restrict__ g, long long total, int cols, long long ld,
                                 const float* __restrict__ mask) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / cols;
    const int c = (int)(i - r * cols);
    if (mask[c] == 0.0f) g[r * ld + c] = 0.0f;
  }
}

// src bf16 [N][HW][ldc] -> dst fp32 [N][C][HW]
__global__ void nhwc_to_nchw_kernel(const bf16* __restrict__ src, float* __restrict__ dst, int N, int C, int HW,
                                    int ldc) {
  __shared__ float tile[32][33];
  const int n = blockIdx.z;
  const int hw0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const bf16* s = src + (long long)n * HW * ldc;
  float* d = dst + (long long)n * C * HW;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int hw = hw0 + i, c = c0 + threadIdx.x;
    tile[i][threadIdx.x] = (c < C && hw < HW) ? __bfloat162float(s[(long long)hw * ldc + c]) : 0.0f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = c0 + i, hw = hw0 + threadIdx.x;
    if (hw < HW && c < C) d[(long long)c * HW + hw] = tile[threadIdx.x][i];
  }
}
// bf16 [N][HW][ldc_in] <-> bf16 [N][C][ld_hw] (flatten in reference (c,h,w) order and its inverse)
template <bool kToChw>
__global__ void bf16_transpose_kernel(const bf16* __restrict__ src, bf16* __restrict__ dst, int C, int HW,
                                      long long src_img, long long dst_img, int ld_hwc, int ld_chw) {
  __shared__ bf16 tile[32][34];
  const int n = blockIdx.z;
  const int hw0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const bf16* s = src + (long long)n * src_img;
  bf16* d = dst + (long long)n * dst_img;
  if (kToChw) {
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
      const int hw = hw0 + i, c = c0 + threadIdx.x;
      tile[i][threadIdx.x] = (c < C && hw < HW) ? s[(long long)hw * ld_hwc + c] : __float2bfloat16(0.0f);
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
      const int c = c0 + i, hw = hw0 + threadIdx.x;
      if (hw < HW && c < C) d[(long long)c * ld_chw + hw] = tile[threadIdx.x][i];
    }
  } else {
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
      const int c = c0 + i, hw = hw0 + threadIdx.x;
      tile[i][threadIdx.x] = (c < C && hw < HW) ? s[(long long)c * ld_chw + hw] : __float2bfloat16(0.0f);
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
      const int hw = hw0 + i, c = c0 + threadIdx.x;
      if (hw < HW && c < C) d[(long long)hw * ld_hwc + c] = tile[threadIdx.x][i];
    }
  }
}

// ---------------------------------------------------------------- 2-D strided helpers
// Activations are 2-D [rows][ld] with `cols` valid entries per row.
template <typename TS, typename TD>
__global__ void copy2d_kernel(const TS* __restrict__ src, long long lds, TD* __restrict__ dst, long long ldd,
                              long long rows, int cols, int accumulate) {
  const long long total = rows * cols;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / cols;
    const int c = (int)(i - r * cols);
    float v = (float)src[r * lds + c];
    if (accumulate) v += (float)dst[r * ldd + c];
    dst[r * ldd + c] = (TD)v;
  }
}

// ---------------------------------------------------------------- activations
// op: 0 relu, 1 sigmoid, 2 tanh, 3 xelu(b)          (reference src/layer/op.h:27-73)
__device__ __forceinline__ float act_fwd(int op, float a, float b) {
  switch (op) {
    case 0: return fmaxf(a, 0.0f);
    case 1: return 1.0f / (1.0f + __expf(-a));
    case 2: return tanhf(a);
    default: return a > 0.0f ? a : a / b;
  }
}
// gradient expressed in terms of the ACTIVATED value y (reference keeps y in the input node)
__device__ __forceinline__ float act_bwd(int op, float y, float b) {
  switch (op) {
    case 0: return y > 0.0f ? 1.0f : 0.0f;
    case 1: return y * (1.0f - y);
    case 2: return 1.0f - y * y;
    default: return y > 0.0f ? 1.0f : 1.0f / b;
  }
}
__global__ void act_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, long long rows, int cols,
                               long long ld, int op, float b) {
  const long long total = rows * cols;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / cols;
    const long long o = r * ld + (i - r * cols);
    y[o] = __float2bfloat16(act_fwd(op, __bfloat162float(x[o]), b));
  }
}
// dx = act'(y) * dy
__global__ void act_bwd_kernel(const bf16* __restrict__ y, const bf16* __restrict__ dy, bf16* __restrict__ dx,
                               long long rows, int cols, long long ld, int op, float b) {
  const long long total = rows * cols;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / cols;
    const long long o = r * ld + (i - r * cols);
    dx[o] = __float2bfloat16(act_bwd(op, __bfloat162float(y[o]), b) * __bfloat162float(dy[o]));
  }
}
// contiguous fast path, 8 bf16 per thread
__global__ void relu_bwd_vec_kernel(const uint4* __restrict__ y, const uint4* __restrict__ dy, uint4* __restrict__ dx,
                                    long long n8) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
    uint4 a = y[i], g = dy[i], o;
    const __nv_bfloat162* ya = reinterpret_cast<const __nv_bfloat162*>(&a);
    const __nv_bfloat162* ga = reinterpret_cast<const __nv_bfloat162*>(&g);
    __nv_bfloat162* oa = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float2 yf = __bfloat1622float2(ya[j]), gf = __bfloat1622float2(ga[j]);
      oa[j] = __floats2bfloat162_rn(yf.x > 0.0f ? gf.x : 0.0f, yf.y > 0.0f ? gf.y : 0.0f);
    }
    dx[i] = o;
  }
}

// ---------------------------------------------------------------- dropout
// mask = (U[0,1) < pkeep) / pkeep, regenerated from (seed, element index) in
// backward instead of being stored (reference src/layer/dropout_layer-inl.hpp:44-58).
__global__ void dropout_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, long long rows, int cols,
                               long long ld, float pkeep, unsigned long long seed,
                               const unsigned long long* __restrict__ counter) {
  // `counter` is a device-resident step counter so that a captured CUDA graph
  // draws a fresh mask on every replay.
  const unsigned long long offset = counter ? *counter : 0ull;
  const Philox rng(seed);
  const long long total = rows * cols;
  const long long nquads = (total + 3) / 4;
  const float scale = 1.0f / pkeep;
  for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < nquads;
       q += (long long)gridDim.x * blockDim.x) {
    uint32_t r4[4];
    rng((unsigned long long)q, offset, r4);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const long long i = q * 4 + j;
      if (i < total) {
        const long long r = i / cols;
        const long long o = r * ld + (i - r * cols);
        const float m = (u01(r4[j]) < pkeep) ? scale : 0.0f;
        y[o] = __float2bfloat16(__bfloat162float(x[o]) * m);
      }
    }
  }
}

// ---------------------------------------------------------------- reductions
// out[c] (+)= sum_r x[r][c]   (bias gradients: reference fullc_layer-inl.hpp:123-125,
// convolution_layer-inl.hpp:117-119).  One CTA handles a strip of 32 columns.
__global__ void colsum_kernel(const bf16* __restrict__ x, long long rows, int cols, long long ld,
                              float* __restrict__ out, int accumulate) {
  __shared__ float red[8][33];
  const int c = blockIdx.x * 32 + threadIdx.x;
  float acc = 0.0f;
  if (c < cols) {
    for (long long r = blockIdx.y * 8 + threadIdx.y; r < rows; r += (long long)gridDim.y * 8)
      acc += __bfloat162float(x[r * ld + c]);
  }
  red[threadIdx.y][threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.y == 0 && c < cols) {
    float s = 0.0f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += red[j][threadIdx.x];
    if (gridDim.y > 1 || accumulate) atomicAdd(out + c, s);
    else out[c] = s;
  }
}

__global__ void counter_add_kernel(unsigned long long* c, unsigned long long inc) { *c += inc; }

__global__ void fill_f32_kernel(float* __restrict__ p, long long n, float v) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    p[i] = v;
}

}  // namespace cxn

using namespace cxn;

extern "C" { unsigned long long cxn_launch_counter_ = 0; }

// ---- launch-site histogram (CXN_LAUNCH_LOG=1): which kernels make up a step, without a profiler attached.  Launches
// replayed from a CUDA graph are counted once, at capture - i.e. the composition of one captured step.
#include <map>
#include <mutex>
#include <string>
#include <stdlib.h>
namespace {
std::mutex g_ll_mu;
std::map<std::string, unsigned long long>* g_ll = nullptr;
void ll_dump() {
  if (!g_ll) return;
  fprintf(stderr, "[cxxnet_b200] launches per call site (graph replays not included):\n");
  for (const auto& kv : *g_ll) fprintf(stderr, "  %8llu  %s\n", kv.second, kv.first.c_str());
}
int ll_init() {
  const char* e = getenv("CXN_LAUNCH_LOG");
  if (e && atoi(e) > 0) {
    g_ll = new std::map<std::string, unsigned long long>();
    atexit(ll_dump);
    return 1;
  }
  return 0;
}
}  // namespace
extern "C" {
int cxn_launch_log_on_ = ll_init();
void cxxnet_note_launch_(const char* file, int line) {
  std::lock_guard<std::mutex> g(g_ll_mu);
  if (!g_ll) return;
  const char* base = file;
  for (const char* q = file; *q; ++q) if (*q == '/') base = q + 1;
  ++(*g_ll)[std::string(base) + ":" + std::to_string(line)];
}
}
CXN_API unsigned long long cxn_launch_count() { return cxn_launch_counter_; }
// launches replayed from a captured CUDA graph never pass through this library's entry points
CXN_API unsigned long long cxn_launch_count_add(unsigned long long n) { return cxn_launch_counter_ += n; }

CXN_API int cxn_nchw_f32_to_nhwc_bf16(const float* src, void* dst, int N, int C, int HW, int ldc, cudaStream_t st) {
  if (N <= 0 || C <= 0 || HW <= 0) return 0;
  if (C <= 4 && ldc == 4 && (reinterpret_cast<uintptr_t>(dst) & 7) == 0) {
    const long long total = (long long)N * HW;
    nchw_to_nhwc4_kernel<<<ew_grid(total, 256, 16), 256, 0, st>>>(src, (bf16*)dst, total, C, HW);
    CXN_CHECK_LAUNCH();
    return 0;
  }
  dim3 grid(ceil_div(HW, 32), ceil_div(C, 32), N), block(32, 8);
  nchw_to_nhwc_kernel<<<grid, block, 0, st>>>(src, (bf16*)dst, N, C, HW, ldc);
  CXN_CHECK_LAUNCH();
  return 0;
}
CXN_API int cxn_nchw_f32_to_nhwc4_pitched(const float* src, void* dst, int N, int C, int H, int W, int Wp, cudaStream_t st) {
  if (N <= 0 || C <= 0 || H <= 0 || W <= 0) return 0;
  if (C > 4 || (Wp & 1) || Wp < W || (reinterpret_cast<uintptr_t>(dst) & 15)) return -1;
  const long long total = (long long)N * H * (Wp / 2);
  nchw_to_nhwc4_pitched_kernel<float><<<ew_grid(total, 256, 16), 256, 0, st>>>(src, (bf16*)dst, total, C, H, W, Wp);
  CXN_CHECK_LAUNCH();
  return 0;
}
// the same from a bf16 NCHW source (a data pipeline that ships half the bytes over PCIe)
CXN_API int cxn_nchw_bf16_to_nhwc4_pitched(const void* src, void* dst, int N, int C, int H, int W, int Wp, cudaStream_t st) {
  if (N <= 0 || C <= 0 || H <= 0 || W <= 0) return 0;
  if (C > 4 || (Wp & 1) || Wp < W || (reinterpret_cast<uintptr_t>(dst) & 15)) return -1;
  const long long total = (long long)N * H * (Wp / 2);
  nchw_to_nhwc4_pitched_kernel<bf16><<<ew_grid(total, 256, 16), 256, 0, st>>>((const bf16*)src, (bf16*)dst, total, C, H, W, Wp);
  CXN_CHECK_LAUNCH();
  return 0;
}
// ... and from 8-bit pixels with the input normalisation (pixel - mean[c]) * scale folded in
CXN_API int cxn_nchw_u8_to_nhwc4_pitched(const void* src, void* dst, int N, int C, int H, int W, int Wp, float mean0,
                                         float mean1, float mean2, float mean3, float scale, cudaStream_t st) {
  if (N <= 0 || C <= 0 || H <= 0 || W <= 0) return 0;
  if (C > 4 || (Wp & 1) || Wp < W || (reinterpret_cast<uintptr_t>(dst) & 15)) return -1;
  const long long total = (long long)N * H * (Wp / 2);
  nchw_u8_to_nhwc4_pitched_kernel<<<ew_grid(total, 256, 16), 256, 0, st>>>((const uint8_t*)src, (bf16*)dst, total, C, H, W,
                                                                           Wp, mean0, mean1, mean2, mean3, scale);
  CXN_CHECK_LAUNCH();
  return 0;
}
// NCHW (C <= 4) -> space-to-depth NHWC bf16 [N][Hs][Ws][64], Hs = ceil(H / 4), Ws = ceil(W / 4); kind 0 fp32,
// 1 bf16, 2 8-bit pixels normalised as (pixel - mean[c]) * scale.  See nchw_to_s2d4_kernel.
CXN_API int cxn_nchw_to_s2d4(const void* src, int kind, void* dst, int N, int C, int H, int W, float mean0, float mean1,
                             float mean2, float mean3, float scale, cudaStream_t st) {
  if (N <= 0 || C <= 0 || H <= 0 || W <= 0) return 0;
  if (C > 4 || kind < 0 || kind > 2 || (reinterpret_cast<uintptr_t>(dst) & 15)) return -1;
  const int Hs = (H + 3) / 4, Ws = (W + 3) / 4;
  const long long total = (long long)N * Hs * 4 * Ws;
  const int grid = ew_grid(total, 256, 16);
  if (kind == 0)
    nchw_to_s2d4_kernel<float, false><<<grid, 256, 0, st>>>((const float*)src, (bf16*)dst, total, C, H, W, Hs, Ws, 0, 0, 0, 0, 1);
  else if (kind == 1)
    nchw_to_s2d4_kernel<bf16, false><<<grid, 256, 0, st>>>((const bf16*)src, (bf16*)dst, total, C, H, W, Hs, Ws, 0, 0, 0, 0, 1);
  else
    nchw_to_s2d4_kernel<uint8_t, true><<<grid, 256, 0, st>>>((const uint8_t*)src, (bf16*)dst, total, C, H, W, Hs, Ws, mean0,
                                                             mean1, mean2, mean3, scale);
  CXN_CHECK_LAUNCH();
  return 0;
}
// g[r * ld + c] = 0 where mask[c] == 0 (rows x cols fp32)
CXN_API int cxn_mask_cols_f32(float* g, long long rows, int cols, long long ld, const float* mask, cudaStream_t st) {
  const long long total = rows * cols;
  if (total <= 0) return 0;
  mask_cols_kernel<<<ew_grid(total, 256, 8), 256, 0, st>>>(g, total, cols, ld, mask);
  CXN_CHECK_LAUNCH();
  return 0;
}
CXN_API int cxn_nhwc_bf16_to_nchw_f32(const void* src, float* dst, int N, int C, int HW, int ldc, cudaStream_t st) {
  if (N <= 0 || C <= 0 || HW <= 0) return 0;
  dim3 grid(ceil_div(HW, 32), ceil_div(C, 32), N), block(32, 8);
  nhwc_to_nchw_kernel<<<grid, block, 0, st>>>((const bf16*)src, dst, N, C, HW, ldc);
  CXN_CHECK_LAUNCH();
  return 0;
}
// to_chw=1: src [N][HW][ld_hwc] -> dst [N][C][ld_chw];  to_chw=0: the inverse.
CXN_API int cxn_bf16_hwc_chw(const void* src, void* dst, int N, int C, int HW, long long src_img, long long dst_img,
                             int ld_hwc, int ld_chw, int to_chw, cudaStream_t st) {
  if (N <= 0 || C <= 0 || HW <= 0) return 0;
  dim3 grid(ceil_div(HW, 32), ceil_div(C, 32), N), block(32, 8);
  if (to_chw)
    bf16_transpose_kernel<true><<<grid, block, 0, st>>>((const bf16*)src, (bf16*)dst, C, HW, src_img, dst_img, ld_hwc, ld_chw);
  else
    bf16_transpose_kernel<false><<<grid, block, 0, st>>>((const bf16*)src, (bf16*)dst, C, HW, src_img, dst_img, ld_hwc, ld_chw);
  CXN_CHECK_LAUNCH();
  return 0;
}
// kind: 0 bf16->bf16, 1 f32->bf16, 2 bf16->f32, 3 f32->f32
CXN_API int cxn_copy2d(const void* src, long long lds, void* dst, long long ldd, long long rows, int cols, int kind,
                       int accumulate, cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return 0;
  const int g = ew_grid(rows * cols, 256);
  switch (kind) {
    case 0: copy2d_kernel<bf16, bf16><<<g, 256, 0, st>>>((const bf16*)src, lds, (bf16*)dst, ldd, rows, cols, accumulate); break;
    case 1: copy2d_kernel<float, bf16><<<g, 256, 0, st>>>((const float*)src, lds, (bf16*)dst, ldd, rows, cols, accumulate); break;
    case 2: copy2d_kernel<bf16, float><<<g, 256, 0, st>>>((const bf16*)src, lds, (float*)dst, ldd, rows, cols, accumulate); break;
    default: copy2d_kernel<float, float><<<g, 256, 0, st>>>((const float*)src, lds, (float*)dst, ldd, rows, cols, accumulate); break;
  }
  CXN_CHECK_LAUNCH();
  return 0;
}
CXN_API int cxn_act_fwd(const void* x, void* y, long long rows, int cols, long long ld, int op, float b,
                        cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return 0;
  act_fwd_kernel<<<ew_grid(rows * cols, 256), 256, 0, st>>>((const bf16*)x, (bf16*)y, rows, cols, ld, op, b);
  CXN_CHECK_LAUNCH();
  return 0;
}
CXN_API int cxn_act_bwd(const void* y, const void* dy, void* dx, long long rows, int cols, long long ld, int op,
                        float b, cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return 0;
  const long long total = rows * cols;
  const bool aligned = ((reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(dy) |
                         reinterpret_cast<uintptr_t>(dx)) & 15) == 0;
  if (op == 0 && ld == cols && (total % 8) == 0 && aligned) {
    relu_bwd_vec_kernel<<<ew_grid(total / 8, 256), 256, 0, st>>>((const uint4*)y, (const uint4*)dy, (uint4*)dx, total / 8);
  } else {
    act_bwd_kernel<<<ew_grid(total, 256), 256, 0, st>>>((const bf16*)y, (const bf16*)dy, (bf16*)dx, rows, cols, ld, op, b);
  }
  CXN_CHECK_LAUNCH();
  return 0;
}
CXN_API int cxn_dropout(const void* x, void* y, long long rows, int cols, long long ld, float pkeep,
                        unsigned long long seed, const unsigned long long* counter, cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return 0;
  dropout_kernel<<<ew_grid((rows * cols + 3) / 4, 256), 256, 0, st>>>((const bf16*)x, (bf16*)y, rows, cols, ld, pkeep, seed, counter);
  CXN_CHECK_LAUNCH();
  return 0;
}
extern "C" int cxn_colsum_bf16_vec(const void* x, long long rows, int cols, long long ld, float* out, cudaStream_t st);

CXN_API int cxn_colsum_bf16(const void* x, long long rows, int cols, long long ld, float* out, int accumulate,
                            cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return 0;
  if (cols % 8 == 0 && cols <= 2048 && ld % 8 == 0 && rows >= 1024 && (reinterpret_cast<uintptr_t>(x) & 15) == 0 &&
      (256 / (cols / 8)) >= 1 && (size_t)(256 / (cols / 8)) * cols * 4 <= 48 * 1024) {
    if (!accumulate) {
      fill_f32_kernel<<<ew_grid(cols, 256), 256, 0, st>>>(out, cols, 0.0f);
      CXN_CHECK_LAUNCH();
    }
    return cxn_colsum_bf16_vec(x, rows, cols, ld, out, st);
  }
  int gy = (int)((rows + 255) / 256);
  const int cap = ceil_div(num_sms() * 4, ceil_div(cols, 32));
  if (gy > cap) gy = cap;
  if (gy < 1) gy = 1;
  if (gy > 1 && !accumulate) {
    fill_f32_kernel<<<ew_grid(cols, 256), 256, 0, st>>>(out, cols, 0.0f);
  }
  dim3 grid(ceil_div(cols, 32), gy), block(32, 8);
  colsum_kernel<<<grid, block, 0, st>>>((const bf16*)x, rows, cols, ld, out, accumulate);
  CXN_CHECK_LAUNCH();
  return 0;
}
CXN_API int cxn_fill_f32(float* p, long long n, float v, cudaStream_t st) {
  if (n <= 0) return 0;
  fill_f32_kernel<<<ew_grid(n, 256), 256, 0, st>>>(p, n, v);
  CXN_CHECK_LAUNCH();
  return 0;
}
// Step trace: a one-thread kernel that stores the GPU's global nanosecond timer; placed between the layers of a step
// (also inside a captured graph) it gives a device-side timeline without a profiler - the only kind available for
// multi-rank runs.  buf[slot] is overwritten at every execution.
__global__ void timestamp_kernel(unsigned long long* buf, int slot) {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  buf[slot] = t;
}
CXN_API int cxn_timestamp(unsigned long long* buf, int slot, cudaStream_t st) {
  timestamp_kernel<<<1, 1, 0, st>>>(buf, slot);
  CXN_CHECK_LAUNCH();
  return 0;
}
CXN_API int cxn_counter_add(unsigned long long* counter, unsigned long long inc, cudaStream_t st) {
  counter_add_kernel<<<1, 1, 0, st>>>(counter, inc);
  CXN_CHECK_LAUNCH();
  return 0;
}
