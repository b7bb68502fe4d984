// Native kernels of the less common layers on NHWC bf16 ([rows = pixels][C] matrices):
//   bias (reference src/layer/bias_layer-inl.hpp:13-82), prelu (prelu_layer-inl.hpp:47-173),
//   batch_norm (batch_norm_layer-inl.hpp: mini-batch statistics in training AND at test time).
// All of them are per-channel, i.e. per-column, operations: a column reduction (one template,
// three element functors) plus element-wise kernels with per-column coefficients.
#include "common.cuh"

namespace cxn {

__device__ __forceinline__ void ml_unpack8(const uint4& v, float* f) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 t = __bfloat1622float2(h[j]);
    f[2 * j] = t.x; f[2 * j + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 ml_pack8(const float* f) {
  uint4 o;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
  for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(f[2 * j], f[2 * j + 1]);
  return o;
}

// ---------------------------------------------------------------- column reductions
// MODE 0: out0 += sum a, out1 += sum a*a            (batch-norm statistics)
// MODE 1: out0 += sum a, out1 += sum a*b            (batch-norm backward: a = dy, b = xhat)
// MODE 2: out0 += sum (a <= 0 ? a*b : 0)            (prelu slope gradient: a = x, b = dy)
template <int MODE>
__device__ __forceinline__ void red_elem(float a, float b, float& v0, float& v1) {
  if (MODE == 0) { v0 = a; v1 = a * a; }
  else if (MODE == 1) { v0 = a; v1 = a * b; }
  else { v0 = a <= 0.0f ? a * b : 0.0f; v1 = 0.0f; }
}

// C % 8 == 0: a thread owns 8 columns (16-byte loads) and strides over rows; partial sums meet in
// shared memory, one atomic per column per CTA.
template <int MODE>
__global__ void colreduce_vec_kernel(const bf16* __restrict__ a, const bf16* __restrict__ b, long long rows, int cols,
                                     long long lda, long long ldb, float* __restrict__ out0, float* __restrict__ out1) {
  extern __shared__ float red[];            // [2][rows_per_block][cols]
  const int tpr = cols >> 3;
  const int rpb = blockDim.x / tpr;
  const int tr = threadIdx.x / tpr, tc = threadIdx.x - tr * tpr;
  float s0[8], s1[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s0[j] = s1[j] = 0.f;
  if (tr < rpb) {
    for (long long r = (long long)blockIdx.x * rpb + tr; r < rows; r += (long long)gridDim.x * rpb) {
      float fa[8], fb[8];
      ml_unpack8(*reinterpret_cast<const uint4*>(a + r * lda + tc * 8), fa);
      if (MODE != 0) ml_unpack8(*reinterpret_cast<const uint4*>(b + r * ldb + tc * 8), fb);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float v0, v1;
        red_elem<MODE>(fa[j], MODE != 0 ? fb[j] : 0.0f, v0, v1);
        s0[j] += v0; s1[j] += v1;
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      red[tr * cols + tc * 8 + j] = s0[j];
      if (MODE != 2) red[(rpb + tr) * cols + tc * 8 + j] = s1[j];
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < cols; c += blockDim.x) {
    float t0 = 0.f, t1 = 0.f;
    for (int r = 0; r < rpb; ++r) {
      t0 += red[r * cols + c];
      if (MODE != 2) t1 += red[(rpb + r) * cols + c];
    }
    atomicAdd(out0 + c, t0);
    if (MODE != 2) atomicAdd(out1 + c, t1);
  }
}

// any C: a thread sums a strip of rows of one column
template <int MODE>
__global__ void colreduce_kernel(const bf16* __restrict__ a, const bf16* __restrict__ b, long long rows, int cols,
                                 long long lda, long long ldb, float* __restrict__ out0, float* __restrict__ out1,
                                 int strip) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cols) return;
  const long long r0 = (long long)blockIdx.y * strip, r1 = min(rows, r0 + strip);
  float t0 = 0.f, t1 = 0.f;
  for (long long r = r0; r < r1; ++r) {
    float v0, v1;
    red_elem<MODE>(__bfloat162float(a[r * lda + c]), MODE != 0 ? __bfloat162float(b[r * ldb + c]) : 0.0f, v0, v1);
    t0 += v0; t1 += v1;
  }
  atomicAdd(out0 + c, t0);
  if (MODE != 2) atomicAdd(out1 + c, t1);
}

template <int MODE>
static int colreduce(const void* a, const void* b, long long rows, int cols, long long lda, long long ldb, float* out0,
                     float* out1, cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return 0;
  const bool al = ((reinterpret_cast<uintptr_t>(a) | (MODE ? reinterpret_cast<uintptr_t>(b) : 0)) & 15) == 0;
  if (cols % 8 == 0 && cols / 8 <= 256 && lda % 8 == 0 && (MODE == 0 || ldb % 8 == 0) && al) {
    const int tpr = cols / 8, threads = 256, rpb = threads / tpr;
    const size_t smem = (size_t)2 * rpb * cols * sizeof(float);
    if (rpb >= 1 && smem <= 48 * 1024) {
      long long blocks = ceil_div_ll(rows, (long long)rpb * 8);
      const long long cap = (long long)num_sms() * 4;
      if (blocks > cap) blocks = cap;
      if (blocks < 1) blocks = 1;
      colreduce_vec_kernel<MODE><<<(int)blocks, threads, smem, st>>>((const bf16*)a, (const bf16*)b, rows, cols, lda, ldb,
                                                                    out0, out1);
      CXN_CHECK_LAUNCH();
      return 0;
    }
  }
  const int strip = 256;
  dim3 grid(ceil_div(cols, 128), (unsigned)ceil_div_ll(rows, strip));
  colreduce_kernel<MODE><<<grid, 128, 0, st>>>((const bf16*)a, (const bf16*)b, rows, cols, lda, ldb, out0, out1, strip);
  CXN_CHECK_LAUNCH();
  return 0;
}

// ---------------------------------------------------------------- element-wise with per-column coefficients
// KIND 0 bias:        y = x + p0[c]
// KIND 1 bn forward:  xhat = (x - p0[c]) * p1[c] (stored to aux if non-null), y = xhat * p2[c] + p3[c]
// KIND 2 bn backward: dx = p0[c] * (g - p1[c] * xhat - p2[c])        (x = g, aux = xhat)
template <int KIND>
__global__ void colwise_kernel(const bf16* __restrict__ x, bf16* __restrict__ aux, bf16* __restrict__ y, long long rows,
                               int cols, long long ldx, long long ldaux, long long ldy, const float* __restrict__ p0,
                               const float* __restrict__ p1, const float* __restrict__ p2,
                               const float* __restrict__ p3) {
  const long long total = rows * cols;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / cols;
    const int c = (int)(i - r * cols);
    const float v = __bfloat162float(x[r * ldx + c]);
    float o;
    if (KIND == 0) {
      o = v + p0[c];
    } else if (KIND == 1) {
      const float xh = (v - p0[c]) * p1[c];
      if (aux != nullptr) aux[r * ldaux + c] = __float2bfloat16(xh);
      o = xh * p2[c] + p3[c];
    } else {
      o = p0[c] * (v - p1[c] * __bfloat162float(aux[r * ldaux + c]) - p2[c]);
    }
    y[r * ldy + c] = __float2bfloat16(o);
  }
}

// batch-norm coefficient kernels (C threads)
__global__ void bn_stats_kernel(const float* __restrict__ sum, const float* __restrict__ sumsq, float cnt, float eps,
                                float* __restrict__ mean, float* __restrict__ istd, int C) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const double m = (double)sum[c] / cnt;
  double var = (double)sumsq[c] / cnt - m * m;
  if (var < 0.0) var = 0.0;
  mean[c] = (float)m;
  istd[c] = (float)(1.0 / sqrt(var + (double)eps));
}
// after the backward reduction (s1 = sum dy, s2 = sum dy*xhat): gradients of slope / bias and the
// coefficients of dx = a * (dy - b * xhat - c)
__global__ void bn_bwd_coef_kernel(const float* __restrict__ s1, const float* __restrict__ s2,
                                   const float* __restrict__ slope, const float* __restrict__ istd, float cnt,
                                   float* __restrict__ gslope, float* __restrict__ gbias, float* __restrict__ ca,
                                   float* __restrict__ cb, float* __restrict__ cc, int C) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  gslope[c] += s2[c];
  gbias[c] += s1[c];
  ca[c] = slope[c] * istd[c];
  cb[c] = s2[c] / cnt;
  cc[c] = s1[c] / cnt;
}

// ---------------------------------------------------------------- prelu
// mask = clamp(slope[c] * (1 + u * 2r - r), 0, 1), u ~ U[0,1) regenerated from (seed, step counter) in
// backward; y = x > 0 ? x : x * mask;  dx = (x > 0 ? 1 : mask) * dy
template <bool BWD>
__global__ void prelu_kernel(const bf16* __restrict__ x, const bf16* __restrict__ dy, bf16* __restrict__ out,
                             long long rows, int cols, long long ldx, long long ldo, const float* __restrict__ slope,
                             float noise, unsigned long long seed, const unsigned long long* __restrict__ counter) {
  const unsigned long long offset = counter ? *counter : 0ull;
  const Philox rng(seed);
  const long long total = rows * cols;
  const long long nquads = (total + 3) / 4;
  for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < nquads;
       q += (long long)gridDim.x * blockDim.x) {
    uint32_t r4[4] = {0, 0, 0, 0};
    if (noise > 0.0f) rng((unsigned long long)q, offset, r4);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const long long i = q * 4 + j;
      if (i >= total) break;
      const long long r = i / cols;
      const int c = (int)(i - r * cols);
      float m = slope[c];
      if (noise > 0.0f) m *= 1.0f + u01(r4[j]) * 2.0f * noise - noise;
      m = fminf(fmaxf(m, 0.0f), 1.0f);
      const float xv = __bfloat162float(x[r * ldx + c]);
      float o;
      if (!BWD) o = xv > 0.0f ? xv : xv * m;
      else o = (xv > 0.0f ? 1.0f : m) * __bfloat162float(dy[r * ldo + c]);
      out[r * ldo + c] = __float2bfloat16(o);
    }
  }
}

// ---------------------------------------------------------------- insanity (randomised leaky ReLU)
// reference src/layer/insanity_layer-inl.hpp:13-102: y = x > 0 ? x : x / s, s ~ U[lb, ub) per element while
// training, (lb + ub) / 2 otherwise; backward multiplies dy by 1 or 1 / s with the same s (Philox stream
// regenerated from (seed, step counter), nothing stored).
template <bool BWD>
__global__ void insanity_kernel(const bf16* __restrict__ x, const bf16* __restrict__ dy, bf16* __restrict__ out,
                                long long rows, int cols, long long ldx, long long ldo, float lb, float ub, int random,
                                unsigned long long seed, const unsigned long long* __restrict__ counter) {
  const unsigned long long offset = counter ? *counter : 0ull;
  const Philox rng(seed);
  const long long total = rows * cols;
  const long long nquads = (total + 3) / 4;
  for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < nquads;
       q += (long long)gridDim.x * blockDim.x) {
    uint32_t r4[4] = {0, 0, 0, 0};
    if (random) rng((unsigned long long)q, offset, r4);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const long long i = q * 4 + j;
      if (i >= total) break;
      const long long r = i / cols;
      const int c = (int)(i - r * cols);
      const float s = random ? u01(r4[j]) * (ub - lb) + lb : 0.5f * (lb + ub);
      const float xv = __bfloat162float(x[r * ldx + c]);
      float o;
      if (!BWD) o = xv > 0.0f ? xv : xv / s;
      else o = (xv > 0.0f ? 1.0f : 1.0f / s) * __bfloat162float(dy[r * ldo + c]);
      out[r * ldo + c] = __float2bfloat16(o);
    }
  }
}

// ---------------------------------------------------------------- insanity pooling: random displacement
// reference src/layer/insanity_pooling_layer-inl.hpp:63-94: while training, every element of the pooling
// input is read from itself (probability `keep`) or from its upper / lower / left / right neighbour
// ((1 - keep) / 4 each, clamped at the border); the max pooling then runs on that displaced copy.
// src[n][y][x][c] = in[n][y'][x'][c], (y', x') chosen by u ~ U[0,1) from Philox(seed, *counter).
__global__ void insanity_displace_kernel(const bf16* __restrict__ in, bf16* __restrict__ out, int N, int H, int W, int C,
                                         long long ld, float keep, unsigned long long seed,
                                         const unsigned long long* __restrict__ counter) {
  const unsigned long long offset = counter ? *counter : 0ull;
  const Philox rng(seed);
  const long long total = (long long)N * H * W * C;
  const long long nquads = (total + 3) / 4;
  const float delta = (1.0f - keep) * 0.25f;
  for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < nquads;
       q += (long long)gridDim.x * blockDim.x) {
    uint32_t r4[4];
    rng((unsigned long long)q, offset, r4);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const long long i = q * 4 + j;
      if (i >= total) break;
      long long r = i;
      const int c = (int)(r % C); r /= C;
      const int x = (int)(r % W); r /= W;
      const int y = (int)(r % H);
      const long long n = r / H;
      const float u = u01(r4[j]);
      int ys = y, xs = x;
      if (u >= keep) {
        if (u < keep + delta) ys = max(y - 1, 0);
        else if (u < keep + 2.0f * delta) ys = min(y + 1, H - 1);
        else if (u < keep + 3.0f * delta) xs = max(x - 1, 0);
        else xs = min(x + 1, W - 1);
      }
      out[((n * H + y) * W + x) * ld + c] = in[((n * H + ys) * W + xs) * ld + c];
    }
  }
}

}  // namespace cxn

using namespace cxn;

CXN_API int cxn_insanity_displace(const void* in, void* out, int N, int H, int W, int C, long long ld, float keep,
                                  unsigned long long seed, const unsigned long long* counter, cudaStream_t st) {
  const long long total = (long long)N * H * W * C;
  if (total <= 0) return 0;
  insanity_displace_kernel<<<ew_grid((total + 3) / 4, 256), 256, 0, st>>>((const bf16*)in, (bf16*)out, N, H, W, C, ld, keep,
                                                                           seed, counter);
  CXN_CHECK_LAUNCH();
  return 0;
}

CXN_API int cxn_insanity_fwd(const void* x, void* y, long long rows, int cols, long long ldx, long long ldy, float lb,
                             float ub, int random, unsigned long long seed, const unsigned long long* counter,
                             cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return 0;
  insanity_kernel<false><<<ew_grid((rows * cols + 3) / 4, 256), 256, 0, st>>>((const bf16*)x, nullptr, (bf16*)y, rows, cols,
                                                                               ldx, ldy, lb, ub, random, seed, counter);
  CXN_CHECK_LAUNCH();
  return 0;
}
CXN_API int cxn_insanity_bwd(const void* x, const void* dy, void* dx, long long rows, int cols, long long ldx, long long ldg,
                             float lb, float ub, int random, unsigned long long seed, const unsigned long long* counter,
                             cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return 0;
  insanity_kernel<true><<<ew_grid((rows * cols + 3) / 4, 256), 256, 0, st>>>((const bf16*)x, (const bf16*)dy, (bf16*)dx, rows,
                                                                              cols, ldx, ldg, lb, ub, random, seed, counter);
  CXN_CHECK_LAUNCH();
  return 0;
}

// x[r][c] += bias[c]  (in place allowed)
CXN_API int cxn_bias_add(const void* x, void* y, long long rows, int cols, long long ld, const float* bias, cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return 0;
  colwise_kernel<0><<<ew_grid(rows * cols, 256), 256, 0, st>>>((const bf16*)x, nullptr, (bf16*)y, rows, cols, ld, 0, ld, bias,
                                                                nullptr, nullptr, nullptr);
  CXN_CHECK_LAUNCH();
  return 0;
}

// Batch-norm forward.  scratch: 2*C floats (zeroed here), mean / istd: C floats each (outputs, kept for backward).
// xhat may be NULL (evaluation: nothing is kept).
CXN_API int cxn_batchnorm_fwd(const void* x, void* y, void* xhat, long long rows, int C, long long ldx, long long ldy,
                              long long ldxhat, const float* slope, const float* bias, float eps, float* scratch,
                              float* mean, float* istd, cudaStream_t st) {
  if (rows <= 0 || C <= 0) return 0;
  cudaError_t e = cudaMemsetAsync(scratch, 0, 2 * (size_t)C * sizeof(float), st);
  if (e != cudaSuccess) return (int)e;
  int rc = colreduce<0>(x, nullptr, rows, C, ldx, 0, scratch, scratch + C, st);
  if (rc) return rc;
  bn_stats_kernel<<<ceil_div(C, 128), 128, 0, st>>>(scratch, scratch + C, (float)rows, eps, mean, istd, C);
  CXN_CHECK_LAUNCH();
  colwise_kernel<1><<<ew_grid(rows * C, 256), 256, 0, st>>>((const bf16*)x, (bf16*)xhat, (bf16*)y, rows, C, ldx, ldxhat, ldy,
                                                            mean, istd, slope, bias);
  CXN_CHECK_LAUNCH();
  return 0;
}

// Batch-norm backward: gslope += sum dy*xhat, gbias += sum dy, dx = slope*istd*(dy - mean(dy) - xhat*mean(dy*xhat)).
// scratch: 5*C floats.
CXN_API int cxn_batchnorm_bwd(const void* dy, const void* xhat, void* dx, long long rows, int C, long long lddy,
                              long long ldxhat, long long lddx, const float* slope, const float* istd, float* gslope,
                              float* gbias, float* scratch, int prop_grad, cudaStream_t st) {
  if (rows <= 0 || C <= 0) return 0;
  cudaError_t e = cudaMemsetAsync(scratch, 0, 2 * (size_t)C * sizeof(float), st);
  if (e != cudaSuccess) return (int)e;
  int rc = colreduce<1>(dy, xhat, rows, C, lddy, ldxhat, scratch, scratch + C, st);
  if (rc) return rc;
  float* ca = scratch + 2 * C;
  bn_bwd_coef_kernel<<<ceil_div(C, 128), 128, 0, st>>>(scratch, scratch + C, slope, istd, (float)rows, gslope, gbias, ca,
                                                       ca + C, ca + 2 * C, C);
  CXN_CHECK_LAUNCH();
  if (prop_grad) {
    colwise_kernel<2><<<ew_grid(rows * C, 256), 256, 0, st>>>((const bf16*)dy, (bf16*)const_cast<void*>(xhat), (bf16*)dx, rows,
                                                              C, lddy, ldxhat, lddx, ca, ca + C, ca + 2 * C, nullptr);
    CXN_CHECK_LAUNCH();
  }
  return 0;
}

CXN_API int cxn_prelu_fwd(const void* x, void* y, long long rows, int cols, long long ldx, long long ldy, const float* slope,
                          float noise, unsigned long long seed, const unsigned long long* counter, cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return 0;
  prelu_kernel<false><<<ew_grid((rows * cols + 3) / 4, 256), 256, 0, st>>>((const bf16*)x, nullptr, (bf16*)y, rows, cols, ldx,
                                                                            ldy, slope, noise, seed, counter);
  CXN_CHECK_LAUNCH();
  return 0;
}
// gslope[c] += sum_{x<=0} x*dy ; dx = (x > 0 ? 1 : mask) * dy  (dx may be NULL)
CXN_API int cxn_prelu_bwd(const void* x, const void* dy, void* dx, long long rows, int cols, long long ldx, long long ldg,
                          const float* slope, float* gslope, float noise, unsigned long long seed,
                          const unsigned long long* counter, cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return 0;
  int rc = colreduce<2>(x, dy, rows, cols, ldx, ldg, gslope, nullptr, st);
  if (rc) return rc;
  if (dx != nullptr) {
    prelu_kernel<true><<<ew_grid((rows * cols + 3) / 4, 256), 256, 0, st>>>((const bf16*)x, (const bf16*)dy, (bf16*)dx, rows,
                                                                             cols, ldx, ldg, slope, noise, seed, counter);
    CXN_CHECK_LAUNCH();
  }
  return 0;
}
