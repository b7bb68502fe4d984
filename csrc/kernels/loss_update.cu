// Loss layers (K13) and parameter updaters (K15) for sm_100a.
//
// Loss (reference src/layer/loss/*.hpp, loss_layer_base-inl.hpp:52-63): the output node
// holds softmax / sigmoid / identity of the logits; the gradient is
// (p - target) * grad_scale / (batch_size * update_period).  The reference
// round-trips through the host for this; here it is one device kernel.
//
// Updaters (reference src/updater/{sgd,nag,adam}_updater-inl.hpp): all weights, gradients
// and optimiser state of a net live in flat fp32 arenas cut into segments
// (one per (layer, tag)); one launch updates every segment, zeroes the
// gradient and refreshes the bf16 shadow copy the GEMMs read.
#include "common.cuh"

namespace cxn {

// ------------------------------------------------------------------ softmax
// in-place row softmax on fp32 [rows][ld], one warp per row
__global__ void softmax_rows_kernel(float* __restrict__ x, int rows, int cols, long long ld) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= rows) return;
  float* r = x + (long long)warp * ld;
  float m = -INFINITY;
  for (int c = lane; c < cols; c += 32) m = fmaxf(m, r[c]);
  m = warp_max(m);
  float s = 0.0f;
  for (int c = lane; c < cols; c += 32) {
    const float e = __expf(r[c] - m);
    r[c] = e;
    s += e;
  }
  s = warp_sum(s);
  const float inv = 1.0f / s;
  for (int c = lane; c < cols; c += 32) r[c] *= inv;
}
// Error metric on the device (reference src/utils/metric.h:92-121 runs this loop on the host after a
// D2H copy): one warp per row finds the first maximum of the scores (strict '>' scan), compares it with
// the label and adds the number of wrong rows to a double accumulator.  cols == 1: score > 0 is class 1.
__global__ void metric_error_kernel(const float* __restrict__ p, long long ld, const float* __restrict__ label,
                                    long long label_ld, int rows, int cols, double* __restrict__ accum) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  float wrong = 0.0f;
  if (warp < rows) {
    const float* r = p + (long long)warp * ld;
    int best;
    if (cols == 1) {
      best = r[0] > 0.0f ? 1 : 0;
    } else {
      float m = -INFINITY;
      int mi = 0x7fffffff;
      for (int c = lane; c < cols; c += 32) {
        const float v = r[c];
        if (v > m) { m = v; mi = c; }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float om = __shfl_xor_sync(0xffffffffu, m, o);
        const int oi = __shfl_xor_sync(0xffffffffu, mi, o);
        if (om > m || (om == m && oi < mi)) { m = om; mi = oi; }
      }
      best = mi;
    }
    wrong = (best != (int)label[(long long)warp * label_ld]) ? 1.0f : 0.0f;
  }
  // one atomic per block
  __shared__ float part[8];
  if (lane == 0) part[threadIdx.x >> 5] = wrong;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.0f;
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) t += part[i];
    if (t != 0.0f) atomicAdd(accum, (double)t);
  }
}
__global__ void sigmoid_rows_kernel(float* __restrict__ x, long long rows, int cols, long long ld) {
  const long long total = rows * cols;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / cols;
    float* p = x + r * ld + (i - r * cols);
    *p = 1.0f / (1.0f + __expf(-*p));
  }
}
// kind 0: softmax  (label = class index, one float per row)
// kind 1: l2 / multi_logistic (label matrix [rows][label_ld], same width as cols)
__global__ void loss_grad_kernel(const float* __restrict__ p, long long ld, const float* __restrict__ label,
                                 long long label_ld, bf16* __restrict__ g, long long ldg, long long rows, int cols,
                                 float scale, int kind) {
  const long long total = rows * cols;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / cols;
    const int c = (int)(i - r * cols);
    float v = p[r * ld + c];
    if (kind == 0) {
      const int k = (int)label[r * label_ld];
      if (c == k) v -= 1.0f;
    } else {
      v -= label[r * label_ld + c];
    }
    g[r * ldg + c] = __float2bfloat16(v * scale);
  }
}

// ------------------------------------------------------------------ updaters
struct Segment {
  long long start;        // element offset in the fp32 arenas (multiple of kChunk)
  long long n;            // elements
  long long shadow_off;   // element offset in the bf16 shadow arena, -1 = none
  int cols;               // row length of the 2-D view (for the padded shadow)
  int shadow_ld;
  float lr, wd, momentum, clip;   // refreshed every step
  float adam_lr_t, beta1, beta2;
  int flags;              // bit 0: the producer overwrites this gradient every step - the updater need not zero it
                          // bit 1: this step's update was applied by the producer (cxn_gemm_wgrad_update): skip
};
constexpr int kChunk = 1024;

__device__ __forceinline__ float clip_grad(float g, float c) {
  // reference sgd_updater-inl.hpp:15-22: NaN -> 0, clamp to [-c, c]
  if (g != g) return 0.0f;
  return fminf(fmaxf(g, -c), c);
}

// kind: 0 sgd, 1 nag, 2 adam.   m2 only used by adam.
__global__ void update_kernel(float* __restrict__ w, float* __restrict__ g, float* __restrict__ m,
                              float* __restrict__ m2, bf16* __restrict__ shadow, const Segment* __restrict__ segs,
                              const int* __restrict__ chunk_seg, long long chunk_begin, long long chunk_end,
                              int kind) {
  for (long long ch = chunk_begin + blockIdx.x; ch < chunk_end; ch += gridDim.x) {
    const Segment s = segs[chunk_seg[ch]];
    if (s.flags & 2) continue;
    const long long base = ch * kChunk;
#pragma unroll
    for (int it = 0; it < kChunk / 256 / 4; ++it) {
      const long long e = base + (it * 256 + threadIdx.x) * 4;
      const long long rel = e - s.start;
      if (rel >= s.n) continue;
      float4 wv = *reinterpret_cast<float4*>(w + e);
      float4 gv = *reinterpret_cast<float4*>(g + e);
      float4 mv = *reinterpret_cast<float4*>(m + e);
      float* wp = reinterpret_cast<float*>(&wv);
      float* gp = reinterpret_cast<float*>(&gv);
      float* mp = reinterpret_cast<float*>(&mv);
      float4 m2v = make_float4(0, 0, 0, 0);
      if (kind == 2) m2v = *reinterpret_cast<float4*>(m2 + e);
      float* m2p = reinterpret_cast<float*>(&m2v);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (rel + j >= s.n) { wp[j] = 0.0f; mp[j] = 0.0f; continue; }
        float gr = gp[j];
        if (kind == 0) {
          if (s.clip != 0.0f) gr = clip_grad(gr, s.clip);
          mp[j] = s.momentum * mp[j] - s.lr * (gr + s.wd * wp[j]);
          wp[j] += mp[j];
        } else if (kind == 1) {
          const float old = mp[j];
          mp[j] = s.momentum * old - s.lr * (gr + s.wd * wp[j]);
          wp[j] += (1.0f + s.momentum) * mp[j] - s.momentum * old;
        } else {
          if (s.wd > 0.0f) gr -= s.wd * wp[j];   // sign as in the reference (adam_updater-inl.hpp:74)
          mp[j] += s.beta1 * (gr - mp[j]);
          m2p[j] += s.beta2 * (gr * gr - m2p[j]);
          wp[j] -= s.adam_lr_t * (mp[j] / (sqrtf(m2p[j]) + 1e-8f));
        }
      }
      *reinterpret_cast<float4*>(w + e) = wv;
      *reinterpret_cast<float4*>(m + e) = mv;
      if (kind == 2) *reinterpret_cast<float4*>(m2 + e) = m2v;
      if (!(s.flags & 1)) *reinterpret_cast<float4*>(g + e) = make_float4(0, 0, 0, 0);
      if (s.shadow_off >= 0) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const long long k = rel + j;
          if (k < s.n) {
            const long long row = k / s.cols;
            shadow[s.shadow_off + row * s.shadow_ld + (k - row * s.cols)] = __float2bfloat16(wp[j]);
          }
        }
      }
    }
  }
}

// refresh every bf16 shadow from the fp32 master (after init / load / set_weight)
__global__ void shadow_kernel(const float* __restrict__ w, bf16* __restrict__ shadow, const Segment* __restrict__ segs,
                              const int* __restrict__ chunk_seg, long long nchunks) {
  for (long long ch = blockIdx.x; ch < nchunks; ch += gridDim.x) {
    const Segment s = segs[chunk_seg[ch]];
    if (s.shadow_off < 0) continue;
    for (int t = threadIdx.x; t < kChunk; t += blockDim.x) {
      const long long k = ch * kChunk + t - s.start;
      if (k < s.n) {
        const long long row = k / s.cols;
        shadow[s.shadow_off + row * s.shadow_ld + (k - row * s.cols)] = __float2bfloat16(w[ch * kChunk + t]);
      }
    }
  }
}

}  // namespace cxn

using namespace cxn;

CXN_API int cxn_softmax_rows(float* x, int rows, int cols, long long ld, cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return 0;
  softmax_rows_kernel<<<ceil_div(rows * 32, 256), 256, 0, st>>>(x, rows, cols, ld);
  CXN_CHECK_LAUNCH();
  return 0;
}
// accum (double, device) += number of rows whose first-maximum class differs from label[row][0]
CXN_API int cxn_metric_error(const float* p, long long ld, const float* label, long long label_ld, int rows, int cols,
                             double* accum, cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return 0;
  metric_error_kernel<<<ceil_div(rows * 32, 256), 256, 0, st>>>(p, ld, label, label_ld, rows, cols, accum);
  CXN_CHECK_LAUNCH();
  return 0;
}
CXN_API int cxn_sigmoid_rows(float* x, long long rows, int cols, long long ld, cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return 0;
  sigmoid_rows_kernel<<<ew_grid(rows * cols, 256), 256, 0, st>>>(x, rows, cols, ld);
  CXN_CHECK_LAUNCH();
  return 0;
}
CXN_API int cxn_loss_grad(const float* p, long long ld, const float* label, long long label_ld, void* g,
                          long long ldg, long long rows, int cols, float scale, int kind, cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return 0;
  loss_grad_kernel<<<ew_grid(rows * cols, 256), 256, 0, st>>>(p, ld, label, label_ld, (bf16*)g, ldg, rows, cols, scale, kind);
  CXN_CHECK_LAUNCH();
  return 0;
}
CXN_API int cxn_segment_bytes() { return (int)sizeof(Segment); }
CXN_API int cxn_update_chunk() { return kChunk; }

// Applies the updater to chunks [chunk_begin, chunk_end) of the arenas (a layer bucket or everything).
CXN_API int cxn_update(float* w, float* g, float* m, float* m2, void* shadow, const void* segs, const int* chunk_seg,
                       long long chunk_begin, long long chunk_end, int kind, cudaStream_t st) {
  const long long nchunks = chunk_end - chunk_begin;
  if (nchunks <= 0) return 0;
  const long long cap = (long long)num_sms() * 8;
  update_kernel<<<(int)(nchunks < cap ? nchunks : cap), 256, 0, st>>>(w, g, m, m2, (bf16*)shadow, (const Segment*)segs,
                                                                      chunk_seg, chunk_begin, chunk_end, kind);
  CXN_CHECK_LAUNCH();
  return 0;
}
CXN_API int cxn_refresh_shadow(const float* w, void* shadow, const void* segs, const int* chunk_seg, long long nchunks,
                               cudaStream_t st) {
  if (nchunks <= 0) return 0;
  const long long cap = (long long)num_sms() * 8;
  shadow_kernel<<<(int)(nchunks < cap ? nchunks : cap), 256, 0, st>>>(w, (bf16*)shadow, (const Segment*)segs, chunk_seg, nchunks);
  CXN_CHECK_LAUNCH();
  return 0;
}
