// Pooling (K8/K9) and local response normalisation (K10/K11) on NHWC bf16.
//
// Pooling semantics (reference src/layer/pooling_layer-inl.hpp:45-106): windows start at
// o*stride and are clipped at the border, output size is
// min(H-k+s-1, H-1)/s+1, avg divides by the full k*k, and max-unpool sends the
// gradient to EVERY input equal to the pooled maximum.
// LRN (reference src/layer/lrn_layer-inl.hpp:53-75):
//   norm = knorm + alpha/n * sum_{|j-c|<=n/2} x_j^2 ;  y = x * norm^-beta
//   dx = g*norm^-beta - 2*beta*alpha/n * x * sum_window( g*x*norm^(-beta-1) )
#include "common.cuh"

namespace cxn {

struct PoolGeom {
  int N, H, W, C, K_y, K_x, stride, OH, OW;
  int mode;   // 0 max, 1 sum, 2 avg
  int relu;   // relu_max_pooling: relu applied to the input first
};

template <int V>
__global__ void pool_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, PoolGeom g) {
  const int cv = g.C / V;
  const long long total = (long long)g.N * g.OH * g.OW * cv;
  const float scale = g.mode == 2 ? 1.0f / (g.K_y * g.K_x) : 1.0f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    long long r = i;
    const int c = (int)(r % cv) * V; r /= cv;
    const int ow = (int)(r % g.OW); r /= g.OW;
    const int oh = (int)(r % g.OH);
    const int n = (int)(r / g.OH);
    const int y0 = oh * g.stride, y1 = min(y0 + g.K_y, g.H);
    const int x0 = ow * g.stride, x1 = min(x0 + g.K_x, g.W);
    float acc[V];
#pragma unroll
    for (int j = 0; j < V; ++j) acc[j] = g.mode == 0 ? -INFINITY : 0.0f;
    for (int yy = y0; yy < y1; ++yy)
      for (int xx = x0; xx < x1; ++xx) {
        const bf16* s = x + (((long long)n * g.H + yy) * g.W + xx) * g.C + c;
        float f[V];
        if (V == 8) {
          const uint4 v = *reinterpret_cast<const uint4*>(s);
          const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float2 t = __bfloat1622float2(h[j]);
            f[2 * j] = t.x; f[2 * j + 1] = t.y;
          }
        } else {
          f[0] = __bfloat162float(*s);
        }
#pragma unroll
        for (int j = 0; j < V; ++j) {
          const float a = g.relu ? fmaxf(f[j], 0.0f) : f[j];
          acc[j] = g.mode == 0 ? fmaxf(acc[j], a) : acc[j] + a;
        }
      }
    bf16* d = y + (((long long)n * g.OH + oh) * g.OW + ow) * g.C + c;
    if (V == 8) {
      uint4 o;
      __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
      for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(acc[2 * j] * scale, acc[2 * j + 1] * scale);
      *reinterpret_cast<uint4*>(d) = o;
    } else {
      *d = __float2bfloat16(acc[0] * scale);
    }
  }
}

// ---- fast paths: 8 channels per thread, one image per blockIdx.y, windows fully unrolled so
// that every load of a thread is in flight before the first use (the generic kernels below walk
// the window with dependent loads and 64-bit index divisions).
__device__ __forceinline__ void unpack8(const uint4& v, float* f) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 t = __bfloat1622float2(h[j]);
    f[2 * j] = t.x; f[2 * j + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 o;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
  for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(f[2 * j], f[2 * j + 1]);
  return o;
}

// forward, window <= KMAX x KMAX
template <int KMAX>
__global__ void __launch_bounds__(256) pool_fwd_img_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, PoolGeom g) {
  const int cv = g.C >> 3;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= g.OH * g.OW * cv) return;
  const int pix = idx / cv, c = (idx - pix * cv) << 3;
  const int oh = pix / g.OW, ow = pix - oh * g.OW;
  const int n = blockIdx.y;
  const int y0 = oh * g.stride, x0 = ow * g.stride;
  const bf16* base = x + (((long long)n * g.H + y0) * g.W + x0) * g.C + c;
  uint4 v[KMAX][KMAX];
  bool ok[KMAX][KMAX];
#pragma unroll
  for (int a = 0; a < KMAX; ++a)
#pragma unroll
    for (int b = 0; b < KMAX; ++b) {
      ok[a][b] = a < g.K_y && b < g.K_x && y0 + a < g.H && x0 + b < g.W;
      v[a][b] = ok[a][b] ? *reinterpret_cast<const uint4*>(base + ((long long)a * g.W + b) * g.C) : make_uint4(0, 0, 0, 0);
    }
  bf16* dst = y + (((long long)n * g.OH + oh) * g.OW + ow) * g.C + c;
  if (g.mode == 0) {
    // max pooling in packed bf16 (exact): 4 instructions per window instead of 8 converts + 8 max
    const __nv_bfloat162 ninf = __float2bfloat162_rn(-INFINITY);
    __nv_bfloat162 m[4] = {ninf, ninf, ninf, ninf};
#pragma unroll
    for (int a = 0; a < KMAX; ++a)
#pragma unroll
      for (int b = 0; b < KMAX; ++b) {
        if (!ok[a][b]) continue;
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v[a][b]);
#pragma unroll
        for (int j = 0; j < 4; ++j) m[j] = __hmax2(m[j], h[j]);
      }
    if (g.relu) {
      const __nv_bfloat162 z = __float2bfloat162_rn(0.0f);
#pragma unroll
      for (int j = 0; j < 4; ++j) m[j] = __hmax2(m[j], z);
    }
    uint4 o;
    __nv_bfloat162* ho = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
    for (int j = 0; j < 4; ++j) ho[j] = m[j];
    *reinterpret_cast<uint4*>(dst) = o;
    return;
  }
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.0f;
#pragma unroll
  for (int a = 0; a < KMAX; ++a)
#pragma unroll
    for (int b = 0; b < KMAX; ++b) {
      if (!ok[a][b]) continue;
      float f[8];
      unpack8(v[a][b], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += g.relu ? fmaxf(f[j], 0.0f) : f[j];
    }
  const float scale = g.mode == 2 ? 1.0f / (g.K_y * g.K_x) : 1.0f;
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] *= scale;
  *reinterpret_cast<uint4*>(dst) = pack8(acc);
}

// backward, at most 2 x 2 windows cover an input pixel (K <= 2 * stride).
// A thread owns U vectors a block-width apart and issues their x loads (the DRAM stream) up
// front: one 16-byte miss per thread in flight is ~12 KB per SM, a quarter of what HBM needs.
template <int U>
__global__ void __launch_bounds__(256) pool_bwd_img_kernel(const bf16* __restrict__ x, const bf16* __restrict__ y,
                                                           const bf16* __restrict__ dy, bf16* __restrict__ dx, PoolGeom g) {
  const int cv = g.C >> 3;
  const int per_img = g.H * g.W * cv;
  const int n = blockIdx.y;
  const int idx0 = blockIdx.x * (blockDim.x * U) + threadIdx.x;
  const long long img = (long long)n * g.H * g.W * g.C;
  uint4 xr[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int idx = idx0 + u * blockDim.x;
    xr[u] = idx < per_img ? *reinterpret_cast<const uint4*>(x + img + (long long)idx * 8) : make_uint4(0, 0, 0, 0);
  }
  const float scale = g.mode == 2 ? 1.0f / (g.K_y * g.K_x) : 1.0f;
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int idx = idx0 + u * blockDim.x;
    if (idx >= per_img) break;
    const int pix = idx / cv, c = (idx - pix * cv) << 3;
    const int yy = pix / g.W, xx = pix - yy * g.W;
    const int py0 = yy < g.K_y ? 0 : (yy - g.K_y + g.stride) / g.stride;
    const int px0 = xx < g.K_x ? 0 : (xx - g.K_x + g.stride) / g.stride;
    const int py1 = min(yy / g.stride + 1, g.OH);
    const int px1 = min(xx / g.stride + 1, g.OW);
    const long long pbase = (((long long)n * g.OH + py0) * g.OW + px0) * g.C + c;
    uint4 pv[2][2], gv[2][2];
    bool ok[2][2];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        ok[a][b] = py0 + a < py1 && px0 + b < px1;
        const long long o = pbase + ((long long)a * g.OW + b) * g.C;
        pv[a][b] = ok[a][b] && g.mode == 0 ? *reinterpret_cast<const uint4*>(y + o) : make_uint4(0, 0, 0, 0);
        gv[a][b] = ok[a][b] ? *reinterpret_cast<const uint4*>(dy + o) : make_uint4(0, 0, 0, 0);
      }
    if (g.mode == 0) {
      // max pooling in packed bf16: mask = (x == pooled) as 1.0 / 0.0, acc += mask * dy.  At most
      // four terms, almost always one, so bf16 accumulation costs nothing measurable.
      const __nv_bfloat162 z = __float2bfloat162_rn(0.0f);
      const __nv_bfloat162* hx = reinterpret_cast<const __nv_bfloat162*>(&xr[u]);
      __nv_bfloat162 xm[4], am[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        xm[j] = g.relu ? __hmax2(hx[j], z) : hx[j];
        am[j] = z;
      }
#pragma unroll
      for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int b = 0; b < 2; ++b) {
          if (!ok[a][b]) continue;
          const __nv_bfloat162* hp = reinterpret_cast<const __nv_bfloat162*>(&pv[a][b]);
          const __nv_bfloat162* hg = reinterpret_cast<const __nv_bfloat162*>(&gv[a][b]);
#pragma unroll
          for (int j = 0; j < 4; ++j) am[j] = __hfma2(__heq2(xm[j], hp[j]), hg[j], am[j]);
        }
      if (g.relu) {
#pragma unroll
        for (int j = 0; j < 4; ++j) am[j] = __hmul2(am[j], __hgt2(hx[j], z));
      }
      uint4 o;
      __nv_bfloat162* ho = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
      for (int j = 0; j < 4; ++j) ho[j] = am[j];
      *reinterpret_cast<uint4*>(dx + img + (long long)idx * 8) = o;
      continue;
    }
    float xv[8], acc[8];
    unpack8(xr[u], xv);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (g.relu) xv[j] = fmaxf(xv[j], 0.0f);
      acc[j] = 0.0f;
    }
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        if (!ok[a][b]) continue;
        float p[8], gr[8];
        unpack8(pv[a][b], p);
        unpack8(gv[a][b], gr);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += (g.mode != 0 || xv[j] == p[j]) ? gr[j] : 0.0f;
      }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      acc[j] *= scale;
      if (g.relu) acc[j] = xv[j] > 0.0f ? acc[j] : 0.0f;
    }
    *reinterpret_cast<uint4*>(dx + img + (long long)idx * 8) = pack8(acc);
  }
}

// Max-pool forward with the same division-free 2-D thread indexing (window <= KMAX x KMAX)
template <int KMAX>
__global__ void __launch_bounds__(256) pool_fwd_row_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, PoolGeom g) {
  const int ow = blockIdx.x * blockDim.y + threadIdx.y;
  if (ow >= g.OW) return;
  const int oh = blockIdx.y, n = blockIdx.z, c = threadIdx.x << 3;
  const int y0 = oh * g.stride, x0 = ow * g.stride;
  const bf16* base = x + (((long long)n * g.H + y0) * g.W + x0) * g.C + c;
  const __nv_bfloat162 ninf = __float2bfloat162_rn(-INFINITY);
  uint4 v[KMAX][KMAX];
#pragma unroll
  for (int a = 0; a < KMAX; ++a)
#pragma unroll
    for (int b = 0; b < KMAX; ++b) {
      const bool ok = a < g.K_y && b < g.K_x && y0 + a < g.H && x0 + b < g.W;
      if (ok) {
        v[a][b] = *reinterpret_cast<const uint4*>(base + ((long long)a * g.W + b) * g.C);
      } else {
        __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&v[a][b]);
        h[0] = h[1] = h[2] = h[3] = ninf;
      }
    }
  __nv_bfloat162 m[4] = {ninf, ninf, ninf, ninf};
#pragma unroll
  for (int a = 0; a < KMAX; ++a)
#pragma unroll
    for (int b = 0; b < KMAX; ++b) {
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v[a][b]);
#pragma unroll
      for (int j = 0; j < 4; ++j) m[j] = __hmax2(m[j], h[j]);
    }
  if (g.relu) {
    const __nv_bfloat162 z = __float2bfloat162_rn(0.0f);
#pragma unroll
    for (int j = 0; j < 4; ++j) m[j] = __hmax2(m[j], z);
  }
  uint4 o;
  __nv_bfloat162* ho = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
  for (int j = 0; j < 4; ++j) ho[j] = m[j];
  *reinterpret_cast<uint4*>(y + (((long long)n * g.OH + oh) * g.OW + ow) * g.C + c) = o;
}

// Max-pool backward, at most 2 x 2 covering windows, division-free indexing: threadIdx.x = channel
// vector, threadIdx.y = pixel inside the row segment, blockIdx = (row segment, row, image).  SH = log2 of
// the stride when it is a power of two (else -1: generic division).
template <int SH>
__global__ void __launch_bounds__(256) pool_bwd_row_kernel(const bf16* __restrict__ x, const bf16* __restrict__ y,
                                                           const bf16* __restrict__ dy, bf16* __restrict__ dx, PoolGeom g) {
  const int xx = blockIdx.x * blockDim.y + threadIdx.y;
  if (xx >= g.W) return;
  const int yy = blockIdx.y, n = blockIdx.z, c = threadIdx.x << 3;
  auto div_s = [&](int v) { return SH >= 0 ? (v >> SH) : v / g.stride; };
  const int py0 = yy < g.K_y ? 0 : div_s(yy - g.K_y + g.stride);
  const int px0 = xx < g.K_x ? 0 : div_s(xx - g.K_x + g.stride);
  const int py1 = min(div_s(yy) + 1, g.OH);
  const int px1 = min(div_s(xx) + 1, g.OW);
  const long long xoff = (((long long)n * g.H + yy) * g.W + xx) * g.C + c;
  const uint4 xr = *reinterpret_cast<const uint4*>(x + xoff);
  const long long pbase = (((long long)n * g.OH + py0) * g.OW + px0) * g.C + c;
  const bool ok01 = px0 + 1 < px1, ok10 = py0 + 1 < py1;
  const bool okv[4] = {py0 < py1 && px0 < px1, py0 < py1 && ok01, ok10 && px0 < px1, ok10 && ok01};
  const long long offv[4] = {0, (long long)g.C, (long long)g.OW * g.C, (long long)(g.OW + 1) * g.C};
  uint4 pv[4], gv[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    pv[k] = okv[k] ? *reinterpret_cast<const uint4*>(y + pbase + offv[k]) : make_uint4(0, 0, 0, 0);
    gv[k] = okv[k] ? *reinterpret_cast<const uint4*>(dy + pbase + offv[k]) : make_uint4(0, 0, 0, 0);
  }
  const __nv_bfloat162 z = __float2bfloat162_rn(0.0f);
  const __nv_bfloat162* hx = reinterpret_cast<const __nv_bfloat162*>(&xr);
  __nv_bfloat162 xm[4], am[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    xm[j] = g.relu ? __hmax2(hx[j], z) : hx[j];
    am[j] = z;
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    // an invalid window contributes dy = 0 (its loads were replaced by zeros), so no branch is needed
    const __nv_bfloat162* hp = reinterpret_cast<const __nv_bfloat162*>(&pv[k]);
    const __nv_bfloat162* hg = reinterpret_cast<const __nv_bfloat162*>(&gv[k]);
#pragma unroll
    for (int j = 0; j < 4; ++j) am[j] = __hfma2(__heq2(xm[j], hp[j]), hg[j], am[j]);
  }
  if (g.relu) {
#pragma unroll
    for (int j = 0; j < 4; ++j) am[j] = __hmul2(am[j], __hgt2(hx[j], z));
  }
  uint4 o;
  __nv_bfloat162* ho = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
  for (int j = 0; j < 4; ++j) ho[j] = am[j];
  *reinterpret_cast<uint4*>(dx + xoff) = o;
}

// Stride-2 max-pool backward, one thread per 2 x 2 block of input pixels (x 8 channels).  The four pixels of a block
// lie in at most the same four windows - (i-1, j-1), (i-1, j), (i, j-1), (i, j) for the block (i, j), window sizes up to
// 4 x 4 - so the pooled value and its gradient are fetched once per block: 12 16-byte loads for four results where
// the per-pixel kernel issues 9 per result (its y / dy re-reads, not DRAM, bound it: 43 % of the copy roofline).
// Same arithmetic and the same window order as pool_bwd_row_kernel.
__global__ void __launch_bounds__(256) pool_bwd_quad_kernel(const bf16* __restrict__ x, const bf16* __restrict__ y,
                                                            const bf16* __restrict__ dy, bf16* __restrict__ dx, PoolGeom g) {
  const int j = blockIdx.x * blockDim.y + threadIdx.y;
  if (2 * j >= g.W) return;
  const int i = blockIdx.y, n = blockIdx.z, c = threadIdx.x << 3;
  uint4 pv[2][2], gv[2][2];
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      const int p = i - 1 + a, q = j - 1 + b;
      const bool ok = p >= 0 && p < g.OH && q >= 0 && q < g.OW;
      const long long off = (((long long)n * g.OH + p) * g.OW + q) * g.C + c;
      pv[a][b] = ok ? *reinterpret_cast<const uint4*>(y + off) : make_uint4(0, 0, 0, 0);
      gv[a][b] = ok ? *reinterpret_cast<const uint4*>(dy + off) : make_uint4(0, 0, 0, 0);   // no window: no gradient
    }
  const __nv_bfloat162 z = __float2bfloat162_rn(0.0f);
  uint4 xr[2][2];
#pragma unroll
  for (int dr = 0; dr < 2; ++dr)
#pragma unroll
    for (int dc = 0; dc < 2; ++dc) {
      const int r = 2 * i + dr, cc = 2 * j + dc;
      xr[dr][dc] = (r < g.H && cc < g.W)
                       ? *reinterpret_cast<const uint4*>(x + (((long long)n * g.H + r) * g.W + cc) * g.C + c)
                       : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
  for (int dr = 0; dr < 2; ++dr)
#pragma unroll
    for (int dc = 0; dc < 2; ++dc) {
      const int r = 2 * i + dr, cc = 2 * j + dc;
      if (r >= g.H || cc >= g.W) continue;
      const __nv_bfloat162* hx = reinterpret_cast<const __nv_bfloat162*>(&xr[dr][dc]);
      __nv_bfloat162 xm[4], am[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        xm[k] = g.relu ? __hmax2(hx[k], z) : hx[k];
        am[k] = z;
      }
#pragma unroll
      for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int b = 0; b < 2; ++b) {
          // window (i - 1 + a, j - 1 + b) starts at row 2 (i - 1 + a): the pixel is dr + 2 (1 - a) rows below its top
          if (dr + 2 * (1 - a) >= g.K_y || dc + 2 * (1 - b) >= g.K_x) continue;     // (uniform over the grid)
          const __nv_bfloat162* hp = reinterpret_cast<const __nv_bfloat162*>(&pv[a][b]);
          const __nv_bfloat162* hg = reinterpret_cast<const __nv_bfloat162*>(&gv[a][b]);
#pragma unroll
          for (int k = 0; k < 4; ++k) am[k] = __hfma2(__heq2(xm[k], hp[k]), hg[k], am[k]);
        }
      if (g.relu) {
#pragma unroll
        for (int k = 0; k < 4; ++k) am[k] = __hmul2(am[k], __hgt2(hx[k], z));
      }
      uint4 o;
      __nv_bfloat162* ho = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
      for (int k = 0; k < 4; ++k) ho[k] = am[k];
      *reinterpret_cast<uint4*>(dx + (((long long)n * g.H + r) * g.W + cc) * g.C + c) = o;
    }
}

// The same with the bias gradient of the layer in front fused in: when the pooled node is the (ReLU-masked) output
// of a convolution, dx IS that convolution's output gradient and its bias gradient is the column sum of dx over all
// pixels - a separate pass over the largest gradient tensors of the net otherwise (reference: the convolution's
// gbias = sum_rows(grad), src/layer/convolution_layer-inl.hpp:127-129).  A block walks `rows_per_block` image rows
// and keeps per-thread fp32 sums of the bf16 values it stores; one shared-memory reduction and 8 atomics per channel
// vector per block end it.  bias_grad must be zeroed (or hold the value to accumulate onto) before the launch.
template <int SH>
__global__ void __launch_bounds__(256) pool_bwd_row_bias_kernel(const bf16* __restrict__ x, const bf16* __restrict__ y,
                                                                const bf16* __restrict__ dy, bf16* __restrict__ dx,
                                                                float* __restrict__ bias_grad, int rows_per_block,
                                                                PoolGeom g) {
  __shared__ float red[256 * 8];
  const int xx = blockIdx.x * blockDim.y + threadIdx.y;
  const bool active = xx < g.W;
  const int n = blockIdx.z, c = threadIdx.x << 3;
  auto div_s = [&](int v) { return SH >= 0 ? (v >> SH) : v / g.stride; };
  const int px0 = xx < g.K_x ? 0 : div_s(xx - g.K_x + g.stride);
  const int px1 = min(div_s(xx) + 1, g.OW);
  const bool ok01 = px0 + 1 < px1;
  const __nv_bfloat162 z = __float2bfloat162_rn(0.0f);
  float bs[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) bs[j] = 0.0f;
  const int y_begin = blockIdx.y * rows_per_block, y_end = min(g.H, y_begin + rows_per_block);
  for (int yy = y_begin; active && yy < y_end; ++yy) {
    const int py0 = yy < g.K_y ? 0 : div_s(yy - g.K_y + g.stride);
    const int py1 = min(div_s(yy) + 1, g.OH);
    const long long xoff = (((long long)n * g.H + yy) * g.W + xx) * g.C + c;
    const uint4 xr = *reinterpret_cast<const uint4*>(x + xoff);
    const long long pbase = (((long long)n * g.OH + py0) * g.OW + px0) * g.C + c;
    const bool ok10 = py0 + 1 < py1;
    const bool okv[4] = {py0 < py1 && px0 < px1, py0 < py1 && ok01, ok10 && px0 < px1, ok10 && ok01};
    const long long offv[4] = {0, (long long)g.C, (long long)g.OW * g.C, (long long)(g.OW + 1) * g.C};
    uint4 pv[4], gv[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      pv[k] = okv[k] ? *reinterpret_cast<const uint4*>(y + pbase + offv[k]) : make_uint4(0, 0, 0, 0);
      gv[k] = okv[k] ? *reinterpret_cast<const uint4*>(dy + pbase + offv[k]) : make_uint4(0, 0, 0, 0);
    }
    const __nv_bfloat162* hx = reinterpret_cast<const __nv_bfloat162*>(&xr);
    __nv_bfloat162 xm[4], am[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      xm[j] = g.relu ? __hmax2(hx[j], z) : hx[j];
      am[j] = z;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const __nv_bfloat162* hp = reinterpret_cast<const __nv_bfloat162*>(&pv[k]);
      const __nv_bfloat162* hg = reinterpret_cast<const __nv_bfloat162*>(&gv[k]);
#pragma unroll
      for (int j = 0; j < 4; ++j) am[j] = __hfma2(__heq2(xm[j], hp[j]), hg[j], am[j]);
    }
    if (g.relu) {
#pragma unroll
      for (int j = 0; j < 4; ++j) am[j] = __hmul2(am[j], __hgt2(hx[j], z));
    }
    uint4 o;
    __nv_bfloat162* ho = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      ho[j] = am[j];
      const float2 f = __bfloat1622float2(am[j]);
      bs[2 * j] += f.x;
      bs[2 * j + 1] += f.y;
    }
    *reinterpret_cast<uint4*>(dx + xoff) = o;
  }
  // block reduction over the pixel dimension, then one atomic per channel
  const int cw = blockDim.x * 8;
#pragma unroll
  for (int j = 0; j < 8; ++j) red[threadIdx.y * cw + c + j] = bs[j];
  __syncthreads();
  if (threadIdx.y == 0) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float sum = 0.0f;
      for (int r = 0; r < (int)blockDim.y; ++r) sum += red[r * cw + c + j];
      atomicAdd(bias_grad + c + j, sum);
    }
  }
}

// dx[n][y][x][c] = sum over covering windows of  partial(x, pooled) * dy   (* 1/k^2 for avg)
// With relu fused, x is compared after relu and the result is masked by x>0.
template <int V>
__global__ void pool_bwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ y, const bf16* __restrict__ dy,
                                bf16* __restrict__ dx, PoolGeom g) {
  const int cv = g.C / V;
  const long long total = (long long)g.N * g.H * g.W * cv;
  const float scale = g.mode == 2 ? 1.0f / (g.K_y * g.K_x) : 1.0f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    long long r = i;
    const int c = (int)(r % cv) * V; r /= cv;
    const int xx = (int)(r % g.W); r /= g.W;
    const int yy = (int)(r % g.H);
    const int n = (int)(r / g.H);
    const int py0 = yy < g.K_y ? 0 : (yy - g.K_y + g.stride) / g.stride;
    const int px0 = xx < g.K_x ? 0 : (xx - g.K_x + g.stride) / g.stride;
    const int py1 = min((yy + g.stride) / g.stride, g.OH);
    const int px1 = min((xx + g.stride) / g.stride, g.OW);
    const long long xoff = (((long long)n * g.H + yy) * g.W + xx) * g.C + c;
    float xv[V], acc[V];
    if (V == 8) {
      const uint4 v = *reinterpret_cast<const uint4*>(x + xoff);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 t = __bfloat1622float2(h[j]);
        xv[2 * j] = t.x; xv[2 * j + 1] = t.y;
      }
    } else {
      xv[0] = __bfloat162float(x[xoff]);
    }
#pragma unroll
    for (int j = 0; j < V; ++j) {
      if (g.relu) xv[j] = fmaxf(xv[j], 0.0f);
      acc[j] = 0.0f;
    }
    for (int py = py0; py < py1; ++py)
      for (int px = px0; px < px1; ++px) {
        const long long poff = (((long long)n * g.OH + py) * g.OW + px) * g.C + c;
        float pv[V], gv[V];
        if (V == 8) {
          const uint4 a = *reinterpret_cast<const uint4*>(y + poff);
          const uint4 b = *reinterpret_cast<const uint4*>(dy + poff);
          const __nv_bfloat162* ha = reinterpret_cast<const __nv_bfloat162*>(&a);
          const __nv_bfloat162* hb = reinterpret_cast<const __nv_bfloat162*>(&b);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float2 t = __bfloat1622float2(ha[j]), u = __bfloat1622float2(hb[j]);
            pv[2 * j] = t.x; pv[2 * j + 1] = t.y;
            gv[2 * j] = u.x; gv[2 * j + 1] = u.y;
          }
        } else {
          pv[0] = __bfloat162float(y[poff]);
          gv[0] = __bfloat162float(dy[poff]);
        }
#pragma unroll
        for (int j = 0; j < V; ++j) {
          if (g.mode == 0) acc[j] += (xv[j] == pv[j]) ? gv[j] : 0.0f;
          else acc[j] += gv[j];
        }
      }
#pragma unroll
    for (int j = 0; j < V; ++j) {
      acc[j] *= scale;
      if (g.relu) acc[j] = xv[j] > 0.0f ? acc[j] : 0.0f;
    }
    if (V == 8) {
      uint4 o;
      __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
      for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(acc[2 * j], acc[2 * j + 1]);
      *reinterpret_cast<uint4*>(dx + xoff) = o;
    } else {
      dx[xoff] = __float2bfloat16(acc[0]);
    }
  }
}

// ------------------------------------------------------------------ LRN
// One warp per pixel; the pixel's channel vector lives in shared memory.
__global__ void lrn_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, long long pixels, int C, int nsize,
                               float salpha, float beta, float knorm) {
  extern __shared__ float sm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
  float* xs = sm + warp * C;
  const int half = nsize / 2;
  for (long long p = (long long)blockIdx.x * wpb + warp; p < pixels; p += (long long)gridDim.x * wpb) {
    const bf16* xp = x + p * C;
    for (int c = lane; c < C; c += 32) xs[c] = __bfloat162float(xp[c]);
    __syncwarp();
    for (int c = lane; c < C; c += 32) {
      const int lo = max(0, c - half), hi = min(C, c + half + 1);
      float s = 0.0f;
      for (int j = lo; j < hi; ++j) s += xs[j] * xs[j];
      const float norm = s * salpha + knorm;
      y[p * C + c] = __float2bfloat16(xs[c] * __powf(norm, -beta));
    }
    __syncwarp();
  }
}
__global__ void lrn_bwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ dy, bf16* __restrict__ dx,
                               long long pixels, int C, int nsize, float salpha, float beta, float knorm) {
  extern __shared__ float sm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
  float* xs = sm + warp * 3 * C;
  float* nb = xs + C;   // norm^-beta
  float* ts = nb + C;   // g * x * norm^(-beta-1)
  const int half = nsize / 2;
  for (long long p = (long long)blockIdx.x * wpb + warp; p < pixels; p += (long long)gridDim.x * wpb) {
    for (int c = lane; c < C; c += 32) xs[c] = __bfloat162float(x[p * C + c]);
    __syncwarp();
    for (int c = lane; c < C; c += 32) {
      const int lo = max(0, c - half), hi = min(C, c + half + 1);
      float s = 0.0f;
      for (int j = lo; j < hi; ++j) s += xs[j] * xs[j];
      const float norm = s * salpha + knorm;
      const float pb = __powf(norm, -beta);
      const float g = __bfloat162float(dy[p * C + c]);
      nb[c] = pb;
      ts[c] = g * xs[c] * pb / norm;
    }
    __syncwarp();
    for (int c = lane; c < C; c += 32) {
      const int lo = max(0, c - half), hi = min(C, c + half + 1);
      float s = 0.0f;
      for (int j = lo; j < hi; ++j) s += ts[j];
      const float g = __bfloat162float(dy[p * C + c]);
      dx[p * C + c] = __float2bfloat16(g * nb[c] - 2.0f * beta * salpha * s * xs[c]);
    }
    __syncwarp();
  }
}

// ------------------------------------------------------------------ LRN, register version
// C % 8 == 0 and C <= 256: TPP = C/8 lanes own one pixel, 8 channels (one 16-byte load) each;
// the +-2 channel halo (local_size <= 5) comes from the neighbouring lanes by shuffle.
// e[0..11] = {left lane's v[6], v[7], own v[0..7], right lane's v[0], v[1]} (zero outside the pixel)
__device__ __forceinline__ void halo2(const float* v, float* e, int lane_in_pix, int tpp) {
  const float l6 = __shfl_up_sync(0xffffffffu, v[6], 1), l7 = __shfl_up_sync(0xffffffffu, v[7], 1);
  const float r0 = __shfl_down_sync(0xffffffffu, v[0], 1), r1 = __shfl_down_sync(0xffffffffu, v[1], 1);
  const bool hl = lane_in_pix > 0, hr = lane_in_pix < tpp - 1;
  e[0] = hl ? l6 : 0.f; e[1] = hl ? l7 : 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) e[2 + j] = v[j];
  e[10] = hr ? r0 : 0.f; e[11] = hr ? r1 : 0.f;
}
template <bool BWD>
__global__ void lrn_reg_kernel(const bf16* __restrict__ x, const bf16* __restrict__ dy, bf16* __restrict__ out,
                               long long pixels, int C, int half, float salpha, float beta, float knorm) {
  const int tpp = C >> 3;                 // lanes per pixel
  const int ppw = 32 / tpp;               // pixels per warp
  const int lane = threadIdx.x & 31;
  const int sub = lane / tpp, lip = lane - sub * tpp;
  const bool active = sub < ppw;
  const long long warp0 = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  // two pixels per thread and iteration, both loads issued before the arithmetic of the first: one
  // 16-byte load in flight per thread is not enough outstanding bytes to cover HBM latency
  constexpr int U = 2;
  for (long long p0 = warp0 * ppw; p0 < pixels; p0 += U * nwarps * ppw) {
    uint4 raw[U], rawg[U];
    bool ok[U];
    long long pixu[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      pixu[u] = p0 + u * nwarps * ppw + sub;
      ok[u] = active && pixu[u] < pixels;
      raw[u] = make_uint4(0, 0, 0, 0);
      rawg[u] = make_uint4(0, 0, 0, 0);
      if (ok[u]) {
        raw[u] = *reinterpret_cast<const uint4*>(x + pixu[u] * C + lip * 8);
        if (BWD) rawg[u] = *reinterpret_cast<const uint4*>(dy + pixu[u] * C + lip * 8);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (p0 + u * nwarps * ppw >= pixels) break;      // warp-uniform: the shuffles below stay converged
      float xv[8], gv[8];
      unpack8(raw[u], xv);
      if (BWD) unpack8(rawg[u], gv);
      float sq[8], e[12];
#pragma unroll
      for (int j = 0; j < 8; ++j) sq[j] = xv[j] * xv[j];
      halo2(sq, e, lip, tpp);
      float nb[8], o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float s = e[2 + j];
        if (half >= 1) s += e[1 + j] + e[3 + j];
        if (half >= 2) s += e[j] + e[4 + j];
        const float norm = s * salpha + knorm;
        const float l2 = __log2f(norm);
        nb[j] = exp2f(-beta * l2);                                      // norm^-beta
        if (!BWD) o[j] = xv[j] * nb[j];
        else sq[j] = gv[j] * xv[j] * exp2f((-beta - 1.0f) * l2);        // t = g * x * norm^(-beta-1)
      }
      if (BWD) {
        halo2(sq, e, lip, tpp);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float s = e[2 + j];
          if (half >= 1) s += e[1 + j] + e[3 + j];
          if (half >= 2) s += e[j] + e[4 + j];
          o[j] = gv[j] * nb[j] - 2.0f * beta * salpha * s * xv[j];
        }
      }
      if (ok[u]) *reinterpret_cast<uint4*>(out + pixu[u] * C + lip * 8) = pack8(o);
    }
  }
}

// ------------------------------------------------------------------ column sums, vectorised
// out[c] (+)= sum_r x[r][c] for C % 8 == 0: each thread owns 8 channels (16-byte loads) and
// strides over rows; partial sums are combined through shared memory, one atomic per column per CTA.
__global__ void colsum_vec_kernel(const bf16* __restrict__ x, long long rows, int cols, long long ld,
                                  float* __restrict__ out) {
  extern __shared__ float red[];            // [rows_per_block][cols]
  const int tpr = cols >> 3;                // threads per row
  const int rpb = blockDim.x / tpr;         // rows per block iteration
  const int tr = threadIdx.x / tpr, tc = threadIdx.x - tr * tpr;
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  if (tr < rpb) {
    // four rows per iteration, all four loads issued first: one 16-byte load in flight per thread does not cover
    // the memory latency (measured 53-57 % of the copy roofline with one)
    constexpr int U = 4;
    const long long step = (long long)gridDim.x * rpb;
    long long r = (long long)blockIdx.x * rpb + tr;
    for (; r + (U - 1) * step < rows; r += U * step) {
      uint4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) v[u] = *reinterpret_cast<const uint4*>(x + (r + u * step) * ld + tc * 8);
#pragma unroll
      for (int u = 0; u < U; ++u) {
        float f[8];
        unpack8(v[u], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    }
    for (; r < rows; r += step) {
      const uint4 v = *reinterpret_cast<const uint4*>(x + r * ld + tc * 8);
      float f[8];
      unpack8(v, f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += f[j];
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) red[tr * cols + tc * 8 + j] = acc[j];
  }
  __syncthreads();
  for (int c = threadIdx.x; c < cols; c += blockDim.x) {
    float s = 0.f;
    for (int r = 0; r < rpb; ++r) s += red[r * cols + c];
    atomicAdd(out + c, s);
  }
}

}  // namespace cxn

using namespace cxn;

static inline bool pool_vec(const void* a, const void* b, int C) {
  return (C % 8 == 0) && (((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15) == 0);
}

CXN_API int cxn_pool_fwd(const void* x, void* y, int N, int H, int W, int C, int ky, int kx, int stride, int OH,
                         int OW, int mode, int relu, cudaStream_t st) {
  PoolGeom g{N, H, W, C, ky, kx, stride, OH, OW, mode, relu};
  if ((long long)N * OH * OW * C <= 0) return 0;
  const bool small = pool_vec(x, y, C) && N <= 65535 && (long long)OH * OW * (C / 8) < (1LL << 30);
  if (small && mode == 0 && ky <= 3 && kx <= 3 && OH <= 65535 && C / 8 <= 256) {
    const int cv = C / 8;
    const int py_max = 256 / cv > 0 ? 256 / cv : 1;
    const int py = ceil_div(OW, ceil_div(OW, py_max));       // equal shares of an output row for its blocks
    dim3 block(cv, py), grid(ceil_div(OW, py), OH, N);
    if (ky <= 2 && kx <= 2) pool_fwd_row_kernel<2><<<grid, block, 0, st>>>((const bf16*)x, (bf16*)y, g);
    else pool_fwd_row_kernel<3><<<grid, block, 0, st>>>((const bf16*)x, (bf16*)y, g);
  } else if (small && ky <= 2 && kx <= 2)
    pool_fwd_img_kernel<2><<<dim3((unsigned)ceil_div_ll((long long)OH * OW * (C / 8), 256), N), 256, 0, st>>>((const bf16*)x, (bf16*)y, g);
  else if (small && ky <= 3 && kx <= 3)
    pool_fwd_img_kernel<3><<<dim3((unsigned)ceil_div_ll((long long)OH * OW * (C / 8), 256), N), 256, 0, st>>>((const bf16*)x, (bf16*)y, g);
  else if (pool_vec(x, y, C))
    pool_fwd_kernel<8><<<ew_grid((long long)N * OH * OW * (C / 8), 256, 16), 256, 0, st>>>((const bf16*)x, (bf16*)y, g);
  else
    pool_fwd_kernel<1><<<ew_grid((long long)N * OH * OW * C, 256, 16), 256, 0, st>>>((const bf16*)x, (bf16*)y, g);
  CXN_CHECK_LAUNCH();
  return 0;
}
// Max-pool backward that also accumulates the column sums of dx (over all N*H*W pixels) into bias_grad[C] (fp32,
// zeroed or holding the value to add to): the bias gradient of a convolution whose ReLU-masked output is the pooled
// node.  Returns -8 when the shape is outside the fused kernel's range (the caller then runs cxn_pool_bwd and a
// column sum).
CXN_API int cxn_pool_bwd_bias(const void* x, const void* y, const void* dy, void* dx, float* bias_grad, int N, int H, int W,
                              int C, int ky, int kx, int stride, int OH, int OW, int mode, int relu, cudaStream_t st) {
  PoolGeom g{N, H, W, C, ky, kx, stride, OH, OW, mode, relu};
  if ((long long)N * H * W * C <= 0) return 0;
  const bool v = pool_vec(x, y, C) && pool_vec(dy, dx, C);
  if (!(v && mode == 0 && N <= 65535 && H <= 65535 && ky <= 2 * stride && kx <= 2 * stride && C / 8 <= 256) || !bias_grad)
    return -8;
  const int cv = C / 8;
  const int py = 256 / cv > 0 ? 256 / cv : 1;
  const int segs = ceil_div(W, py);
  // about eight blocks per SM, each walking a band of image rows
  long long bands = ceil_div_ll(8LL * num_sms(), (long long)segs * N);
  if (bands < 1) bands = 1;
  if (bands > H) bands = H;
  const int rpb = ceil_div(H, (int)bands);
  dim3 block(cv, py), grid(segs, ceil_div(H, rpb), N);
  const bf16 *xp = (const bf16*)x, *yp = (const bf16*)y, *gp = (const bf16*)dy;
  if (stride == 1) pool_bwd_row_bias_kernel<0><<<grid, block, 0, st>>>(xp, yp, gp, (bf16*)dx, bias_grad, rpb, g);
  else if (stride == 2) pool_bwd_row_bias_kernel<1><<<grid, block, 0, st>>>(xp, yp, gp, (bf16*)dx, bias_grad, rpb, g);
  else if (stride == 4) pool_bwd_row_bias_kernel<2><<<grid, block, 0, st>>>(xp, yp, gp, (bf16*)dx, bias_grad, rpb, g);
  else pool_bwd_row_bias_kernel<-1><<<grid, block, 0, st>>>(xp, yp, gp, (bf16*)dx, bias_grad, rpb, g);
  CXN_CHECK_LAUNCH();
  return 0;
}

CXN_API int cxn_pool_bwd(const void* x, const void* y, const void* dy, void* dx, int N, int H, int W, int C, int ky,
                         int kx, int stride, int OH, int OW, int mode, int relu, cudaStream_t st) {
  PoolGeom g{N, H, W, C, ky, kx, stride, OH, OW, mode, relu};
  if ((long long)N * H * W * C <= 0) return 0;
  const bool v = pool_vec(x, y, C) && pool_vec(dy, dx, C);
  static const bool quad_on = []() { const char* e = getenv("CXN_POOL_QUAD"); return e == nullptr || e[0] != '0'; }();
  if (quad_on && v && mode == 0 && stride == 2 && ky <= 4 && kx <= 4 && N <= 65535 && H <= 2 * 65535 && C / 8 <= 256) {
    // stride-2 max pooling: one thread per 2 x 2 input block and 8 channels
    const int cv = C / 8, qw = (W + 1) / 2;
    const int py_max = 256 / cv > 0 ? 256 / cv : 1;
    const int py = ceil_div(qw, ceil_div(qw, py_max));       // equal shares of a row for its blocks
    dim3 block(cv, py), grid(ceil_div(qw, py), (H + 1) / 2, N);
    pool_bwd_quad_kernel<<<grid, block, 0, st>>>((const bf16*)x, (const bf16*)y, (const bf16*)dy, (bf16*)dx, g);
  } else if (v && mode == 0 && N <= 65535 && H <= 65535 && ky <= 2 * stride && kx <= 2 * stride && C / 8 <= 256) {
    // max pooling: one thread per (pixel, 8 channels) with 2-D thread indexing
    const int cv = C / 8;
    const int py = 256 / cv > 0 ? 256 / cv : 1;
    dim3 block(cv, py), grid(ceil_div(W, py), H, N);
    const bf16 *xp = (const bf16*)x, *yp = (const bf16*)y, *gp = (const bf16*)dy;
    if (stride == 1) pool_bwd_row_kernel<0><<<grid, block, 0, st>>>(xp, yp, gp, (bf16*)dx, g);
    else if (stride == 2) pool_bwd_row_kernel<1><<<grid, block, 0, st>>>(xp, yp, gp, (bf16*)dx, g);
    else if (stride == 4) pool_bwd_row_kernel<2><<<grid, block, 0, st>>>(xp, yp, gp, (bf16*)dx, g);
    else pool_bwd_row_kernel<-1><<<grid, block, 0, st>>>(xp, yp, gp, (bf16*)dx, g);
  } else if (v && N <= 65535 && ky <= 2 * stride && kx <= 2 * stride && (long long)H * W * (C / 8) < (1LL << 30))
    pool_bwd_img_kernel<4><<<dim3((unsigned)ceil_div_ll((long long)H * W * (C / 8), 256 * 4), N), 256, 0, st>>>(
        (const bf16*)x, (const bf16*)y, (const bf16*)dy, (bf16*)dx, g);
  else if (v)
    pool_bwd_kernel<8><<<ew_grid((long long)N * H * W * (C / 8), 256, 16), 256, 0, st>>>(
        (const bf16*)x, (const bf16*)y, (const bf16*)dy, (bf16*)dx, g);
  else
    pool_bwd_kernel<1><<<ew_grid((long long)N * H * W * C, 256, 16), 256, 0, st>>>(
        (const bf16*)x, (const bf16*)y, (const bf16*)dy, (bf16*)dx, g);
  CXN_CHECK_LAUNCH();
  return 0;
}
static inline bool lrn_reg_ok(const void* a, const void* b, int C, int nsize) {
  return (C % 8 == 0) && C <= 256 && nsize <= 5 && (nsize & 1) &&
         (((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15) == 0);
}
static inline int lrn_reg_grid(long long pixels, int C) {
  const int ppw = 32 / (C / 8);
  const long long warps = ceil_div_ll(pixels, ppw);
  const long long blocks = ceil_div_ll(warps, 8);
  const long long cap = (long long)num_sms() * 8;
  return (int)(blocks < cap ? blocks : cap);
}

CXN_API int cxn_lrn_fwd(const void* x, void* y, long long pixels, int C, int nsize, float alpha, float beta,
                        float knorm, cudaStream_t st) {
  if (pixels <= 0 || C <= 0) return 0;
  if (lrn_reg_ok(x, y, C, nsize)) {
    lrn_reg_kernel<false><<<lrn_reg_grid(pixels, C), 256, 0, st>>>((const bf16*)x, nullptr, (bf16*)y, pixels, C, nsize / 2,
                                                                   alpha / nsize, beta, knorm);
    CXN_CHECK_LAUNCH();
    return 0;
  }
  const int threads = 256, wpb = threads / 32;
  const size_t smem = (size_t)wpb * C * sizeof(float);
  if (smem > 48 * 1024) return -7;
  long long blocks = ceil_div_ll(pixels, wpb);
  const long long cap = (long long)num_sms() * 8;
  lrn_fwd_kernel<<<(int)(blocks < cap ? blocks : cap), threads, smem, st>>>((const bf16*)x, (bf16*)y, pixels, C, nsize,
                                                                              alpha / nsize, beta, knorm);
  CXN_CHECK_LAUNCH();
  return 0;
}
CXN_API int cxn_lrn_bwd(const void* x, const void* dy, void* dx, long long pixels, int C, int nsize, float alpha,
                        float beta, float knorm, cudaStream_t st) {
  if (pixels <= 0 || C <= 0) return 0;
  if (lrn_reg_ok(x, dy, C, nsize) && (reinterpret_cast<uintptr_t>(dx) & 15) == 0) {
    lrn_reg_kernel<true><<<lrn_reg_grid(pixels, C), 256, 0, st>>>((const bf16*)x, (const bf16*)dy, (bf16*)dx, pixels, C,
                                                                  nsize / 2, alpha / nsize, beta, knorm);
    CXN_CHECK_LAUNCH();
    return 0;
  }
  const int threads = 128, wpb = threads / 32;
  const size_t smem = (size_t)wpb * 3 * C * sizeof(float);
  if (smem > 48 * 1024) return -7;
  long long blocks = ceil_div_ll(pixels, wpb);
  const long long cap = (long long)num_sms() * 8;
  lrn_bwd_kernel<<<(int)(blocks < cap ? blocks : cap), threads, smem, st>>>((const bf16*)x, (const bf16*)dy, (bf16*)dx,
                                                                              pixels, C, nsize, alpha / nsize, beta, knorm);
  CXN_CHECK_LAUNCH();
  return 0;
}

// vectorised column sum (bias gradients of conv layers); falls back to the generic kernel in elementwise.cu
CXN_API int cxn_colsum_bf16_vec(const void* x, long long rows, int cols, long long ld, float* out, cudaStream_t st) {
  const int tpr = cols / 8;
  const int threads = 256;
  const int rpb = threads / tpr;
  const size_t smem = (size_t)rpb * cols * sizeof(float);
  long long blocks = ceil_div_ll(rows, (long long)rpb * 8);
  const long long cap = (long long)num_sms() * 4;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  colsum_vec_kernel<<<(int)blocks, threads, smem, st>>>((const bf16*)x, rows, cols, ld, out);
  CXN_CHECK_LAUNCH();
  return 0;
}
