// NHWC bf16 patch gather / scatter used by the convolution layer (K1-K3).
// col layout: [group][n*OH*OW + oh*OW + ow][(kh*KW + kw)*Cg + c], row stride ldk.
// Semantics follow the reference's unpack_patch2col / pack_col2patch use in
// src/layer/convolution_layer-inl.hpp:79-153 (cross-correlation, zero padding),
// with a channels-last ordering of the reduction index.
#include "common.cuh"

namespace cxn {

struct ConvGeom {
  int N, H, W, C;       // input
  int G, Cg;            // groups, channels per group
  int KH, KW, stride, pad_y, pad_x;
  int OH, OW;
  int ldk;              // col row stride (elements)
  long long gstride;    // col group stride (elements)
};

template <int V>  // V = 8 (uint4) or 1
__global__ void im2col_kernel(const bf16* __restrict__ x, bf16* __restrict__ col, ConvGeom g) {
  const int cv = g.Cg / V;
  const long long rows = (long long)g.N * g.OH * g.OW;
  const long long per_row = (long long)g.KH * g.KW * cv;
  const long long total = rows * per_row * g.G;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    long long r = i;
    const int c = (int)(r % cv) * V; r /= cv;
    const int kw = (int)(r % g.KW); r /= g.KW;
    const int kh = (int)(r % g.KH); r /= g.KH;
    const int grp = (int)(r % g.G); r /= g.G;
    const long long row = r;
    const int ow = (int)(row % g.OW);
    const long long t = row / g.OW;
    const int oh = (int)(t % g.OH);
    const int n = (int)(t / g.OH);
    const int ih = oh * g.stride - g.pad_y + kh;
    const int iw = ow * g.stride - g.pad_x + kw;
    const bool in = (ih >= 0 && ih < g.H && iw >= 0 && iw < g.W);
    bf16* d = col + grp * g.gstride + row * g.ldk + (kh * g.KW + kw) * g.Cg + c;
    const bf16* s = x + (((long long)n * g.H + ih) * g.W + iw) * g.C + grp * g.Cg + c;
    if (V == 8) {
      uint4 v = make_uint4(0, 0, 0, 0);
      if (in) v = *reinterpret_cast<const uint4*>(s);
      *reinterpret_cast<uint4*>(d) = v;
    } else {
      *d = in ? *s : __float2bfloat16(0.0f);
    }
  }
}

// dx[n][ih][iw][grp*Cg + c] = sum over (kh,kw) of dcol entries that read this pixel.
template <int V>
__global__ void col2im_kernel(const bf16* __restrict__ col, bf16* __restrict__ dx, ConvGeom g) {
  const int cv = g.Cg / V;
  const long long total = (long long)g.N * g.H * g.W * g.G * cv;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    long long r = i;
    const int c = (int)(r % cv) * V; r /= cv;
    const int grp = (int)(r % g.G); r /= g.G;
    const int iw = (int)(r % g.W); r /= g.W;
    const int ih = (int)(r % g.H);
    const int n = (int)(r / g.H);
    float acc[V];
#pragma unroll
    for (int j = 0; j < V; ++j) acc[j] = 0.0f;
    for (int kh = 0; kh < g.KH; ++kh) {
      const int th = ih + g.pad_y - kh;
      if (th < 0 || th % g.stride) continue;
      const int oh = th / g.stride;
      if (oh >= g.OH) continue;
      for (int kw = 0; kw < g.KW; ++kw) {
        const int tw = iw + g.pad_x - kw;
        if (tw < 0 || tw % g.stride) continue;
        const int ow = tw / g.stride;
        if (ow >= g.OW) continue;
        const long long row = ((long long)n * g.OH + oh) * g.OW + ow;
        const bf16* s = col + grp * g.gstride + row * g.ldk + (kh * g.KW + kw) * g.Cg + c;
        if (V == 8) {
          const uint4 v = *reinterpret_cast<const uint4*>(s);
          const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float2 f = __bfloat1622float2(h[j]);
            acc[2 * j] += f.x;
            acc[2 * j + 1] += f.y;
          }
        } else {
          acc[0] += __bfloat162float(*s);
        }
      }
    }
    bf16* d = dx + (((long long)n * g.H + ih) * g.W + iw) * g.C + grp * g.Cg + c;
    if (V == 8) {
      uint4 o;
      __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
      for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(acc[2 * j], acc[2 * j + 1]);
      *reinterpret_cast<uint4*>(d) = o;
    } else {
      *d = __float2bfloat16(acc[0]);
    }
  }
}

static inline bool vec_ok(const void* a, const void* b, const ConvGeom& g) {
  return (g.Cg % 8 == 0) && (g.C % 8 == 0) && (g.ldk % 8 == 0) && (g.gstride % 8 == 0) &&
         (((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15) == 0);
}

}  // namespace cxn

using namespace cxn;

CXN_API int cxn_im2col_nhwc(const void* x, void* col, int N, int H, int W, int C, int G, int KH, int KW, int stride,
                            int pad_y, int pad_x, int OH, int OW, int ldk, long long gstride, cudaStream_t st) {
  ConvGeom g{N, H, W, C, G, C / G, KH, KW, stride, pad_y, pad_x, OH, OW, ldk, gstride};
  const long long rows = (long long)N * OH * OW;
  if (rows <= 0) return 0;
  if (vec_ok(x, col, g)) {
    const long long total = rows * KH * KW * (g.Cg / 8) * G;
    im2col_kernel<8><<<ew_grid(total, 256, 16), 256, 0, st>>>((const bf16*)x, (bf16*)col, g);
  } else {
    const long long total = rows * KH * KW * g.Cg * G;
    im2col_kernel<1><<<ew_grid(total, 256, 16), 256, 0, st>>>((const bf16*)x, (bf16*)col, g);
  }
  CXN_CHECK_LAUNCH();
  return 0;
}

CXN_API int cxn_col2im_nhwc(const void* col, void* dx, int N, int H, int W, int C, int G, int KH, int KW, int stride,
                            int pad_y, int pad_x, int OH, int OW, int ldk, long long gstride, cudaStream_t st) {
  ConvGeom g{N, H, W, C, G, C / G, KH, KW, stride, pad_y, pad_x, OH, OW, ldk, gstride};
  if ((long long)N * H * W <= 0) return 0;
  if (vec_ok(col, dx, g)) {
    const long long total = (long long)N * H * W * G * (g.Cg / 8);
    col2im_kernel<8><<<ew_grid(total, 256, 16), 256, 0, st>>>((const bf16*)col, (bf16*)dx, g);
  } else {
    const long long total = (long long)N * H * W * C;
    col2im_kernel<1><<<ew_grid(total, 256, 16), 256, 0, st>>>((const bf16*)col, (bf16*)dx, g);
  }
  CXN_CHECK_LAUNCH();
  return 0;
}
