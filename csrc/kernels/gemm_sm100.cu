// tcgen05 / TMEM / TMA GEMM for sm_100a.
//
//   D[z] (M x N) (+)= A[z] (M x K) * B[z]^T (N x K)  (+ bias[z][n]) (relu)
//
// bf16 operands, fp32 accumulation in tensor memory.  Either operand may be
// K-major (row-major [MN][K]) or MN-major (row-major [K][MN]); this covers
//   fullc / conv forward     : A K-major,  B K-major        (cf. reference
//   fullc / conv backward-data: A K-major,  B MN-major        src/layer/fullc_layer-inl.hpp:108-129,
//   fullc / conv backward-wgt : A MN-major, B MN-major         src/layer/convolution_layer-inl.hpp:79-153)
// without any transposed copies.  The reference does these with cuBLAS SGEMM
// behind mshadow dot(); nothing here is derived from that code.
//
// Structure (persistent, warp specialised, one CTA per SM):
//   warp 0      : TMA producer  (cp.async.bulk.tensor.3d -> 128B-swizzled smem ring)
//   warp 1      : tcgen05.mma issuer (one elected lane), owns the TMEM allocation
//   warps 2..5  : epilogue, tcgen05.ld -> bias/relu/convert -> global
// Two accumulator stages in TMEM let the epilogue of tile i overlap the
// main loop of tile i+1.  Split-K tiles reduce with fp32 red.global.add.
#include "tc_common.cuh"
#include <mutex>
#include <unordered_map>
#include <string.h>

namespace cxn {
namespace gemm {

constexpr int BM = 128;
constexpr int BK = 64;           // 64 bf16 = 128 bytes = one swizzle span
constexpr int UMMA_K = 16;
constexpr int kThreads = 192;        // pair / cluster kernels: TMA warp, MMA warp, 4 epilogue warps
constexpr int kThreadsWide = 320;    // gemm_kernel: 8 epilogue warps (two per TMEM lane quadrant, half of the columns each)
constexpr int kAccStages = 2;

enum : int {
  kFlagRelu = 1,      // max(x,0) after bias
  kFlagOutBf16 = 2,   // D is bf16 (else fp32)
  kFlagAtomic = 4,    // fp32 D += partial via red.global.add (split-K or accumulate)
  kFlagBeta1 = 8,     // fp32 D = D + result (non atomic read-modify-write)
  kFlagUpdate = 16,   // the result is a weight gradient that is consumed on the spot: the epilogue applies the updater
                      // to the fp32 master / momentum at D's coordinates and refreshes the bf16 copy; D is not written
};

// Operands of the fused updater epilogue (kFlagUpdate).  w / m / m2 are addressed like D (row * ldd + column);
// hyper points at {lr, wd, momentum, clip, adam_lr_t, beta1, beta2} of the layer's entry in the per-step device
// table (loss_update.cu Segment, refreshed before the backward pass) - read at run time, so a captured step follows
// the schedules.  kind: 0 sgd, 1 nag, 2 adam, the arithmetic of update_kernel.
struct UpdArgs {
  float* w;
  float* m;
  float* m2;
  bf16* shadow;
  long long shadow_ld;
  const float* hyper;
  int kind;
};

struct Params {
  int M, N, K;
  int batch, splits, kb_per_split;
  void* D;
  long long ldd, strideD;
  const float* bias;
  long long strideBias;
  int flags;
  UpdArgs u;
};

struct UpdHyper {
  float lr, wd, momentum, clip, adam_lr_t, beta1, beta2;
};

__device__ __forceinline__ void update_one(float& w, float& m, float& m2, float g, const UpdHyper& h, int kind) {
  if (kind == 0) {
    if (h.clip != 0.0f) g = (g != g) ? 0.0f : fminf(fmaxf(g, -h.clip), h.clip);
    m = h.momentum * m - h.lr * (g + h.wd * w);
    w += m;
  } else if (kind == 1) {
    const float old = m;
    m = h.momentum * old - h.lr * (g + h.wd * w);
    w += (1.0f + h.momentum) * m - h.momentum * old;
  } else {
    if (h.wd > 0.0f) g -= h.wd * w;
    m += h.beta1 * (g - m);
    m2 += h.beta2 * (g * g - m2);
    w -= h.adam_lr_t * (m / (sqrtf(m2) + 1e-8f));
  }
}

// WIDE: the variant for short-K fp32 problems (the fully connected weight gradients: K = batch rows): a two / four
// stage ring leaves room for every epilogue warp to stage ALL of its columns of a tile (32 rows x BN / 2), so a
// result row leaves in one 512-byte (BN = 256) or 256-byte (BN = 128) piece instead of four / two 128-byte pieces
// written microseconds apart.
template <int BN, bool WIDE = false>
struct Cfg {
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = BN * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = WIDE ? (BN >= 256 ? 2 : 4) : ((BN >= 256) ? 4 : (BN >= 128 ? 6 : 8));
  static constexpr int kTmemCols = kAccStages * BN;  // 128 / 256 / 512
  // epilogue staging: per epilogue warp a 32 x 32 fp32 chunk (WIDE: 32 x BN / 2) whose 16-byte columns are
  // XOR-swizzled with the row (conflict-free 128-bit shared accesses both ways, no padding); 8 epilogue warps
  static constexpr int kWarpStageFloats = 32 * (WIDE ? BN / 2 : 32);
  static constexpr int kStageOutBytes = 8 * kWarpStageFloats * 4;
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align*/ + 256 /*barriers*/ + kStageOutBytes;
  static_assert(kSmemBytes <= 232448, "shared memory budget of one CTA");
};

using namespace cxn::tc;

// One 32-column chunk of an accumulator tile, held one row per lane (v[j] = column n0 + j of row row0 + lane):
// bias / ReLU / conversion and the store, reduction or read-modify-write into D.  `warp_stage` is this warp's
// 32 x 32 float staging area (fp32 results pass through a shared-memory transpose so that every store instruction
// covers complete 128-byte row segments).
template <bool UPD = false>
__device__ __forceinline__ void epilogue_chunk(const Params& p, uint32_t (&v)[32], int row0, int lane, int n0,
                                               bool has_work, const float* bias, int z, float* warp_stage) {
  if constexpr (UPD) {
    // fused updater: the accumulator IS the weight gradient of rows [row0, row0 + 32) x columns [n0, n0 + 32)
    UpdHyper h;
    h.lr = __ldg(p.u.hyper); h.wd = __ldg(p.u.hyper + 1); h.momentum = __ldg(p.u.hyper + 2);
    h.clip = __ldg(p.u.hyper + 3); h.adam_lr_t = __ldg(p.u.hyper + 4); h.beta1 = __ldg(p.u.hyper + 5);
    h.beta2 = __ldg(p.u.hyper + 6);
    const int kind = p.u.kind;
    if (n0 + 32 <= p.N && (p.ldd & 3) == 0 && (p.u.shadow_ld & 3) == 0) {
      float* sm = warp_stage;
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        float4 o4;
        o4.x = has_work ? __uint_as_float(v[4 * q]) : 0.0f;
        o4.y = has_work ? __uint_as_float(v[4 * q + 1]) : 0.0f;
        o4.z = has_work ? __uint_as_float(v[4 * q + 2]) : 0.0f;
        o4.w = has_work ? __uint_as_float(v[4 * q + 3]) : 0.0f;
        *reinterpret_cast<float4*>(sm + lane * 32 + 4 * (q ^ (lane & 7))) = o4;
      }
      __syncwarp();
      const int rsub = lane >> 3, cq = (lane & 7) * 4;
      const long long col = n0 + cq;
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        // all loads of four row segments first (w, m: 8 x 16 bytes in flight per thread), then the arithmetic
        float4 wv[4], mv[4], m2v[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int r = (half * 4 + i) * 4 + rsub;
          wv[i] = mv[i] = m2v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (row0 + r < p.M) {
            const long long off = (long long)(row0 + r) * p.ldd + col;
            wv[i] = *reinterpret_cast<const float4*>(p.u.w + off);
            mv[i] = *reinterpret_cast<const float4*>(p.u.m + off);
            if (kind == 2) m2v[i] = *reinterpret_cast<const float4*>(p.u.m2 + off);
          }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int r = (half * 4 + i) * 4 + rsub;
          if (row0 + r < p.M) {
            const float4 g4 = *reinterpret_cast<const float4*>(sm + r * 32 + 4 * ((lane & 7) ^ (r & 7)));
            update_one(wv[i].x, mv[i].x, m2v[i].x, g4.x, h, kind);
            update_one(wv[i].y, mv[i].y, m2v[i].y, g4.y, h, kind);
            update_one(wv[i].z, mv[i].z, m2v[i].z, g4.z, h, kind);
            update_one(wv[i].w, mv[i].w, m2v[i].w, g4.w, h, kind);
            const long long off = (long long)(row0 + r) * p.ldd + col;
            *reinterpret_cast<float4*>(p.u.w + off) = wv[i];
            *reinterpret_cast<float4*>(p.u.m + off) = mv[i];
            if (kind == 2) *reinterpret_cast<float4*>(p.u.m2 + off) = m2v[i];
            if (p.u.shadow != nullptr) {
              const __nv_bfloat162 h0 = __floats2bfloat162_rn(wv[i].x, wv[i].y);
              const __nv_bfloat162 h1 = __floats2bfloat162_rn(wv[i].z, wv[i].w);
              uint2 u2;
              u2.x = *reinterpret_cast<const uint32_t*>(&h0);
              u2.y = *reinterpret_cast<const uint32_t*>(&h1);
              *reinterpret_cast<uint2*>(p.u.shadow + (long long)(row0 + r) * p.u.shadow_ld + col) = u2;
            }
          }
        }
      }
      __syncwarp();
      return;
    }
    // ragged edge: one row per lane, element by element
    const int row = row0 + lane;
    if (row >= p.M) return;
#pragma unroll 1
    for (int j = 0; j < 32; ++j) {
      if (n0 + j >= p.N) break;
      const long long off = (long long)row * p.ldd + n0 + j;
      float w = p.u.w[off], m = p.u.m[off], m2 = (kind == 2) ? p.u.m2[off] : 0.0f;
      float g = 0.0f;
#pragma unroll
      for (int jj = 0; jj < 32; ++jj)
        if (jj == j) g = has_work ? __uint_as_float(v[jj]) : 0.0f;
      update_one(w, m, m2, g, h, kind);
      p.u.w[off] = w;
      p.u.m[off] = m;
      if (kind == 2) p.u.m2[off] = m2;
      if (p.u.shadow != nullptr) p.u.shadow[(long long)row * p.u.shadow_ld + n0 + j] = __float2bfloat16(w);
    }
    return;
  }
  const bool relu = p.flags & kFlagRelu;
  const bool out_bf16 = p.flags & kFlagOutBf16;
  const bool atomic = p.flags & kFlagAtomic;
  const bool beta1 = p.flags & kFlagBeta1;
  const int row = row0 + lane;
  const bool row_ok = row < p.M;
  const bool full = (n0 + 32 <= p.N);
  // fp32 results (weight gradients) go through a shared-memory transpose: a thread owns a row of the
  // accumulator, so direct stores put 32 rows x 16 bytes into every instruction; staged, each
  // instruction writes four complete 128-byte row segments.
  if (!out_bf16 && !atomic && full && bias == nullptr && !relu && (p.ldd & 3) == 0 &&
      ((reinterpret_cast<uintptr_t>(p.D) & 15) == 0) && (p.strideD & 3) == 0) {
    float* sm = warp_stage;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      float4 o4;
      o4.x = has_work ? __uint_as_float(v[4 * q]) : 0.0f;
      o4.y = has_work ? __uint_as_float(v[4 * q + 1]) : 0.0f;
      o4.z = has_work ? __uint_as_float(v[4 * q + 2]) : 0.0f;
      o4.w = has_work ? __uint_as_float(v[4 * q + 3]) : 0.0f;
      *reinterpret_cast<float4*>(sm + lane * 32 + 4 * (q ^ (lane & 7))) = o4;
    }
    __syncwarp();
    const int rsub = lane >> 3, cq = (lane & 7) * 4;
    float* dbase = reinterpret_cast<float*>(p.D) + (long long)z * p.strideD + n0 + cq;
    
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int r = it * 4 + rsub;
      if (row0 + r < p.M) {
        float4 o4 = *reinterpret_cast<const float4*>(sm + r * 32 + 4 * ((lane & 7) ^ (r & 7)));
        float4* dp = reinterpret_cast<float4*>(dbase + (long long)(row0 + r) * p.ldd);
        if (beta1) {
          const float4 c4 = *dp;
          o4.x += c4.x; o4.y += c4.y; o4.z += c4.z; o4.w += c4.w;
        }
        *dp = o4;
      }
    }
    __syncwarp();
    return;
  }
  if (!row_ok) return;
  float f[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) f[j] = has_work ? __uint_as_float(v[j]) : 0.0f;
  if (bias != nullptr) {
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (full || n0 + j < p.N) f[j] += __ldg(bias + n0 + j);
  }
  if (relu) {
#pragma unroll
    for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.0f);
  }
  const long long off = (long long)z * p.strideD + (long long)row * p.ldd + n0;
  if (out_bf16) {
    bf16* d = reinterpret_cast<bf16*>(p.D) + off;
    if (full && ((reinterpret_cast<uintptr_t>(d) & 15) == 0)) {
      uint4* d4 = reinterpret_cast<uint4*>(d);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        __nv_bfloat162 h0 = __floats2bfloat162_rn(f[q * 8 + 0], f[q * 8 + 1]);
        __nv_bfloat162 h1 = __floats2bfloat162_rn(f[q * 8 + 2], f[q * 8 + 3]);
        __nv_bfloat162 h2 = __floats2bfloat162_rn(f[q * 8 + 4], f[q * 8 + 5]);
        __nv_bfloat162 h3 = __floats2bfloat162_rn(f[q * 8 + 6], f[q * 8 + 7]);
        uint4 u;
        u.x = *reinterpret_cast<uint32_t*>(&h0);
        u.y = *reinterpret_cast<uint32_t*>(&h1);
        u.z = *reinterpret_cast<uint32_t*>(&h2);
        u.w = *reinterpret_cast<uint32_t*>(&h3);
        d4[q] = u;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (n0 + j < p.N) d[j] = __float2bfloat16(f[j]);
    }
  } else {
    float* d = reinterpret_cast<float*>(p.D) + off;
    if (atomic && full && ((reinterpret_cast<uintptr_t>(d) & 15) == 0)) {
      // 128-bit reductions: a quarter of the L2 atomic operations of the scalar form
#pragma unroll
      for (int q = 0; q < 8; ++q)
        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(d + 4 * q), "f"(f[4 * q]),
                     "f"(f[4 * q + 1]), "f"(f[4 * q + 2]), "f"(f[4 * q + 3])
                     : "memory");
    } else if (atomic) {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (full || n0 + j < p.N) atomicAdd(d + j, f[j]);
    } else if (full && ((reinterpret_cast<uintptr_t>(d) & 15) == 0)) {
      float4* d4 = reinterpret_cast<float4*>(d);
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        float4 o = make_float4(f[q * 4], f[q * 4 + 1], f[q * 4 + 2], f[q * 4 + 3]);
        if (beta1) {
          const float4 c = d4[q];
          o.x += c.x; o.y += c.y; o.z += c.z; o.w += c.w;
        }
        d4[q] = o;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (n0 + j < p.N) d[j] = beta1 ? d[j] + f[j] : f[j];
    }
  }
}

// WIDE epilogue of one warp: its 32 rows x W (= BN / 2) columns of the accumulator go through shared memory and leave
// row by row - one instruction writes (or, with the fused updater, reads and rewrites) W * 4 contiguous bytes of one
// or two result rows.  Returns false (nothing done) when the fast path does not apply; the caller then walks the
// 32-column chunks.  taddr: tensor-memory address of the warp's first column in its lane quadrant.
template <int W, bool UPD>
__device__ __forceinline__ bool epilogue_wide(const Params& p, uint32_t taddr, int row0, int lane, int n0w, bool has_work,
                                              int z, float* sm) {
  constexpr int C16 = W / 4;            // 16-byte columns per row: 32 or 16
  constexpr int RPI = 32 / C16;         // rows per instruction: 1 or 2
  constexpr int ITERS = 32 / RPI;
  if (n0w + W > p.N || (p.ldd & 3) != 0 || (p.strideD & 3) != 0 || (reinterpret_cast<uintptr_t>(p.D) & 15) != 0 ||
      (p.flags & (kFlagRelu | kFlagOutBf16 | kFlagAtomic)) != 0 || p.bias != nullptr)
    return false;
  if constexpr (UPD) {
    // (adam keeps a third array: it takes the 32-column chunks)
    if ((p.u.shadow_ld & 3) != 0 || p.u.kind == 2) return false;
  }
#pragma unroll
  for (int c = 0; c < W / 32; ++c) {
    uint32_t v[32];
    tmem_ld32(taddr + (uint32_t)(c * 32), v);
    tmem_ld_wait();
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      float4 o4;
      o4.x = has_work ? __uint_as_float(v[4 * q]) : 0.0f;
      o4.y = has_work ? __uint_as_float(v[4 * q + 1]) : 0.0f;
      o4.z = has_work ? __uint_as_float(v[4 * q + 2]) : 0.0f;
      o4.w = has_work ? __uint_as_float(v[4 * q + 3]) : 0.0f;
      *reinterpret_cast<float4*>(sm + lane * W + 4 * ((c * 8 + q) ^ (lane & 7))) = o4;
    }
  }
  __syncwarp();
  const int rsub = lane / C16, c16 = lane % C16;
  const long long col = n0w + 4 * c16;
  if constexpr (!UPD) {
    const bool beta1 = p.flags & kFlagBeta1;
    float* dbase = reinterpret_cast<float*>(p.D) + (long long)z * p.strideD + col;
#pragma unroll 8
    for (int it = 0; it < ITERS; ++it) {
      const int r = it * RPI + rsub;
      if (row0 + r < p.M) {
        float4 o4 = *reinterpret_cast<const float4*>(sm + r * W + 4 * (c16 ^ (r & 7)));
        float4* dp = reinterpret_cast<float4*>(dbase + (long long)(row0 + r) * p.ldd);
        if (beta1) {
          const float4 c4 = *dp;
          o4.x += c4.x; o4.y += c4.y; o4.z += c4.z; o4.w += c4.w;
        }
        *dp = o4;
      }
    }
  } else {
    UpdHyper h;
    h.lr = __ldg(p.u.hyper); h.wd = __ldg(p.u.hyper + 1); h.momentum = __ldg(p.u.hyper + 2);
    h.clip = __ldg(p.u.hyper + 3); h.adam_lr_t = __ldg(p.u.hyper + 4); h.beta1 = __ldg(p.u.hyper + 5);
    h.beta2 = __ldg(p.u.hyper + 6);
    const int kind = p.u.kind;
    // groups of G result rows: all (w, m) loads of a group are issued before its arithmetic - 2 x G 16-byte loads
    // in flight per thread.  Measured on fc6 (37.7 M parameters): G = 8 145 us; a double-buffered G = 4 pipeline 167 us;
    // the same with streaming (evict-first) hints 172 us - what counts is the number of loads in flight.
    constexpr int G = (ITERS < 16) ? ITERS : 16;
#pragma unroll 1
    for (int g0 = 0; g0 < ITERS; g0 += G) {
      float4 wv[G], mv[G];
#pragma unroll
      for (int i = 0; i < G; ++i) {
        const int r = (g0 + i) * RPI + rsub;
        wv[i] = mv[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row0 + r < p.M) {
          const long long off = (long long)(row0 + r) * p.ldd + col;
          wv[i] = *reinterpret_cast<const float4*>(p.u.w + off);
          mv[i] = *reinterpret_cast<const float4*>(p.u.m + off);
        }
      }
#pragma unroll
      for (int i = 0; i < G; ++i) {
        const int r = (g0 + i) * RPI + rsub;
        if (row0 + r < p.M) {
          const float4 g4 = *reinterpret_cast<const float4*>(sm + r * W + 4 * (c16 ^ (r & 7)));
          float unused = 0.0f;
          update_one(wv[i].x, mv[i].x, unused, g4.x, h, kind);
          update_one(wv[i].y, mv[i].y, unused, g4.y, h, kind);
          update_one(wv[i].z, mv[i].z, unused, g4.z, h, kind);
          update_one(wv[i].w, mv[i].w, unused, g4.w, h, kind);
          const long long off = (long long)(row0 + r) * p.ldd + col;
          *reinterpret_cast<float4*>(p.u.w + off) = wv[i];
          *reinterpret_cast<float4*>(p.u.m + off) = mv[i];
          if (p.u.shadow != nullptr) {
            const __nv_bfloat162 h0 = __floats2bfloat162_rn(wv[i].x, wv[i].y);
            const __nv_bfloat162 h1 = __floats2bfloat162_rn(wv[i].z, wv[i].w);
            uint2 u2;
            u2.x = *reinterpret_cast<const uint32_t*>(&h0);
            u2.y = *reinterpret_cast<const uint32_t*>(&h1);
            *reinterpret_cast<uint2*>(p.u.shadow + (long long)(row0 + r) * p.u.shadow_ld + col) = u2;
          }
        }
      }
    }
  }
  __syncwarp();
  return true;
}

template <int BN, bool A_MN, bool B_MN, bool UPD = false, bool WIDE = false>
__global__ void __launch_bounds__(kThreadsWide, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const Params p) {
  using C = Cfg<BN, WIDE>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t bar_base = smem_base + C::kStages * C::kStageBytes;
  // barrier layout (8 bytes each): full[S], empty[S], tfull[2], tempty[2], then tmem ptr
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (C::kStages + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * C::kStages + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * C::kStages + kAccStages + a); };
  volatile uint32_t* tmem_ptr_smem =
      reinterpret_cast<volatile uint32_t*>(smem_gen + C::kStages * C::kStageBytes + 8 * (2 * C::kStages + 2 * kAccStages));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  float* stage_out = reinterpret_cast<float*>(smem_gen + C::kStages * C::kStageBytes + 256);

  const int tiles_m = ceil_div(p.M, BM);
  const int tiles_n = ceil_div(p.N, BN);
  const int num_kb = ceil_div(p.K, BK);
  const long long total_tiles = (long long)tiles_m * tiles_n * p.batch * p.splits;

  if (threadIdx.x == 0) {
    for (int s = 0; s < C::kStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < kAccStages; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), 8);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32((const void*)tmem_ptr_smem)),
                 "n"(C::kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    // =========================== TMA producer ===========================
    // the whole warp walks the schedule, one elected lane issues (see elect_one in tc_common.cuh)
    int stage = 0;
    uint32_t phase = 0;
    for (long long t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      long long r = t;
      const int n_blk = (int)(r % tiles_n); r /= tiles_n;
      const int m_blk = (int)(r % tiles_m); r /= tiles_m;
      const int split = (int)(r % p.splits);
      const int z = (int)(r / p.splits);
      const int kb0 = split * p.kb_per_split;
      const int kb1 = min(num_kb, kb0 + p.kb_per_split);
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(empty_bar(stage), phase ^ 1u);
        const uint32_t sa = smem_base + stage * C::kStageBytes;
        const uint32_t sb = sa + C::kABytes;
        if (elect_one()) {
          mbar_expect_tx(full_bar(stage), C::kStageBytes);
          if (!A_MN) {
            tma_load_3d(sa, &tmA, full_bar(stage), kb * BK, m_blk * BM, z);
          } else {
#pragma unroll
            for (int h = 0; h < BM / 64; ++h)
              tma_load_3d(sa + h * (64 * BK * 2), &tmA, full_bar(stage), m_blk * BM + h * 64, kb * BK, z);
          }
          if (!B_MN) {
            tma_load_3d(sb, &tmB, full_bar(stage), kb * BK, n_blk * BN, z);
          } else {
#pragma unroll
            for (int h = 0; h < BN / 64; ++h)
              tma_load_3d(sb + h * (64 * BK * 2), &tmB, full_bar(stage), n_blk * BN + h * 64, kb * BK, z);
          }
        }
        __syncwarp();
        if (++stage == C::kStages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer =============================
    // warp-uniform loop, one elected lane issues the MMAs and their commits
    // instruction descriptor: D=f32, A=B=bf16, majors, N, M
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((A_MN ? 1u : 0u) << 15) |
                           ((B_MN ? 1u : 0u) << 16) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    constexpr uint32_t kLboMN = 64 * BK * 2;  // next 64-wide MN atom (one TMA box)
    constexpr uint32_t kSbo = 1024;           // 8 rows x 128 bytes
    constexpr uint32_t kAStep = A_MN ? (UMMA_K * 128) : (UMMA_K * 2);
    constexpr uint32_t kBStep = B_MN ? (UMMA_K * 128) : (UMMA_K * 2);
    // descriptor = constant high part | (address >> 4): stepping along K only adds to the low word
    const uint64_t adesc_hi = make_desc(0, A_MN ? kLboMN : 0u, kSbo);
    const uint64_t bdesc_hi = make_desc(0, B_MN ? kLboMN : 0u, kSbo);
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (long long t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      long long r = t / ((long long)tiles_n * tiles_m);
      const int split = (int)(r % p.splits);
      const int kb0 = split * p.kb_per_split;
      const int kb1 = min(num_kb, kb0 + p.kb_per_split);
      mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(full_bar(stage), phase);
        tc_fence_after();
        const uint32_t sa = smem_base + stage * C::kStageBytes;
        const uint32_t sb = sa + C::kABytes;
        if (elect_one()) {
          const uint64_t a0 = adesc_hi | (uint64_t)((sa >> 4) & 0x3FFF), b0 = bdesc_hi | (uint64_t)((sb >> 4) & 0x3FFF);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k)
            umma_bf16(d_tmem, a0 + (uint64_t)((k * kAStep) >> 4), b0 + (uint64_t)((k * kBStep) >> 4), idesc,
                      (kb > kb0 || k > 0) ? 1u : 0u);
          umma_commit(empty_bar(stage));                   // frees the smem slot when these MMAs retire
          if (kb == kb1 - 1) umma_commit(tfull_bar(acc));  // accumulator complete -> epilogue
        }
        __syncwarp();
        if (++stage == C::kStages) { stage = 0; phase ^= 1u; }
      }
      if (++acc == kAccStages) { acc = 0; acc_phase ^= 1u; }
    }
  } else {
    // =========================== epilogue ===============================
    // eight warps: warp w reads TMEM lane quadrant w & 3 (a hardware rule) and takes the lower or upper half of the
    // tile's columns - the accumulator drains twice as fast as with one warp per quadrant
    const int quad = warp & 3;
    const int chalf = (warp - 2) >> 2;
    constexpr int kColsPerWarp = BN / 2;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (long long t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      long long r = t;
      const int n_blk = (int)(r % tiles_n); r /= tiles_n;
      const int m_blk = (int)(r % tiles_m); r /= tiles_m;
      const int split = (int)(r % p.splits);
      const int z = (int)(r / p.splits);
      const int kb0 = split * p.kb_per_split;
      const int kb1 = min(num_kb, kb0 + p.kb_per_split);
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const bool has_work = kb1 > kb0;
      const float* bias = (p.bias != nullptr && split == 0) ? p.bias + (long long)z * p.strideBias : nullptr;
      bool done = false;
      if constexpr (WIDE) {
        const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * BN + chalf * kColsPerWarp);
        done = epilogue_wide<kColsPerWarp, UPD>(p, taddr, m_blk * BM + quad * 32, lane, n_blk * BN + chalf * kColsPerWarp,
                                                has_work, z, stage_out + (warp - 2) * C::kWarpStageFloats);
      }
      if (!done) {
#pragma unroll 1
        for (int c0 = chalf * kColsPerWarp; c0 < (chalf + 1) * kColsPerWarp; c0 += 32) {
          const int n0 = n_blk * BN + c0;
          if (n0 >= p.N) break;  // warp-uniform
          uint32_t v[32];
          const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * BN + c0);
          tmem_ld32(taddr, v);
          tmem_ld_wait();
          epilogue_chunk<UPD>(p, v, m_blk * BM + quad * 32, lane, n0, has_work, bias, z,
                              stage_out + (warp - 2) * C::kWarpStageFloats);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar(acc));
      if (++acc == kAccStages) { acc = 0; acc_phase ^= 1u; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(C::kTmemCols)
                 : "memory");
  }
}

// ===================================================================================================
// CTA-pair variant (tcgen05 cta_group::2) for large problems: the two CTAs of a cluster - the two SMs of one TPC -
// compute one 256 x 256 tile with ONE stream of M = 256 MMAs.  Each CTA loads its own 128 rows of A and HALF of the
// B tile (128 of the 256 columns) and receives its 128 rows of the accumulator in its own tensor memory: per SM and
// k-block 32 KB arrive from L2 for 128 x 256 outputs where the single-CTA 128 x 256 tile needs 48 KB.  The large
// layers sit on the chip-wide L2 -> SM throughput ceiling (profiles/), so bytes per FLOP is what this buys.
//
// Protocol (CTA 0 of the pair = leader):
//   full[s]   lives in the leader.  Its producer arms it with the bytes of BOTH CTAs' loads; every load of either CTA
//             (cp.async.bulk.tensor ... .cta_group::2, barrier address with the peer bit cleared) reports there.
//   empty[s]  one per CTA, count 1: the leader's tcgen05.commit.cta_group::2 ... multicast arrives on both.
//   tfull[a]  one per CTA, count 1: the same multicast commit after a tile's last MMA; each CTA's epilogue warps read
//             their own tensor memory.
//   tempty[a] in the leader, count 8: four epilogue warps of each CTA arrive (the peer's through shared::cluster).
// Tensor memory is allocated and freed by warp 1 of both CTAs (cta_group::2); cluster barriers bracket the kernel
// so that neither CTA's shared memory, barriers or tensor memory is touched after it has gone.
template <bool A_MN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
gemm_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const Params p) {
  constexpr int BN = 256;            // tile width of the pair; each CTA stages BN / 2 columns of B
  constexpr int BNH = BN / 2;
  constexpr int BM2 = 2 * BM;
  using C = Cfg<BNH>;                // per-CTA ring: 128 x 64 of A + 128 x 64 of B per stage
  constexpr int kTmemCols = kAccStages * BN;     // 512: two 256-column accumulators
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t bar_base = smem_base + C::kStages * C::kStageBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (C::kStages + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * C::kStages + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * C::kStages + kAccStages + a); };
  volatile uint32_t* tmem_ptr_smem =
      reinterpret_cast<volatile uint32_t*>(smem_gen + C::kStages * C::kStageBytes + 8 * (2 * C::kStages + 2 * kAccStages));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t crank = cluster_ctarank();
  const bool leader = crank == 0;
  float* stage_out = reinterpret_cast<float*>(smem_gen + C::kStages * C::kStageBytes + 256);

  const int tiles_m = ceil_div(p.M, BM2);
  const int tiles_n = ceil_div(p.N, BN);
  const int num_kb = ceil_div(p.K, BK);
  const long long total_tiles = (long long)tiles_m * tiles_n * p.batch * p.splits;
  const long long pair0 = blockIdx.x >> 1, npairs = gridDim.x >> 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < C::kStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < kAccStages; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), 8);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32((const void*)tmem_ptr_smem)),
                 "n"(kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  cluster_sync_all();              // both CTAs' barriers exist before anything arrives on them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    // =========================== TMA producer (both CTAs) ===========================
    int stage = 0;
    uint32_t phase = 0;
    for (long long t = pair0; t < total_tiles; t += npairs) {
      long long r = t;
      const int n_blk = (int)(r % tiles_n); r /= tiles_n;
      const int m_blk = (int)(r % tiles_m); r /= tiles_m;
      const int split = (int)(r % p.splits);
      const int z = (int)(r / p.splits);
      const int kb0 = split * p.kb_per_split;
      const int kb1 = min(num_kb, kb0 + p.kb_per_split);
      const int m0 = m_blk * BM2 + (int)crank * BM;
      const int n0 = n_blk * BN + (int)crank * BNH;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(empty_bar(stage), phase ^ 1u);
        const uint32_t sa = smem_base + stage * C::kStageBytes;
        const uint32_t sb = sa + C::kABytes;
        if (elect_one()) {
          if (leader) mbar_expect_tx(full_bar(stage), 2 * C::kStageBytes);
          if (!A_MN) {
            tma_load_3d_2cta(sa, &tmA, full_bar(stage), kb * BK, m0, z);
          } else {
#pragma unroll
            for (int h = 0; h < BM / 64; ++h)
              tma_load_3d_2cta(sa + h * (64 * BK * 2), &tmA, full_bar(stage), m0 + h * 64, kb * BK, z);
          }
          if (!B_MN) {
            tma_load_3d_2cta(sb, &tmB, full_bar(stage), kb * BK, n0, z);
          } else {
#pragma unroll
            for (int h = 0; h < BNH / 64; ++h)
              tma_load_3d_2cta(sb + h * (64 * BK * 2), &tmB, full_bar(stage), n0 + h * 64, kb * BK, z);
          }
        }
        __syncwarp();
        if (++stage == C::kStages) { stage = 0; phase ^= 1u; }
      }
    }
    // tail: every stage this CTA filled has been consumed (the multicast commits have arrived here) before the
    // CTA may leave - a commit must not find this CTA's barriers gone
    for (int i = 0; i < C::kStages; ++i) {
      mbar_wait(empty_bar(stage), phase ^ 1u);
      if (++stage == C::kStages) { stage = 0; phase ^= 1u; }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer (leader only) =============================
    if (leader) {
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((A_MN ? 1u : 0u) << 15) |
                             ((B_MN ? 1u : 0u) << 16) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM2 >> 4) << 24);
      constexpr uint32_t kLboMN = 64 * BK * 2;
      constexpr uint32_t kSbo = 1024;
      constexpr uint32_t kAStep = A_MN ? (UMMA_K * 128) : (UMMA_K * 2);
      constexpr uint32_t kBStep = B_MN ? (UMMA_K * 128) : (UMMA_K * 2);
      const uint64_t adesc_hi = make_desc(0, A_MN ? kLboMN : 0u, kSbo);
      const uint64_t bdesc_hi = make_desc(0, B_MN ? kLboMN : 0u, kSbo);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (long long t = pair0; t < total_tiles; t += npairs) {
        long long r = t / ((long long)tiles_n * tiles_m);
        const int split = (int)(r % p.splits);
        const int kb0 = split * p.kb_per_split;
        const int kb1 = min(num_kb, kb0 + p.kb_per_split);
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sa = smem_base + stage * C::kStageBytes;
          const uint32_t sb = sa + C::kABytes;
          if (elect_one()) {
            const uint64_t a0 = adesc_hi | (uint64_t)((sa >> 4) & 0x3FFF), b0 = bdesc_hi | (uint64_t)((sb >> 4) & 0x3FFF);
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k)
              umma_bf16_2cta(d_tmem, a0 + (uint64_t)((k * kAStep) >> 4), b0 + (uint64_t)((k * kBStep) >> 4), idesc,
                             (kb > kb0 || k > 0) ? 1u : 0u);
            umma_commit_2cta(empty_bar(stage), 3);                   // frees the slot in both CTAs
            if (kb == kb1 - 1) umma_commit_2cta(tfull_bar(acc), 3);  // accumulator halves complete in both CTAs
          }
          __syncwarp();
          if (++stage == C::kStages) { stage = 0; phase ^= 1u; }
        }
        if (kb1 <= kb0 && elect_one()) umma_commit_2cta(tfull_bar(acc), 3);   // empty split: nothing to wait for
        __syncwarp();
        if (++acc == kAccStages) { acc = 0; acc_phase ^= 1u; }
      }
    }
  } else {
    // =========================== epilogue (both CTAs, own accumulator rows) ===============================
    const int quad = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (long long t = pair0; t < total_tiles; t += npairs) {
      long long r = t;
      const int n_blk = (int)(r % tiles_n); r /= tiles_n;
      const int m_blk = (int)(r % tiles_m); r /= tiles_m;
      const int split = (int)(r % p.splits);
      const int z = (int)(r / p.splits);
      const int kb0 = split * p.kb_per_split;
      const int kb1 = min(num_kb, kb0 + p.kb_per_split);
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const bool has_work = kb1 > kb0;
      const float* bias = (p.bias != nullptr && split == 0) ? p.bias + (long long)z * p.strideBias : nullptr;
      const int row0 = m_blk * BM2 + (int)crank * BM + quad * 32;
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 32) {
        const int n0 = n_blk * BN + c0;
        if (n0 >= p.N) break;  // warp-uniform
        uint32_t v[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * BN + c0);
        tmem_ld32(taddr, v);
        tmem_ld_wait();
        if (row0 < p.M) epilogue_chunk(p, v, row0, lane, n0, has_work, bias, z, stage_out + (warp - 2) * (32 * 32));
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_leader(tempty_bar(acc));
      if (++acc == kAccStages) { acc = 0; acc_phase ^= 1u; }
    }
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kTmemCols) : "memory");
  }
}

// ===================================================================================================
// Cluster variant for skinny K-major-A problems (opt-in: CXN_GEMM_CLUSTER=2|4).  The fully connected forward and
// backward-data GEMMs have M = batch rows and thousands of output columns: with 128 x BN tiles every column tile
// re-reads the same activation tile (fc6 forward fetches its 4.7 MB activation matrix 64 times from L2).  Here the CL
// CTAs of a cluster take CL neighbouring column tiles of one row tile; each fetches 128 / CL rows of the activation
// tile and multicasts them into every CTA's ring, and fetches its own weight tile.  Same barrier protocol as
// conv_fprop_cluster_kernel (full = own loads + the peers' slices, empty = multicast tcgen05.commit of all CL MMA
// warps, producer tail + cluster barrier before exit), which has run on hardware.  No split-K, one batch, no atomics.
template <int BN, bool B_MN, int CL>
__global__ void __cluster_dims__(CL, 1, 1) __launch_bounds__(kThreads, 1)
gemm_cluster_kernel(const __grid_constant__ CUtensorMap tmASlice, const __grid_constant__ CUtensorMap tmB, const Params p) {
  using C = Cfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t bar_base = smem_base + C::kStages * C::kStageBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (C::kStages + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * C::kStages + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * C::kStages + kAccStages + a); };
  volatile uint32_t* tmem_ptr_smem =
      reinterpret_cast<volatile uint32_t*>(smem_gen + C::kStages * C::kStageBytes + 8 * (2 * C::kStages + 2 * kAccStages));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t crank = cluster_ctarank();
  constexpr uint16_t kMask = (uint16_t)((1u << CL) - 1u);
  constexpr int kSliceRows = BM / CL;

  const int tiles_m = ceil_div(p.M, BM);
  const int tiles_n = ceil_div(p.N, BN);
  const int groups_n = ceil_div(tiles_n, CL);
  const int num_kb = ceil_div(p.K, BK);
  const long long total_groups = (long long)tiles_m * groups_n;
  const long long cluster_id = blockIdx.x / CL, num_clusters = gridDim.x / CL;

  if (threadIdx.x == 0) {
    for (int s = 0; s < C::kStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), CL);
    }
    for (int a = 0; a < kAccStages; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), 4);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmASlice) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32((const void*)tmem_ptr_smem)),
                 "n"(C::kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  auto decode = [&](long long t, int& m_blk, int& n_blk) {
    const int ng = (int)(t % groups_n);
    m_blk = (int)(t / groups_n);
    n_blk = ng * CL + (int)crank;            // may be >= tiles_n in the last group: computed, not stored
  };

  if (warp == 0) {
    // =========================== TMA producer ===========================
    int stage = 0;
    uint32_t phase = 0;
    for (long long t = cluster_id; t < total_groups; t += num_clusters) {
      int m_blk, n_blk;
      decode(t, m_blk, n_blk);
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait_warp(empty_bar(stage), phase ^ 1u);
        const uint32_t sa = smem_base + stage * C::kStageBytes;
        const uint32_t sb = sa + C::kABytes;
        if (elect_one()) {
          mbar_expect_tx(full_bar(stage), C::kStageBytes);
          // this CTA's slice of the activation tile, into every CTA of the cluster
          tma_load_3d_mc(sa + crank * (uint32_t)(kSliceRows * BK * 2), &tmASlice, full_bar(stage), kb * BK,
                         m_blk * BM + (int)crank * kSliceRows, 0, kMask);
          if (!B_MN) {
            tma_load_3d(sb, &tmB, full_bar(stage), kb * BK, n_blk * BN, 0);
          } else {
#pragma unroll
            for (int h = 0; h < BN / 64; ++h)
              tma_load_3d(sb + h * (64 * BK * 2), &tmB, full_bar(stage), n_blk * BN + h * 64, kb * BK, 0);
          }
        }
        __syncwarp();
        if (++stage == C::kStages) { stage = 0; phase ^= 1u; }
      }
    }
    for (int i = 0; i < C::kStages; ++i) {        // tail: every stage released by all MMA warps of the cluster
      mbar_wait_warp(empty_bar(stage), phase ^ 1u);
      if (++stage == C::kStages) { stage = 0; phase ^= 1u; }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer =============================
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((B_MN ? 1u : 0u) << 16) | ((uint32_t)(BN >> 3) << 17) |
                           ((uint32_t)(BM >> 4) << 24);
    constexpr uint32_t kLboMN = 64 * BK * 2;
    constexpr uint32_t kSbo = 1024;
    constexpr uint32_t kAStep = UMMA_K * 2;
    constexpr uint32_t kBStep = B_MN ? (UMMA_K * 128) : (UMMA_K * 2);
    const uint64_t adesc_hi = make_desc(0, 0u, kSbo);
    const uint64_t bdesc_hi = make_desc(0, B_MN ? kLboMN : 0u, kSbo);
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (long long t = cluster_id; t < total_groups; t += num_clusters) {
      mbar_wait_warp(tempty_bar(acc), acc_phase ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait_warp(full_bar(stage), phase);
        tc_fence_after();
        const uint32_t sa = smem_base + stage * C::kStageBytes;
        const uint32_t sb = sa + C::kABytes;
        if (elect_one()) {
          const uint64_t a0 = adesc_hi | (uint64_t)((sa >> 4) & 0x3FFF), b0 = bdesc_hi | (uint64_t)((sb >> 4) & 0x3FFF);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k)
            umma_bf16(d_tmem, a0 + (uint64_t)((k * kAStep) >> 4), b0 + (uint64_t)((k * kBStep) >> 4), idesc,
                      (kb > 0 || k > 0) ? 1u : 0u);
          umma_commit_mc(empty_bar(stage), kMask);          // frees the slot in every CTA of the cluster
          if (kb == num_kb - 1) umma_commit(tfull_bar(acc));
        }
        __syncwarp();
        if (++stage == C::kStages) { stage = 0; phase ^= 1u; }
      }
      if (++acc == kAccStages) { acc = 0; acc_phase ^= 1u; }
    }
  } else {
    // =========================== epilogue ===============================
    const int quad = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    const bool relu = p.flags & kFlagRelu;
    const bool out_bf16 = p.flags & kFlagOutBf16;
    for (long long t = cluster_id; t < total_groups; t += num_clusters) {
      int m_blk, n_blk;
      decode(t, m_blk, n_blk);
      mbar_wait_warp(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const int row = m_blk * BM + quad * 32 + lane;
      const bool row_ok = row < p.M;
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 32) {
        const int n0 = n_blk * BN + c0;
        if (n0 >= p.N) break;  // warp-uniform (covers the tiles past the last column tile)
        uint32_t v[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * BN + c0);
        tmem_ld32(taddr, v);
        tmem_ld_wait();
        if (!row_ok) continue;
        const bool full = (n0 + 32 <= p.N);
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
        if (p.bias != nullptr) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (full || n0 + j < p.N) f[j] += __ldg(p.bias + n0 + j);
        }
        if (relu) {
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.0f);
        }
        const long long off = (long long)row * p.ldd + n0;
        if (out_bf16) {
          bf16* d = reinterpret_cast<bf16*>(p.D) + off;
          if (full && ((reinterpret_cast<uintptr_t>(d) & 15) == 0)) {
            uint4* d4 = reinterpret_cast<uint4*>(d);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              __nv_bfloat162 h0 = __floats2bfloat162_rn(f[q * 8 + 0], f[q * 8 + 1]);
              __nv_bfloat162 h1 = __floats2bfloat162_rn(f[q * 8 + 2], f[q * 8 + 3]);
              __nv_bfloat162 h2 = __floats2bfloat162_rn(f[q * 8 + 4], f[q * 8 + 5]);
              __nv_bfloat162 h3 = __floats2bfloat162_rn(f[q * 8 + 6], f[q * 8 + 7]);
              uint4 u;
              u.x = *reinterpret_cast<uint32_t*>(&h0);
              u.y = *reinterpret_cast<uint32_t*>(&h1);
              u.z = *reinterpret_cast<uint32_t*>(&h2);
              u.w = *reinterpret_cast<uint32_t*>(&h3);
              d4[q] = u;
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (n0 + j < p.N) d[j] = __float2bfloat16(f[j]);
          }
        } else {
          float* d = reinterpret_cast<float*>(p.D) + off;
          if (full && ((reinterpret_cast<uintptr_t>(d) & 15) == 0)) {
            float4* d4 = reinterpret_cast<float4*>(d);
#pragma unroll
            for (int q = 0; q < 8; ++q) d4[q] = make_float4(f[q * 4], f[q * 4 + 1], f[q * 4 + 2], f[q * 4 + 3]);
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (n0 + j < p.N) d[j] = f[j];
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar(acc));
      if (++acc == kAccStages) { acc = 0; acc_phase ^= 1u; }
    }
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(C::kTmemCols)
                 : "memory");
  }
}

// Second half of an internally split GEMM: D = act(ws + bias) (+ D), ws (fp32 [M][ldw]) is left zeroed
// for the next call.
__global__ void splitk_finish_kernel(float* __restrict__ ws, int ldw, void* __restrict__ D, long long ldd,
                                     const float* __restrict__ bias, int M, int N, int flags) {
  const int nq = ldw >> 2;
  const long long total = (long long)M * nq;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int row = (int)(i / nq), c = (int)(i - (long long)row * nq) << 2;
    float4* w4 = reinterpret_cast<float4*>(ws + (long long)row * ldw + c);
    const float4 v = *w4;
    *w4 = make_float4(0.f, 0.f, 0.f, 0.f);
    float f[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (c + j >= N) continue;
      if (bias != nullptr) f[j] += __ldg(bias + c + j);
      if (flags & kFlagRelu) f[j] = fmaxf(f[j], 0.0f);
      const long long o = (long long)row * ldd + c + j;
      if (flags & kFlagOutBf16) {
        reinterpret_cast<bf16*>(D)[o] = __float2bfloat16(f[j]);
      } else {
        float* d = reinterpret_cast<float*>(D) + o;
        *d = (flags & kFlagBeta1) ? *d + f[j] : f[j];
      }
    }
  }
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

struct MapKey {
  const void* ptr;
  long long ld, bstride;
  int rows, cols, batch, mn_major, box_rows;
  bool operator==(const MapKey& o) const { return memcmp(this, &o, sizeof(MapKey)) == 0; }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    const uint64_t* w = reinterpret_cast<const uint64_t*>(&k);
    size_t h = 1469598103934665603ull;
    for (size_t i = 0; i < sizeof(MapKey) / 8; ++i) h = (h ^ w[i]) * 1099511628211ull;
    return h;
  }
};

// Tensor map for one operand.  `mn` = extent along M or N, `k` = reduction extent.
//   K-major : memory [batch][mn][ld>=k]   box {64 (k), box_rows (mn)}
//   MN-major: memory [batch][k][ld>=mn]   box {64 (mn), 64 (k)}
static int make_map(CUtensorMap* out, const void* ptr, int mn, int k, long long ld, int batch, long long bstride,
                    bool mn_major, int box_rows) {
  static std::mutex mu;
  static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
  MapKey key;
  memset(&key, 0, sizeof(key));
  key.ptr = ptr; key.ld = ld; key.bstride = bstride; key.rows = mn; key.cols = k; key.batch = batch;
  key.mn_major = mn_major ? 1 : 0; key.box_rows = box_rows;
  {
    std::lock_guard<std::mutex> g(mu);
    auto it = cache.find(key);
    if (it != cache.end()) { *out = it->second; return 0; }
  }
  EncodeTiledFn enc = get_encode();
  if (!enc) { fprintf(stderr, "[cxxnet_b200] cuTensorMapEncodeTiled unavailable\n"); return -2; }
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) || (ld % 8) || (batch > 1 && (bstride % 8))) {
    fprintf(stderr, "[cxxnet_b200] gemm operand not 16B aligned (ptr=%p ld=%lld bstride=%lld)\n", ptr, ld, bstride);
    return -3;
  }
  cuuint64_t gdim[3], gstr[2];
  cuuint32_t box[3], estr[3] = {1, 1, 1};
  if (!mn_major) {
    gdim[0] = (cuuint64_t)k; gdim[1] = (cuuint64_t)mn;
    box[0] = BK; box[1] = (cuuint32_t)box_rows;
  } else {
    gdim[0] = (cuuint64_t)mn; gdim[1] = (cuuint64_t)k;
    box[0] = 64; box[1] = BK;
  }
  gdim[2] = (cuuint64_t)batch;
  box[2] = 1;
  gstr[0] = (cuuint64_t)ld * 2;
  gstr[1] = (cuuint64_t)(batch > 1 ? bstride : (long long)ld * (long long)gdim[1]) * 2;
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    fprintf(stderr, "[cxxnet_b200] cuTensorMapEncodeTiled failed: %d (mn=%d k=%d ld=%lld batch=%d)\n", (int)r, mn, k,
            ld, batch);
    return -4;
  }
  {
    std::lock_guard<std::mutex> g(mu);
    if (cache.size() > 4096) cache.clear();
    cache[key] = *out;
  }
  return 0;
}

template <int BN, bool A_MN, bool B_MN, bool UPD = false, bool WIDE = false>
static int launch(const CUtensorMap& ta, const CUtensorMap& tb, const Params& p, int grid, cudaStream_t st) {
  using C = Cfg<BN, WIDE>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_kernel<BN, A_MN, B_MN, UPD, WIDE>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes);
    if (e != cudaSuccess) return (int)e;
    attr_set = true;
  }
  gemm_kernel<BN, A_MN, B_MN, UPD, WIDE><<<grid, kThreadsWide, C::kSmemBytes, st>>>(ta, tb, p);
  CXN_CHECK_LAUNCH();
  return 0;
}

template <bool A_MN, bool B_MN>
static int launch_pair(const CUtensorMap& ta, const CUtensorMap& tb, const Params& p, int grid, cudaStream_t st) {
  using C = Cfg<128>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_pair_kernel<A_MN, B_MN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         C::kSmemBytes);
    if (e != cudaSuccess) return (int)e;
    attr_set = true;
  }
  gemm_pair_kernel<A_MN, B_MN><<<grid, kThreads, C::kSmemBytes, st>>>(ta, tb, p);
  CXN_CHECK_LAUNCH();
  return 0;
}
static int g_gemm_pair = -1;       // CTA pairs (cta_group::2, 256 x 256 tiles): 1 auto (default), 0 off, 2 whenever legal
static int gemm_pair_mode() {
  if (g_gemm_pair < 0) {
    const char* e = getenv("CXN_GEMM_PAIR");
    g_gemm_pair = e ? atoi(e) : 1;
  }
  return g_gemm_pair;
}

template <int BN, bool B_MN, int CL>
static int launch_cluster(const CUtensorMap& ta, const CUtensorMap& tb, const Params& p, int grid, cudaStream_t st) {
  using C = Cfg<BN>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_cluster_kernel<BN, B_MN, CL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         C::kSmemBytes);
    if (e != cudaSuccess) return (int)e;
    attr_set = true;
  }
  gemm_cluster_kernel<BN, B_MN, CL><<<grid, kThreads, C::kSmemBytes, st>>>(ta, tb, p);
  CXN_CHECK_LAUNCH();
  return 0;
}
static int g_gemm_cluster = -1;    // 2 | 4: cluster size of the activation-multicast route; -1 = read CXN_GEMM_CLUSTER
static int gemm_cluster_mode() {
  if (g_gemm_cluster < 0) {
    const char* e = getenv("CXN_GEMM_CLUSTER");
    g_gemm_cluster = e ? atoi(e) : 0;
  }
  return g_gemm_cluster;
}

// fp32 scratch of the internally split GEMMs: zeroed buffers per device, grown on demand (never while a stream is
// capturing - the first, eager, steps size them).  A device has kWsLanes independent buffers ("lanes"); the calling
// thread picks one with cxn_gemm_workspace_lane (default 0).  Work that may run CONCURRENTLY - e.g. the main-stream
// and exchange-stream branches of a captured step - must use different lanes, otherwise two split GEMMs would
// accumulate into, and re-zero, the same buffer.  Within a lane, different streams are ordered through an event;
// inside a stream capture neither the wait nor the record is issued (sequential use of one lane inside a graph is
// ordered by the graph's own dependencies).
struct SplitWs {
  float* p = nullptr;
  size_t n = 0;
  cudaEvent_t ev = nullptr;
  cudaStream_t last = nullptr;
  bool used = false;
};
constexpr int kWsLanes = 4;
static std::mutex g_ws_mu;
static SplitWs g_ws[64][kWsLanes];
static thread_local int t_ws_lane = 0;
static thread_local UpdArgs t_upd = {};   // operands of the next kFlagUpdate call (set by cxn_gemm_wgrad_update)

static bool stream_capturing(cudaStream_t st) {
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(st, &cs) != cudaSuccess) return false;
  return cs != cudaStreamCaptureStatusNone;
}

static float* splitk_workspace(size_t bytes, cudaStream_t st) {
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return nullptr;
  std::lock_guard<std::mutex> g(g_ws_mu);
  SplitWs& w = g_ws[dev][t_ws_lane];
  const bool capturing = stream_capturing(st);
  if (w.n < bytes) {
    if (capturing) return nullptr;            // cannot allocate now: the caller falls back to the unsplit kernel
    if (w.p) { cudaDeviceSynchronize(); cudaFree(w.p); }
    w.p = nullptr; w.n = 0;
    const size_t want = bytes < (16u << 20) ? (16u << 20) : bytes;
    if (cudaMalloc(&w.p, want) != cudaSuccess) { w.p = nullptr; cudaGetLastError(); return nullptr; }
    if (cudaMemsetAsync(w.p, 0, want, st) != cudaSuccess) return nullptr;
    w.n = want;
    if (!w.ev) cudaEventCreateWithFlags(&w.ev, cudaEventDisableTiming);
    // the zero fill counts as a use: another stream must wait for it
    cudaEventRecord(w.ev, st);
    w.last = st; w.used = true;
  }
  if (!capturing && w.used && w.last != st && w.ev) cudaStreamWaitEvent(st, w.ev, 0);
  return w.p;
}
static void splitk_release(cudaStream_t st) {
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return;
  std::lock_guard<std::mutex> g(g_ws_mu);
  SplitWs& w = g_ws[dev][t_ws_lane];
  if (!stream_capturing(st) && w.ev) cudaEventRecord(w.ev, st);
  w.last = st; w.used = true;
}

}  // namespace gemm
}  // namespace cxn

// CTA-pair route (cta_group::2): 1 = automatic (large problems), 0 = never, 2 = whenever the shape allows,
// -1 = re-read CXN_GEMM_PAIR
CXN_API int cxn_gemm_pair(int mode) {
  if (mode < -1 || mode > 2) return -1;
  cxn::gemm::g_gemm_pair = mode;
  return 0;
}

// Split-K scratch lane (0..3) used by this thread's following cxn_gemm_bf16 calls; returns the previous lane or -1.
CXN_API int cxn_gemm_workspace_lane(int lane) {
  if (lane < 0 || lane >= cxn::gemm::kWsLanes) return -1;
  const int prev = cxn::gemm::t_ws_lane;
  cxn::gemm::t_ws_lane = lane;
  return prev;
}

// experiment knob: cluster size (0 | 2 | 4) of the activation-multicast route; -1 re-reads CXN_GEMM_CLUSTER
CXN_API int cxn_gemm_cluster(int cl) {
  if (cl != -1 && cl != 0 && cl != 2 && cl != 4) return -1;
  cxn::gemm::g_gemm_cluster = cl;
  return 0;
}

// D[z] = A[z] * B[z]^T.  All leading dimensions / strides are in elements.
//   a_mn / b_mn: 0 = operand stored [rows=MN][K] (K contiguous), 1 = stored [K][MN] (MN contiguous)
//   flags: cxn::gemm::kFlag*;  splits: split-K factor (>1 requires kFlagAtomic, fp32 D)
//   bn: tile width 64/128/256 or 0 = choose.
CXN_API int cxn_gemm_bf16(const void* A, long long lda, long long strideA, int a_mn, const void* B, long long ldb,
                          long long strideB, int b_mn, void* D, long long ldd, long long strideD, const float* bias,
                          long long strideBias, int M, int N, int K, int batch, int splits, int flags, int bn,
                          cudaStream_t stream) {
  using namespace cxn;
  using namespace cxn::gemm;
  if (M <= 0 || N <= 0 || K <= 0 || batch <= 0) return 0;
  const int num_kb = ceil_div(K, BK);
  if (splits < 1) splits = 1;
  if (splits > num_kb) splits = num_kb;
  if (splits > 1 && !(flags & kFlagAtomic)) return -5;
  if ((flags & kFlagAtomic) && (flags & kFlagOutBf16)) return -6;
  const int sms = num_sms();
  // Skinny problems (few output tiles, long K - the fully connected layers at any batch size) are
  // split along K internally: partial tiles meet in a zeroed fp32 scratch through vector
  // reductions and a second small kernel applies bias / activation / conversion.  More CTAs pull
  // on the weight stream, and wider tiles re-read the activations less often.
  static const bool auto_split = getenv("CXN_GEMM_NO_AUTOSPLIT") == nullptr;
  // (measured: worth it only while even 64-wide tiles would leave a quarter of the SMs idle)
  const long long tiles64 = (long long)ceil_div(M, BM) * ceil_div(N, 64);
  if (auto_split && splits == 1 && batch == 1 && bn == 0 && !(flags & (kFlagAtomic | kFlagUpdate)) && num_kb >= 16 &&
      tiles64 * 4 < 3LL * sms) {
    int best_bn = 0, best_splits = 1;
    const int cand[3] = {256, 128, 64};
    for (int ci = 0; ci < 3; ++ci) {
      const int b = cand[ci];
      if (b > 64 && N <= b / 2) continue;
      const long long t = (long long)ceil_div(M, BM) * ceil_div(N, b);
      if (t >= sms) break;                       // enough tiles without splitting
      int sp = (int)(sms / t);
      if (sp > num_kb / 8) sp = num_kb / 8;      // keep at least 8 k-blocks per split
      if (sp < 2) continue;
      if (t * sp >= (3 * sms) / 4 || ci == 2) { best_bn = b; best_splits = sp; break; }
    }
    const int ldw = (N + 3) & ~3;
    const size_t ws_bytes = (size_t)M * ldw * sizeof(float);
    if (best_bn && ws_bytes <= (64u << 20)) {
      float* ws = splitk_workspace(ws_bytes, stream);
      if (ws != nullptr) {
        int rc = cxn_gemm_bf16(A, lda, strideA, a_mn, B, ldb, strideB, b_mn, ws, ldw, 0, nullptr, 0, M, N, K, 1, best_splits,
                               kFlagAtomic, best_bn, stream);
        if (rc) return rc;
        const long long quads = (long long)M * (ldw >> 2);
        splitk_finish_kernel<<<ew_grid(quads, 256), 256, 0, stream>>>(ws, ldw, D, ldd, bias, M, N, flags);
        CXN_CHECK_LAUNCH();
        splitk_release(stream);
        return 0;
      }
    }
  }
  int kbps = ceil_div(num_kb, splits);
  splits = ceil_div(num_kb, kbps);
  // CTA pairs: 256 x 256 tiles computed by two SMs with M = 256 MMAs (gemm_pair_kernel) when the problem has at least
  // one tile per pair of SMs - two thirds of the L2 -> SM bytes of the 128 x 256 single-CTA tile
  {
    const int pm = gemm_pair_mode();
    const long long t256 = (long long)ceil_div(M, 256) * ceil_div(N, 256) * batch * splits;
    const bool legal = (bn == 0 || bn == 256) && M > 128 && N > 128 && sms >= 2 && !(flags & kFlagUpdate);
    // (automatic: only long reductions - with few k-blocks the tile is epilogue-bound and the single-CTA kernel's eight
    // epilogue warps win; measured in profiles/r2_gemm_sweep.txt)
    if (legal && (pm == 2 || (pm == 1 && t256 >= sms / 2 && num_kb >= 16))) {
      Params pp;
      pp.M = M; pp.N = N; pp.K = K; pp.batch = batch; pp.splits = splits; pp.kb_per_split = kbps;
      pp.D = D; pp.ldd = ldd; pp.strideD = strideD; pp.bias = bias; pp.strideBias = strideBias; pp.flags = flags;
      pp.u = UpdArgs{};
      CUtensorMap pa, pb;
      int rc2 = make_map(&pa, A, M, K, lda, batch, strideA, a_mn != 0, BM);
      if (rc2) return rc2;
      rc2 = make_map(&pb, B, N, K, ldb, batch, strideB, b_mn != 0, 128);
      if (rc2) return rc2;
      const long long max_pairs = sms / 2;
      const int gridp = 2 * (int)(t256 < max_pairs ? t256 : max_pairs);
      if (!a_mn && !b_mn) return launch_pair<false, false>(pa, pb, pp, gridp, stream);
      if (!a_mn && b_mn) return launch_pair<false, true>(pa, pb, pp, gridp, stream);
      if (a_mn && b_mn) return launch_pair<true, true>(pa, pb, pp, gridp, stream);
      return launch_pair<true, false>(pa, pb, pp, gridp, stream);
    }
  }
  if (bn == 0) {
    auto tiles = [&](int b) { return (long long)ceil_div(M, BM) * ceil_div(N, b) * batch * splits; };
    if (N > 128 && tiles(256) >= sms) bn = 256;
    else if (N > 64 && tiles(128) * 4 >= 3LL * sms) bn = 128;
    else bn = (N > 64 && tiles(64) > 2LL * sms) ? 128 : 64;
  }
  Params p;
  p.M = M; p.N = N; p.K = K; p.batch = batch; p.splits = splits; p.kb_per_split = kbps;
  p.D = D; p.ldd = ldd; p.strideD = strideD; p.bias = bias; p.strideBias = strideBias; p.flags = flags;
  p.u = UpdArgs{};
  if (flags & kFlagUpdate) {
    // fused updater epilogue: weight-gradient operand layout only (both operands stored [K][MN]), one pass over K
    if (!a_mn || !b_mn || batch != 1 || splits != 1 || bias != nullptr || (flags & ~kFlagUpdate)) return -7;
    p.u = t_upd;
    if (p.u.w == nullptr || p.u.m == nullptr || p.u.hyper == nullptr || (p.u.kind == 2 && p.u.m2 == nullptr)) return -7;
  }
  CUtensorMap ta, tb;
  const int cl = gemm_cluster_mode();
  if ((cl == 2 || cl == 4) && !a_mn && batch == 1 && splits == 1 && !(flags & (kFlagAtomic | kFlagBeta1)) && bn <= 128 &&
      ceil_div(N, bn) >= cl && sms >= cl &&
      memory_is_local(A, ((size_t)(M - 1) * (size_t)lda + (size_t)K) * sizeof(bf16))) {
    // skinny problem, many column tiles: clusters of `cl` column tiles share the activation tile by multicast.
    // The multicast tensor map must cover this GPU's own memory only (memory_is_local, common.cuh); any other
    // operand takes the ordinary kernel below.
    int rc2 = make_map(&ta, A, M, K, lda, 1, strideA, false, BM / cl);
    if (rc2) return rc2;
    rc2 = make_map(&tb, B, N, K, ldb, 1, strideB, b_mn != 0, bn);
    if (rc2) return rc2;
    const long long groups = (long long)ceil_div(M, BM) * ceil_div(ceil_div(N, bn), cl);
    const long long max_clusters = sms / cl;
    const int gridc = cl * (int)(groups < max_clusters ? groups : max_clusters);
#define CXN_CL_DISPATCH(BN_, CL_)                                                          \
  return b_mn ? launch_cluster<BN_, true, CL_>(ta, tb, p, gridc, stream)                  \
              : launch_cluster<BN_, false, CL_>(ta, tb, p, gridc, stream);
    if (bn == 128) { if (cl == 4) { CXN_CL_DISPATCH(128, 4) } CXN_CL_DISPATCH(128, 2) }
    if (cl == 4) { CXN_CL_DISPATCH(64, 4) }
    CXN_CL_DISPATCH(64, 2)
#undef CXN_CL_DISPATCH
  }
  int rc = make_map(&ta, A, M, K, lda, batch, strideA, a_mn != 0, BM);
  if (rc) return rc;
  rc = make_map(&tb, B, N, K, ldb, batch, strideB, b_mn != 0, bn);
  if (rc) return rc;
  const long long total = (long long)ceil_div(M, BM) * ceil_div(N, bn) * batch * splits;
  const int grid = (int)(total < sms ? total : sms);
  // short-K fp32 problems in the weight-gradient layout (the fully connected weight gradients): WIDE epilogue - whole
  // result rows of a warp's columns per instruction (Cfg<BN, true>).  CXN_GEMM_WIDE: 1 (default) = with the fused
  // updater only (measured: 174 -> 145 us on fc6; the plain fp32 store gets slower, 42 -> 44 us), 2 = also plain
  // stores, 0 = never.
  static const int wide_mode = []() { const char* e = getenv("CXN_GEMM_WIDE"); return e == nullptr ? 1 : atoi(e); }();
  const bool wide = (wide_mode >= 2 || (wide_mode == 1 && (flags & kFlagUpdate))) && a_mn && b_mn && bn >= 128 &&
                    kbps <= 8 && bias == nullptr && !(flags & (kFlagOutBf16 | kFlagAtomic | kFlagRelu));
  if (flags & kFlagUpdate) {
    if (wide && bn == 256) return launch<256, true, true, true, true>(ta, tb, p, grid, stream);
    if (wide && bn == 128) return launch<128, true, true, true, true>(ta, tb, p, grid, stream);
    if (bn == 256) return launch<256, true, true, true>(ta, tb, p, grid, stream);
    if (bn == 128) return launch<128, true, true, true>(ta, tb, p, grid, stream);
    return launch<64, true, true, true>(ta, tb, p, grid, stream);
  }
  if (wide && bn == 256) return launch<256, true, true, false, true>(ta, tb, p, grid, stream);
  if (wide && bn == 128) return launch<128, true, true, false, true>(ta, tb, p, grid, stream);
#define CXN_DISPATCH(BN_)                                                                      \
  if (!a_mn && !b_mn) return launch<BN_, false, false>(ta, tb, p, grid, stream);              \
  if (!a_mn && b_mn) return launch<BN_, false, true>(ta, tb, p, grid, stream);                \
  if (a_mn && b_mn) return launch<BN_, true, true>(ta, tb, p, grid, stream);                  \
  return launch<BN_, true, false>(ta, tb, p, grid, stream);
  if (bn == 256) { CXN_DISPATCH(256) }
  if (bn == 128) { CXN_DISPATCH(128) }
  { CXN_DISPATCH(64) }
#undef CXN_DISPATCH
}

// Weight gradient of a fully connected layer with the updater applied in the epilogue (K6 + K15 in one kernel):
//   g = A^T-stored . B^T-stored as in cxn_gemm_bf16(a_mn = 1, b_mn = 1), M x N over K (= batch rows), then per element
//   the sgd / nag / adam step of update_kernel on w / m (/ m2) at [row * ldw + column] and the bf16 copy at
//   [row * shadow_ld + column].  The gradient itself is never stored: 18 bytes of traffic per parameter instead of the
//   26 of GEMM + multi-tensor update.  `hyper` = &segment.lr of the layer's entry in the device table.
//   Returns -7 for arguments the fused path does not take (the caller then runs GEMM + cxn_update).
CXN_API int cxn_gemm_wgrad_update(const void* A, long long lda, const void* B, long long ldb, int M, int N, int K,
                                  float* w, float* m, float* m2, void* shadow, long long ldw, long long shadow_ld,
                                  const float* hyper, int kind, cudaStream_t stream) {
  using namespace cxn::gemm;
  if (kind < 0 || kind > 2) return -7;
  t_upd.w = w; t_upd.m = m; t_upd.m2 = m2; t_upd.shadow = (cxn::bf16*)shadow; t_upd.shadow_ld = shadow_ld;
  t_upd.hyper = hyper; t_upd.kind = kind;
  const int rc = cxn_gemm_bf16(A, lda, 0, 1, B, ldb, 0, 1, w, ldw, 0, nullptr, 0, M, N, K, 1, 1, kFlagUpdate, 0, stream);
  t_upd = UpdArgs{};
  return rc;
}
