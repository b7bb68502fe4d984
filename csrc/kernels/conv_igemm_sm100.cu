// Implicit-GEMM convolution on tcgen05 / TMEM for sm_100a (K1-K3 of SURVEY.md section 2.9).
//
// No im2col buffer exists: the patch matrix goes straight from the NHWC bf16 activation tensor
// into 128B-swizzled shared memory, the dense operand arrives by (tiled) TMA.  Two ways to fetch
// the patches:
//   ATMA   im2col-mode TMA (cuTensorMapEncodeIm2col): one instruction per k-block loads 128/256
//          output positions x 64 channels through one filter tap, padding = out-of-bounds zero
//          fill.  Needs channels-per-group % 64 == 0.  No thread touches the data.
//   gather four producer warps issue cp.async (zero-fill for padding) from a per-CTA gather table;
//          any channels-per-group % 8 == 0, and the 4-channel-pitch RGB input.
//
//   FPROP  y[n,oh,ow, g*Ng + j]  = sum_{kh,kw,c} x[n, oh*s-p+kh, ow*s-p+kw, g*Cg + c] * W[g][j][kh][kw][c]
//          A = gathered patches (K-major, rows = output pixels), B = W via TMA (K-major)
//          epilogue: + bias, optional ReLU, bf16 NHWC store.
//          The same kernel computes backward-data of stride-1 convolutions: gather from dy with
//          the spatially flipped, channel-transposed filter (cxn_conv_flip_weights).
//   WGRAD  dW[g][j][kh][kw][c] += sum_{n,oh,ow} dy[n,oh,ow, g*Ng + j] * x[n, oh*s-p+kh, ow*s-p+kw, g*Cg + c]
//          A = dy via TMA (MN-major), B = gathered patches (MN-major: rows = pixels = reduction),
//          split-K over pixels, fp32 red.global.add epilogue.
//
// Reference behaviour being reproduced: src/layer/convolution_layer-inl.hpp:79-153 (im2col in
// chunks + one SGEMM per group + col2im); semantics (cross-correlation, zero padding, groups)
// are identical, the algorithm is not.
//
// Warp roles (10 warps, persistent, one CTA per SM):
//   0 TMA producer | 1 MMA issuer + TMEM owner | 2-5 epilogue | 6-9 gather producers
// The gather granule is 16 bytes (8 channels) when Cg % 8 == 0, or 8 bytes (one 4-channel
// pixel) for the 3-channel input layer stored with a channel pitch of 4.
#include "tc_common.cuh"
#include <mutex>
#include <string.h>

namespace cxn {
namespace igemm {

using namespace cxn::tc;

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int UMMA_K = 16;
constexpr int kThreads = 320;
constexpr int kGatherWarp0 = 6;
constexpr int kGatherThreads = 128;
constexpr int kMaxGranules = 1024;  // gather table entries (K / granule)

struct Params {
  // gathered tensor (NHWC, pixel stride Cs elements)
  const bf16* src;
  int N, H, W, Cs;
  int Cg;          // channels per group as laid out in K (4 in the padded-RGB mode)
  int KH, KW, stride_y, stride_x, pad_y, pad_x;
  int OH, OW;      // pixel grid enumerated by the GEMM rows
  int G;
  int Kg;          // KH*KW*Cg
  int gran;        // elements per gather granule: 8 or 4
  int Ng;          // dense-operand extent per group (output channels)
  int bn;          // tile width actually used (multiple of 16, <= BNMAX)
  // output
  void* D;
  long long ldd, strideD;   // strideD: per-group offset (elements)
  const float* bias;        // [G*Ng] or null
  int relu;
  const bf16* mask;         // optional [rows][ldd] tensor: outputs whose mask element is <= 0 are written as zero
                            // (backward-data folding the ReLU that precedes this layer in the forward pass)
  // wgrad
  int splits, kb_per_split;
  int mask_period, mask_valid;   // wgrad: columns n with n % period >= valid are not accumulated (0 = off)
  // fprop: whole filter panel resident in shared memory (single group, single n-tile, small K*N)
  int b_resident;
  int stages;      // depth of the shared-memory ring actually used (<= Cfg::kMaxStages)
  int ring_bytes;  // shared memory before the barriers: stages * stage bytes, or (resident filter) A ring + panel
  // optional strided view for the im2col tensor map (elements; 0 = the dense NHWC tensor): channel
  // extent per position, row stride, image stride.  Cs stays the stride between positions.
  long long view_cdim, view_row, view_img;
  int Cgs;         // channels between the first channels of consecutive groups: Cg, or less when a view's groups are
                   // channel windows padded up to a multiple of 64 (cxn_conv_view_group_stride)
  // optional strided view of the OUTPUT (fprop only; 0 = rows are consecutive): GEMM row r = (image, j, i) of an
  // o_h x o_w grid lands at element o_base + image * o_img + j * o_row + i * o_pix (+ group / column offsets).  Used by
  // backward-data of strided convolutions: each stride phase is a stride-1 problem whose results interleave in dx.
  int o_w, o_hw;
  long long o_base, o_pix, o_row, o_img;
  int o_vh, o_vw;  // valid extent of the view's grid (0 = all of it): positions outside are computed and discarded
  // FLAT: stride-1 forward problem without padding computed over the INPUT's pixel grid (H x W instead of OH x OW):
  // there a filter tap is a constant row offset, so a tile's patch rows through tap (kh, kw) are one contiguous 2-D box
  // fetched in TILED mode at row m0 + kh * W + kw (im2col-mode fetches bound layers with few output columns,
  // profiles/r2_experiments.md).  The rows with ow >= OW or oh >= OH read across row / image ends and are discarded by
  // the epilogue through the output view above.
  int flat;
  long long* prof; // [grid][16] cycle counters of the warp roles (builds with -DCXN_IGEMM_PROF only)
  int dbg;         // timing experiments only (results are garbage): 2 skip the TMA loads, 4 skip the MMAs
};

#ifdef CXN_IGEMM_PROF
constexpr bool kProf = true;
#else
constexpr bool kProf = false;
#endif

struct GatherEntry { int off; short kh, kw; };
constexpr int kHostFixedBytes = 1024 + 256 + kMaxGranules * 8 + 4 * 32 * 36 * 4;   // = Cfg::kFixedBytes

// MH = number of 128-row accumulator halves of one tile (fprop only): with MH = 2 a CTA computes a
// 256-row tile with two MMAs per k-step that share the filter tile, halving the filter's L2 traffic.
template <int BNMAX, int MH>
struct Cfg {
  static constexpr int kAHalfBytes = BM * BK * 2;
  static constexpr int kABytes = MH * kAHalfBytes;
  static constexpr int kBBytes = BNMAX * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kTableBytes = kMaxGranules * 8;
  // epilogue staging (weight gradient): per epilogue warp a 32 x 32 fp32 chunk, rows padded to 36 floats
  static constexpr int kStageOutBytes = 4 * 32 * 36 * 4;
  static constexpr int kFixedBytes = 1024 + 256 + kTableBytes + kStageOutBytes;
  static constexpr int kMaxStagesFit = (232448 - kFixedBytes) / kStageBytes;
  static constexpr int kMaxStages = kMaxStagesFit > 6 ? 6 : kMaxStagesFit;
  static constexpr int kAccStages = (2 * MH * BNMAX <= 512) ? 2 : 1;
  static constexpr int kTmemCols = kAccStages * MH * BNMAX;
  static constexpr int smem_bytes(int stages) { return stages * kStageBytes + kFixedBytes; }
};

template <int BNMAX, bool WGRAD, int GRAN, int MH, bool ATMA>
__global__ void __launch_bounds__(kThreads, 1)
conv_igemm_kernel(const __grid_constant__ CUtensorMap tmDense, const __grid_constant__ CUtensorMap tmPatch,
                  const Params p) {
  using C = Cfg<BNMAX, MH>;
  static_assert(!(WGRAD && MH != 1), "row-paired tiles are a forward-convolution feature");
  static_assert(!(ATMA && GRAN != 8), "im2col TMA needs 64-channel chunks");
  // With ATMA the TMA / MMA warps run warp-uniform and elect the issuing lane (as the GEMM does);
  // with the gather producers alive the single-lane form is faster (fewer pollers on the barriers
  // the 128 gather threads arrive on).
  constexpr bool UNI = ATMA;
  constexpr int kAccStages = C::kAccStages;
  constexpr int BMT = BM * MH;               // tile rows
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const int nst = p.stages;
  const uint32_t bar_base = smem_base + (uint32_t)p.ring_bytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (nst + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * nst + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * nst + kAccStages + a); };
  const uint32_t bres_bar = bar_base + 8u * (2 * nst + 2 * kAccStages);
  volatile uint32_t* tmem_ptr_smem =
      reinterpret_cast<volatile uint32_t*>(smem_gen + p.ring_bytes + 8 * (2 * nst + 2 * kAccStages + 1));
  // smem addressing: ring of {A,B} stages, or (resident filter) ring of A stages followed by the panel
  const bool bres = !WGRAD && p.b_resident;
  const uint32_t a_stride = bres ? (uint32_t)C::kABytes : (uint32_t)C::kStageBytes;
  const uint32_t bres_base = smem_base + nst * C::kABytes;
  const uint32_t bres_kb_bytes = (uint32_t)p.bn * BK * 2;
  GatherEntry* table = reinterpret_cast<GatherEntry*>(smem_gen + p.ring_bytes + 256);
  float* stage_out = reinterpret_cast<float*>(smem_gen + p.ring_bytes + 256 + C::kTableBytes);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const long long rows = p.flat ? (long long)p.N * p.H * p.W : (long long)p.N * p.OH * p.OW;

  // GEMM view.  FPROP: M = rows (gathered), N = Ng, K = Kg.   WGRAD: M = Ng, N = Kg (gathered), K = rows.
  const long long gm = WGRAD ? p.Ng : rows;
  const int gn = WGRAD ? p.Kg : p.Ng;
  const long long gk = WGRAD ? rows : p.Kg;
  const int tiles_m = (int)ceil_div_ll(gm, BMT);
  const int tiles_n = ceil_div(gn, p.bn);
  const int num_kb = (int)ceil_div_ll(gk, BK);
  const int splits = WGRAD ? p.splits : 1;
  const int kbps = WGRAD ? p.kb_per_split : num_kb;
  const long long total_tiles = (long long)tiles_m * tiles_n * p.G * splits;

  // gather table: granule index -> (element offset of the tap inside the image, kh, kw)
  const int ngran = p.Kg / GRAN;
  for (int i = threadIdx.x; !ATMA && i < ngran; i += blockDim.x) {
    const int k = i * GRAN;
    const int pos = k / p.Cg, c0 = k - pos * p.Cg;
    const int kh = pos / p.KW, kw = pos - kh * p.KW;
    GatherEntry e;
    e.off = (kh * p.W + kw) * p.Cs + c0;
    e.kh = (short)kh;
    e.kw = (short)kw;
    table[i] = e;
  }

  if (threadIdx.x == 0) {
    for (int s = 0; s < nst; ++s) {
      mbar_init(full_bar(s), ATMA ? 1 : 1 + kGatherThreads);
      mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < kAccStages; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), 4);
    }
    mbar_init(bres_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmDense) : "memory");
    if (ATMA) asm volatile("prefetch.tensormap [%0];" ::"l"(&tmPatch) : "memory");
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

  const int cpk = p.Cg >> 6;   // 64-channel chunks per filter tap (ATMA)

  // (all lanes of a warp poll: with one polling lane and the rest parked at the reconvergence point the AlexNet step was
  // 2.242 ms instead of 2.202 - measured at the end of round 2)
  auto wait_bar = [&](uint32_t bar, uint32_t parity) { mbar_wait(bar, parity); };
  auto issue = [&]() -> bool { return UNI ? elect_one() : true; };   // UNI: all lanes run the loop, one issues
  // role profiling: cycles spent inside waits / phases, accumulated per CTA
  const bool prof_on = kProf && p.prof != nullptr;
  long long pw0 = 0, pw1 = 0, pw2 = 0;
  const long long prof_t0 = prof_on ? clock64() : 0;
  auto timed_wait = [&](uint32_t bar, uint32_t parity, long long& accum) {
    if (prof_on) {
      const long long t = clock64();
      wait_bar(bar, parity);
      accum += clock64() - t;
    } else {
      wait_bar(bar, parity);
    }
  };
  auto prof_store = [&](int slot, long long v) {
    if (prof_on && lane == 0) p.prof[(long long)blockIdx.x * 16 + slot] = v;
  };

  auto decode_tile = [&](long long t, int& n_blk, int& m_blk, int& split, int& g) {
    long long r = t;
    n_blk = (int)(r % tiles_n); r /= tiles_n;
    m_blk = (int)(r % tiles_m); r /= tiles_m;
    split = (int)(r % splits);
    g = (int)(r / splits);
  };

  if (warp == 0) {
    // =========================== TMA producer ===========================
    // dense operand always; with ATMA also the patch operand (im2col mode)
    if (UNI || lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      if (bres && issue()) {
        // the whole filter panel, once per CTA
        mbar_expect_tx(bres_bar, (uint32_t)num_kb * bres_kb_bytes);
        for (int kb = 0; kb < num_kb; ++kb) tma_load_3d(bres_base + kb * bres_kb_bytes, &tmDense, bres_bar, kb * BK, 0, 0);
      }
      if (UNI) __syncwarp();
      const int ohw = p.OH * p.OW;
      const int step_oh = BK / p.OW, step_ow = BK - step_oh * p.OW;     // 64 pixels = step_oh rows + step_ow columns
      for (long long t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        int n_blk, m_blk, split, g;
        decode_tile(t, n_blk, m_blk, split, g);
        const int kb0 = split * kbps, kb1 = min(num_kb, kb0 + kbps);
        // ATMA fprop: base output position of the tile -> (w, h, n) of its tap-(0,0) input pixel
        int pn = 0, pw = 0, ph = 0;
        if (ATMA && !WGRAD) {
          const long long m0 = (long long)m_blk * BMT;
          pn = (int)(m0 / ohw);
          const int r = (int)(m0 - (long long)pn * ohw);
          const int oh = r / p.OW;
          pw = (r - oh * p.OW) * p.stride_x - p.pad_x;
          ph = oh * p.stride_y - p.pad_y;
        }
        // weight gradient on the im2col route: the tile's (up to four) 64-column atoms -> (channel, kw, kh), and the
        // first pixel of the split -> (ow, oh, n), stepped by 64 pixels per k-block without divisions
        int wg_c[4] = {0, 0, 0, 0}, wg_kw[4] = {0, 0, 0, 0}, wg_kh[4] = {0, 0, 0, 0};
        int wg_n = 0, wg_oh = 0, wg_ow = 0;
        int fp_chunk = 0, fp_kw = 0, fp_kh = 0;     // forward: (64-channel chunk, kw, kh) of k-block kb0 = 0
        if (ATMA && WGRAD) {
#pragma unroll
          for (int a = 0; a < 4; ++a) {
            const int n0 = n_blk * p.bn + a * 64;
            const int tap = n0 / p.Cg, c0 = n0 - tap * p.Cg;
            wg_kh[a] = tap / p.KW;
            wg_kw[a] = tap - wg_kh[a] * p.KW;
            wg_c[a] = g * p.Cgs + c0;
          }
          const long long pix0 = (long long)kb0 * BK;
          wg_n = (int)(pix0 / ohw);
          const int r = (int)(pix0 - (long long)wg_n * ohw);
          wg_oh = r / p.OW;
          wg_ow = r - wg_oh * p.OW;
        }
        for (int kb = kb0; kb < kb1; ++kb) {
          timed_wait(empty_bar(stage), phase ^ 1u, pw0);
          const uint32_t sa = smem_base + stage * a_stride;
          const uint32_t sb = sa + C::kABytes;
          if (issue()) {
            if (p.dbg & 2) {
              mbar_arrive(full_bar(stage));
            } else if (!WGRAD) {
              const int pk = kb;
              uint32_t tx = bres ? 0u : (uint32_t)p.bn * BK * 2;
              if (ATMA) tx += (uint32_t)C::kABytes;
              if (tx) mbar_expect_tx(full_bar(stage), tx); else mbar_arrive(full_bar(stage));
              if (ATMA && p.flat) {
                // A = BMT consecutive pixels of the input grid, shifted by the tap: one tiled-mode box
                tma_load_3d(sa, &tmPatch, full_bar(stage), g * p.Cgs + fp_chunk * 64,
                            (int)((long long)m_blk * BMT) + fp_kh * p.W + fp_kw, 0);
              } else if (ATMA) {
                // A = patches: BMT output positions x 64 channels through tap (kh, kw)
                tma_load_im2col_4d(sa, &tmPatch, full_bar(stage), g * p.Cgs + fp_chunk * 64, pw, ph, pn, fp_kw, fp_kh);
              }
              // B = W[g] : K-major box {64 (k), bn (n)}
              if (!bres) tma_load_3d(sb, &tmDense, full_bar(stage), pk * BK, n_blk * p.bn, g);
            } else {
              // A = dy[:, g*Ng + m] : MN-major, two boxes {64 (m), 64 (k)}
              uint32_t tx = (uint32_t)C::kABytes;
              int natoms_ld = 0;
              if (ATMA) {
                natoms_ld = min((p.bn + 63) >> 6, (p.Kg - n_blk * p.bn + 63) >> 6);
                tx += (uint32_t)natoms_ld * (64 * BK * 2);
              }
              mbar_expect_tx(full_bar(stage), tx);
#pragma unroll
              for (int h = 0; h < BM / 64; ++h)
                tma_load_3d(sa + h * (64 * BK * 2), &tmDense, full_bar(stage), m_blk * BM + h * 64, kb * BK, g);
              if (ATMA) {
                // B = patches, MN-major: per 64-column atom (one tap, 64 channels) the 64 pixels of this k-block.
                // (Everything this needs was decoded outside the k loop: the elected lane's instruction stream is
                // what paces the loads of this kernel - bench/wgrad_roles.py.)
                const int w0 = wg_ow * p.stride_x - p.pad_x, h0 = wg_oh * p.stride_y - p.pad_y;
#pragma unroll
                for (int a = 0; a < 4; ++a)
                  if (a < natoms_ld)
                    tma_load_im2col_4d(sb + a * (64 * BK * 2), &tmPatch, full_bar(stage), wg_c[a], w0, h0, wg_n, wg_kw[a],
                                       wg_kh[a]);
              }
            }
          }
          if (ATMA && !WGRAD) {
            // the next k-block: next 64-channel chunk of the tap, then the next tap
            if (++fp_chunk == cpk) {
              fp_chunk = 0;
              if (++fp_kw == p.KW) { fp_kw = 0; ++fp_kh; }
            }
          }
          if (ATMA && WGRAD) {
            // the next k-block starts 64 pixels on
            wg_ow += step_ow;
            wg_oh += step_oh;
            if (wg_ow >= p.OW) { wg_ow -= p.OW; ++wg_oh; }
            while (wg_oh >= p.OH) { wg_oh -= p.OH; ++wg_n; }
          }
          if (UNI) __syncwarp();
          if (++stage == nst) { stage = 0; phase ^= 1u; }
        }
      }
      prof_store(0, pw0);
      prof_store(1, prof_on ? clock64() - prof_t0 : 0);
    }
    __syncwarp();
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    if (UNI || lane == 0) {
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((WGRAD ? 1u : 0u) << 15) |
                             ((WGRAD ? 1u : 0u) << 16) | ((uint32_t)(p.bn >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
      constexpr uint32_t kLboMN = 64 * BK * 2;
      constexpr uint32_t kSbo = 1024;
      constexpr uint32_t kStep = WGRAD ? (UMMA_K * 128) : (UMMA_K * 2);
      // descriptor = constant high word | (address >> 4): stepping along K only touches the low word
      const uint64_t desc_hi = make_desc(0, WGRAD ? kLboMN : 0u, kSbo);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      if (bres) wait_bar(bres_bar, 0);
      for (long long t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        int n_blk, m_blk, split, g;
        decode_tile(t, n_blk, m_blk, split, g);
        const int kb0 = split * kbps, kb1 = min(num_kb, kb0 + kbps);
        timed_wait(tempty_bar(acc), acc_phase ^ 1u, pw1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * MH * BNMAX);
        for (int kb = kb0; kb < kb1; ++kb) {
          timed_wait(full_bar(stage), phase, pw0);
          if (!ATMA) fence_proxy_async();   // cp.async (generic proxy) writes -> tensor-core (async proxy) reads
          tc_fence_after();
          const uint32_t sa = smem_base + stage * a_stride;
          const uint32_t sb = bres ? bres_base + kb * bres_kb_bytes : sa + C::kABytes;
          if (issue()) {
            const uint64_t a0 = desc_hi | (uint64_t)((sa >> 4) & 0x3FFF), b0 = desc_hi | (uint64_t)((sb >> 4) & 0x3FFF);
            if (!(p.dbg & 4)) {
#pragma unroll
              for (int k = 0; k < BK / UMMA_K; ++k) {
#pragma unroll
                for (int h = 0; h < MH; ++h)
                  umma_bf16(d_tmem + (uint32_t)(h * BNMAX), a0 + (uint64_t)((h * C::kAHalfBytes + k * kStep) >> 4),
                            b0 + (uint64_t)((k * kStep) >> 4), idesc, (kb > kb0 || k > 0) ? 1u : 0u);
              }
            }
            umma_commit(empty_bar(stage));
            if (kb == kb1 - 1) umma_commit(tfull_bar(acc));
          }
          if (UNI) __syncwarp();
          if (++stage == nst) { stage = 0; phase ^= 1u; }
        }
        if (++acc == kAccStages) { acc = 0; acc_phase ^= 1u; }
      }
      prof_store(2, pw0);
      prof_store(3, pw1);
      prof_store(4, prof_on ? clock64() - prof_t0 : 0);
    }
    __syncwarp();
  } else if (ATMA && warp >= kGatherWarp0) {
    // im2col TMA fetches the patches: nothing to gather
  } else if (warp >= kGatherWarp0) {
    // =========================== gather producers ===========================
    // Lane mapping: the lanes of one cp.async instruction cover whole 128-byte rows of the
    // tile (GPR granules per row, RPI = 32/GPR rows per instruction), so every instruction reads
    // a few contiguous runs of global memory instead of 32 scattered lines.
    constexpr int GPR = 64 / GRAN;            // granules per 128-byte row: 8 or 16
    constexpr int RPI = 32 / GPR;             // rows per instruction: 4 or 2
    constexpr int GBYTES = GRAN * 2;          // 16 or 8
    const int wl = warp - kGatherWarp0;       // 0..3
    const int q = lane & (GPR - 1);           // granule inside the row
    const int rsub = lane / GPR;              // row inside the instruction
    const uint32_t qoff16 = (GRAN == 8) ? (uint32_t)q : (uint32_t)(q >> 1);
    const uint32_t qsub = (GRAN == 8) ? 0u : (uint32_t)((q & 1) << 3);
    int stage = 0;
    uint32_t phase = 0;
    const int ohw = p.OH * p.OW;
    // validity of the KH x KW taps of one output pixel as two bit ranges (rows | cols << 16)
    auto tap_mask = [&](int ih0, int iw0) -> uint32_t {
      const int h_lo = max(0, -ih0), h_hi = min(p.KH, p.H - ih0);
      const int w_lo = max(0, -iw0), w_hi = min(p.KW, p.W - iw0);
      const uint32_t mh = h_hi > h_lo ? ((1u << h_hi) - 1u) & ~((1u << h_lo) - 1u) : 0u;
      const uint32_t mw = w_hi > w_lo ? ((1u << w_hi) - 1u) & ~((1u << w_lo) - 1u) : 0u;
      return mh | (mw << 16);
    };
    auto pixel_state = [&](int n, int oh, int ow, bool ok, long long& off, uint32_t& mask) {
      const int ih0 = oh * p.stride_y - p.pad_y, iw0 = ow * p.stride_x - p.pad_x;
      mask = ok ? tap_mask(ih0, iw0) : 0u;
      off = (((long long)n * p.H + ih0) * p.W + iw0) * p.Cs;
    };
    auto decode_pixel = [&](long long pix, bool ok, long long& off, uint32_t& mask) {
      int n = 0, oh = 0, ow = 0;
      if (ok) {
        n = (int)(pix / ohw);
        const int r = (int)(pix - (long long)n * ohw);
        oh = r / p.OW;
        ow = r - oh * p.OW;
      }
      pixel_state(n, oh, ow, ok, off, mask);
    };
    for (long long t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      int n_blk, m_blk, split, g;
      decode_tile(t, n_blk, m_blk, split, g);
      const int kb0 = split * kbps, kb1 = min(num_kb, kb0 + kbps);
      const bf16* gsrc = p.src + (long long)g * p.Cgs;
      const long long tset = prof_on ? clock64() : 0;
      if (!WGRAD) {
        // A tile: MH halves of 128 output pixels x 64 reduction elements; warp wl owns rows
        // [32*wl, 32*wl+32) of every half
        constexpr int ITERS1 = 32 / RPI;      // 8 or 16
        constexpr int ITERS = ITERS1 * MH;
        long long roff[ITERS];
        uint32_t rmask[ITERS];
#pragma unroll
        for (int i = 0; i < ITERS; ++i) {
          const int rl = (i / ITERS1) * BM + wl * 32 + (i % ITERS1) * RPI + rsub;
          const long long row = (long long)m_blk * BMT + rl;
          decode_pixel(row, row < rows, roff[i], rmask[i]);
        }
        if (prof_on) pw1 += clock64() - tset;
        for (int kb = kb0; kb < kb1; ++kb) {
          const int gi = kb * GPR + q;
          const bool gok = gi < ngran;
          const GatherEntry e = table[gok ? gi : 0];
          const uint32_t sel = gok ? ((1u << e.kh) | (1u << (16 + e.kw))) : 0xFFFFFFFFu;
          timed_wait(empty_bar(stage), phase ^ 1u, pw0);
          const uint32_t abase = smem_base + stage * a_stride;
#pragma unroll
          for (int i = 0; i < ITERS; ++i) {
            const int rl = (i / ITERS1) * BM + wl * 32 + (i % ITERS1) * RPI + rsub;   // half*128 + row: 128 B per row
            const bool ok = gok && ((rmask[i] & sel) == sel);
            const uint32_t dst = abase + rl * 128 + ((qoff16 ^ (uint32_t)(rl & 7)) << 4) + qsub;
            const void* sp = ok ? (const void*)(gsrc + roff[i] + e.off) : (const void*)p.src;
            if (GRAN == 8) cp_async_16(dst, sp, ok); else cp_async_8(dst, sp, ok);
          }
          cp_async_arrive_noinc(full_bar(stage));
          if (++stage == nst) { stage = 0; phase ^= 1u; }
        }
      } else {
        // B tile: 64 pixels (reduction rows) x bn reduction-columns, MN-major atoms of [64 rows][128 B];
        // warp wl owns pixel rows [16*wl, 16*wl+16)
        constexpr int ITERS = 16 / RPI;       // 4 or 8
        constexpr int MAXATOMS = BNMAX / 64;
        const int natoms = (p.bn + 63) >> 6;
        GatherEntry ent[MAXATOMS];
        uint32_t sel[MAXATOMS];
#pragma unroll
        for (int a = 0; a < MAXATOMS; ++a) {
          const int gi = (n_blk * p.bn + a * 64) / GRAN + q;
          const bool gok = (a < natoms) && gi < ngran;
          ent[a] = table[gok ? gi : 0];
          sel[a] = gok ? ((1u << ent[a].kh) | (1u << (16 + ent[a].kw))) : 0xFFFFFFFFu;
        }
        // pixel coordinates of this thread's rows, advanced by 64 pixels per k-block without divisions
        int pn[ITERS], poh[ITERS], pow_[ITERS];
#pragma unroll
        for (int i = 0; i < ITERS; ++i) {
          const long long pix = (long long)kb0 * BK + wl * 16 + i * RPI + rsub;
          pn[i] = (int)(pix / ohw);
          const int rr = (int)(pix - (long long)pn[i] * ohw);
          poh[i] = rr / p.OW;
          pow_[i] = rr - poh[i] * p.OW;
        }
        const int adv_n = BK / ohw, adv_r = BK - adv_n * ohw;
        const int adv_oh = adv_r / p.OW, adv_ow = adv_r - adv_oh * p.OW;
        for (int kb = kb0; kb < kb1; ++kb) {
          long long roff[ITERS];
          uint32_t rmask[ITERS];
#pragma unroll
          for (int i = 0; i < ITERS; ++i) {
            pixel_state(pn[i], poh[i], pow_[i], pn[i] < p.N, roff[i], rmask[i]);
            // advance to the pixel 64 rows further
            pow_[i] += adv_ow;
            int c = pow_[i] >= p.OW;
            pow_[i] -= c * p.OW;
            poh[i] += adv_oh + c;
            c = poh[i] >= p.OH;
            poh[i] -= c * p.OH;
            pn[i] += adv_n + c;
          }
          timed_wait(empty_bar(stage), phase ^ 1u, pw0);
          const uint32_t sb = smem_base + stage * C::kStageBytes + C::kABytes;
#pragma unroll
          for (int a = 0; a < MAXATOMS; ++a) {
            if (a < natoms) {
#pragma unroll
              for (int i = 0; i < ITERS; ++i) {
                const int rl = wl * 16 + i * RPI + rsub;
                const bool ok = (rmask[i] & sel[a]) == sel[a];
                const uint32_t dst = sb + a * (64 * BK * 2) + rl * 128 + ((qoff16 ^ (uint32_t)(rl & 7)) << 4) + qsub;
                const void* sp = ok ? (const void*)(gsrc + roff[i] + ent[a].off) : (const void*)p.src;
                if (GRAN == 8) cp_async_16(dst, sp, ok); else cp_async_8(dst, sp, ok);
              }
            }
          }
          cp_async_arrive_noinc(full_bar(stage));
          if (++stage == nst) { stage = 0; phase ^= 1u; }
        }
      }
    }
    if (wl == 0) {
      prof_store(5, pw0);
      prof_store(6, pw1);
      prof_store(7, prof_on ? clock64() - prof_t0 : 0);
    }
  } else {
    // =========================== epilogue ===========================
    const int quad = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (long long t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      int n_blk, m_blk, split, g;
      decode_tile(t, n_blk, m_blk, split, g);
      timed_wait(tfull_bar(acc), acc_phase, pw0);
      const long long tbody = prof_on ? clock64() : 0;
      tc_fence_after();
      const float* bias = (!WGRAD && p.bias != nullptr) ? p.bias + (long long)g * p.Ng : nullptr;
#pragma unroll 1
      for (int hc = 0; hc < MH * BNMAX; hc += 32) {
        const int h = hc / BNMAX, c0 = hc - h * BNMAX;
        if (c0 >= p.bn) continue;
        const long long row = (long long)m_blk * BMT + h * BM + quad * 32 + lane;
        const bool row_ok = row < gm;
        const int n0 = n_blk * p.bn + c0;
        if (n0 >= gn) continue;
        uint32_t v[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * MH * BNMAX + hc);
        tmem_ld32(taddr, v);
        tmem_ld_wait();
        const int lim = min(32, min(p.bn - c0, gn - n0));   // valid columns in this chunk
        if (WGRAD && lim == 32 && (p.ldd & 3) == 0 && (p.strideD & 3) == 0 &&
            (reinterpret_cast<uintptr_t>(p.D) & 15) == 0) {
          // Transpose through shared memory: a thread owns an accumulator row, so direct reductions put
          // 32 rows x 16 bytes into every instruction; staged, an instruction covers four complete
          // 128-byte row segments.
          float* sm = stage_out + quad * (32 * 36);
#pragma unroll
          for (int q = 0; q < 8; ++q)
            *reinterpret_cast<float4*>(sm + lane * 36 + 4 * q) =
                make_float4(__uint_as_float(v[4 * q]), __uint_as_float(v[4 * q + 1]), __uint_as_float(v[4 * q + 2]),
                            __uint_as_float(v[4 * q + 3]));
          __syncwarp();
          const int rsub = lane >> 3, cq = (lane & 7) * 4;
          const long long row0 = (long long)m_blk * BMT + h * BM + quad * 32;
          float* dbase = reinterpret_cast<float*>(p.D) + (long long)g * p.strideD + n0 + cq;
          const bool col_ok = !(p.mask_period && (n0 + cq) % p.mask_period >= p.mask_valid);   // structural zeros (4-aligned)
#pragma unroll
          for (int it = 0; it < 8; ++it) {
            const int r = it * 4 + rsub;
            if (col_ok && row0 + r < gm) {
              const float4 o4 = *reinterpret_cast<const float4*>(sm + r * 36 + cq);
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dbase + (row0 + r) * p.ldd), "f"(o4.x),
                           "f"(o4.y), "f"(o4.z), "f"(o4.w)
                           : "memory");
            }
          }
          __syncwarp();
          continue;
        }
        if (!row_ok) continue;
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
        long long pix_off = row * p.ldd;
        if (!WGRAD && p.o_w) {
          const long long img = row / p.o_hw;
          const int rr = (int)(row - img * p.o_hw);
          const int jj = rr / p.o_w;
          if (p.o_vw && (jj >= p.o_vh || rr - jj * p.o_w >= p.o_vw)) continue;     // outside the valid part of the grid
          pix_off = p.o_base + img * p.o_img + (long long)jj * p.o_row + (long long)(rr - jj * p.o_w) * p.o_pix;
        }
        const long long off = (long long)g * p.strideD + pix_off + n0;
        if (!WGRAD) {
          if (bias != nullptr) {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (j < lim) f[j] += __ldg(bias + n0 + j);
          }
          if (p.relu) {
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.0f);
          }
          if (p.mask != nullptr) {
            const bf16* mk = p.mask + off;
            if (lim == 32 && ((reinterpret_cast<uintptr_t>(mk) & 15) == 0)) {
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                const uint4 mv = *reinterpret_cast<const uint4*>(mk + 8 * q);
                const __nv_bfloat162* mh2 = reinterpret_cast<const __nv_bfloat162*>(&mv);
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                  const float2 m2 = __bfloat1622float2(mh2[t]);
                  if (!(m2.x > 0.0f)) f[q * 8 + 2 * t] = 0.0f;
                  if (!(m2.y > 0.0f)) f[q * 8 + 2 * t + 1] = 0.0f;
                }
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (j < lim && !(__bfloat162float(mk[j]) > 0.0f)) f[j] = 0.0f;
            }
          }
          bf16* d = reinterpret_cast<bf16*>(p.D) + off;
          if (lim == 32 && ((reinterpret_cast<uintptr_t>(d) & 15) == 0)) {
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
              if (j < lim) d[j] = __float2bfloat16(f[j]);
          }
        } else {
          float* d = reinterpret_cast<float*>(p.D) + off;
          if (lim == 32 && ((reinterpret_cast<uintptr_t>(d) & 15) == 0)) {
            // 128-bit reductions: a quarter of the L2 atomic operations of the scalar form
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              if (p.mask_period && (n0 + 4 * q) % p.mask_period >= p.mask_valid) continue;   // structural zeros (4-aligned)
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(d + 4 * q), "f"(f[4 * q]),
                           "f"(f[4 * q + 1]), "f"(f[4 * q + 2]), "f"(f[4 * q + 3])
                           : "memory");
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (j < lim && !(p.mask_period && (n0 + j) % p.mask_period >= p.mask_valid)) atomicAdd(d + j, f[j]);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar(acc));
      if (prof_on) pw1 += clock64() - tbody;
      if (++acc == kAccStages) { acc = 0; acc_phase ^= 1u; }
    }
    if (quad == 2) {   // warp 2
      prof_store(8, pw0);
      prof_store(9, pw1);
      prof_store(10, prof_on ? clock64() - prof_t0 : 0);
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
// Cluster variant of the forward kernel (opt-in: cxn_conv_tune(7, 2) or CXN_CONV_CLUSTER=2).
// Two CTAs of a cluster work on two different 128/256-row pixel tiles of the SAME output columns, so they
// need the same filter tile: each fetches one half of it (bn/2 filter rows x 64) and multicasts it into
// both CTAs' rings - the filter crosses L2->SM once per pair instead of once per CTA.  The patches come
// by im2col-mode TMA into the CTA's own ring (no gather warps: 6 warps).  The tensor map of the multicast
// load covers this GPU's weights only.
//
// Protocol per ring stage: full barrier (1 arrival + bytes) = own patch tile + both filter halves;
// empty barrier (2 arrivals) = tcgen05.commit of BOTH CTAs' MMA warps (multicast commit), because a
// producer writes into its peer's stage as well as its own.  Both CTAs run the same (n, group, k) loop,
// so the barriers of a pair advance in lock step; every wait is bounded (mbar_wait traps).
constexpr int kClusterThreads = 192;   // warp 0 TMA | warp 1 MMA + TMEM | warps 2-5 epilogue

template <int BNMAX, int MH>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kClusterThreads, 1)
conv_fprop_cluster_kernel(const __grid_constant__ CUtensorMap tmHalf, const __grid_constant__ CUtensorMap tmPatch,
                          const Params p) {
  using C = Cfg<BNMAX, MH>;
  constexpr int kAccStages = C::kAccStages;
  constexpr int BMT = BM * MH;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const int nst = p.stages;
  const uint32_t bar_base = smem_base + (uint32_t)p.ring_bytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (nst + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * nst + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * nst + kAccStages + a); };
  volatile uint32_t* tmem_ptr_smem =
      reinterpret_cast<volatile uint32_t*>(smem_gen + p.ring_bytes + 8 * (2 * nst + 2 * kAccStages + 1));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t crank = cluster_ctarank();
  const long long rows = (long long)p.N * p.OH * p.OW;
  const int tiles_m = (int)ceil_div_ll(rows, BMT);
  const int pairs_m = (tiles_m + 1) >> 1;
  const int tiles_n = ceil_div(p.Ng, p.bn);
  const int num_kb = ceil_div(p.Kg, BK);
  const long long total_pairs = (long long)pairs_m * tiles_n * p.G;
  const long long cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
  const int half_rows = p.bn >> 1;               // filter rows each CTA of the pair fetches
  const int cpk = p.Cg >> 6;

  if (threadIdx.x == 0) {
    for (int s = 0; s < nst; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 2);
    }
    for (int a = 0; a < kAccStages; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), 4);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmHalf) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmPatch) : "memory");
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
  cluster_sync_all();          // the peer's barriers exist before anything of ours can reach them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  auto decode_pair = [&](long long t, int& n_blk, int& m_blk, int& g) {
    n_blk = (int)(t % tiles_n);
    const long long r = t / tiles_n;
    m_blk = 2 * (int)(r % pairs_m) + (int)crank;
    g = (int)(r / pairs_m);
  };

  if (warp == 0) {
    // =========================== TMA producer ===========================
    int stage = 0;
    uint32_t phase = 0;
    const int ohw = p.OH * p.OW;
    for (long long t = cluster_id; t < total_pairs; t += num_clusters) {
      int n_blk, m_blk, g;
      decode_pair(t, n_blk, m_blk, g);
      // an odd tile count leaves the last pair with one real tile: its partner recomputes that tile
      // (same loads, so the pair stays in lock step) and stores nothing
      const int m_ld = m_blk < tiles_m ? m_blk : tiles_m - 1;
      const long long m0 = (long long)m_ld * BMT;
      const int pn = (int)(m0 / ohw);
      const int r = (int)(m0 - (long long)pn * ohw);
      const int oh = r / p.OW;
      const int pw = (r - oh * p.OW) * p.stride_x - p.pad_x;
      const int ph = oh * p.stride_y - p.pad_y;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait_warp(empty_bar(stage), phase ^ 1u);
        const uint32_t sa = smem_base + stage * (uint32_t)C::kStageBytes;
        const uint32_t sb = sa + C::kABytes;
        if (elect_one()) {
          mbar_expect_tx(full_bar(stage), (uint32_t)C::kABytes + (uint32_t)p.bn * BK * 2);
          const int tap = kb / cpk, chunk = kb - tap * cpk;
          const int kh = tap / p.KW, kw = tap - kh * p.KW;
          tma_load_im2col_4d(sa, &tmPatch, full_bar(stage), g * p.Cgs + chunk * 64, pw, ph, pn, kw, kh);
          tma_load_3d_mc(sb + crank * (uint32_t)half_rows * 128u, &tmHalf, full_bar(stage), kb * BK,
                         n_blk * p.bn + (int)crank * half_rows, g, (uint16_t)0x3);
        }
        __syncwarp();
        if (++stage == nst) { stage = 0; phase ^= 1u; }
      }
    }
    // tail: every stage we (or the peer) filled has been released by BOTH MMA warps before this CTA may
    // leave - the peer's last commits target our barriers
    for (int i = 0; i < nst; ++i) {
      mbar_wait_warp(empty_bar(stage), phase ^ 1u);
      if (++stage == nst) { stage = 0; phase ^= 1u; }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.bn >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    constexpr uint32_t kSbo = 1024;
    constexpr uint32_t kStep = UMMA_K * 2;
    const uint64_t desc_hi = make_desc(0, 0u, kSbo);
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (long long t = cluster_id; t < total_pairs; t += num_clusters) {
      mbar_wait_warp(tempty_bar(acc), acc_phase ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * MH * BNMAX);
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait_warp(full_bar(stage), phase);
        tc_fence_after();
        const uint32_t sa = smem_base + stage * (uint32_t)C::kStageBytes;
        const uint32_t sb = sa + C::kABytes;
        if (elect_one()) {
          const uint64_t a0 = desc_hi | (uint64_t)((sa >> 4) & 0x3FFF), b0 = desc_hi | (uint64_t)((sb >> 4) & 0x3FFF);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
#pragma unroll
            for (int h = 0; h < MH; ++h)
              umma_bf16(d_tmem + (uint32_t)(h * BNMAX), a0 + (uint64_t)((h * C::kAHalfBytes + k * kStep) >> 4),
                        b0 + (uint64_t)((k * kStep) >> 4), idesc, (kb > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit_mc(empty_bar(stage), (uint16_t)0x3);     // frees this stage in both CTAs
          if (kb == num_kb - 1) umma_commit(tfull_bar(acc));
        }
        __syncwarp();
        if (++stage == nst) { stage = 0; phase ^= 1u; }
      }
      if (++acc == kAccStages) { acc = 0; acc_phase ^= 1u; }
    }
  } else {
    // =========================== epilogue ===========================
    const int quad = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (long long t = cluster_id; t < total_pairs; t += num_clusters) {
      int n_blk, m_blk, g;
      decode_pair(t, n_blk, m_blk, g);
      mbar_wait_warp(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const float* bias = p.bias != nullptr ? p.bias + (long long)g * p.Ng : nullptr;
#pragma unroll 1
      for (int hc = 0; hc < MH * BNMAX; hc += 32) {
        const int h = hc / BNMAX, c0 = hc - h * BNMAX;
        if (c0 >= p.bn) continue;
        const long long row = (long long)m_blk * BMT + h * BM + quad * 32 + lane;
        const int n0 = n_blk * p.bn + c0;
        if (n0 >= p.Ng) continue;
        uint32_t v[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * MH * BNMAX + hc);
        tmem_ld32(taddr, v);
        tmem_ld_wait();
        if (row >= rows) continue;
        const int lim = min(32, min(p.bn - c0, p.Ng - n0));
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
        const long long off = (long long)g * p.strideD + row * p.ldd + n0;
        if (bias != nullptr) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j < lim) f[j] += __ldg(bias + n0 + j);
        }
        if (p.relu) {
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.0f);
        }
        if (p.mask != nullptr) {
          const bf16* mk = p.mask + off;
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j < lim && !(__bfloat162float(mk[j]) > 0.0f)) f[j] = 0.0f;
        }
        bf16* d = reinterpret_cast<bf16*>(p.D) + off;
        if (lim == 32 && ((reinterpret_cast<uintptr_t>(d) & 15) == 0)) {
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
            if (j < lim) d[j] = __float2bfloat16(f[j]);
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
  cluster_sync_all();          // nobody leaves while the peer may still signal our barriers or fill our ring
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(C::kTmemCols)
                 : "memory");
  }
}

// ===================================================================================================
// CTA-pair variant of the forward kernel (tcgen05 cta_group::2; cxn_conv_tune(10, 1) or CXN_CONV_PAIR2=1).
// The large stride-1 layers run at the chip-wide L2 -> SM throughput ceiling (profiles/): per k-block a 128 x bn
// tile pulls 16 KB of patches and bn x 128 B of filter into its SM.  Here the two CTAs of a cluster (the two SMs of
// a TPC) compute ONE 256-row x bn tile with M = 256 MMAs: each CTA fetches its own 128 output positions (im2col TMA)
// and HALF of the filter tile (bn / 2 rows), the tensor core reads both halves across the pair, and each CTA
// receives its 128 rows of the accumulator in its own tensor memory.  Filter bytes per SM halve; nothing is
// multicast.  Protocol as gemm_pair_kernel (gemm_sm100.cu): full barriers in the leader armed with both CTAs'
// bytes, empty / tfull by multicast commit, tempty in the leader with arrivals from both CTAs' epilogue warps.
template <int BNMAX>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kClusterThreads, 1)
conv_fprop_pair_kernel(const __grid_constant__ CUtensorMap tmHalf, const __grid_constant__ CUtensorMap tmPatch,
                       const Params p) {
  constexpr int kAccStages = 2;
  constexpr int kABytes = BM * BK * 2;
  constexpr int kBHalfBytes = (BNMAX / 2) * BK * 2;
  constexpr int kStageBytes = kABytes + kBHalfBytes;
  constexpr int kTmemCols = kAccStages * BNMAX;
  constexpr int BMT = 2 * BM;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const int nst = p.stages;
  const uint32_t bar_base = smem_base + (uint32_t)p.ring_bytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (nst + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * nst + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * nst + kAccStages + a); };
  volatile uint32_t* tmem_ptr_smem =
      reinterpret_cast<volatile uint32_t*>(smem_gen + p.ring_bytes + 8 * (2 * nst + 2 * kAccStages + 1));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t crank = cluster_ctarank();
  const bool leader = crank == 0;
  const long long rows = (long long)p.N * p.OH * p.OW;
  const int tiles_m = (int)ceil_div_ll(rows, BMT);
  const int tiles_n = ceil_div(p.Ng, p.bn);
  const int num_kb = ceil_div(p.Kg, BK);
  const long long total_tiles = (long long)tiles_m * tiles_n * p.G;
  const long long pair0 = blockIdx.x >> 1, npairs = gridDim.x >> 1;
  const int half_rows = p.bn >> 1;               // filter rows each CTA stages
  const int cpk = p.Cg >> 6;

  if (threadIdx.x == 0) {
    for (int s = 0; s < nst; ++s) {
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
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmHalf) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmPatch) : "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32((const void*)tmem_ptr_smem)),
                 "n"(kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  auto decode_tile = [&](long long t, int& n_blk, int& m_blk, int& g) {
    n_blk = (int)(t % tiles_n);
    const long long r = t / tiles_n;
    m_blk = (int)(r % tiles_m);
    g = (int)(r / tiles_m);
  };

  if (warp == 0) {
    // =========================== TMA producer (both CTAs) ===========================
    int stage = 0;
    uint32_t phase = 0;
    const int ohw = p.OH * p.OW;
    for (long long t = pair0; t < total_tiles; t += npairs) {
      int n_blk, m_blk, g;
      decode_tile(t, n_blk, m_blk, g);
      // first output position of this CTA's 128 rows -> (w, h, n) of its tap-(0,0) input pixel; a position past the
      // last image reads out of bounds = zeros
      const long long m0 = (long long)m_blk * BMT + (long long)crank * BM;
      const int pn = (int)(m0 / ohw);
      const int r = (int)(m0 - (long long)pn * ohw);
      const int oh = r / p.OW;
      const int pw = (r - oh * p.OW) * p.stride_x - p.pad_x;
      const int ph = oh * p.stride_y - p.pad_y;
      int chunk = 0, kw = 0, kh = 0;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait_warp(empty_bar(stage), phase ^ 1u);
        const uint32_t sa = smem_base + stage * (uint32_t)kStageBytes;
        const uint32_t sb = sa + kABytes;
        if (elect_one()) {
          if (leader) mbar_expect_tx(full_bar(stage), 2u * ((uint32_t)kABytes + (uint32_t)half_rows * BK * 2));
          tma_load_im2col_4d_2cta(sa, &tmPatch, full_bar(stage), g * p.Cgs + chunk * 64, pw, ph, pn, kw, kh);
          tma_load_3d_2cta(sb, &tmHalf, full_bar(stage), kb * BK, n_blk * p.bn + (int)crank * half_rows, g);
        }
        if (++chunk == cpk) {       // next 64-channel chunk of the tap, then the next tap (no divisions in this loop)
          chunk = 0;
          if (++kw == p.KW) { kw = 0; ++kh; }
        }
        __syncwarp();
        if (++stage == nst) { stage = 0; phase ^= 1u; }
      }
    }
    for (int i = 0; i < nst; ++i) {     // tail: the multicast commits for everything this CTA staged have arrived
      mbar_wait_warp(empty_bar(stage), phase ^ 1u);
      if (++stage == nst) { stage = 0; phase ^= 1u; }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer (leader only) ===========================
    if (leader) {
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.bn >> 3) << 17) | ((uint32_t)(BMT >> 4) << 24);
      constexpr uint32_t kSbo = 1024;
      constexpr uint32_t kStep = UMMA_K * 2;
      const uint64_t desc_hi = make_desc(0, 0u, kSbo);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (long long t = pair0; t < total_tiles; t += npairs) {
        mbar_wait_warp(tempty_bar(acc), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BNMAX);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait_warp(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sa = smem_base + stage * (uint32_t)kStageBytes;
          const uint32_t sb = sa + kABytes;
          if (elect_one()) {
            const uint64_t a0 = desc_hi | (uint64_t)((sa >> 4) & 0x3FFF), b0 = desc_hi | (uint64_t)((sb >> 4) & 0x3FFF);
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k)
              umma_bf16_2cta(d_tmem, a0 + (uint64_t)((k * kStep) >> 4), b0 + (uint64_t)((k * kStep) >> 4), idesc,
                             (kb > 0 || k > 0) ? 1u : 0u);
            umma_commit_2cta(empty_bar(stage), (uint16_t)0x3);
            if (kb == num_kb - 1) umma_commit_2cta(tfull_bar(acc), (uint16_t)0x3);
          }
          __syncwarp();
          if (++stage == nst) { stage = 0; phase ^= 1u; }
        }
        if (++acc == kAccStages) { acc = 0; acc_phase ^= 1u; }
      }
    }
  } else {
    // =========================== epilogue (both CTAs, own 128 rows) ===========================
    const int quad = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (long long t = pair0; t < total_tiles; t += npairs) {
      int n_blk, m_blk, g;
      decode_tile(t, n_blk, m_blk, g);
      mbar_wait_warp(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const float* bias = p.bias != nullptr ? p.bias + (long long)g * p.Ng : nullptr;
      const long long row = (long long)m_blk * BMT + (long long)crank * BM + quad * 32 + lane;
#pragma unroll 1
      for (int c0 = 0; c0 < BNMAX; c0 += 32) {
        if (c0 >= p.bn) break;
        const int n0 = n_blk * p.bn + c0;
        if (n0 >= p.Ng) break;
        uint32_t v[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * BNMAX + c0);
        tmem_ld32(taddr, v);
        tmem_ld_wait();
        if (row >= rows) continue;
        const int lim = min(32, min(p.bn - c0, p.Ng - n0));
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
        const long long off = (long long)g * p.strideD + row * p.ldd + n0;
        if (bias != nullptr) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j < lim) f[j] += __ldg(bias + n0 + j);
        }
        if (p.relu) {
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.0f);
        }
        if (p.mask != nullptr) {
          const bf16* mk = p.mask + off;
          if (lim == 32 && ((reinterpret_cast<uintptr_t>(mk) & 15) == 0)) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const uint4 mv = *reinterpret_cast<const uint4*>(mk + 8 * q);
              const __nv_bfloat162* mh2 = reinterpret_cast<const __nv_bfloat162*>(&mv);
#pragma unroll
              for (int tt = 0; tt < 4; ++tt) {
                const float2 m2 = __bfloat1622float2(mh2[tt]);
                if (!(m2.x > 0.0f)) f[q * 8 + 2 * tt] = 0.0f;
                if (!(m2.y > 0.0f)) f[q * 8 + 2 * tt + 1] = 0.0f;
              }
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (j < lim && !(__bfloat162float(mk[j]) > 0.0f)) f[j] = 0.0f;
          }
        }
        bf16* d = reinterpret_cast<bf16*>(p.D) + off;
        if (lim == 32 && ((reinterpret_cast<uintptr_t>(d) & 15) == 0)) {
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
            if (j < lim) d[j] = __float2bfloat16(f[j]);
        }
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
// Halo-tile variant of the forward kernel for stride-1 convolutions with few output columns (opt-in:
// cxn_conv_tune(9, 1) or CXN_CONV_HALO=1; built for conv2's backward-data problem: 5x5, 128 -> 48 channels).
// The im2col routes fetch every input element once per filter tap.  Here the GEMM rows enumerate the output over the
// PADDED-width grid (Wp = W + 2 pad_x columns per output row), where tap (kh, kw) is a constant shift of kh Wp + kw
// pixels: one tiled TMA load brings the patch of a tile - (rows + KH - 1) x Wp pixels x 64 channels, zero-filled
// outside the image - and every tap reads it through a descriptor whose start address is moved by that many
// 128-byte rows (a 128B-swizzled operand may start at any row: profiles/r1_desc_probe.txt).  A tile is MH halves of
// R = 128 / Wp output rows; the Wp - OW junk columns and the rows past the image are computed and not stored.
// Patch traffic drops by KH KW at the price of Wp / OW more tensor work; the filter tiles stream through their own
// ring, one {64 x bn} tile per (tap, 64-channel chunk).
//
// warps: 0 TMA producer | 1 MMA issuer + TMEM owner | 2-5 epilogue.
struct HaloGeom {
  int wp;        // padded width = GEMM rows per output row
  int rows_half; // output rows per 128-row half
  int tiles_img; // tiles per image
  int patch_h;   // input rows of one patch: MH * rows_half + KH - 1
  int a_stage_bytes, na, nb;   // patch stage size (1024-aligned), patch ring depth, filter ring depth
};

constexpr int kHaloThreads = 192;
constexpr int kHaloBN = 64;          // accumulator columns per half (bn <= 64)

__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2,
                                            int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst),
      "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

template <int MH>
__global__ void __launch_bounds__(kHaloThreads, 1)
conv_fprop_halo_kernel(const __grid_constant__ CUtensorMap tmDense, const __grid_constant__ CUtensorMap tmTile,
                       const Params p, const HaloGeom hg) {
  constexpr int kAcc = 2;
  constexpr int kTmemCols = kAcc * MH * kHaloBN;       // 128 or 256
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const int na = hg.na, nb = hg.nb;
  const uint32_t b_bytes = (uint32_t)p.bn * BK * 2;                 // one filter tile (bn % 8 == 0: 1024-aligned)
  const uint32_t a_base = smem_base, b_base = smem_base + (uint32_t)na * hg.a_stage_bytes;
  const uint32_t bar_base = b_base + (uint32_t)nb * ((b_bytes + 1023u) & ~1023u);
  const uint32_t b_stride = (b_bytes + 1023u) & ~1023u;
  auto fulla = [&](int s) { return bar_base + 8u * s; };
  auto emptya = [&](int s) { return bar_base + 8u * (na + s); };
  auto fullb = [&](int s) { return bar_base + 8u * (2 * na + s); };
  auto emptyb = [&](int s) { return bar_base + 8u * (2 * na + nb + s); };
  auto tfull = [&](int a) { return bar_base + 8u * (2 * na + 2 * nb + a); };
  auto tempty = [&](int a) { return bar_base + 8u * (2 * na + 2 * nb + kAcc + a); };
  volatile uint32_t* tmem_ptr_smem =
      reinterpret_cast<volatile uint32_t*>(smem_gen + (bar_base - smem_base) + 8 * (2 * na + 2 * nb + 2 * kAcc));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int cpk = p.Cg >> 6;                         // 64-channel chunks
  const int taps = p.KH * p.KW;
  const long long total_tiles = (long long)p.N * hg.tiles_img * p.G;
  const int rows_tile = MH * hg.rows_half;           // output rows per tile

  if (threadIdx.x == 0) {
    for (int s = 0; s < na; ++s) { mbar_init(fulla(s), 1); mbar_init(emptya(s), 1); }
    for (int s = 0; s < nb; ++s) { mbar_init(fullb(s), 1); mbar_init(emptyb(s), 1); }
    for (int a = 0; a < kAcc; ++a) { mbar_init(tfull(a), 1); mbar_init(tempty(a), 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmDense) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmTile) : "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32((const void*)tmem_ptr_smem)),
                 "n"(kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  auto decode_tile = [&](long long t, int& n, int& ti, int& g) {
    ti = (int)(t % hg.tiles_img);
    const long long r = t / hg.tiles_img;
    n = (int)(r % p.N);
    g = (int)(r / p.N);
  };

  if (warp == 0) {
    // =========================== TMA producer ===========================
    // items: (tile, chunk) pairs of this CTA in order; the patch of item q + 1 is requested before the filter
    // tiles of item q, so that it is in flight while q's taps run
    long long ntile_cta = 0;
    for (long long t = blockIdx.x; t < total_tiles; t += gridDim.x) ++ntile_cta;
    const long long items = ntile_cta * cpk;
    int sb = 0;
    uint32_t pb = 0;
    auto issue_a = [&](long long q) {
      const int s = (int)(q % na);
      const uint32_t ph = (uint32_t)((q / na) & 1);
      mbar_wait_warp(emptya(s), ph ^ 1u);
      if (elect_one()) {
        const long long t = blockIdx.x + (q / cpk) * (long long)gridDim.x;
        const int chunk = (int)(q % cpk);
        int n, ti, g;
        decode_tile(t, n, ti, g);
        mbar_expect_tx(fulla(s), (uint32_t)(64 * hg.wp * hg.patch_h * 2));
        tma_load_4d(a_base + (uint32_t)s * hg.a_stage_bytes, &tmTile, fulla(s), g * p.Cgs + chunk * 64, -p.pad_x,
                    ti * rows_tile - p.pad_y, n);
      }
      __syncwarp();
    };
    if (items > 0) issue_a(0);
    for (long long q = 0; q < items; ++q) {
      if (q + 1 < items) issue_a(q + 1);
      const long long t = blockIdx.x + (q / cpk) * (long long)gridDim.x;
      const int chunk = (int)(q % cpk);
      const int g = (int)(t / ((long long)hg.tiles_img * p.N));
      for (int tap = 0; tap < taps; ++tap) {
        mbar_wait_warp(emptyb(sb), pb ^ 1u);
        if (elect_one()) {
          mbar_expect_tx(fullb(sb), b_bytes);
          tma_load_3d(b_base + (uint32_t)sb * b_stride, &tmDense, fullb(sb), tap * p.Cg + chunk * 64, 0, g);
        }
        __syncwarp();
        if (++sb == nb) { sb = 0; pb ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.bn >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    const uint64_t desc_hi = make_desc(0, 0u, 1024u);
    int sb = 0;
    uint32_t pb = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    long long q = 0;
    for (long long t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      mbar_wait_warp(tempty(acc), acc_phase ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * MH * kHaloBN);
      for (int chunk = 0; chunk < cpk; ++chunk, ++q) {
        const int sa = (int)(q % na);
        mbar_wait_warp(fulla(sa), (uint32_t)((q / na) & 1));
        const uint32_t patch = a_base + (uint32_t)sa * hg.a_stage_bytes;
        for (int tap = 0; tap < taps; ++tap) {
          mbar_wait_warp(fullb(sb), pb);
          tc_fence_after();
          if (elect_one()) {
            const int kh = tap / p.KW, kw = tap - kh * p.KW;
            const uint32_t sbt = b_base + (uint32_t)sb * b_stride;
            const uint64_t b0 = desc_hi | (uint64_t)((sbt >> 4) & 0x3FFF);
#pragma unroll
            for (int h = 0; h < MH; ++h) {
              // rows of half h start (h R + kh) Wp + kw pixels into the patch; one pixel = one 128-byte row
              const uint32_t sat = patch + (uint32_t)(((h * hg.rows_half + kh) * hg.wp + kw) * 128);
              const uint64_t a0 = desc_hi | (uint64_t)((sat >> 4) & 0x3FFF);
#pragma unroll
              for (int k = 0; k < BK / UMMA_K; ++k)
                umma_bf16(d_tmem + (uint32_t)(h * kHaloBN), a0 + (uint64_t)((k * UMMA_K * 2) >> 4),
                          b0 + (uint64_t)((k * UMMA_K * 2) >> 4), idesc, (chunk > 0 || tap > 0 || k > 0) ? 1u : 0u);
            }
            umma_commit(emptyb(sb));
            if (tap == taps - 1) {
              umma_commit(emptya(sa));
              if (chunk == cpk - 1) umma_commit(tfull(acc));
            }
          }
          __syncwarp();
          if (++sb == nb) { sb = 0; pb ^= 1u; }
        }
      }
      if (++acc == kAcc) { acc = 0; acc_phase ^= 1u; }
    }
  } else {
    // =========================== epilogue ===========================
    const int quad = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    const int r = quad * 32 + lane;                 // GEMM row inside a half = position on the padded grid
    const int orow = r / hg.wp, ow = r - orow * hg.wp;
    for (long long t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      int n, ti, g;
      decode_tile(t, n, ti, g);
      mbar_wait_warp(tfull(acc), acc_phase);
      tc_fence_after();
      const float* bias = p.bias != nullptr ? p.bias + (long long)g * p.Ng : nullptr;
#pragma unroll 1
      for (int hc = 0; hc < MH * kHaloBN; hc += 32) {
        const int h = hc / kHaloBN, c0 = hc - h * kHaloBN;
        if (c0 >= p.bn || c0 >= p.Ng) continue;                        // warp-uniform
        uint32_t v[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * MH * kHaloBN + hc);
        tmem_ld32(taddr, v);
        tmem_ld_wait();
        const int oh = ti * rows_tile + h * hg.rows_half + orow;
        if (orow >= hg.rows_half || ow >= p.OW || oh >= p.OH) continue;   // junk column / row of the padded grid
        const int lim = min(32, p.Ng - c0);
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
        const long long row = ((long long)n * p.OH + oh) * p.OW + ow;
        const long long off = (long long)g * p.strideD + row * p.ldd + c0;
        if (bias != nullptr) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j < lim) f[j] += __ldg(bias + c0 + j);
        }
        if (p.relu) {
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.0f);
        }
        if (p.mask != nullptr) {
          const bf16* mk = p.mask + off;
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j < lim && !(__bfloat162float(mk[j]) > 0.0f)) f[j] = 0.0f;
        }
        bf16* d = reinterpret_cast<bf16*>(p.D) + off;
        if ((reinterpret_cast<uintptr_t>(d) & 15) == 0 && (lim & 7) == 0) {
          uint4* d4 = reinterpret_cast<uint4*>(d);
#pragma unroll
          for (int q8 = 0; q8 < 4; ++q8) {
            if (q8 * 8 < lim) {
              __nv_bfloat162 h0 = __floats2bfloat162_rn(f[q8 * 8 + 0], f[q8 * 8 + 1]);
              __nv_bfloat162 h1 = __floats2bfloat162_rn(f[q8 * 8 + 2], f[q8 * 8 + 3]);
              __nv_bfloat162 h2 = __floats2bfloat162_rn(f[q8 * 8 + 4], f[q8 * 8 + 5]);
              __nv_bfloat162 h3 = __floats2bfloat162_rn(f[q8 * 8 + 6], f[q8 * 8 + 7]);
              uint4 u;
              u.x = *reinterpret_cast<uint32_t*>(&h0);
              u.y = *reinterpret_cast<uint32_t*>(&h1);
              u.z = *reinterpret_cast<uint32_t*>(&h2);
              u.w = *reinterpret_cast<uint32_t*>(&h3);
              d4[q8] = u;
            }
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j < lim) d[j] = __float2bfloat16(f[j]);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty(acc));
      if (++acc == kAcc) { acc = 0; acc_phase ^= 1u; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kTmemCols) : "memory");
  }
}

// W[g][oc][kh][kw][ic] (ld_in) -> Wt[g][ic][KH-1-kh][KW-1-kw][oc] (ld_out): the filter of the equivalent
// forward convolution that computes backward-data of a stride-1 convolution.  Per (group, tap) this is
// the transpose of an ocg x icg block: 32 x 32 tiles through shared memory, both sides coalesced.
// (cpitch: elements between the taps of the source layout - icg, or more when its channels are padded)
__global__ void flip_weights_kernel(const bf16* __restrict__ w, bf16* __restrict__ wt, int ocg, int icg, int taps,
                                    int ld_in, int ld_out, int cpitch) {
  __shared__ bf16 tile[32][33];
  const int g = blockIdx.z / taps, t = blockIdx.z - g * taps;
  const int oc0 = blockIdx.y * 32, ic0 = blockIdx.x * 32;
  const bf16* src = w + (long long)g * ocg * ld_in + (long long)t * cpitch;
  bf16* dst = wt + (long long)g * icg * ld_out + (long long)(taps - 1 - t) * ocg;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int oc = oc0 + i, ic = ic0 + threadIdx.x;
    tile[i][threadIdx.x] = (oc < ocg && ic < icg) ? src[(long long)oc * ld_in + ic] : __float2bfloat16(0.0f);
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int ic = ic0 + i, oc = oc0 + threadIdx.x;
    if (ic < icg && oc < ocg) dst[(long long)ic * ld_out + oc] = tile[threadIdx.x][i];
  }
}

// The sub-filter of one stride phase, flipped and channel-transposed (backward-data of a strided convolution):
//   Wt[g][ic][ty][tx][oc] = W[g][oc][ph + s (KHa - 1 - ty)][pw + s (KWb - 1 - tx)][ic]        ty < KHa, tx < KWb
// blockIdx.z enumerates (group, ty, tx); per tap a 32 x 32 transpose through shared memory as above.
__global__ void flip_weights_phase_kernel(const bf16* __restrict__ w, bf16* __restrict__ wt, int ocg, int icg, int KW,
                                          int s, int ph, int pw, int KHa, int KWb, int ld_in, int ld_out) {
  __shared__ bf16 tile[32][33];
  const int taps = KHa * KWb;
  const int g = blockIdx.z / taps, t = blockIdx.z - g * taps;
  const int ty = t / KWb, tx = t - ty * KWb;
  const int kh = ph + s * (KHa - 1 - ty), kw = pw + s * (KWb - 1 - tx);
  const int oc0 = blockIdx.y * 32, ic0 = blockIdx.x * 32;
  const bf16* src = w + (long long)g * ocg * ld_in + (long long)(kh * KW + kw) * icg;
  bf16* dst = wt + (long long)g * icg * ld_out + (long long)t * ocg;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int oc = oc0 + i, ic = ic0 + threadIdx.x;
    tile[i][threadIdx.x] = (oc < ocg && ic < icg) ? src[(long long)oc * ld_in + ic] : __float2bfloat16(0.0f);
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int ic = ic0 + i, oc = oc0 + threadIdx.x;
    if (ic < icg && oc < ocg) dst[(long long)ic * ld_out + oc] = tile[threadIdx.x][i];
  }
}

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

static int encode3d(CUtensorMap* out, const void* ptr, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t s1_elems,
                    uint64_t s2_elems, uint32_t b0, uint32_t b1) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return -2;
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) || (s1_elems % 8) || (d2 > 1 && (s2_elems % 8))) {
    fprintf(stderr, "[cxxnet_b200] conv operand not 16B aligned (ptr=%p s1=%llu s2=%llu)\n", ptr,
            (unsigned long long)s1_elems, (unsigned long long)s2_elems);
    return -3;
  }
  cuuint64_t gdim[3] = {d0, d1, d2};
  cuuint64_t gstr[2] = {s1_elems * 2, (d2 > 1 ? s2_elems : s1_elems * d1) * 2};
  cuuint32_t box[3] = {b0, b1, 1}, estr[3] = {1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    fprintf(stderr, "[cxxnet_b200] conv cuTensorMapEncodeTiled failed: %d\n", (int)r);
    return -4;
  }
  return 0;
}

// tuning knobs (cxn_conv_tune): 0 = automatic
static int g_tune_mh = 0;          // 1 / 2: force the number of 128-row halves per forward tile
static int g_tune_stages = 0;      // force the ring depth
static int g_tune_bnmax = 0;       // 128 / 256: force the forward tile-width class
static int g_tune_dbg = 0;         // timing experiments (Params::dbg)
static int g_tune_atma = -1;       // 0 / 1: fetch the patches with im2col-mode TMA
static int g_tune_cluster = -1;    // 2: forward layers on the im2col-TMA route run as CTA pairs that share the
                                   // filter tile by multicast (conv_fprop_cluster_kernel); -1 = read CXN_CONV_CLUSTER
static int g_tune_fill = -1;       // 1: forward layers with fewer tiles than SMs take narrower column tiles until
                                   // the grid fills the machine (small per-GPU batches); -1 = read CXN_CONV_FILL
static int g_tune_halo = -1;       // 1: stride-1 forward problems with <= 64 output columns per group run as halo
                                   // tiles (conv_fprop_halo_kernel); 2: the same with single-half tiles; -1 = CXN_CONV_HALO
static long long* g_prof = nullptr;

typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeIm2colFn get_encode_im2col() {
  static EncodeIm2colFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* q = nullptr;
    cudaDriverEntryPointQueryResult r;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &q, cudaEnableDefault, &r) == cudaSuccess &&
        r == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeIm2colFn>(q);
  });
  return fn;
}

// Can the patch operand of this problem come through an im2col tensor map?
static bool atma_ok(const Params& p) {
  if (p.gran != 8 || p.Cg % 64 || p.KH * p.KW < 1) return false;
  if (p.pad_x > 127 || p.pad_y > 127 || p.KW - 1 - p.pad_x > 127 || p.KH - 1 - p.pad_y > 127) return false;
  if (p.stride_x < 1 || p.stride_x > 8 || p.stride_y < 1 || p.stride_y > 8) return false;
  // the map enumerates floor((W + 2 pad - KW) / stride) + 1 positions per row: must be the caller's grid
  if ((p.W + 2 * p.pad_x - p.KW) / p.stride_x + 1 != p.OW || (p.H + 2 * p.pad_y - p.KH) / p.stride_y + 1 != p.OH) return false;
  if (p.W + 2 * p.pad_x < p.KW || p.H + 2 * p.pad_y < p.KH) return false;
  if ((reinterpret_cast<uintptr_t>(p.src) & 15) || (p.Cs % 8)) return false;
  // (a view of ONE group may expose fewer channels than its last 64-channel chunk: the tensor map zero-fills the rest)
  if (p.view_cdim && ((p.view_row % 8) || (p.view_img % 8) || (p.view_cdim % 8) || (p.Cgs % 8) ||
                      p.view_cdim + 63 < (long long)(p.G - 1) * p.Cgs + p.Cg))
    return false;
  return get_encode_im2col() != nullptr;
}

// NHWC activation tensor as an im2col tensor map: `pixels` output positions x 64 channels per load
static int encode_im2col(CUtensorMap* out, const Params& p, int pixels) {
  EncodeIm2colFn enc = get_encode_im2col();
  if (!enc) return -2;
  cuuint64_t gdim[4] = {(cuuint64_t)p.Cs, (cuuint64_t)p.W, (cuuint64_t)p.H, (cuuint64_t)p.N};
  cuuint64_t gstr[3] = {(cuuint64_t)p.Cs * 2, (cuuint64_t)p.W * p.Cs * 2, (cuuint64_t)p.H * p.W * p.Cs * 2};
  if (p.view_cdim) {
    // positions may overlap in memory (channel extent larger than the stride between positions)
    gdim[0] = (cuuint64_t)p.view_cdim;
    gstr[1] = (cuuint64_t)p.view_row * 2;
    gstr[2] = (cuuint64_t)p.view_img * 2;
  }
  // bounding box of the tap-(0,0) pixel: [-pad, extent + pad - (K-1))
  int lower[2] = {-p.pad_x, -p.pad_y};
  int upper[2] = {p.pad_x - (p.KW - 1), p.pad_y - (p.KH - 1)};
  cuuint32_t estr[4] = {1, (cuuint32_t)p.stride_x, (cuuint32_t)p.stride_y, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<bf16*>(p.src), gdim, gstr, lower, upper, 64,
                   (cuuint32_t)pixels, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    fprintf(stderr, "[cxxnet_b200] cuTensorMapEncodeIm2col failed: %d\n", (int)r);
    return -5;
  }
  return 0;
}

template <int BNMAX, bool WGRAD, int GRAN, int MH, bool ATMA>
static int launch_g(const CUtensorMap& tm, const CUtensorMap& tmp, const Params& p, int grid, cudaStream_t st) {
  using C = Cfg<BNMAX, MH>;
  if (p.stages < 2 || p.stages > C::kMaxStages) return -20;
  if (p.ring_bytes < p.stages * (p.b_resident ? C::kABytes : C::kStageBytes) || (p.ring_bytes & 1023)) return -22;
  const int smem = p.ring_bytes + C::kFixedBytes;
  if (smem > 232448) return -23;
  static int max_set = 0;
  if (smem > max_set) {
    cudaError_t e = cudaFuncSetAttribute(conv_igemm_kernel<BNMAX, WGRAD, GRAN, MH, ATMA>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return (int)e;
    max_set = smem;
  }
  conv_igemm_kernel<BNMAX, WGRAD, GRAN, MH, ATMA><<<grid, kThreads, smem, st>>>(tm, tmp, p);
  CXN_CHECK_LAUNCH();
  return 0;
}
template <int BNMAX, int MH>
static int launch_cluster(const CUtensorMap& tm_half, const CUtensorMap& tmp, const Params& p, int grid, cudaStream_t st) {
  using C = Cfg<BNMAX, MH>;
  if (p.stages < 2 || p.stages > C::kMaxStages || (grid & 1)) return -20;
  if (p.ring_bytes < p.stages * C::kStageBytes || (p.ring_bytes & 1023)) return -22;
  const int smem = p.ring_bytes + C::kFixedBytes;
  if (smem > 232448) return -23;
  static int max_set = 0;
  if (smem > max_set) {
    cudaError_t e = cudaFuncSetAttribute(conv_fprop_cluster_kernel<BNMAX, MH>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return (int)e;
    max_set = smem;
  }
  conv_fprop_cluster_kernel<BNMAX, MH><<<grid, kClusterThreads, smem, st>>>(tm_half, tmp, p);
  CXN_CHECK_LAUNCH();
  return 0;
}
template <int BNMAX>
static int launch_pair(const CUtensorMap& tm_half, const CUtensorMap& tmp, const Params& p, int grid, cudaStream_t st) {
  constexpr int kStageBytes = BM * BK * 2 + (BNMAX / 2) * BK * 2;
  if (p.stages < 2 || p.stages > 8 || (grid & 1)) return -20;
  if (p.ring_bytes < p.stages * kStageBytes || (p.ring_bytes & 1023)) return -22;
  const int smem = p.ring_bytes + 1024 + 256;
  if (smem > 232448) return -23;
  static int max_set = 0;
  if (smem > max_set) {
    cudaError_t e = cudaFuncSetAttribute(conv_fprop_pair_kernel<BNMAX>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return (int)e;
    max_set = smem;
  }
  conv_fprop_pair_kernel<BNMAX><<<grid, kClusterThreads, smem, st>>>(tm_half, tmp, p);
  CXN_CHECK_LAUNCH();
  return 0;
}
static int g_tune_flat = -1;       // forward problems without padding over the input's pixel grid with tiled-mode loads
                                   // (Params::flat); -1 = CXN_CONV_FLAT, default 0
static long long g_flat_launches = 0;      // forward launches that took the FLAT route (cxn_conv_flat_launches: tests)
static int flat_mode() {
  if (g_tune_flat < 0) {
    const char* e = getenv("CXN_CONV_FLAT");
    g_tune_flat = e ? atoi(e) : 0;
  }
  return g_tune_flat;
}
static thread_local int t_view_gstride = 0;   // group stride of the NEXT *_view call of this thread (cxn_conv_view_group_stride)
static int g_tune_pair = -1;       // 1: CTA pairs (cta_group::2) for forward problems on the im2col-TMA route with
                                   // more than 128 output columns per group; -1 = read CXN_CONV_PAIR2
static int pair_mode() {
  if (g_tune_pair < 0) {
    const char* e = getenv("CXN_CONV_PAIR2");
    g_tune_pair = e ? atoi(e) : 1;
  }
  return g_tune_pair;
}
// NHWC tensor as a tiled 4-D map: box {64 channels, box_w pixels, box_h rows, 1 image}, zero fill outside
static int encode_tile4d(CUtensorMap* out, const Params& p, int box_w, int box_h) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return -2;
  if (box_w > 256 || box_h > 256 || box_w < 1 || box_h < 1) return -6;
  cuuint64_t gdim[4] = {(cuuint64_t)p.Cs, (cuuint64_t)p.W, (cuuint64_t)p.H, (cuuint64_t)p.N};
  cuuint64_t gstr[3] = {(cuuint64_t)p.Cs * 2, (cuuint64_t)p.W * p.Cs * 2, (cuuint64_t)p.H * p.W * p.Cs * 2};
  cuuint32_t box[4] = {64, (cuuint32_t)box_w, (cuuint32_t)box_h, 1}, estr[4] = {1, 1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<bf16*>(p.src), gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    fprintf(stderr, "[cxxnet_b200] cuTensorMapEncodeTiled (4-D) failed: %d\n", (int)r);
    return -5;
  }
  return 0;
}
static int halo_mode() {
  if (g_tune_halo < 0) {
    const char* e = getenv("CXN_CONV_HALO");
    g_tune_halo = e ? atoi(e) : 0;
  }
  return g_tune_halo;
}
// Does the problem fit the halo-tile kernel?  Fills the geometry for `mh` halves per tile.
static bool halo_plan(const Params& p, int mh, HaloGeom* hg, int* smem_bytes) {
  if (p.stride_x != 1 || p.stride_y != 1 || p.gran != 8 || p.Cg % 64 || p.view_cdim) return false;
  if (p.KH * p.KW < 4 || p.Ng > kHaloBN || p.bn > kHaloBN || (p.bn & 7)) return false;
  const int wp = p.W + 2 * p.pad_x;
  if (wp > 128 || wp < p.OW + p.KW - 1 || p.OW != wp - p.KW + 1 || p.OH != p.H + 2 * p.pad_y - p.KH + 1) return false;
  if ((reinterpret_cast<uintptr_t>(p.src) & 15) || (p.Cs % 8)) return false;
  hg->wp = wp;
  hg->rows_half = 128 / wp;
  if (hg->rows_half < 1) return false;
  const int rows_tile = mh * hg->rows_half;
  hg->tiles_img = ceil_div(p.OH, rows_tile);
  hg->patch_h = rows_tile + p.KH - 1;
  if (hg->patch_h > 256) return false;
  // pixels a descriptor may touch: the last half's last tap starts ((mh-1) R + KH-1) Wp + KW-1 rows in and spans 128
  int pixels = ((mh - 1) * hg->rows_half + p.KH - 1) * wp + p.KW - 1 + 128;
  if (pixels < wp * hg->patch_h) pixels = wp * hg->patch_h;
  hg->a_stage_bytes = (pixels * 128 + 1023) & ~1023;
  const int b_stride = (p.bn * BK * 2 + 1023) & ~1023;
  const int budget = 232448 - 1024 - 512;
  hg->nb = 8;
  hg->na = (budget - hg->nb * b_stride) / hg->a_stage_bytes;
  if (hg->na > 3) hg->na = 3;
  if (hg->na < 2) return false;
  *smem_bytes = hg->na * hg->a_stage_bytes + hg->nb * b_stride + 512 + 1024;
  return true;
}
template <int MH>
static int launch_halo(const CUtensorMap& tm, const CUtensorMap& tmt, const Params& p, const HaloGeom& hg, int smem, int grid,
                       cudaStream_t st) {
  static int max_set = 0;
  if (smem > max_set) {
    cudaError_t e = cudaFuncSetAttribute(conv_fprop_halo_kernel<MH>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return (int)e;
    max_set = smem;
  }
  conv_fprop_halo_kernel<MH><<<grid, kHaloThreads, smem, st>>>(tm, tmt, p, hg);
  CXN_CHECK_LAUNCH();
  return 0;
}
static int cluster_mode() {
  if (g_tune_cluster < 0) {
    const char* e = getenv("CXN_CONV_CLUSTER");
    g_tune_cluster = e ? atoi(e) : 0;
  }
  return g_tune_cluster;
}
// atma: the patch operand comes through `tmp` (im2col map), otherwise through the gather warps
template <int BNMAX, bool WGRAD, int MH>
static int launch(const CUtensorMap& tm, const CUtensorMap& tmp, bool atma, const Params& p, int grid, cudaStream_t st) {
  if (p.gran == 8) {
    if (atma) return launch_g<BNMAX, WGRAD, 8, MH, true>(tm, tmp, p, grid, st);
    return launch_g<BNMAX, WGRAD, 8, MH, false>(tm, tmp, p, grid, st);
  }
  if (MH == 1 && !atma) return launch_g<BNMAX, WGRAD, 4, 1, false>(tm, tmp, p, grid, st);
  return -21;
}

static int check_geom(const Params& p) {
  if (p.KH > 16 || p.KW > 16) return -10;
  if (p.gran != 8 && p.gran != 4) return -11;
  if (p.gran == 4 && p.Cg != 4) return -12;
  if (p.Cg % p.gran) return -13;
  if (p.Kg / p.gran > kMaxGranules) return -14;
  if (p.Cs % p.gran || (reinterpret_cast<uintptr_t>(p.src) & (p.gran * 2 - 1))) return -15;
  return 0;
}

// pick the tile width: the smallest multiple of 16 that covers n in the fewest tiles
static int pick_bn(int n, int bnmax) {
  const int tiles = ceil_div(n, bnmax);
  int bn = ceil_div(ceil_div(n, tiles), 16) * 16;
  return bn < 16 ? 16 : bn;
}

}  // namespace igemm
}  // namespace cxn

using namespace cxn;
using namespace cxn::igemm;

// Forward convolution (and stride-1 backward-data with flipped filters).
//   src  NHWC bf16 [N][H][W][Cs]; group g reads channels [g*Cg, (g+1)*Cg)   (gran 8: Cg % 8 == 0;
//        gran 4: single group, Cg = Cs = 4 = RGB padded to 4)
//   w    bf16 [G][Ng][ldw >= KH*KW*Cg] with the reduction ordered (kh, kw, c)
//   y    NHWC bf16 [N*OH*OW][ldy], group g writes columns [g*Ng, (g+1)*Ng)
// CTA cap of this thread's weight-gradient launches (cxn_conv_wgrad_grid_cap); 0 = the whole GPU
static thread_local int t_wgrad_cap = 0;

// output view of the next conv_fprop_impl call of this thread (set by cxn_conv_dgrad_strided)
struct OutView { int w, h; long long base, pix, row, img; };
static thread_local OutView t_oview = {0, 0, 0, 0, 0, 0};

static int conv_fprop_impl(const void* src, int N, int H, int W, int Cs, int Cg, int G, int gran, const void* w,
                           long long ldw, int Ng, int KH, int KW, int stride, int pad_y, int pad_x, int OH, int OW,
                           void* y, long long ldy, const float* bias, int relu, const void* mask, long long view_cdim,
                           long long view_row, long long view_img, cudaStream_t st) {
  Params p;
  memset(&p, 0, sizeof(p));
  p.view_cdim = view_cdim; p.view_row = view_row; p.view_img = view_img;
  p.src = (const bf16*)src; p.N = N; p.H = H; p.W = W; p.Cs = Cs; p.Cg = Cg; p.KH = KH; p.KW = KW;
  p.Cgs = (view_cdim > 0 && t_view_gstride > 0) ? t_view_gstride : Cg;
  t_view_gstride = 0;
  p.stride_y = stride & 0xFFFF; p.stride_x = (stride >> 16) ? (stride >> 16) : p.stride_y; p.pad_y = pad_y; p.pad_x = pad_x; p.OH = OH; p.OW = OW; p.G = G; p.Kg = KH * KW * Cg;
  p.gran = gran; p.Ng = Ng; p.D = y; p.ldd = ldy; p.strideD = Ng; p.bias = bias; p.relu = relu;
  p.mask = (const bf16*)mask;
  p.splits = 1; p.kb_per_split = 0;
  const bool oview = t_oview.w > 0;
  if (oview) {
    if (t_oview.w != OW || t_oview.h != OH) return -19;
    p.o_w = OW; p.o_hw = OH * OW; p.o_base = t_oview.base; p.o_pix = t_oview.pix; p.o_row = t_oview.row; p.o_img = t_oview.img;
  }
  int rc = check_geom(p);
  if (rc) return rc;
  // FLAT route (see Params::flat): stride 1, no padding, few output columns per group (such layers are bound by the
  // im2col-mode fetch rate), at most 15 % of discarded rows
  bool flat = false;
  if (flat_mode() && !oview && !p.view_cdim && gran == 8 && p.stride_x == 1 && p.stride_y == 1 && pad_x == 0 && pad_y == 0 &&
      Ng <= 128 && (long long)H * W * 100 <= (long long)OH * OW * 115 && (long long)N * H * W < (1LL << 31) - 65536) {
    bool atma_e = atma_ok(p);
    if (g_tune_atma >= 0) atma_e = atma_e && g_tune_atma;
    flat = atma_e;
  }
  if (flat) {
    ++g_flat_launches;
    p.flat = 1;
    p.o_w = W; p.o_hw = H * W; p.o_base = 0; p.o_pix = ldy; p.o_row = (long long)OW * ldy; p.o_img = (long long)OH * OW * ldy;
    p.o_vh = OH; p.o_vw = OW;
  }
  const long long rows = flat ? (long long)N * H * W : (long long)N * OH * OW;
  if (rows <= 0) return 0;
  // Tile shape: row-paired tiles (256 x <=128, two accumulators sharing the filter tile) when one
  // n-tile covers the group's outputs and there is enough work to fill the machine with them -
  // the filter tile is then read half as often; otherwise 128 x <=256.
  int mh = 1, bnmax = Ng > 128 ? 256 : 128;
  if (gran == 8 && Ng <= 128) {
    const long long t2 = ceil_div_ll(rows, 2 * BM) * G;
    if (t2 >= 2LL * num_sms()) mh = 2;
  }
  if (g_tune_mh && gran == 8) { mh = g_tune_mh; if (mh == 1) bnmax = Ng > 128 ? 256 : 128; }
  if (g_tune_bnmax) bnmax = g_tune_bnmax;
  p.bn = pick_bn(Ng, bnmax);
  if (g_tune_fill < 0) {
    const char* e = getenv("CXN_CONV_FILL");
    g_tune_fill = e ? atoi(e) : 0;
  }
  if (g_tune_fill == 1) {
    // under one wave of tiles: halve the tile width (not below 64 columns) while that adds CTAs
    while (p.bn >= 128 && ceil_div_ll(rows, mh * BM) * ceil_div(Ng, p.bn) * G < num_sms()) {
      const int nb = pick_bn(Ng, p.bn / 2);
      if (nb >= p.bn || nb < 64) break;
      p.bn = nb;
    }
  }
  p.dbg = g_tune_dbg;
  p.prof = g_prof;
  const int stage_bytes = (mh * BM + bnmax) * BK * 2;
  int max_st = (232448 - kHostFixedBytes) / stage_bytes;
  if (max_st > 6) max_st = 6;
  p.stages = g_tune_stages ? g_tune_stages : max_st;
  CUtensorMap tm;
  rc = encode3d(&tm, w, (uint64_t)p.Kg, (uint64_t)Ng, (uint64_t)G, (uint64_t)ldw, (uint64_t)Ng * ldw, BK, p.bn);
  if (rc) return rc;
  const long long tiles = ceil_div_ll(rows, mh * BM) * ceil_div(Ng, p.bn) * G;
  const int grid = (int)(tiles < num_sms() ? tiles : num_sms());
  // Small single-group filters stay resident in shared memory for the whole (persistent) kernel: the
  // ring then holds only the patch operand.  Prefer that over row-paired tiles when it only fits with
  // 128-row tiles - no filter traffic at all beats half the filter traffic.
  p.b_resident = 0;
  p.ring_bytes = p.stages * stage_bytes;
  if (G == 1 && Ng <= p.bn && tiles > 2LL * grid) {
    const long long panel = ((long long)ceil_div(p.Kg, BK) * p.bn * BK * 2 + 1023) & ~1023LL;
    const long long budget = 232448 - kHostFixedBytes;
    for (int try_mh = mh; try_mh >= 1 && !p.b_resident; --try_mh) {
      if (g_tune_mh && try_mh != mh) break;
      const long long a_bytes = (long long)try_mh * BM * BK * 2;
      int st_fit = (int)((budget - panel) / a_bytes);
      if (st_fit > 6) st_fit = 6;
      if (g_tune_stages && st_fit > g_tune_stages) st_fit = g_tune_stages;
      if (panel < budget && st_fit >= 3) {
        mh = try_mh;
        p.stages = st_fit;
        p.b_resident = 1;
        p.ring_bytes = (int)(st_fit * a_bytes + panel);
      }
    }
  }
  const long long tiles_final = ceil_div_ll(rows, mh * BM) * ceil_div(Ng, p.bn) * G;
  const int grid_final = (int)(tiles_final < num_sms() ? tiles_final : num_sms());
  bool atma = atma_ok(p) && !oview;        // (only the generic kernel knows the output view)
  if (g_tune_atma >= 0 && !p.view_cdim) atma = atma && g_tune_atma;
  if (p.view_cdim && !atma) return -17;    // a strided view only exists as a tensor map
  CUtensorMap tmp = tm;
  if (atma && flat) {
    rc = encode3d(&tmp, src, (uint64_t)Cs, (uint64_t)rows, 1, (uint64_t)Cs, (uint64_t)rows * Cs, 64, mh * BM);
    if (rc) return rc;
  } else if (atma) {
    rc = encode_im2col(&tmp, p, mh * BM);
    if (rc) return rc;
  }
  if (flat && !atma) return -24;
  if (halo_mode() > 0 && atma && !flat) {     // (takes precedence over the resident-filter variant when asked for)
    HaloGeom hg;
    int smem_h = 0;
    const int hmh = halo_mode() == 2 ? 1 : 2;
    if (halo_plan(p, hmh, &hg, &smem_h)) {
      CUtensorMap tmt;
      rc = encode_tile4d(&tmt, p, hg.wp, hg.patch_h);
      if (rc) return rc;
      const long long htiles = (long long)N * hg.tiles_img * G;
      const int hgrid = (int)(htiles < num_sms() ? htiles : num_sms());
      return hmh == 2 ? launch_halo<2>(tm, tmt, p, hg, smem_h, hgrid, st) : launch_halo<1>(tm, tmt, p, hg, smem_h, hgrid, st);
    }
  }
  if (pair_mode() >= 1 && atma && !flat && !(p.bn & 15) && num_sms() >= 2 &&
      (pair_mode() == 2 || (!p.b_resident && Ng > 128 && ceil_div_ll(rows, 2 * BM) * ceil_div(Ng, p.bn) * G >= num_sms() / 2))) {
    // CTA pairs: one 256-row x bn tile per pair, every CTA stages its own 128 positions and half of the filter tile
    Params pp = p;
    const int pbn = p.bn;
    const int bnmax_p = pbn > 128 ? 256 : 128;
    const int stage_b = BM * BK * 2 + (bnmax_p / 2) * BK * 2;
    int st_fit = (232448 - 1024 - 256) / stage_b;
    if (st_fit > 8) st_fit = 8;
    pp.stages = g_tune_stages ? (g_tune_stages < st_fit ? g_tune_stages : st_fit) : st_fit;
    pp.ring_bytes = pp.stages * stage_b;
    CUtensorMap tmh, tmp1;
    rc = encode3d(&tmh, w, (uint64_t)p.Kg, (uint64_t)Ng, (uint64_t)G, (uint64_t)ldw, (uint64_t)Ng * ldw, BK, pbn / 2);
    if (rc) return rc;
    rc = encode_im2col(&tmp1, pp, BM);
    if (rc) return rc;
    const long long ptiles = ceil_div_ll(rows, 2 * BM) * ceil_div(Ng, pbn) * G;
    const long long max_pairs = num_sms() / 2;
    const int grid_p = 2 * (int)(ptiles < max_pairs ? ptiles : max_pairs);
    return bnmax_p == 256 ? launch_pair<256>(tmh, tmp1, pp, grid_p, st) : launch_pair<128>(tmh, tmp1, pp, grid_p, st);
  }
  const long long tiles_m = ceil_div_ll(rows, mh * BM);
  if (cluster_mode() == 2 && atma && !flat && !p.b_resident && tiles_m >= 2 && num_sms() >= 2 &&
      memory_is_local(w, (size_t)G * (size_t)Ng * (size_t)ldw * sizeof(bf16))) {   // multicast map: own memory only
    // CTA pairs on two pixel tiles of the same output columns: each fetches half of the filter tile and
    // multicasts it to both (tensor map over this GPU's weights)
    CUtensorMap tmh;
    rc = encode3d(&tmh, w, (uint64_t)p.Kg, (uint64_t)Ng, (uint64_t)G, (uint64_t)ldw, (uint64_t)Ng * ldw, BK, p.bn / 2);
    if (rc) return rc;
    const long long pairs = ((tiles_m + 1) / 2) * ceil_div(Ng, p.bn) * G;
    const long long max_clusters = num_sms() / 2;
    const int grid_c = 2 * (int)(pairs < max_clusters ? pairs : max_clusters);
    if (mh == 2) return bnmax == 256 ? launch_cluster<256, 2>(tmh, tmp, p, grid_c, st) : launch_cluster<128, 2>(tmh, tmp, p, grid_c, st);
    return bnmax == 256 ? launch_cluster<256, 1>(tmh, tmp, p, grid_c, st) : launch_cluster<128, 1>(tmh, tmp, p, grid_c, st);
  }
  if (mh == 2) return bnmax == 256 ? launch<256, false, 2>(tm, tmp, atma, p, grid_final, st) : launch<128, false, 2>(tm, tmp, atma, p, grid_final, st);
  return bnmax == 256 ? launch<256, false, 1>(tm, tmp, atma, p, grid_final, st) : launch<128, false, 1>(tm, tmp, atma, p, grid_final, st);
}

// Host-only: would the halo-tile kernel take this forward problem, and with which geometry?  Returns 1 and fills
// out[0..7] = {padded width, output rows per half, tiles per image, patch rows, patch stage bytes, patch stages,
// filter stages, dynamic shared memory} or returns 0.  (Lets the planning be tested without a GPU.)
CXN_API int cxn_conv_halo_plan(int N, int H, int W, int Cs, int Cg, int G, int Ng, int KH, int KW, int stride, int pad_y,
                               int pad_x, int OH, int OW, int halves, int* out) {
  Params p;
  memset(&p, 0, sizeof(p));
  static bf16 aligned_dummy[8] __attribute__((aligned(16)));
  p.src = aligned_dummy; p.N = N; p.H = H; p.W = W; p.Cs = Cs; p.Cg = Cg; p.Cgs = Cg; p.KH = KH; p.KW = KW;
  p.stride_y = p.stride_x = stride; p.pad_y = pad_y; p.pad_x = pad_x; p.OH = OH; p.OW = OW; p.G = G; p.Kg = KH * KW * Cg;
  p.gran = 8; p.Ng = Ng;
  p.bn = pick_bn(Ng, Ng > 128 ? 256 : 128);
  HaloGeom hg;
  int smem = 0;
  if (halves < 1 || halves > 2 || !halo_plan(p, halves, &hg, &smem)) return 0;
  out[0] = hg.wp; out[1] = hg.rows_half; out[2] = hg.tiles_img; out[3] = hg.patch_h; out[4] = hg.a_stage_bytes;
  out[5] = hg.na; out[6] = hg.nb; out[7] = smem;
  return 1;
}

// Tuning / experiment knobs; value 0 (or -1 for the switches) restores the automatic choice.
//   0: row halves per forward tile   1: ring stages   2: gather through L1   3: tap-inner order   4: tile-width class
//   5: timing experiments   6: im2col TMA on / off   7: 2 = CTA-pair clusters with a multicast filter tile (forward)
//   8: 1 = narrower column tiles when the forward grid would not fill the machine
//   9: 1 | 2 = halo-tile kernel (two | one 128-row halves per tile) for stride-1 problems with <= 64 output columns
//  10: 1 = CTA pairs (cta_group::2 MMAs, each CTA stages half of the filter tile) for im2col-TMA forward problems with
//      more than 128 output columns per group and at least one tile per SM pair; 2 = whenever legal; -1 = CXN_CONV_PAIR2
// Role profiling of the forward kernel: buf = device array of [148][16] int64 (null switches it off).
// slots: 0 TMA wait-empty, 1 TMA total | 2 MMA wait-full, 3 MMA wait-tmem-empty, 4 MMA total |
// 5 gather wait-empty, 6 gather tile setup, 7 gather total | 8 epilogue wait-tmem-full, 9 epilogue body, 10 total
CXN_API int cxn_conv_set_prof(void* buf) {
  g_prof = (long long*)buf;
  return 0;
}

CXN_API int cxn_conv_tune(int key, int value) {
  switch (key) {
    case 0: g_tune_mh = value; break;
    case 1: g_tune_stages = value; break;
    case 2: break;   // (was: gather through L1; removed - no gain and it cost the gather loop a branch)
    case 3: break;   // (was: tap-inner k-block order; removed - measured slower)
    case 4: g_tune_bnmax = value; break;
    case 5: g_tune_dbg = value; break;
    case 6: g_tune_atma = value; break;
    case 7: g_tune_cluster = value; break;
    case 8: g_tune_fill = value; break;
    case 9: g_tune_halo = value; break;
    case 10: g_tune_pair = value; break;
    case 12: g_tune_flat = value; break;
    default: return -1;
  }
  return 0;
}

// Weight gradient: dw[g][j][(kh,kw,c)] += sum_pixels dy[pixel][g*Ng + j] * patch(pixel)[(kh,kw,c)]  (fp32, atomics)
static int conv_wgrad_impl(const void* src, int N, int H, int W, int Cs, int Cg, int G, int gran, const void* dy,
                           long long lddy, int Ng, int KH, int KW, int stride, int pad_y, int pad_x, int OH, int OW,
                           float* dw, long long lddw, int splits, int mask_period, int mask_valid, long long view_cdim,
                           long long view_row, long long view_img, cudaStream_t st) {
  Params p;
  memset(&p, 0, sizeof(p));
  p.view_cdim = view_cdim; p.view_row = view_row; p.view_img = view_img;
  if (mask_period && ((mask_period | mask_valid) & 3)) return -16;
  p.mask_period = mask_period; p.mask_valid = mask_valid;
  p.src = (const bf16*)src; p.N = N; p.H = H; p.W = W; p.Cs = Cs; p.Cg = Cg; p.KH = KH; p.KW = KW;
  p.Cgs = (view_cdim > 0 && t_view_gstride > 0) ? t_view_gstride : Cg;
  t_view_gstride = 0;
  p.stride_y = stride & 0xFFFF; p.stride_x = (stride >> 16) ? (stride >> 16) : p.stride_y; p.pad_y = pad_y; p.pad_x = pad_x; p.OH = OH; p.OW = OW; p.G = G; p.Kg = KH * KW * Cg;
  p.gran = gran; p.Ng = Ng; p.D = dw; p.ldd = lddw; p.strideD = (long long)Ng * lddw; p.bias = nullptr; p.relu = 0;
  int rc = check_geom(p);
  if (rc) return rc;
  const long long rows = (long long)N * OH * OW;
  if (rows <= 0) return 0;
  const int bnmax = 256;
  p.bn = pick_bn(p.Kg, bnmax);
  if (p.bn % 64) p.bn = ceil_div(p.bn, 64) * 64;   // gathered MN-major operand is built from 64-wide atoms
  const int num_kb = (int)ceil_div_ll(rows, BK);
  const long long base_tiles = (long long)ceil_div(Ng, BM) * ceil_div(p.Kg, p.bn) * G;
  if (splits <= 0) {
    // aim at ~4 tiles per SM so the round-robin tile assignment stays balanced and the fp32 reduction of one tile
    // overlaps the main loop of the next (two accumulator stages), but keep at least 8 k-blocks per tile.  (Measured:
    // fewer, deeper splits - one tile per SM at batch 32..128 - are 1.5-3 % slower per step, profiles/r2_experiments.md.)
    splits = (int)((4LL * num_sms()) / base_tiles);
    const int max_splits = num_kb / 8 > 0 ? num_kb / 8 : 1;
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
  }
  if (splits > num_kb) splits = num_kb;
  if (splits < 1) splits = 1;
  p.kb_per_split = ceil_div(num_kb, splits);
  p.splits = ceil_div(num_kb, p.kb_per_split);
  CUtensorMap tm;
  rc = encode3d(&tm, dy, (uint64_t)Ng, (uint64_t)rows, (uint64_t)G, (uint64_t)lddy, (uint64_t)Ng, 64, BK);
  if (rc) return rc;
  const long long tiles = base_tiles * p.splits;
  // (t_wgrad_cap: the caller runs this launch beside another kernel - a sub-wave backward-data problem on another
  // stream - and leaves that kernel its SMs: both are persistent with one CTA per SM)
  const int sms_here = (t_wgrad_cap > 0 && t_wgrad_cap < num_sms()) ? t_wgrad_cap : num_sms();
  const int grid = (int)(tiles < sms_here ? tiles : sms_here);
  p.stages = g_tune_stages ? g_tune_stages : 4;
  p.ring_bytes = p.stages * (BM + 256) * BK * 2;
  p.dbg = g_tune_dbg;
  p.prof = g_prof;
  bool atma = atma_ok(p);
  if (g_tune_atma >= 0 && !p.view_cdim) atma = atma && g_tune_atma;
  if (p.view_cdim && !atma) return -17;
  CUtensorMap tmp = tm;
  if (atma) {
    rc = encode_im2col(&tmp, p, BK);
    if (rc) return rc;
  }
  return launch<256, true, 1>(tm, tmp, atma, p, grid, st);
}

// ---- public entry points
CXN_API int cxn_conv_fprop(const void* src, int N, int H, int W, int Cs, int Cg, int G, int gran, const void* w,
                           long long ldw, int Ng, int KH, int KW, int stride, int pad_y, int pad_x, int OH, int OW,
                           void* y, long long ldy, const float* bias, int relu, cudaStream_t st) {
  return conv_fprop_impl(src, N, H, W, Cs, Cg, G, gran, w, ldw, Ng, KH, KW, stride, pad_y, pad_x, OH, OW, y, ldy, bias, relu,
                         nullptr, 0, 0, 0, st);
}
// cxn_conv_fprop whose outputs are zeroed where mask[row][col] <= 0 (mask: bf16, laid out like y).  Backward-data of a
// layer whose input went through a ReLU: dx = (x > 0) * (dy conv flipped W) in one kernel.
CXN_API int cxn_conv_fprop_masked(const void* src, int N, int H, int W, int Cs, int Cg, int G, int gran, const void* w,
                                  long long ldw, int Ng, int KH, int KW, int stride, int pad_y, int pad_x, int OH, int OW,
                                  void* y, long long ldy, const void* mask, cudaStream_t st) {
  return conv_fprop_impl(src, N, H, W, Cs, Cg, G, gran, w, ldw, Ng, KH, KW, stride, pad_y, pad_x, OH, OW, y, ldy, nullptr, 0,
                         mask, 0, 0, 0, st);
}
// The same over a strided view of the input: position (n, h, w) starts at src + n*img_stride + h*row_stride + w*Cs and
// exposes `cdim` >= G*Cg contiguous elements (positions may overlap in memory).  Only the im2col-TMA route can read
// such a view (Cg % 64 == 0 etc., else -17).  Used by the first layer: a filter row of a 4-channel-pitch image is
// one contiguous run, so "16 pixels x 4 channels" is a 64-channel position every `stride` pixels.
CXN_API int cxn_conv_fprop_view(const void* src, int N, int H, int W, int Cs, int Cg, int G, const void* w, long long ldw,
                                int Ng, int KH, int KW, int stride, int pad_y, int pad_x, int OH, int OW, void* y,
                                long long ldy, const float* bias, int relu, long long cdim, long long row_stride,
                                long long img_stride, cudaStream_t st) {
  if (cdim <= 0) return -18;
  return conv_fprop_impl(src, N, H, W, Cs, Cg, G, 8, w, ldw, Ng, KH, KW, stride, pad_y, pad_x, OH, OW, y, ldy, bias, relu,
                         nullptr, cdim, row_stride, img_stride, st);
}
CXN_API int cxn_conv_wgrad_masked(const void* src, int N, int H, int W, int Cs, int Cg, int G, int gran, const void* dy,
                                  long long lddy, int Ng, int KH, int KW, int stride, int pad_y, int pad_x, int OH, int OW,
                                  float* dw, long long lddw, int splits, int mask_period, int mask_valid,
                                  cudaStream_t st) {
  return conv_wgrad_impl(src, N, H, W, Cs, Cg, G, gran, dy, lddy, Ng, KH, KW, stride, pad_y, pad_x, OH, OW, dw, lddw, splits,
                         mask_period, mask_valid, 0, 0, 0, st);
}
CXN_API int cxn_conv_wgrad_view(const void* src, int N, int H, int W, int Cs, int Cg, int G, const void* dy, long long lddy,
                                int Ng, int KH, int KW, int stride, int pad_y, int pad_x, int OH, int OW, float* dw,
                                long long lddw, int splits, int mask_period, int mask_valid, long long cdim,
                                long long row_stride, long long img_stride, cudaStream_t st) {
  if (cdim <= 0) return -18;
  return conv_wgrad_impl(src, N, H, W, Cs, Cg, G, 8, dy, lddy, Ng, KH, KW, stride, pad_y, pad_x, OH, OW, dw, lddw, splits,
                         mask_period, mask_valid, cdim, row_stride, img_stride, st);
}

CXN_API int cxn_conv_flip_weights(const void* w, void* wt, int G, int ocg, int icg, int KH, int KW, int ld_in,
                                  int ld_out, cudaStream_t st) {
  const long long total = (long long)G * ocg * KH * KW * icg;
  if (total <= 0) return 0;
  if ((long long)G * KH * KW > 65535) return -1;
  dim3 grid(ceil_div(icg, 32), ceil_div(ocg, 32), G * KH * KW), block(32, 8);
  flip_weights_kernel<<<grid, block, 0, st>>>((const bf16*)w, (bf16*)wt, ocg, icg, KH * KW, ld_in, ld_out, icg);
  CXN_CHECK_LAUNCH();
  return 0;
}
// The same from a filter whose taps are `cpitch` >= icg elements apart (channel-padded internal layout: the padding
// channels are not part of the backward-data problem).
CXN_API int cxn_conv_flip_weights_pitch(const void* w, void* wt, int G, int ocg, int icg, int KH, int KW, int ld_in,
                                        int ld_out, int cpitch, cudaStream_t st) {
  const long long total = (long long)G * ocg * KH * KW * icg;
  if (total <= 0) return 0;
  if ((long long)G * KH * KW > 65535 || cpitch < icg) return -1;
  dim3 grid(ceil_div(icg, 32), ceil_div(ocg, 32), G * KH * KW), block(32, 8);
  flip_weights_kernel<<<grid, block, 0, st>>>((const bf16*)w, (bf16*)wt, ocg, icg, KH * KW, ld_in, ld_out, cpitch);
  CXN_CHECK_LAUNCH();
  return 0;
}

// K2 for strided convolutions, without an im2col buffer: backward-data as stride*stride implicit-GEMM problems.
// Input positions with (ih + pad_y) % s == ph, (iw + pad_x) % s == pw only meet the filter taps kh = ph (mod s),
// kw = pw (mod s), and over that sub-grid the layer is a STRIDE-1 correlation of dy with the flipped sub-filter - the
// forward kernel, run once per phase on dy, writing through a strided view of dx.  No multiplication by an inserted
// zero is ever made (the dilated-dy formulation does s*s times the work).  Reference: the col2im path of
// src/layer/convolution_layer-inl.hpp:140-153.
//   dy [N][OH][OW][lddy] (group g: channels [g*ocg, (g+1)*ocg)), w: the forward filter [G][ocg][KH][KW][icg] (ldw),
//   dx [N][H][W][lddx], mask: optional tensor laid out like dx (outputs whose mask element is <= 0 become zero),
//   scratch: bf16 room for the phase filters, cxn_conv_dgrad_strided_scratch() elements.
// Returns -24 when a phase would need a negative padding (pad >= stride with short filters): use the explicit path.
CXN_API long long cxn_conv_dgrad_strided_scratch(int G, int ocg, int icg, int KH, int KW, int stride) {
  long long total = 0;
  for (int ph = 0; ph < stride; ++ph)
    for (int pw = 0; pw < stride; ++pw) {
      const int KHa = ph < KH ? ceil_div(KH - ph, stride) : 0, KWb = pw < KW ? ceil_div(KW - pw, stride) : 0;
      if (KHa * KWb) total += (long long)G * icg * (((long long)KHa * KWb * ocg + 7) / 8 * 8);
    }
  return total;
}
// 1 when every stride phase of this geometry has a non-negative padding (cxn_conv_dgrad_strided would not return -24)
CXN_API int cxn_conv_dgrad_strided_ok(int KH, int KW, int stride, int pad_y, int pad_x) {
  if (stride < 2) return 0;
  for (int ph = 0; ph < stride; ++ph) {
    const int KHa = ph < KH ? ceil_div(KH - ph, stride) : 0;
    const int ih0 = ((ph - pad_y) % stride + stride) % stride;
    if (KHa && KHa - 1 - (ih0 + pad_y - ph) / stride < 0) return 0;
  }
  for (int pw = 0; pw < stride; ++pw) {
    const int KWb = pw < KW ? ceil_div(KW - pw, stride) : 0;
    const int iw0 = ((pw - pad_x) % stride + stride) % stride;
    if (KWb && KWb - 1 - (iw0 + pad_x - pw) / stride < 0) return 0;
  }
  return 1;
}
CXN_API int cxn_conv_dgrad_strided(const void* dy, int N, int OH, int OW, long long lddy, int ocg, int G, const void* w,
                                   long long ldw, int icg, int KH, int KW, int stride, int pad_y, int pad_x, int H, int W,
                                   void* dx, long long lddx, const void* mask, void* scratch, cudaStream_t st) {
  if (stride < 2 || (ocg % 8) || (lddy % 8)) return -25;
  bool any_empty = false;
  for (int ph = 0; ph < stride; ++ph) {
    const int KHa = ph < KH ? ceil_div(KH - ph, stride) : 0;
    const int ih0 = ((ph - pad_y) % stride + stride) % stride;
    if (KHa && KHa - 1 - (ih0 + pad_y - ph) / stride < 0) return -24;
    if (!KHa) any_empty = true;
  }
  for (int pw = 0; pw < stride; ++pw) {
    const int KWb = pw < KW ? ceil_div(KW - pw, stride) : 0;
    const int iw0 = ((pw - pad_x) % stride + stride) % stride;
    if (KWb && KWb - 1 - (iw0 + pad_x - pw) / stride < 0) return -24;
    if (!KWb) any_empty = true;
  }
  if (any_empty) {   // stride larger than the filter: some input positions receive nothing
    cudaError_t e = cudaMemsetAsync(dx, 0, (size_t)N * H * W * lddx * sizeof(bf16), st);
    if (e != cudaSuccess) return (int)e;
  }
  bf16* wt = (bf16*)scratch;
  for (int ph = 0; ph < stride; ++ph)
    for (int pw = 0; pw < stride; ++pw) {
      const int KHa = ph < KH ? ceil_div(KH - ph, stride) : 0, KWb = pw < KW ? ceil_div(KW - pw, stride) : 0;
      if (KHa * KWb == 0) continue;
      const int ih0 = ((ph - pad_y) % stride + stride) % stride, iw0 = ((pw - pad_x) % stride + stride) % stride;
      const int ld_out = (KHa * KWb * ocg + 7) / 8 * 8;
      bf16* wt_phase = wt;
      wt += (long long)G * icg * ld_out;
      if (ih0 >= H || iw0 >= W) continue;
      const int IHp = ceil_div(H - ih0, stride), IWp = ceil_div(W - iw0, stride);
      if ((long long)G * KHa * KWb > 65535) return -1;
      // rows of the phase filter are padded to 8 elements; the padding must read as zero
      if (ld_out != KHa * KWb * ocg) {
        cudaError_t e = cudaMemsetAsync(wt_phase, 0, (size_t)G * icg * ld_out * sizeof(bf16), st);
        if (e != cudaSuccess) return (int)e;
      }
      dim3 grid(ceil_div(icg, 32), ceil_div(ocg, 32), G * KHa * KWb), block(32, 8);
      flip_weights_phase_kernel<<<grid, block, 0, st>>>((const bf16*)w, wt_phase, ocg, icg, KW, stride, ph, pw, KHa, KWb,
                                                        (int)ldw, ld_out);
      CXN_CHECK_LAUNCH();
      const int pad_py = KHa - 1 - (ih0 + pad_y - ph) / stride, pad_px = KWb - 1 - (iw0 + pad_x - pw) / stride;
      t_oview.w = IWp; t_oview.h = IHp;
      t_oview.base = ((long long)ih0 * W + iw0) * lddx;
      t_oview.pix = (long long)stride * lddx;
      t_oview.row = (long long)stride * W * lddx;
      t_oview.img = (long long)H * W * lddx;
      const int rc = conv_fprop_impl(dy, N, OH, OW, (int)lddy, ocg, G, 8, wt_phase, ld_out, icg, KHa, KWb, 1, pad_py, pad_px, IHp,
                                     IWp, dx, lddx, nullptr, 0, mask, 0, 0, 0, st);
      t_oview.w = 0;
      if (rc) return rc;
    }
  return 0;
}

// Limit the grid of this thread's following weight-gradient launches to `ctas` CTAs (0 = no limit) and return the
// previous value.  The kernels here are persistent with one CTA per SM; a weight gradient launched beside a
// backward-data problem that fills only part of the GPU (small per-GPU batches) takes the remaining SMs, and the two
// run side by side instead of one after the other.
// The next cxn_conv_fprop_view / cxn_conv_wgrad_view call of this thread reads group g's channels from channel
// g * stride on (instead of g * Cg): groups as overlapping channel windows, each padded up to a multiple of 64 channels
// that meet structural zeros of the filter (AlexNet conv2: Cg = 64, stride = 48).  One launch for all groups.
CXN_API int cxn_conv_view_group_stride(int stride) {
  if (stride < 0 || (stride % 8)) return -1;
  t_view_gstride = stride;
  return 0;
}

// how many forward launches of this process took the FLAT route (Params::flat) so far
CXN_API long long cxn_conv_flat_launches() { return g_flat_launches; }

CXN_API int cxn_conv_wgrad_grid_cap(int ctas) {
  const int prev = t_wgrad_cap;
  t_wgrad_cap = ctas < 0 ? 0 : ctas;
  return prev;
}

CXN_API int cxn_conv_wgrad(const void* src, int N, int H, int W, int Cs, int Cg, int G, int gran, const void* dy,
                           long long lddy, int Ng, int KH, int KW, int stride, int pad_y, int pad_x, int OH, int OW,
                           float* dw, long long lddw, int splits, cudaStream_t st) {
  return cxn_conv_wgrad_masked(src, N, H, W, Cs, Cg, G, gran, dy, lddy, Ng, KH, KW, stride, pad_y, pad_x, OH, OW, dw, lddw,
                               splits, 0, 0, st);
}
