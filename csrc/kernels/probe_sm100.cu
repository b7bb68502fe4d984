// Hardware probe (not on any training path): what does tcgen05.mma read when the start address of a
// 128B-swizzled K-major operand is moved down by whole 128-byte rows, i.e. is no longer aligned to the
// 1024-byte swizzle pattern?  The planned "halo tile" convolution wants to feed the
// filter taps of a stride-1 convolution as row-shifted views of one shared-memory patch; whether that needs
// the descriptor's base-offset field, or works at all, is a property of the hardware to be measured.
//
//   cxn_probe_shifted_desc(out[128][16] fp32, shift_rows, base_offset_mode, stream)
//     shared memory holds A[r][k] = a(r, k) for r < 160, k < 64 written with the address-based swizzle the
//     gather producers use; B[n][k] = (k == 4 n), so D[m][n] = "A row m as the hardware sees it", column 4 n.
//     The A descriptor starts at row `shift_rows`; base_offset_mode 0 leaves the base-offset field 0, 1 sets it
//     to (start address >> 7) & 7.  The caller compares out[m][n] with a(m + shift_rows, 4 n).
#include "tc_common.cuh"

namespace cxn {
namespace probe {

using namespace cxn::tc;

__host__ __device__ inline float probe_value(int r, int k) { return (float)(((r * 7 + k * 3) % 17) - 8); }

__global__ void __launch_bounds__(128, 1) shifted_desc_kernel(float* out, int shift_rows, int bo_mode) {
  constexpr int kRowsA = 160, kRowsB = 16;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t a_base = base, b_base = base + kRowsA * 128;          // 160 * 128 = 20480: 1024-aligned
  const uint32_t bar = b_base + 2048;
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(gen + kRowsA * 128 + 2048 + 16);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // fill A and B: element (r, k) lives at r*128 + ((k/8 ^ (r & 7)) * 16) + (k % 8) * 2
  for (int i = threadIdx.x; i < (kRowsA + kRowsB) * 64; i += blockDim.x) {
    const int r = i >> 6, k = i & 63;
    const bool isb = r >= kRowsA;
    const int rr = isb ? r - kRowsA : r;
    const float v = isb ? ((k == 4 * rr) ? 1.0f : 0.0f) : probe_value(rr, k);
    uint8_t* tile = gen + (isb ? kRowsA * 128 : 0);
    *reinterpret_cast<bf16*>(tile + rr * 128 + (((k >> 3) ^ (rr & 7)) << 4) + (k & 7) * 2) = __float2bfloat16(v);
  }
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32((const void*)tmem_slot)),
                 "n"(32)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  if (threadIdx.x == 0) {
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(16 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint32_t a_start = a_base + (uint32_t)shift_rows * 128u;
    uint64_t adesc = make_desc(a_start, 0u, 1024u);
    if (bo_mode == 1) adesc |= (uint64_t)((a_start >> 7) & 7u) << 49;
    const uint64_t bdesc = make_desc(b_base, 0u, 1024u);
#pragma unroll
    for (int k = 0; k < 4; ++k)
      umma_bf16(tmem, adesc + (uint64_t)((k * 32) >> 4), bdesc + (uint64_t)((k * 32) >> 4), idesc, k > 0 ? 1u : 0u);
    umma_commit(bar);
  }
  mbar_wait(bar, 0);
  tc_fence_after();
  uint32_t v[32];
  tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16), v);
  tmem_ld_wait();
  const int row = warp * 32 + lane;
#pragma unroll
  for (int j = 0; j < 16; ++j) out[row * 16 + j] = __uint_as_float(v[j]);
  tc_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(32) : "memory");
}

}  // namespace probe
}  // namespace cxn

// expected value of out[m][n] if the hardware reads "row m + shift_rows" of the logical matrix
CXN_API float cxn_probe_expected(int m, int n, int shift_rows) { return cxn::probe::probe_value(m + shift_rows, 4 * n); }

CXN_API int cxn_probe_shifted_desc(float* out, int shift_rows, int base_offset_mode, cudaStream_t st) {
  if (shift_rows < 0 || shift_rows > 32) return -1;
  const int smem = 160 * 128 + 2048 + 64 + 1024;
  cxn::probe::shifted_desc_kernel<<<1, 128, smem, st>>>(out, shift_rows, base_offset_mode);
  CXN_CHECK_LAUNCH();
  return 0;
}
