// Collectives of the model-parallel (column-sharded) fully connected layers over NVLink / NVSwitch peer memory.
//
// The reference keeps fully connected layers data parallel and pays for it on the wire: every GPU pushes the whole
// weight gradient to the parameter server and pulls the whole weight back (src/updater/async_updater-inl.hpp:94-127),
// or, with `fullc_gather`, gathers inputs and output gradients so that ONE device computes the weight gradient
// (:67-93,190-199).  Here a fully connected layer may instead be sharded by output columns: rank r owns rows
// [r*ns, (r+1)*ns) of W (fp32 master, bf16 shadow, momentum) and nothing of W ever crosses a link.  What crosses
// are activations, a few MB per step, moved by the two kernels below through every rank's CUDA-IPC mapping of every
// other rank's symmetric block (parallel/symm.py):
//
//   tp_push_kernel         2-D block copy from local memory into every rank's block (peer stores), one descriptor per
//                          destination: row gather (data-parallel rows -> X_all on every rank), column gather (my
//                          output columns for the whole batch -> X_all of the next layer), and the two all-to-all
//                          transposes between the column-sharded and the row-sharded (data-parallel) layouts.
//                          Ends with ONE flag barrier: signal "my stores are complete", wait for every peer's signal.
//   tp_reduce_pull_kernel  sum of the same 2-D block of every rank's partial result (backward-data of a column shard
//                          is a partial sum over the shards): signal "my partial is complete", wait, 128-bit peer
//                          loads of bf16 partials, fp32 sum, bf16 result in local memory (reduce-scatter by pull).
//
// Barrier protocol: per (slot, source rank) one 32-bit flag in the destination's block, written with
// st.release.sys and polled with ld.acquire.sys.  The sequence number of a launch is a per-slot device counter that
// the kernel itself advances, so a launch captured in a CUDA graph replays correctly; all ranks issue the same
// sequence of launches per slot.  No entry barrier is needed: every buffer a peer writes into (or reads from) is
// reused only a whole training step later, and between two uses of a buffer every rank has waited for a later
// signal of every other rank (see parallel/tensor_parallel.py for the argument per buffer).
// Every wait is bounded and traps on timeout.
#include "common.cuh"
#include <stdlib.h>
#include <string.h>

namespace cxn {
namespace tp {

constexpr int kMaxRanks = 16;
constexpr int kSlots = 64;
constexpr int kThreads = 256;

struct PushDesc {          // one destination rank
  long long src_off;       // bytes from the source base to the first byte of the sub-block for this destination
  long long dst_off;       // bytes from the destination BLOCK base to where the sub-block's first byte goes
  int nrows;               // rows to write (rows >= valid_rows are written as zeros)
  int valid_rows;
  int row_bytes;           // multiple of 16
  int pad_;
};

struct Params {
  char* peers[kMaxRanks];          // block base of every rank as mapped here
  PushDesc d[kMaxRanks];
  const char* src;                 // push: local source base;  pull: unused
  long long src_ld, dst_ld;        // bytes between rows
  long long off_flags;             // byte offset of the [kSlots][kMaxRanks] flag words inside a block
  unsigned* counters;              // local [kSlots][2]: {CTAs done, launches so far}
  // pull
  long long part_off;              // byte offset of the partial-sum matrix inside every block
  long long part_ld;               // bytes
  long long pull_off;              // first byte of my sub-block inside the partial matrix (row0 * ld + col0 * 2)
  char* dst;                       // local destination
  int pull_rows, pull_row_bytes;
  int rank, nranks, slot;
  char* mc;                        // NVSwitch multicast address of the symmetric block, or null
  int bcast;                       // push: every destination receives the same sub-block at the same offset
};

__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_peer_16(const void* p) {
  uint4 v;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}
// NVLS (multicast address of the block): one store reaches every rank's block; one load returns the fp32-accumulated
// sum of the same eight bf16 values of every rank's block.  Plain multimem instructions, not TMA.
__device__ __forceinline__ void mc_st_16(void* p, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(__uint_as_float(v.x)),
               "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
               : "memory");
}
__device__ __forceinline__ uint4 mc_ld_reduce_bf16x8(const void* p) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ unsigned* flag_ptr(char* base, long long off_flags, int slot, int src) {
  return reinterpret_cast<unsigned*>(base + off_flags) + slot * kMaxRanks + src;
}
__device__ __forceinline__ void signal_all(const Params& p, unsigned seq) {
  if (threadIdx.x < p.nranks) st_release_sys(flag_ptr(p.peers[threadIdx.x], p.off_flags, p.slot, p.rank), seq);
}
// wait until every rank has published `seq` in this rank's flag row of the slot (threads < nranks poll one flag each)
__device__ __forceinline__ void wait_all(const Params& p, unsigned seq) {
  if (threadIdx.x < p.nranks) {
    const unsigned* f = flag_ptr(p.peers[p.rank], p.off_flags, p.slot, threadIdx.x);
    const long long t0 = clock64();
    unsigned spins = 0;
    while ((int)(ld_acquire_sys(f) - seq) < 0) {
      if (++spins > 64) __nanosleep(32);       // the first polls are back to back: a peer is usually a few us away
      if ((spins & 4095u) == 0 && clock64() - t0 > 20000000000LL) {
        printf("[cxxnet_b200 tp] rank %d timed out waiting for rank %d (slot %d seq %u)\n", p.rank, (int)threadIdx.x,
               p.slot, seq);
        __trap();
      }
    }
  }
  __syncthreads();
}

// ---------------------------------------------------------------- push + barrier
__global__ void __launch_bounds__(kThreads) tp_push_kernel(const Params p) {
  unsigned* done = p.counters + 2 * p.slot;
  const unsigned seq = *reinterpret_cast<volatile unsigned*>(done + 1) + 1u;
  if (p.mc != nullptr && p.bcast) {
    // a gather: the same sub-block goes to the same place in every rank's block - one multicast store per 16 bytes
    const PushDesc d = p.d[p.rank];
    if (d.nrows > 0 && d.row_bytes > 0) {
      const int vpr = d.row_bytes >> 4;
      const long long total = (long long)d.nrows * vpr;
      const char* s = p.src + d.src_off;
      char* o = p.mc + d.dst_off;
      for (long long v = blockIdx.x * (long long)kThreads + threadIdx.x; v < total; v += (long long)gridDim.x * kThreads) {
        const int r = (int)(v / vpr);
        const int c = (int)(v - (long long)r * vpr);
        uint4 val = make_uint4(0u, 0u, 0u, 0u);
        if (r < d.valid_rows) val = *reinterpret_cast<const uint4*>(s + (long long)r * p.src_ld + ((long long)c << 4));
        mc_st_16(o + (long long)r * p.dst_ld + ((long long)c << 4), val);
      }
    }
  } else
  for (int i = 0; i < p.nranks; ++i) {
    const int dst = (p.rank + i) % p.nranks;             // own copy first, peers staggered over the links
    const PushDesc d = p.d[dst];
    if (d.nrows <= 0 || d.row_bytes <= 0) continue;
    const int vpr = d.row_bytes >> 4;                     // 16-byte vectors per row
    const long long total = (long long)d.nrows * vpr;
    const char* s = p.src + d.src_off;
    char* o = p.peers[dst] + d.dst_off;
    for (long long v = blockIdx.x * (long long)kThreads + threadIdx.x; v < total; v += (long long)gridDim.x * kThreads) {
      const int r = (int)(v / vpr);
      const int c = (int)(v - (long long)r * vpr);
      uint4 val = make_uint4(0u, 0u, 0u, 0u);
      if (r < d.valid_rows) val = *reinterpret_cast<const uint4*>(s + (long long)r * p.src_ld + ((long long)c << 4));
      *reinterpret_cast<uint4*>(o + (long long)r * p.dst_ld + ((long long)c << 4)) = val;
    }
  }
  // my stores are complete (system scope) -> the CTA that finishes last signals every peer and waits for theirs;
  // the other CTAs simply end: the kernel, and with it the stream, completes when that last CTA has seen all flags
  __threadfence_system();
  __syncthreads();
  __shared__ unsigned s_last;
  if (threadIdx.x == 0) s_last = (atomicAdd(done, 1u) == gridDim.x - 1) ? 1u : 0u;
  __syncthreads();
  if (!s_last) return;
  __threadfence_system();
  signal_all(p, seq);
  wait_all(p, seq);
  if (threadIdx.x == 0) {
    done[0] = 0u;
    done[1] = seq;
  }
}

// ---------------------------------------------------------------- barrier + pull-reduce
__device__ __forceinline__ void acc_bf16x8(float (&a)[8], const uint4& v) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 f = __bfloat1622float2(h[j]);
    a[2 * j] += f.x;
    a[2 * j + 1] += f.y;
  }
}

template <int NR>     // ranks whose loads are in flight together
__global__ void __launch_bounds__(kThreads) tp_reduce_pull_kernel(const Params p) {
  unsigned* done = p.counters + 2 * p.slot;
  const unsigned seq = *reinterpret_cast<volatile unsigned*>(done + 1) + 1u;
  // my partial was completed by the previous kernel on this stream
  if (blockIdx.x == 0) signal_all(p, seq);
  wait_all(p, seq);
  const int vpr = p.pull_row_bytes >> 4;
  const long long total = (long long)p.pull_rows * vpr;
  for (long long v = blockIdx.x * (long long)kThreads + threadIdx.x; v < total; v += (long long)gridDim.x * kThreads) {
    const int r = (int)(v / vpr);
    const int c = (int)(v - (long long)r * vpr);
    const long long off = p.part_off + p.pull_off + (long long)r * p.part_ld + ((long long)c << 4);
    if (p.mc != nullptr) {
      // the switch adds the ranks' partials (fp32 accumulation, bf16 result)
      *reinterpret_cast<uint4*>(p.dst + (long long)r * p.dst_ld + ((long long)c << 4)) = mc_ld_reduce_bf16x8(p.mc + off);
      continue;
    }
    float a[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = 0.f;
    for (int r0 = 0; r0 < p.nranks; r0 += NR) {
      uint4 buf[NR];
#pragma unroll
      for (int k = 0; k < NR; ++k)
        if (r0 + k < p.nranks) buf[k] = ld_peer_16(p.peers[(p.rank + r0 + k) % p.nranks] + off);
#pragma unroll
      for (int k = 0; k < NR; ++k)
        if (r0 + k < p.nranks) acc_bf16x8(a, buf[k]);
    }
    uint4 o;
    __nv_bfloat162* oh = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
    for (int j = 0; j < 4; ++j) oh[j] = __floats2bfloat162_rn(a[2 * j], a[2 * j + 1]);
    *reinterpret_cast<uint4*>(p.dst + (long long)r * p.dst_ld + ((long long)c << 4)) = o;
  }
  __syncthreads();
  if (threadIdx.x == 0 && atomicAdd(done, 1u) == gridDim.x - 1) {
    done[0] = 0u;
    done[1] = seq;
  }
}

}  // namespace tp
}  // namespace cxn

using namespace cxn;
using namespace cxn::tp;

CXN_API int cxn_tp_flag_bytes() { return kSlots * kMaxRanks * 4; }
CXN_API int cxn_tp_counter_bytes() { return kSlots * 2 * 4; }
CXN_API int cxn_tp_max_slots() { return kSlots; }

// Multicast address of the symmetric block (or null): gathers then use multimem.st, reductions multimem.ld_reduce.
// CXN_TP_NVLS=0 keeps the unicast path.  One GPU per process.
static char* g_tp_mc = nullptr;
CXN_API int cxn_tp_set_multicast(void* mc_base) {
  static const bool off = getenv("CXN_TP_NVLS") != nullptr && atoi(getenv("CXN_TP_NVLS")) == 0;
  g_tp_mc = off ? nullptr : reinterpret_cast<char*>(mc_base);
  return 0;
}

static int fill_common(Params* p, const unsigned long long* peers, int nranks, int rank, long long off_flags, int slot,
                       unsigned* counters) {
  if (nranks < 1 || nranks > kMaxRanks || rank < 0 || rank >= nranks || slot < 0 || slot >= kSlots || !counters) return -40;
  memset(p, 0, sizeof(*p));
  for (int i = 0; i < nranks; ++i) p->peers[i] = reinterpret_cast<char*>(peers[i]);
  p->off_flags = off_flags; p->counters = counters; p->rank = rank; p->nranks = nranks; p->slot = slot;
  p->mc = nranks > 1 ? g_tp_mc : nullptr;
  return 0;
}

// Push one 2-D sub-block per destination rank into every rank's symmetric block, then barrier.
//   desc: HOST array of nranks x {src_off, dst_off, nrows, valid_rows, row_bytes} as 5 long longs per rank
//   src / src_ld: local source base and row pitch in bytes; dst_ld: row pitch at the destination in bytes
// All offsets, pitches and row_bytes are multiples of 16.  `counters`: cxn_tp_counter_bytes() of zeroed local memory.
CXN_API int cxn_tp_push(const unsigned long long* peers, int nranks, int rank, long long off_flags, int slot,
                        unsigned* counters, const void* src, long long src_ld, long long dst_ld, const long long* desc,
                        int grid, cudaStream_t st) {
  Params p;
  int rc = fill_common(&p, peers, nranks, rank, off_flags, slot, counters);
  if (rc) return rc;
  if ((reinterpret_cast<uintptr_t>(src) & 15) || (src_ld & 15) || (dst_ld & 15)) return -41;
  long long work = 0;
  for (int i = 0; i < nranks; ++i) {
    const long long* q = desc + 5 * i;
    if ((q[0] & 15) || (q[1] & 15) || (q[4] & 15) || q[2] < 0 || q[4] < 0) return -41;
    p.d[i].src_off = q[0]; p.d[i].dst_off = q[1]; p.d[i].nrows = (int)q[2]; p.d[i].valid_rows = (int)q[3];
    p.d[i].row_bytes = (int)q[4];
    work += q[2] * (q[4] >> 4);
  }
  p.src = (const char*)src; p.src_ld = src_ld; p.dst_ld = dst_ld;
  p.bcast = 1;
  for (int i = 1; i < nranks; ++i)
    if (memcmp(&p.d[i], &p.d[0], sizeof(PushDesc)) != 0) p.bcast = 0;
  if (p.bcast) work /= nranks;       // one multicast store instead of a store per rank
  long long need = (work + kThreads - 1) / kThreads;
  if (grid < 1) grid = 1;
  if (grid > num_sms()) grid = num_sms();
  if (need < grid) grid = (int)(need < 1 ? 1 : need);
  tp_push_kernel<<<grid, kThreads, 0, st>>>(p);
  CXN_CHECK_LAUNCH();
  return 0;
}

// Barrier, then dst[r][c] = sum over ranks of partial_rank[row0 + r][col0 + c] (bf16 in, fp32 sum, bf16 out).
//   part_off / part_ld: the partial matrix inside every rank's block (byte offset, row pitch in bytes)
//   pull_off: byte offset of my sub-block's first element inside the partial matrix; rows x row_bytes its extent
CXN_API int cxn_tp_reduce_pull(const unsigned long long* peers, int nranks, int rank, long long off_flags, int slot,
                               unsigned* counters, long long part_off, long long part_ld, long long pull_off, int rows,
                               int row_bytes, void* dst, long long dst_ld, int grid, cudaStream_t st) {
  Params p;
  int rc = fill_common(&p, peers, nranks, rank, off_flags, slot, counters);
  if (rc) return rc;
  if ((part_off & 15) || (part_ld & 15) || (pull_off & 15) || (row_bytes & 15) || (dst_ld & 15) ||
      (reinterpret_cast<uintptr_t>(dst) & 15) || rows < 0 || row_bytes < 0)
    return -41;
  p.part_off = part_off; p.part_ld = part_ld; p.pull_off = pull_off; p.pull_rows = rows; p.pull_row_bytes = row_bytes;
  p.dst = (char*)dst; p.dst_ld = dst_ld;
  const long long work = (long long)rows * (row_bytes >> 4);
  long long need = (work + kThreads - 1) / kThreads;
  if (grid < 1) grid = 1;
  if (grid > num_sms()) grid = num_sms();        // every CTA spins on the flags: all must be resident
  if (need < grid) grid = (int)(need < 1 ? 1 : need);
  if (nranks <= 2) tp_reduce_pull_kernel<2><<<grid, kThreads, 0, st>>>(p);
  else if (nranks <= 4) tp_reduce_pull_kernel<4><<<grid, kThreads, 0, st>>>(p);
  else tp_reduce_pull_kernel<8><<<grid, kThreads, 0, st>>>(p);
  CXN_CHECK_LAUNCH();
  return 0;
}
