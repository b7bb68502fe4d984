// K17 / K19: fused gradient reduce-scatter + sgd/nag/adam update + weight all-gather over
// NVLink / NVSwitch peer memory, one launch per layer bucket.
//
// What the reference does for the same step (src/updater/async_updater-inl.hpp:94-127 with
// mshadow-ps "local"): every GPU pushes its gradient to a host-staged parameter server, a
// CPU thread sums, every GPU pulls the sum and runs its own updater kernels (3+ launches per
// tensor).  Here every rank maps every other rank's parameter arena (CUDA IPC), and ONE
// kernel per bucket does, with no NCCL call and no separate update kernel:
//
//   entry barrier  system-scope flags in peer memory: "my gradients for this bucket are final"
//   phase 1        rank r owns chunks [r*per, (r+1)*per) of the bucket: 128-bit peer loads of
//                  every rank's gradient shard, summed in registers (reduce-scatter), the
//                  updater applied to the owned shard (momentum lives only in the owner),
//                  the new weights stored into every rank's arena (all-gather, peer stores)
//   exit barrier   "I have read your gradients and written your weights"
//   phase 2        local: zero the bucket's gradients, refresh the bf16 GEMM shadow
//
// All ranks issue the same bucket sequence on an in-order stream, so the spin waits cannot
// deadlock; every wait is bounded and traps on timeout.
#include "common.cuh"
#include <mutex>
#include <vector>
#include <stdlib.h>
#include <string.h>

namespace cxn {
namespace xchg {

constexpr int kMaxRanks = 16;
constexpr int kMaxSlots = 256;
constexpr int kChunk = 1024;
constexpr int kThreads = 256;
constexpr int kMaxBucketSegs = 8;    // segments cached in shared memory (512 B: the kernel must fit beside the GEMM /
                                     // convolution CTAs it overlaps); buckets may hold more (load_seg)

struct Segment {   // must match loss_update.cu
  long long start, n, shadow_off;
  int cols, shadow_ld;
  float lr, wd, momentum, clip;
  float adam_lr_t, beta1, beta2, pad_;
};

struct Params {
  char* peers[kMaxRanks];   // base of every rank's symmetric block, as mapped in THIS process
  long long off_w, off_g, off_m, off_m2, off_shadow, off_flags;   // byte offsets inside a block
  const Segment* segs;
  const int* chunk_seg;
  long long cb, ce;         // bucket = chunks [cb, ce)
  int rank, nranks, kind;
  unsigned seq;
  const unsigned* seq_dev;  // optional device word added to seq when the kernel starts (cxn_xchg_set_epoch_ptr)
  int slot;
  char* mc;                 // NVSwitch multicast address of the symmetric block (null = unicast peer loads / stores)
  unsigned* counter;        // local scratch: CTA completion counter (zero between launches)
  int dbg;                  // profiling aid (CXN_XCHG_DBG): 1 skip peer loads, 2 skip peer stores, 4 skip phase 2
};

__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_sys_f4(const float* p) {
  float4 v;
  // weak 128-bit load that does not allocate in L1: the data was published before the flag
  // acquire + bar.sync that precede every use, and is read exactly once per launch
  asm volatile("ld.global.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ void st_sys_f4(float* p, float4 v) {
  asm volatile("st.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
// NVLS: one load that returns the SUM of the same 16 bytes of every rank's block (reduced inside the switch), one
// store that lands in every rank's block.  Plain multimem instructions on the multicast address, not TMA.
__device__ __forceinline__ float4 mc_ld_reduce_f4(const float* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void mc_st_f4(float* p, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}
__device__ __forceinline__ unsigned* flag_ptr(char* base, long long off_flags, int which, int slot, int src) {
  return reinterpret_cast<unsigned*>(base + off_flags) + ((which * kMaxSlots + slot) * kMaxRanks + src);
}
// wait until every rank has published `seq` in local flag row (which, slot)
__device__ __forceinline__ void wait_all(const Params& p, int which, unsigned seq) {
  if (threadIdx.x < p.nranks) {
    const unsigned* f = flag_ptr(p.peers[p.rank], p.off_flags, which, p.slot, threadIdx.x);
    const long long t0 = clock64();
    unsigned spins = 0;
    while ((int)(ld_acquire_sys(f) - seq) < 0) {
      if (++spins > 32) __nanosleep(32);      // the first polls back to back: a peer is usually within a few us
      if ((spins & 4095u) == 0 && clock64() - t0 > 20000000000LL) {
        printf("[cxxnet_b200 exchange] rank %d timed out waiting for rank %d (slot %d phase %d seq %u)\n", p.rank,
               (int)threadIdx.x, p.slot, which, seq);
        __trap();
      }
    }
  }
  __syncthreads();
}

// Sequence number of this launch.  The host part is constant per captured CUDA graph; the device
// part is a step counter the training loop bumps on the main stream before a step's first exchange
// launch (and never while one is in flight), so replaying a graph still advances the barriers.
__device__ __forceinline__ unsigned launch_seq(const Params& p) {
  return p.seq + (p.seq_dev != nullptr ? *reinterpret_cast<const volatile unsigned*>(p.seq_dev) : 0u);
}

__device__ __forceinline__ float clip_grad(float g, float c) {
  if (g != g) return 0.0f;
  return fminf(fmaxf(g, -c), c);
}

// Segment of a chunk: segments are aligned to chunks, so the chunk table names it.  The first kMaxBucketSegs
// segments of the bucket are cached in shared memory; a bucket with more (many small layers merged into one
// launch) reads the rest from the global table.  The lookup is off the critical path: the peer loads of a chunk
// do not depend on it (the arenas are padded to whole chunks, padding gradients are zero).
__device__ __forceinline__ Segment load_seg(const Params& p, const Segment* sseg, int seg_lo, long long ch) {
  const int k = __ldg(p.chunk_seg + ch) - seg_lo;
  return k < kMaxBucketSegs ? sseg[k] : p.segs[seg_lo + k];
}

template <int NR, int U>   // NR: compile-time upper bound on ranks per load group, U: chunks in flight
__global__ void __launch_bounds__(kThreads, 4)   // <= 64 registers: must co-reside with the GEMM CTAs it overlaps
fused_rs_update_ag_kernel(const Params p) {
  const unsigned seq = launch_seq(p);
  __shared__ Segment sseg[kMaxBucketSegs];
  char* self = p.peers[p.rank];
  // the segments this bucket covers (contiguous in the table)
  const int seg_lo = p.chunk_seg[p.cb], seg_hi = p.chunk_seg[p.ce - 1];
  const int nseg = min(seg_hi - seg_lo + 1, kMaxBucketSegs);
  if (threadIdx.x < nseg) sseg[threadIdx.x] = p.segs[seg_lo + threadIdx.x];
  // ---------------- entry barrier: every rank's gradients for this bucket are complete
  if (blockIdx.x == 0 && threadIdx.x < p.nranks)
    st_release_sys(flag_ptr(p.peers[threadIdx.x], p.off_flags, 0, p.slot, p.rank), seq);
  wait_all(p, 0, seq);

  // ---------------- phase 1: reduce-scatter + update + all-gather on the owned shard
  const long long nch = p.ce - p.cb;
  const long long per = (nch + p.nranks - 1) / p.nranks;
  const long long my0 = p.cb + (long long)p.rank * per;
  const long long my1 = min(p.ce, my0 + per);
  float* w = reinterpret_cast<float*>(self + p.off_w);
  float* m = reinterpret_cast<float*>(self + p.off_m);
  float* m2 = reinterpret_cast<float*>(self + p.off_m2);
  for (long long ch0 = my0 + (long long)blockIdx.x * U; ch0 < my1; ch0 += (long long)gridDim.x * U) {
    float4 acc[U];
    bool live[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      live[u] = ch0 + u < my1;
    }
    if (p.mc != nullptr && !(p.dbg & 1)) {
      // the switch adds the ranks' gradients: one load per 16 bytes instead of one per rank
      const float* gmc = reinterpret_cast<const float*>(p.mc + p.off_g);
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (live[u]) acc[u] = mc_ld_reduce_f4(gmc + (ch0 + u) * kChunk + threadIdx.x * 4);
    } else
    // issue every load of a group of up to NR ranks before consuming any of them
    for (int r0 = 0; r0 < p.nranks; r0 += NR) {
      float4 buf[NR][U];
#pragma unroll
      for (int rr = 0; rr < NR; ++rr) {
        if (r0 + rr < p.nranks) {
          const int src = (p.dbg & 1) ? p.rank : (p.rank + r0 + rr) % p.nranks;   // local copy first, peers staggered
          const float* gsrc = reinterpret_cast<const float*>(p.peers[src] + p.off_g);
#pragma unroll
          for (int u = 0; u < U; ++u)
            buf[rr][u] = live[u] ? ld_sys_f4(gsrc + (ch0 + u) * kChunk + threadIdx.x * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
#pragma unroll
      for (int rr = 0; rr < NR; ++rr) {
        if (r0 + rr < p.nranks) {
#pragma unroll
          for (int u = 0; u < U; ++u) {
            acc[u].x += buf[rr][u].x; acc[u].y += buf[rr][u].y; acc[u].z += buf[rr][u].z; acc[u].w += buf[rr][u].w;
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (!live[u]) continue;
      const long long e = (ch0 + u) * kChunk + threadIdx.x * 4;
      const Segment s = load_seg(p, sseg, seg_lo, ch0 + u);
      const long long rel = e - s.start;
      if (rel >= s.n) continue;             // chunk padding behind the segment's last element
      float4 wv = *reinterpret_cast<const float4*>(w + e);
      float4 mv = *reinterpret_cast<const float4*>(m + e);
      float4 m2v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (p.kind == 2) m2v = *reinterpret_cast<const float4*>(m2 + e);
      float* wp = reinterpret_cast<float*>(&wv);
      float* gp = reinterpret_cast<float*>(&acc[u]);
      float* mp = reinterpret_cast<float*>(&mv);
      float* m2p = reinterpret_cast<float*>(&m2v);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (rel + j >= s.n) { wp[j] = 0.f; mp[j] = 0.f; continue; }
        float gr = gp[j];
        if (p.kind == 0) {
          if (s.clip != 0.0f) gr = clip_grad(gr, s.clip);
          mp[j] = s.momentum * mp[j] - s.lr * (gr + s.wd * wp[j]);
          wp[j] += mp[j];
        } else if (p.kind == 1) {
          const float old = mp[j];
          mp[j] = s.momentum * old - s.lr * (gr + s.wd * wp[j]);
          wp[j] += (1.0f + s.momentum) * mp[j] - s.momentum * old;
        } else {
          if (s.wd > 0.0f) gr -= s.wd * wp[j];
          mp[j] += s.beta1 * (gr - mp[j]);
          m2p[j] += s.beta2 * (gr * gr - m2p[j]);
          wp[j] -= s.adam_lr_t * (mp[j] / (sqrtf(m2p[j]) + 1e-8f));
        }
      }
      *reinterpret_cast<float4*>(m + e) = mv;
      if (p.kind == 2) *reinterpret_cast<float4*>(m2 + e) = m2v;
      if (p.mc != nullptr && !(p.dbg & 2)) {
        mc_st_f4(reinterpret_cast<float*>(p.mc + p.off_w) + e, wv);      // one store, every rank's arena
      } else {
        for (int r = 0; r < p.nranks; ++r) {
          const int dst = (p.dbg & 2) ? p.rank : (p.rank + r) % p.nranks;
          st_sys_f4(reinterpret_cast<float*>(p.peers[dst] + p.off_w) + e, wv);
        }
      }
    }
  }

  // ---------------- exit barrier: my reads of peer gradients and writes of peer weights are done
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned prev = atomicAdd(p.counter, 1u);
    if (prev == gridDim.x - 1) {
      *p.counter = 0u;
      __threadfence_system();
      for (int r = 0; r < p.nranks; ++r) st_release_sys(flag_ptr(p.peers[r], p.off_flags, 1, p.slot, p.rank), seq);
    }
  }
  wait_all(p, 1, seq);

  // ---------------- phase 2 (local): zero gradients, refresh the bf16 shadow of the bucket
  float* g = reinterpret_cast<float*>(self + p.off_g);
  bf16* shadow = reinterpret_cast<bf16*>(self + p.off_shadow);
  constexpr int V = 4;
  for (long long ch0 = p.cb + (long long)blockIdx.x * V; ch0 < p.ce && !(p.dbg & 4); ch0 += (long long)gridDim.x * V) {
    float4 wv[V];
    bool live[V];
#pragma unroll
    for (int u = 0; u < V; ++u) {
      const long long e = (ch0 + u) * kChunk + threadIdx.x * 4;
      live[u] = ch0 + u < p.ce;
      wv[u] = live[u] ? ld_sys_f4(w + e) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < V; ++u) {
      if (!live[u]) continue;
      const long long e = (ch0 + u) * kChunk + threadIdx.x * 4;
      const Segment s = load_seg(p, sseg, seg_lo, ch0 + u);
      const long long rel = e - s.start;
      if (rel >= s.n) continue;
      *reinterpret_cast<float4*>(g + e) = make_float4(0.f, 0.f, 0.f, 0.f);
      if (s.shadow_off >= 0) {
        const float* wp = reinterpret_cast<const float*>(&wv[u]);
        const long long row = rel / s.cols;
        const long long col = rel - row * s.cols;
        bf16* dst = shadow + s.shadow_off + row * s.shadow_ld + col;
        if (col + 3 < s.cols && rel + 3 < s.n && ((reinterpret_cast<uintptr_t>(dst) & 7) == 0)) {
          __nv_bfloat162 a = __floats2bfloat162_rn(wp[0], wp[1]), b2 = __floats2bfloat162_rn(wp[2], wp[3]);
          uint2 o;
          o.x = *reinterpret_cast<uint32_t*>(&a);
          o.y = *reinterpret_cast<uint32_t*>(&b2);
          *reinterpret_cast<uint2*>(dst) = o;
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const long long k = rel + j;
            if (k < s.n) {
              const long long r2 = k / s.cols;
              shadow[s.shadow_off + r2 * s.shadow_ld + (k - r2 * s.cols)] = __float2bfloat16(wp[j]);
            }
          }
        }
      }
    }
  }
}

// K19: copy [off, off+bytes) of the root's block into every rank's block, then barrier.
__global__ void __launch_bounds__(kThreads) broadcast_kernel(const Params p, long long off, long long bytes, int root) {
  const unsigned seq = launch_seq(p);
  if (p.rank == root) {
    const long long n16 = bytes / 16;
    const uint4* src = reinterpret_cast<const uint4*>(p.peers[root] + off);
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n16;
         i += (long long)gridDim.x * blockDim.x) {
      const uint4 v = src[i];
      for (int r = 0; r < p.nranks; ++r)
        if (r != root) reinterpret_cast<uint4*>(p.peers[r] + off)[i] = v;
    }
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned prev = atomicAdd(p.counter, 1u);
    if (prev == gridDim.x - 1) {
      *p.counter = 0u;
      __threadfence_system();
      for (int r = 0; r < p.nranks; ++r) st_release_sys(flag_ptr(p.peers[r], p.off_flags, 1, p.slot, p.rank), seq);
    }
  }
  wait_all(p, 1, seq);
}

// K18: all-gather of up to two local pieces into every rank's symmetric block, bracketed by the
// same flag barriers as the fused exchange.  Rank r's piece k lands at dst_off[k] + r * slot_bytes[k]
// in every block; bytes beyond valid_bytes[k] (a short last batch) are zero-filled.  Used to
// gather the [activation | output-gradient] rows of a fully connected layer (the reference's
// fullc_gather, src/updater/async_updater-inl.hpp:67-93) and to publish updated weight shards.
struct GatherArgs {
  const char* src[2];
  long long dst_off[2], slot_bytes[2], valid_bytes[2];
  int npieces;
};
__global__ void __launch_bounds__(kThreads) allgather_kernel(const Params p, const GatherArgs a) {
  const unsigned seq = launch_seq(p);
  if (blockIdx.x == 0 && threadIdx.x < p.nranks)
    st_release_sys(flag_ptr(p.peers[threadIdx.x], p.off_flags, 0, p.slot, p.rank), seq);
  wait_all(p, 0, seq);
  for (int k = 0; k < a.npieces; ++k) {
    const long long n16 = a.slot_bytes[k] / 16, v16 = a.valid_bytes[k] / 16;
    const uint4* src = reinterpret_cast<const uint4*>(a.src[k]);
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n16;
         i += (long long)gridDim.x * blockDim.x) {
      const uint4 v = i < v16 ? src[i] : make_uint4(0, 0, 0, 0);
      for (int r = 0; r < p.nranks; ++r) {
        const int dst = (p.rank + r) % p.nranks;
        reinterpret_cast<uint4*>(p.peers[dst] + a.dst_off[k] + (long long)p.rank * a.slot_bytes[k])[i] = v;
      }
    }
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned prev = atomicAdd(p.counter, 1u);
    if (prev == gridDim.x - 1) {
      *p.counter = 0u;
      __threadfence_system();
      for (int r = 0; r < p.nranks; ++r) st_release_sys(flag_ptr(p.peers[r], p.off_flags, 1, p.slot, p.rank), seq);
    }
  }
  wait_all(p, 1, seq);
}

}  // namespace xchg
}  // namespace cxn

using namespace cxn;
using namespace cxn::xchg;

// Device word (4-byte aligned, lives as long as exchanges are launched) whose value every exchange
// kernel adds to its `seq` argument; NULL (the default) = the argument alone.  One GPU per process.
static const unsigned* g_epoch_ptr = nullptr;
// Multicast address of the symmetric block (parallel/symm.py, kind "nvls"), or null: the reduce-scatter and the
// all-gather of the fused exchange then use multimem.ld_reduce / multimem.st.  CXN_XCHG_NVLS=0 keeps the unicast path.
static char* g_mc_base = nullptr;
CXN_API int cxn_xchg_set_multicast(void* mc_base) {
  static const bool off = getenv("CXN_XCHG_NVLS") != nullptr && atoi(getenv("CXN_XCHG_NVLS")) == 0;
  g_mc_base = off ? nullptr : reinterpret_cast<char*>(mc_base);
  return 0;
}
CXN_API int cxn_xchg_set_epoch_ptr(const void* dev_word) {
  g_epoch_ptr = reinterpret_cast<const unsigned*>(dev_word);
  return 0;
}

CXN_API int cxn_xchg_flag_bytes() { return 2 * kMaxSlots * kMaxRanks * 4; }
CXN_API int cxn_xchg_max_ranks() { return kMaxRanks; }
CXN_API int cxn_xchg_max_slots() { return kMaxSlots; }

static int fill_params(Params* p, const unsigned long long* peers, int nranks, int rank, long long off_w,
                       long long off_g, long long off_m, long long off_m2, long long off_shadow, long long off_flags,
                       const void* segs, const int* chunk_seg, long long cb, long long ce, int kind, unsigned seq,
                       int slot, unsigned* counter) {
  if (nranks < 1 || nranks > kMaxRanks || slot < 0 || slot >= kMaxSlots) return -20;
  memset(p, 0, sizeof(*p));
  for (int i = 0; i < nranks; ++i) p->peers[i] = reinterpret_cast<char*>(peers[i]);
  p->off_w = off_w; p->off_g = off_g; p->off_m = off_m; p->off_m2 = off_m2; p->off_shadow = off_shadow;
  p->off_flags = off_flags; p->segs = (const Segment*)segs; p->chunk_seg = chunk_seg; p->cb = cb; p->ce = ce;
  p->rank = rank; p->nranks = nranks; p->kind = kind; p->seq = seq; p->slot = slot; p->counter = counter;
  p->seq_dev = g_epoch_ptr;
  p->mc = nranks > 1 ? g_mc_base : nullptr;
  static const int dbg = getenv("CXN_XCHG_DBG") ? atoi(getenv("CXN_XCHG_DBG")) : 0;
  p->dbg = dbg;
  return 0;
}

// One bucket of the fused gradient exchange.  `peers` is a HOST array of nranks block base
// addresses (this rank's own block at index `rank`).  `grid` CTAs (<= resident capacity; the
// kernel spins on peer flags, so every CTA must be co-resident).
static int launch_fused(const Params& p, int nranks, int grid, cudaStream_t st) {
  if (grid < 1) grid = 1;
  if (grid > 2 * num_sms()) grid = 2 * num_sms();
  // with the multicast address one load returns the reduced gradient: no per-rank load buffers, so more chunks in flight
  // per CTA (the switch round trip is several microseconds; bandwidth = bytes in flight / latency)
  static const int mc_u = getenv("CXN_XCHG_MC_U") ? atoi(getenv("CXN_XCHG_MC_U")) : 1;   // (measured on 8 GPUs, fc6-sized bucket: 1 -> 411 us, 8 -> 456 us)
  if (p.mc != nullptr && mc_u == 8) fused_rs_update_ag_kernel<1, 8><<<grid, kThreads, 0, st>>>(p);
  else if (p.mc != nullptr && mc_u == 4) fused_rs_update_ag_kernel<1, 4><<<grid, kThreads, 0, st>>>(p);
  else if (nranks <= 2) fused_rs_update_ag_kernel<2, 4><<<grid, kThreads, 0, st>>>(p);
  else if (nranks <= 4) fused_rs_update_ag_kernel<4, 2><<<grid, kThreads, 0, st>>>(p);
  else fused_rs_update_ag_kernel<8, 1><<<grid, kThreads, 0, st>>>(p);
  CXN_CHECK_LAUNCH();
  return 0;
}

CXN_API int cxn_fused_rs_update_ag(const unsigned long long* peers, int nranks, int rank, long long off_w,
                                   long long off_g, long long off_m, long long off_m2, long long off_shadow,
                                   long long off_flags, const void* segs, const int* chunk_seg, long long cb,
                                   long long ce, int kind, unsigned seq, int slot, unsigned* counter, int grid,
                                   cudaStream_t st) {
  Params p;
  int rc = fill_params(&p, peers, nranks, rank, off_w, off_g, off_m, off_m2, off_shadow, off_flags, segs, chunk_seg,
                       cb, ce, kind, seq, slot, counter);
  if (rc) return rc;
  if (ce <= cb) return 0;
  return launch_fused(p, nranks, grid, st);
}

// ------------------------------------------------------------------ entry point without allocation
// The fused exchange needs one word of scratch per bucket: the counter its CTAs use to find out
// which of them finished last.  cxn_fused_rs_update_ag takes that word from the caller, which in
// practice means the caller allocated it.  Some hosts cannot: an inference / training server that
// hands all device memory to its own pool at start-up (a later cudaMalloc simply fails), or a
// latency-critical loop that must not hit the implicit device-wide synchronisation of
// cudaMalloc / cudaFree.  For them the counters below live in the library image itself.
//
//   * kNoAllocSlots words = 1 KiB, resident on every GPU that loads the library until exit.
//   * slot must be < kNoAllocSlots (one word per bucket); larger values are refused.
//   * The word is cleared on the stream before every launch, so a call never depends on what an
//     earlier (possibly aborted) call left behind.
//   * All calls share the array: two calls for the same slot that overlap on the GPU race on
//     its word.  Calls on one stream, or for different slots, do not.
//   * The first call on each device resolves the array's device address (cudaGetSymbolAddress
//     with that device current); make that first call outside any stream capture.
constexpr int kNoAllocSlots = 256;
constexpr int kNoAllocMaxDevices = 64;
static __device__ __align__(256) unsigned g_noalloc_counters[kNoAllocSlots];

CXN_API int cxn_fused_rs_update_ag_noalloc_max_slots() { return kNoAllocSlots; }

CXN_API int cxn_fused_rs_update_ag_noalloc(const unsigned long long* peers, int nranks, int rank, long long off_w,
                                           long long off_g, long long off_m, long long off_m2, long long off_shadow,
                                           long long off_flags, const void* segs, const int* chunk_seg, long long cb,
                                           long long ce, int kind, unsigned seq, int slot, int grid,
                                           cudaStream_t st) {
  if (slot < 0 || slot >= kNoAllocSlots) {
    fprintf(stderr, "[cxxnet_b200] cxn_fused_rs_update_ag_noalloc: slot %d is outside the built-in scratch "
                    "(%d slots); use cxn_fused_rs_update_ag with caller-provided scratch\n", slot, kNoAllocSlots);
    return -30;
  }
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= kNoAllocMaxDevices) {
    fprintf(stderr, "[cxxnet_b200] cxn_fused_rs_update_ag_noalloc: device ordinal %d not supported\n", dev);
    return -31;
  }
  // g_noalloc_counters names a host-side handle here, not a device address: resolve it once per
  // device, with that device current, and hand the resolved pointer to the kernel.
  static std::mutex mu;
  static unsigned* dev_ptr[kNoAllocMaxDevices] = {nullptr};
  unsigned* base = nullptr;
  {
    std::lock_guard<std::mutex> lk(mu);
    if (dev_ptr[dev] == nullptr) {
      void* sym = nullptr;
      e = cudaGetSymbolAddress(&sym, g_noalloc_counters);
      if (e != cudaSuccess) {
        fprintf(stderr, "[cxxnet_b200] cudaGetSymbolAddress(g_noalloc_counters) failed on device %d: %s\n", dev,
                cudaGetErrorString(e));
        return (int)e;
      }
      dev_ptr[dev] = static_cast<unsigned*>(sym);
    }
    base = dev_ptr[dev];
  }
  Params p;
  int rc = fill_params(&p, peers, nranks, rank, off_w, off_g, off_m, off_m2, off_shadow, off_flags, segs, chunk_seg,
                       cb, ce, kind, seq, slot, base + slot);
  if (rc) return rc;
  if (ce <= cb) return 0;
  e = cudaMemsetAsync(base + slot, 0, sizeof(unsigned), st);   // the counter must start at zero
  if (e != cudaSuccess) return (int)e;
  return launch_fused(p, nranks, grid, st);
}

CXN_API int cxn_symm_broadcast(const unsigned long long* peers, int nranks, int rank, long long off_flags,
                               long long off, long long bytes, int root, unsigned seq, int slot, unsigned* counter,
                               int grid, cudaStream_t st) {
  Params p;
  int rc = fill_params(&p, peers, nranks, rank, 0, 0, 0, 0, 0, off_flags, nullptr, nullptr, 0, 0, 0, seq, slot, counter);
  if (rc) return rc;
  if (bytes % 16) return -21;
  if (grid < 1) grid = 1;
  if (grid > num_sms()) grid = num_sms();
  broadcast_kernel<<<grid, kThreads, 0, st>>>(p, off, bytes, root);
  CXN_CHECK_LAUNCH();
  return 0;
}

// npieces = 1 or 2; all byte counts multiples of 16; src pointers 16-byte aligned.
CXN_API int cxn_symm_allgather(const unsigned long long* peers, int nranks, int rank, long long off_flags,
                               const void* src0, long long dst_off0, long long slot_bytes0, long long valid_bytes0,
                               const void* src1, long long dst_off1, long long slot_bytes1, long long valid_bytes1,
                               int npieces, unsigned seq, int slot, unsigned* counter, int grid, cudaStream_t st) {
  Params p;
  int rc = fill_params(&p, peers, nranks, rank, 0, 0, 0, 0, 0, off_flags, nullptr, nullptr, 0, 0, 0, seq, slot, counter);
  if (rc) return rc;
  if (npieces < 1 || npieces > 2) return -22;
  GatherArgs a;
  a.src[0] = (const char*)src0; a.dst_off[0] = dst_off0; a.slot_bytes[0] = slot_bytes0; a.valid_bytes[0] = valid_bytes0;
  a.src[1] = (const char*)src1; a.dst_off[1] = dst_off1; a.slot_bytes[1] = slot_bytes1; a.valid_bytes[1] = valid_bytes1;
  a.npieces = npieces;
  for (int k = 0; k < npieces; ++k)
    if ((a.slot_bytes[k] % 16) || (a.valid_bytes[k] % 16) || (reinterpret_cast<uintptr_t>(a.src[k]) & 15)) return -21;
  if (grid < 1) grid = 1;
  if (grid > num_sms()) grid = num_sms();
  allgather_kernel<<<grid, kThreads, 0, st>>>(p, a);
  CXN_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ CUDA IPC helpers
// ---- locality registry: every range this library maps from another GPU is recorded here, and memory_is_local()
// refuses those ranges (as well as anything the driver does not report as plain device memory of the current GPU).
namespace cxn {
namespace {
struct PeerRange { uintptr_t lo, hi; };
std::mutex g_peer_mu;
std::vector<PeerRange> g_peer_ranges;
}  // namespace
static void note_peer_range(const void* p, size_t bytes) {
  std::lock_guard<std::mutex> g(g_peer_mu);
  g_peer_ranges.push_back({reinterpret_cast<uintptr_t>(p), reinterpret_cast<uintptr_t>(p) + bytes});
}
static void forget_peer_range(const void* p) {
  std::lock_guard<std::mutex> g(g_peer_mu);
  for (size_t i = 0; i < g_peer_ranges.size(); ++i)
    if (g_peer_ranges[i].lo == reinterpret_cast<uintptr_t>(p)) { g_peer_ranges.erase(g_peer_ranges.begin() + i); return; }
}
bool memory_is_local(const void* p, size_t bytes) {
  if (p == nullptr) return false;
  const uintptr_t lo = reinterpret_cast<uintptr_t>(p), hi = lo + (bytes ? bytes : 1);
  {
    std::lock_guard<std::mutex> g(g_peer_mu);
    for (const PeerRange& r : g_peer_ranges)
      if (lo < r.hi && r.lo < hi) return false;
  }
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) return false;
  // both ends of the range must be ordinary device allocations of the current device
  const void* ends[2] = {p, reinterpret_cast<const void*>(hi - 1)};
  for (int i = 0; i < 2; ++i) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, ends[i]) != cudaSuccess) { cudaGetLastError(); return false; }
    if (a.type != cudaMemoryTypeDevice || a.device != dev) return false;
    // an IPC / VMM import of a peer allocation reports the exporting device; a mapping the driver attributes to
    // this device but whose base is a recorded peer range was caught above
    CUdeviceptr base = 0; size_t sz = 0;
    typedef CUresult (*RangeFn)(CUdeviceptr*, size_t*, CUdeviceptr);
    static RangeFn range_fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
      void* q = nullptr;
      cudaDriverEntryPointQueryResult r;
      if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &q, cudaEnableDefault, &r) == cudaSuccess &&
          r == cudaDriverEntryPointSuccess)
        range_fn = reinterpret_cast<RangeFn>(q);
    });
    if (range_fn && range_fn(&base, &sz, (CUdeviceptr)ends[i]) == CUDA_SUCCESS) {
      std::lock_guard<std::mutex> g(g_peer_mu);
      for (const PeerRange& r : g_peer_ranges)
        if ((uintptr_t)base < r.hi && r.lo < (uintptr_t)base + sz) return false;
    }
  }
  return true;
}
}  // namespace cxn

// record a range that maps (or multicasts to) another GPU's memory, however it was obtained (VMM imports, multicast
// addresses of torch's symmetric memory): memory_is_local() answers 0 for anything that overlaps it
CXN_API int cxn_note_peer_range(const void* p, unsigned long long bytes) {
  if (p == nullptr || bytes == 0) return -1;
  cxn::note_peer_range(p, (size_t)bytes);
  return 0;
}

// 1 = the range is plain memory of the current device (eligible for the cluster-multicast routes), 0 = not
CXN_API int cxn_memory_is_local(const void* p, unsigned long long bytes) { return cxn::memory_is_local(p, (size_t)bytes) ? 1 : 0; }

CXN_API int cxn_ipc_alloc(unsigned long long bytes, unsigned long long* ptr, unsigned char* handle64) {
  void* d = nullptr;
  cudaError_t e = cudaMalloc(&d, bytes);
  if (e != cudaSuccess) return (int)e;
  e = cudaMemset(d, 0, bytes);
  if (e != cudaSuccess) return (int)e;
  cudaIpcMemHandle_t h;
  e = cudaIpcGetMemHandle(&h, d);
  if (e != cudaSuccess) return (int)e;
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(handle64, &h, 64);
  *ptr = reinterpret_cast<unsigned long long>(d);
  return 0;
}
CXN_API int cxn_ipc_open(const unsigned char* handle64, unsigned long long* ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  void* d = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&d, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) return (int)e;
  // record the mapping's extent so that no multicast tensor map is ever built over it
  size_t sz = 0;
  {
    CUdeviceptr base = 0;
    typedef CUresult (*RangeFn)(CUdeviceptr*, size_t*, CUdeviceptr);
    void* q = nullptr;
    cudaDriverEntryPointQueryResult r;
    if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &q, cudaEnableDefault, &r) == cudaSuccess &&
        r == cudaDriverEntryPointSuccess && q)
      if (reinterpret_cast<RangeFn>(q)(&base, &sz, (CUdeviceptr)d) != CUDA_SUCCESS) sz = 0;
  }
  cxn::note_peer_range(d, sz ? sz : ((size_t)1 << 40));
  *ptr = reinterpret_cast<unsigned long long>(d);
  return 0;
}
CXN_API int cxn_ipc_close(unsigned long long ptr) {
  cxn::forget_peer_range(reinterpret_cast<void*>(ptr));
  return (int)cudaIpcCloseMemHandle(reinterpret_cast<void*>(ptr));
}
CXN_API int cxn_ipc_free(unsigned long long ptr) { return (int)cudaFree(reinterpret_cast<void*>(ptr)); }
