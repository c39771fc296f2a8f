// bn.cu — NHWC SyncBatchNorm for sm_100a: Welford statistics, cross-rank
// exchange by peer stores, fused BN(+residual)+ReLU forward and backward.
//
// Data layout.  An activation is [M, C] row-major (channels_last), M = N*H*W.
//
// Two-kernel path (every layer that does not qualify for the one-kernel path below).
// A CTA of 256 threads is 32 rows x 8 sixteen-byte vectors: it owns one
// 128-byte channel chunk (64 bf16 / 32 fp32 channels) and walks down the rows,
// so every warp request is four full 128-byte lines and every thread keeps the
// same channels for its whole life (their statistics / coefficients stay in
// registers).  grid = (row splits, channel chunks), one resident wave: 4 CTAs
// per SM for the forward kernels (<= 64 registers), 3 for the backward ones.
// Rows are staged through per-thread shared-memory rings filled by cp.async, so
// each thread has 4-8 rows per input in flight (RowRing below).
//
// Reductions (stats, backward sums): thread -> warp shuffles over the rows of a warp ->
// the 8 warp partials through shared memory -> the 8 CTAs of a cluster through DSMEM ->
// one partial per cluster in `scratch` -> the last cluster leader of each chunk (ticket
// counter) merges the partials, writes the rank-local result and, when world > 1, the
// whole CTA stores it into every peer's exchange slot as LL words {value, seq}.  The
// element-wise kernels that follow poll the words of their own chunk, merge the ranks'
// vectors in rank order (so all ranks get bit-identical statistics) and stream.
//
// One-kernel path (b200dp_bn_forward / b200dp_bn_backward on wide, short layers): one
// thread-block cluster per channel chunk does both passes; see "one kernel per direction".
//
// Algorithmic traffic per element (bf16): stats 2 B; fwd apply 4 B (+2 with a
// residual); bwd reduce 4 B; bwd apply 6 B (+2 for d_residual); the ReLU mask
// is recomputed from x (mode 1) or read from y (mode 2, +2 B per bwd pass).
#include <mutex>
#include <stdlib.h>

#include <cooperative_groups.h>

#include "common.cuh"

namespace {

using namespace b200dp;
namespace cg = cooperative_groups;

constexpr int kT = 256;           // threads per CTA
constexpr int kCV = 8;            // 16-byte vectors per chunk row (128 B)
constexpr int kRows = kT / kCV;   // 32 rows per CTA pass
constexpr int kOccFwd = 4;        // resident CTAs per SM, forward kernels
constexpr int kOccBwd = 3;        // resident CTAs per SM, backward kernels
constexpr int kMaxCtas = kOccFwd * B200DP_NUM_SMS;
constexpr int kMaxPartials = kMaxCtas * 64;  // (row split, channel) pairs
constexpr int kUnroll = 4;        // rows per CTA pass assumed by the grid planner
constexpr int kCluster = 8;       // CTAs per thread-block cluster in the reduction kernels

// scratch layout (bytes): tickets[128] u32 | counts[1024] f32 | pa[kMaxPartials] | pb[kMaxPartials]
constexpr size_t kScratchTickets = 0;
constexpr size_t kScratchCounts = 128 * 4;
constexpr size_t kScratchA = kScratchCounts + 1024 * 4;
constexpr size_t kScratchB = kScratchA + (size_t)kMaxPartials * 4;
constexpr size_t kScratchBytes = kScratchB + (size_t)kMaxPartials * 4;

struct Grid2 {
  int gx, gy;
};

template <typename T>
Grid2 plan_grid(int64_t M, int C, int occ) {
  const int chunk = kCV * Vec<T>::N;
  Grid2 g;
  g.gy = (C + chunk - 1) / chunk;
  int64_t want = (M + (int64_t)kRows * kUnroll - 1) / ((int64_t)kRows * kUnroll);
  int cap = occ * B200DP_NUM_SMS / g.gy;
  if (cap < 1) cap = 1;
  g.gx = (int)(want < 1 ? 1 : (want > cap ? cap : want));
  return g;
}

// Reduction kernels run in clusters of kCluster CTAs along x (same channel chunk): the
// row split is a whole number of clusters, and the grid must not exceed the number of
// clusters the GPU can hold at once (GPC boundaries make that a little less than
// occupancy * SMs / kCluster: 71 instead of 74 on B200) or a few clusters run as a second
// wave and double the kernel's duration.
template <typename T>
Grid2 plan_grid_clustered(int64_t M, int C, int max_clusters) {
  Grid2 g = plan_grid<T>(M, C, kOccFwd);
  int cap = max_clusters / g.gy * kCluster;
  g.gx = (g.gx + kCluster - 1) / kCluster * kCluster;
  if (g.gx > cap) g.gx = cap;
  if (g.gx < kCluster) g.gx = kCluster;
  return g;
}

// cudaOccupancyMaxActiveClusters for a kernel of this file with clusters of `cs` CTAs of `threads` threads
// and `dyn_smem` bytes of dynamic shared memory, asked once per (kernel, device, cluster size).  Forward runs
// on the caller's thread and backward on autograd's, so the small cache is guarded.  Returns 0 when the
// runtime cannot place such a cluster at all.  Also opts the kernel in to what the launch will need
// (cluster sizes above 8, more than 48 KB of dynamic shared memory).
std::mutex g_occ_mutex;
template <typename KernelT>
int max_active_clusters(KernelT kernel, int cs = kCluster, int threads = kT, int dyn_smem = 0) {
  struct Entry { const void* fn; int dev; int cs; int n; };
  static Entry table[1024];
  static int used = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 64;
  const void* key = reinterpret_cast<const void*>(kernel);
  std::lock_guard<std::mutex> lock(g_occ_mutex);
  for (int i = 0; i < used; ++i)
    if (table[i].fn == key && table[i].dev == dev && table[i].cs == cs) return table[i].n;
  if (cs > 8) {  // clusters of more than 8 CTAs are an opt-in per kernel
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) cudaGetLastError();
  }
  if (dyn_smem > 48 * 1024) {
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_smem) != cudaSuccess) cudaGetLastError();
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(cs * B200DP_NUM_SMS, 1, 1);
  cfg.blockDim = dim3(threads, 1, 1);
  cfg.dynamicSmemBytes = dyn_smem;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = cs;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, kernel, &cfg) != cudaSuccess || n <= 0) {
    cudaGetLastError();
    n = (cs == kCluster && threads == kT) ? 64 : 0;  // the two-kernel path: be conservative (8 per GPC); else: not available
  }
  if (used < 1024) {
    table[used].fn = key;
    table[used].dev = dev;
    table[used].cs = cs;
    table[used].n = n;
    ++used;
  }
  return n;
}

// ---- Chan / Welford merge of (na, ma, qa) with (nb, mb, qb), N channels -----
template <int N>
__device__ __forceinline__ void welford_merge(float& na, float* ma, float* qa, float nb,
                                              const float* mb, const float* qb) {
  const float n = na + nb;
  if (nb > 0.f) {
    const float fb = nb / n;  // n > 0 here
    const float w = na * fb;
#pragma unroll
    for (int k = 0; k < N; ++k) {
      const float d = mb[k] - ma[k];
      ma[k] = fmaf(d, fb, ma[k]);
      qa[k] = qa[k] + qb[k] + d * d * w;
    }
  }
  na = n;
}

// ---- reduction over the row slots of a CTA ---------------------------------------------------
// Thread tid owns channel group tid % cvw and row slot tid / cvw (cvw = 16-byte vectors per row chunk:
// 8 in the two-kernel path, 2/4/8 in the one-kernel path), so the lanes of a warp that share a channel
// group are cvw apart: log2(32 / cvw) shuffle steps leave the warp's partial in lanes < cvw.
template <int N, bool WELFORD>
__device__ __forceinline__ void warp_fold(int cvw, float& n, float* a, float* b) {
  for (int o = 16; o >= cvw; o >>= 1) {
    float oa[N], ob[N];
    const float on = WELFORD ? __shfl_down_sync(0xffffffffu, n, o) : 0.f;
#pragma unroll
    for (int k = 0; k < N; ++k) {
      oa[k] = __shfl_down_sync(0xffffffffu, a[k], o);
      ob[k] = __shfl_down_sync(0xffffffffu, b[k], o);
    }
    if (WELFORD) {
      welford_merge<N>(n, a, b, on, oa, ob);
    } else {
#pragma unroll
      for (int k = 0; k < N; ++k) {
        a[k] += oa[k];
        b[k] += ob[k];
      }
    }
  }
}

// The 8 warp partials meet in shared memory (wp: kT/32 * cvw * (2N+1) floats) and threads tid < cvw
// fold them in warp order: one barrier instead of the five of a tree over 32 row slots.  On return
// threads tid < cvw hold the CTA's (n, a[], b[]) for their channel group.
template <int N, bool WELFORD>
__device__ __forceinline__ void cta_fold(int cvw, float* wp, float& n, float* a, float* b) {
  constexpr int W = 2 * N + 1;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  warp_fold<N, WELFORD>(cvw, n, a, b);
  if (lane < cvw) {
    float* d = wp + (w * cvw + lane) * W;
#pragma unroll
    for (int k = 0; k < N; ++k) {
      d[k] = a[k];
      d[N + k] = b[k];
    }
    d[2 * N] = n;
  }
  __syncthreads();
  if (tid < cvw) {
    for (int q = 1; q < kT / 32; ++q) {
      const float* src = wp + (q * cvw + tid) * W;
      float oa[N], ob[N];
#pragma unroll
      for (int k = 0; k < N; ++k) {
        oa[k] = src[k];
        ob[k] = src[N + k];
      }
      if (WELFORD) {
        welford_merge<N>(n, a, b, src[2 * N], oa, ob);
      } else {
#pragma unroll
        for (int k = 0; k < N; ++k) {
          a[k] += oa[k];
          b[k] += ob[k];
        }
      }
    }
  }
}

// Same interface as the tree it replaces: sA is the CTA's reduction scratch (>= 8*8*(2N+1) floats).
template <int N, bool WELFORD>
__device__ __forceinline__ void cta_tree(float* sA, float*, float*, float& n, float* a, float* b) {
  cta_fold<N, WELFORD>(kCV, sA, n, a, b);
}

// After the CTA tree, threads with r == 0 hold the CTA's (n, a[], b[]) for their
// channels.  Stage 1: the kCluster CTAs of a cluster hand their results to the cluster's
// first CTA through distributed shared memory.  Stage 2: the cluster leaders publish one
// partial each to `scratch`, elect the last leader of the chunk (ticket), and that CTA
// merges the partials.  Returns true in that one CTA, where r == 0 threads then hold the
// chunk's final values.  (592 CTAs leave 74 partials instead of 592: the serial tail of
// the last CTA is what bounded the reduction kernels before.)
template <int N, bool WELFORD>
__device__ __forceinline__ bool grid_merge(void* scratch, int C, int chbase, bool chan_ok,
                                           float* sA, float* sB, float* sN, float& n, float* a,
                                           float* b, cg::cluster_group::arrival_token&& started) {
  __shared__ bool is_last;
  // the leader's copy is written by all CTAs of the cluster (DSMEM stores), one row per CTA rank
  __shared__ __align__(16) float xchg[kCluster][kCV][2 * N + 4];
  const int tid = threadIdx.x, r = tid / kCV;
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned rank = cluster.block_rank();
  // every CTA of the cluster arrived on this barrier when it started: DSMEM of the leader exists
  cluster.barrier_wait(std::move(started));
  if (r == 0 && rank != 0) {  // push: nobody waits on a remote read, and one cluster barrier is enough
    float* dst = cluster.map_shared_rank(&xchg[rank][tid][0], 0);
#pragma unroll
    for (int k = 0; k < N; ++k) {
      dst[k] = a[k];
      dst[N + k] = b[k];
    }
    dst[2 * N] = n;
  }
  cluster.sync();  // release/acquire at cluster scope: the pushed values are visible to the leader
  if (rank != 0) return false;
  if (r == 0) {
    for (unsigned q = 1; q < kCluster; ++q) {
      const float* src = &xchg[q][tid][0];
      float oa[N], ob[N];
#pragma unroll
      for (int k = 0; k < N; ++k) {
        oa[k] = src[k];
        ob[k] = src[N + k];
      }
      if (WELFORD) {
        welford_merge<N>(n, a, b, src[2 * N], oa, ob);
      } else {
#pragma unroll
        for (int k = 0; k < N; ++k) {
          a[k] += oa[k];
          b[k] += ob[k];
        }
      }
    }
  }

  unsigned int* tickets = reinterpret_cast<unsigned int*>((char*)scratch + kScratchTickets);
  float* counts = reinterpret_cast<float*>((char*)scratch + kScratchCounts);
  float* pa = reinterpret_cast<float*>((char*)scratch + kScratchA);
  float* pb = reinterpret_cast<float*>((char*)scratch + kScratchB);
  const int bx = blockIdx.x / kCluster, by = blockIdx.y, gx = gridDim.x / kCluster;
  // partial index: (by * gx + bx) * chunk + local channel  (< kMaxPartials by construction)
  const int chunk = kCV * N;
  const int lc = (tid % kCV) * N;
  if (gx == 1) return true;  // a single cluster per chunk: its result is final

  if (r == 0) {
    float* da = pa + ((size_t)by * gx + bx) * chunk + lc;
    float* db = pb + ((size_t)by * gx + bx) * chunk + lc;
#pragma unroll
    for (int k = 0; k < N; ++k) {
      da[k] = a[k];
      db[k] = b[k];
    }
    if (WELFORD && tid == 0) counts[by * gx + bx] = n;
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    const unsigned int t = atomicAdd(&tickets[by], 1u);
    is_last = (t == (unsigned)gx - 1);
    if (is_last) tickets[by] = 0;  // self-resetting for the next call
  }
  __syncthreads();
  if (!is_last) return false;
  __threadfence();

  // every thread folds the partials bx = r, r+32, …; then one more CTA tree
  n = 0.f;
#pragma unroll
  for (int k = 0; k < N; ++k) a[k] = b[k] = 0.f;
  for (int p = r; p < gx; p += kRows) {
    const float* sa = pa + ((size_t)by * gx + p) * chunk + lc;
    const float* sb = pb + ((size_t)by * gx + p) * chunk + lc;
    float oa[N], ob[N];
#pragma unroll
    for (int k = 0; k < N; ++k) {
      oa[k] = ld_cg_f32(sa + k);
      ob[k] = ld_cg_f32(sb + k);
    }
    if (WELFORD) {
      welford_merge<N>(n, a, b, ld_cg_f32(counts + by * gx + p), oa, ob);
    } else {
#pragma unroll
      for (int k = 0; k < N; ++k) {
        a[k] += oa[k];
        b[k] += ob[k];
      }
    }
  }
  (void)chan_ok;
  (void)C;
  (void)chbase;
  cta_tree<N, WELFORD>(sA, sB, sN, n, a, b);
  return true;
}

// ---- cross-rank exchange of one channel chunk ---------------------------------------------------
// A chunk is `cw` consecutive channels starting at `cb` (cw <= kChunkMax).  Its exchange vector is
// a[cw] | b[cw] | cnt, staged in shared memory as stage[0..cw) | stage[cw..2cw) | stage[2cw].  In a rank's
// slot (b200dp_sync_area::bn_ll[slot][rank]) a[c] sits at word c, b[c] at C + c, and the count at
// 2*MAX_CHANNELS + cb/64 (every chunk that covers part of a 64-channel group writes the same value
// there), so ranks whose kernels cut the channels into different chunks still understand each other.
// Words are LL words {value, seq}; two adjacent ones travel as one 16-byte store / load.
constexpr int kChunkMax = kCV * 8;                 // channels per chunk at most (bf16, 8 vectors)
constexpr int kXrStride = 2 * kChunkMax + 4;       // floats per rank in the gathered copy

// The whole CTA stores the staged vector into every rank's slot: world * (cw + 1) stores spread over
// its threads (before: 8 threads issuing 17 scalar stores per peer, back to back).
__device__ __forceinline__ void push_chunk(const b200dp_peers& P, uint32_t seq, int C, int cb, int cw,
                                           const float* stage) {
  const int slot = seq % B200DP_XCHG_SLOTS;
  const int valid = (C - cb < cw) ? C - cb : cw;  // multiple of 4: whole pairs
  const int pairs = valid >> 1;
  const int units = 2 * pairs + 1;
  for (int u = threadIdx.x; u < P.world * units; u += blockDim.x) {
    const int p = u / units;
    int j = u - p * units;
    uint2* d = P.sync[p]->bn_ll[slot][P.rank];
    if (j < pairs) {
      st_ll2(d + cb + 2 * j, stage[2 * j], stage[2 * j + 1], seq);
    } else if (j < 2 * pairs) {
      j -= pairs;
      st_ll2(d + C + cb + 2 * j, stage[cw + 2 * j], stage[cw + 2 * j + 1], seq);
    } else {
      st_ll(d + 2 * B200DP_BN_MAX_CHANNELS + (cb >> 6), stage[2 * cw], seq);
    }
  }
}

// The whole CTA polls the ranks' words of this chunk in its own sync area (the peers stored them there)
// and leaves them in shared memory: xr[rank * kXrStride + ...] = a[cw] | b[cw] | cnt.  At most
// ceil(world * (cw + 1) / blockDim.x) polls per thread, side by side (before: 17 one after the other).
__device__ __forceinline__ void wait_and_gather(const b200dp_peers& P, uint32_t seq, int C, int cb, int cw,
                                                float* xr) {
  const int slot = seq % B200DP_XCHG_SLOTS;
  const int valid = (C - cb < cw) ? C - cb : cw;
  const int pairs = valid >> 1;
  const int units = 2 * pairs + 1;
  bool ok = true;
  for (int u = threadIdx.x; u < P.world * units; u += blockDim.x) {
    const int p = u / units;
    int j = u - p * units;
    const uint2* d = P.sync[P.rank]->bn_ll[slot][p];
    float* o = xr + p * kXrStride;
    if (j < pairs) {
      ok &= ld_ll2(d + cb + 2 * j, seq, o + 2 * j, o + 2 * j + 1, P.timeout);
    } else if (j < 2 * pairs) {
      j -= pairs;
      ok &= ld_ll2(d + C + cb + 2 * j, seq, o + cw + 2 * j, o + cw + 2 * j + 1, P.timeout);
    } else {
      ok &= ld_ll(d + 2 * B200DP_BN_MAX_CHANNELS + (cb >> 6), seq, o + 2 * cw, P.timeout);
    }
  }
  if (!ok) atomicExch(&P.sync[P.rank]->error, 1u);
  __syncthreads();
}

// ---- per-thread staging ring in shared memory -------------------------------
// Each thread owns D slots of 16 bytes per input stream and fills them with
// cp.async (LDGSTS): the loads of its next D rows are in flight while it works on
// the current one, at no cost in registers.  Slots are private to the thread, so
// the only synchronisation is cp.async.wait_group.  With 3-4 CTAs per SM this
// keeps 100-150 KB of requests per SM outstanding, which is what the HBM3e
// latency-bandwidth product asks for (~45-65 KB per SM at 6.5 TB/s).
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async16_evict_first(uint32_t dst, const void* src, uint64_t pol) {
  asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;" ::"r"(dst), "l"(src),
               "l"(pol)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "r"(addr));
  return r;
}

// Programmatic dependent launch (sm_90+): the reduction kernel lets its consumer start as soon
// as every CTA has finished streaming, so the consumer's launch, its primed loads and the
// first DRAM round trip overlap the reduction's serial tail; the consumer blocks in
// pdl_wait() until the producer grid has completed and its writes are visible.  Both are
// no-ops when the kernels are launched the ordinary way.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// rows this thread owns: first row, distance between its rows, how many
struct RowRange {
  int64_t first, step;
  int count;
};
__device__ __forceinline__ RowRange my_rows(int64_t M, int r) {
  RowRange rr;
  rr.step = (int64_t)gridDim.x * kRows;
  rr.first = (int64_t)blockIdx.x * kRows + r;
  rr.count = rr.first < M ? (int)((M - rr.first + rr.step - 1) / rr.step) : 0;
  return rr;
}

template <int D, int NS, bool EVICT_FIRST, bool W2 = false, int NT = kT>  // W2: stream 2 carries one 32-bit word per row; NT: threads per CTA
struct RowRing {
  uint32_t base;  // shared-space address of this thread's slot (stream 0, depth 0)
  uint64_t pol;
  __device__ __forceinline__ void init(void* smem) {
    base = (uint32_t)__cvta_generic_to_shared(smem) + threadIdx.x * 16;
    // Keep the slot address in ONE ordinary register.  Left to itself ptxas 12.9 may split it into a
    // uniform shared-window base plus a per-thread offset and emit LDGSTS [R+UR], desc[UR] with the
    // uniform operands unset (seen in bn_bwd_apply_kernel<float, *>): the GPU raises "illegal instruction".
    // (An identity shuffle: the value is unchanged, but the compiler can no longer prove which part of it
    // is uniform across the warp.)
    // init() may run under a divergent branch (threads past the last channel skip it): only the lanes
    // that are here take part.
    base = __shfl_sync(__activemask(), base, threadIdx.x & 31);
    if (EVICT_FIRST) pol = evict_first_policy();
  }
  __device__ __forceinline__ uint32_t slot(int d, int s) const { return base + (s * D + d) * (NT * 16); }
  __device__ __forceinline__ void issue(int d, const void* p0, const void* p1, const void* p2) {
    if (EVICT_FIRST) {
      cp_async16_evict_first(slot(d, 0), p0, pol);
      if (NS > 1) cp_async16_evict_first(slot(d, 1), p1, pol);
      if (NS > 2 && !W2) cp_async16_evict_first(slot(d, 2), p2, pol);
    } else {
      cp_async16(slot(d, 0), p0);
      if (NS > 1) cp_async16(slot(d, 1), p1);
      if (NS > 2 && !W2) cp_async16(slot(d, 2), p2);
    }
    if (NS > 2 && W2)
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(slot(d, 2)), "l"(p2) : "memory");
  }
  // Fill the ring with the first rows; one commit group per row, always D groups.
  template <typename PtrF>
  __device__ __forceinline__ void prime(int nrows, PtrF ptrs) {
#pragma unroll
    for (int k = 0; k < D; ++k) {
      if (k < nrows) {
        const void *p0, *p1, *p2;
        ptrs(k, p0, p1, p2);
        issue(k, p0, p1, p2);
      }
      cp_async_commit();
    }
  }
  // proc(k, v0, v1, v2) for k = 0 … nrows-1 in order; refills the slot it has just drained.
  template <typename PtrF, typename ProcF>
  __device__ __forceinline__ void run(int nrows, PtrF ptrs, ProcF proc) {
    int d = 0;
    for (int k = 0; k < nrows; ++k) {
      cp_async_wait<D - 1>();  // the group of row k has landed
      uint4 v0 = lds128(slot(d, 0)), v1, v2;
      if (NS > 1) v1 = lds128(slot(d, 1));
      if (NS > 2 && !W2) v2 = lds128(slot(d, 2));
      if (NS > 2 && W2) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v2.x) : "r"(slot(d, 2)));
      proc(k, v0, v1, v2);     // consumes the values: the slot may be overwritten after this
      if (k + D < nrows) {
        const void *p0, *p1, *p2;
        ptrs(k + D, p0, p1, p2);
        issue(d, p0, p1, p2);
      }
      cp_async_commit();
      d = (d + 1 == D) ? 0 : d + 1;
    }
    cp_async_wait<0>();
  }
  static constexpr int kBytes = D * NS * NT * 16;
};


// per-channel vectors: N consecutive floats, 16-byte aligned
template <int N>
__device__ __forceinline__ void ldg_vec(const float* p, float* out) {
#pragma unroll
  for (int k = 0; k < N; k += 4) {
    const float4 v = *reinterpret_cast<const float4*>(p + k);
    out[k] = v.x; out[k + 1] = v.y; out[k + 2] = v.z; out[k + 3] = v.w;
  }
}
template <int N>
__device__ __forceinline__ void sts_vec(float* p, const float* in) {
#pragma unroll
  for (int k = 0; k < N; k += 4) *reinterpret_cast<float4*>(p + k) = make_float4(in[k], in[k + 1], in[k + 2], in[k + 3]);
}
// lanes of a packed bf16 pair kept (0xffff) or cleared (0) by bits 2j and 2j+1 of a ReLU mask word
__device__ __forceinline__ uint32_t pair_keep(uint32_t bits, int j) {
  return ((bits >> (2 * j)) & 1u) * 0xffffu | ((bits >> (2 * j + 1)) & 1u) * 0xffff0000u;
}
__device__ __forceinline__ uint4 apply_mask_word(const uint4& g, uint32_t bits) {
  return make_uint4(g.x & pair_keep(bits, 0), g.y & pair_keep(bits, 1), g.z & pair_keep(bits, 2),
                    g.w & pair_keep(bits, 3));
}

template <typename T>
struct is_bf16 { static constexpr bool value = false; };
template <>
struct is_bf16<__nv_bfloat16> { static constexpr bool value = true; };

// =============================== forward: stats ===============================
constexpr int kDepthStats = 8;
constexpr int kDefaultBnOrder = 10;  // the apply kernels walk against the reductions (B200: -1 % step time)

template <typename T>
__global__ void __cluster_dims__(kCluster, 1, 1) __launch_bounds__(kT, kOccFwd)
bn_stats_kernel(const T* __restrict__ x, int64_t M, int C, void* scratch, float* stats_local,
                const __grid_constant__ b200dp_peers P, uint32_t seq, int rev) {
  constexpr int N = Vec<T>::N;
  using Ring = RowRing<kDepthStats, 1, false>;
  constexpr int kTreeBytes = (2 * N * kT + kT) * 4;
  __shared__ __align__(16) unsigned char smem[Ring::kBytes > kTreeBytes ? Ring::kBytes : kTreeBytes];
  float* sA = reinterpret_cast<float*>(smem);  // the reduction tree reuses the ring once it is drained
  float* sB = sA + N * kT;
  float* sN = sB + N * kT;
  const int tid = threadIdx.x, r = tid / kCV, cv = tid % kCV;
  const int c0 = blockIdx.y * (kCV * N) + cv * N;
  const bool chan_ok = c0 < C;
  auto started = cg::this_cluster().barrier_arrive();  // waited on just before the DSMEM push

  float n = 0.f, mean[N], m2[N];
#pragma unroll
  for (int k = 0; k < N; ++k) mean[k] = m2[k] = 0.f;

  pdl_wait();  // no-op unless launched as a programmatic dependent of the kernel that writes x
  if (chan_ok) {
    const RowRange rr = my_rows(M, r);
    // rev: walk this thread's rows from the last to the first (see bn_order())
    const int64_t step = rev ? -rr.step * C : rr.step * C;
    const T* base = x + rr.first * C + c0 + (rev ? (int64_t)(rr.count - 1) * rr.step * C : 0);
    auto ptrs = [&](int k, const void*& p0, const void*& p1, const void*& p2) {
      p0 = base + k * step;
      p1 = p2 = nullptr;
    };
    auto proc = [&](int, const uint4& v, const uint4&, const uint4&) {
      if constexpr (is_bf16<T>::value) {
        // bf16 rows: raw first and second moments with FHADD/FHFMA (2 instructions per
        // element, products exact in fp32); turned into (mean, M2) once per thread below
        // and merged from there on with Chan's formula.  A thread owns at most a few
        // hundred rows, so the single-pass cancellation stays far below bf16's own step.
#pragma unroll
        for (int k = 0; k < N; ++k) {
          const uint16_t h = vec_half(v, k);
          mean[k] = add_bf16_f32(h, mean[k]);      // running sum
          m2[k] = fma_bf16_f32(h, h, m2[k]);       // running sum of squares
        }
      } else {
        float f[N];
        Vec<T>::unpack(v, f);
        n += 1.f;
        const float rn = __frcp_rn(n);
#pragma unroll
        for (int k = 0; k < N; ++k) {
          const float d = f[k] - mean[k];
          mean[k] = fmaf(d, rn, mean[k]);
          m2[k] = fmaf(d, f[k] - mean[k], m2[k]);
        }
      }
    };
    Ring ring;
    ring.init(smem);
    ring.prime(rr.count, ptrs);
    ring.run(rr.count, ptrs, proc);
    if constexpr (is_bf16<T>::value) {
      n = (float)rr.count;
      if (rr.count > 0) {
        const float rn = 1.0f / n;
#pragma unroll
        for (int k = 0; k < N; ++k) {
          const float s = mean[k];
          mean[k] = s * rn;
          m2[k] = fmaxf(fmaf(-s, mean[k], m2[k]), 0.f);
        }
      }
    }
  }
  pdl_launch_dependents();
  __syncthreads();  // every thread is done with its ring slots before the tree overwrites them
  cta_tree<N, true>(sA, sB, sN, n, mean, m2);
  if (!grid_merge<N, true>(scratch, C, c0, chan_ok, sA, sB, sN, n, mean, m2, std::move(started))) return;

  if (r == 0 && chan_ok) {
#pragma unroll
    for (int k = 0; k < N; ++k) {
      stats_local[c0 + k] = mean[k];
      stats_local[C + c0 + k] = m2[k];
    }
    if (tid == 0 && blockIdx.y == 0) stats_local[2 * C] = n;
  }
  if (P.world > 1) {
    constexpr int cw = kCV * N;
    __syncthreads();  // the fold above is done with sA everywhere
    if (r == 0) {
#pragma unroll
      for (int k = 0; k < N; ++k) {
        sA[cv * N + k] = mean[k];
        sA[cw + cv * N + k] = m2[k];
      }
      if (tid == 0) sA[2 * cw] = n;
    }
    __syncthreads();
    push_chunk(P, seq, C, blockIdx.y * cw, cw, sA);
  }
}

// ============================ forward: normalise =============================
struct FwdArgs {
  const void* x;
  const void* res;
  void* y;
  int64_t M;
  int C;
  const float* stats_local;
  const float* weight;
  const float* bias;
  float eps, momentum;
  float* running_mean;
  float* running_var;
  float* save_mean;
  float* save_invstd;
  float* save_count;
  // inference: use these instead of batch statistics
  const float* given_mean;
  const float* given_var;
  // optional output: one 32-bit word per 16-byte vector of y, bit j = (y[j] > 0)
  uint32_t* maskw;
  int rev;  // walk the rows from the last to the first
  long long* batches_tracked;  // optional: the layer's num_batches_tracked counter, incremented once
};

// Per-channel coefficients of one chunk are worked out once per CTA, by its first kCV
// threads (one 16-byte channel group each), and handed to the other 31 rows of threads
// through shared memory: a small layer gives every thread only a handful of rows, and
// 256 threads each redoing the divisions and square roots cost more than the rows.
template <typename T, bool RELU, bool RES>
__global__ void __launch_bounds__(kT, kOccFwd) bn_fwd_apply_kernel(const __grid_constant__ FwdArgs A, const __grid_constant__ b200dp_peers P, uint32_t seq) {
  constexpr int N = Vec<T>::N;
  using Ring = RowRing<RES ? 4 : 8, RES ? 2 : 1, true>;
  __shared__ __align__(16) unsigned char smem[Ring::kBytes];
  __shared__ __align__(16) float coef[3][kChunkMax];  // -mean, a = invstd*weight, b = bias
  const int tid = threadIdx.x, r = tid / kCV, cv = tid % kCV;
  const int C = A.C;
  const int c0 = blockIdx.y * (kCV * N) + cv * N;
  const bool chan_ok = c0 < C;

  // Start streaming now: the first rows fly while the flags are awaited and the
  // coefficients are computed below.
  const T* xb = static_cast<const T*>(A.x);
  const T* rb = static_cast<const T*>(A.res);
  T* yb = static_cast<T*>(A.y);
  const RowRange rr = my_rows(A.M, r);
  const int64_t step = A.rev ? -rr.step * C : rr.step * C;
  const int64_t off0 = rr.first * C + c0 + (A.rev ? (int64_t)(rr.count - 1) * rr.step * C : 0);
  auto ptrs = [&](int k, const void*& p0, const void*& p1, const void*& p2) {
    p0 = xb + off0 + k * step;
    p1 = RES ? (const void*)(rb + off0 + k * step) : nullptr;
    p2 = nullptr;
  };
  Ring ring;
  ring.init(smem);
  const int nrows = chan_ok ? rr.count : 0;
  ring.prime(nrows, ptrs);
  if (A.batches_tracked && blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) *A.batches_tracked += 1;
  pdl_wait();  // statistics of the producer kernel are complete and visible from here on

  __shared__ __align__(16) float xr[B200DP_MAX_WORLD * kXrStride];
  constexpr int cw = kCV * N;
  if (!A.given_mean && P.world > 1) wait_and_gather(P, seq, C, blockIdx.y * cw, cw, xr);
  if (tid < kCV && chan_ok) {  // r == 0: this thread's c0 is the group it publishes
    float mean[N], a[N], b[N];
    if (A.given_mean) {
#pragma unroll
      for (int k = 0; k < N; ++k) {
        mean[k] = A.given_mean[c0 + k];
        const float invstd = 1.0f / sqrtf(A.given_var[c0 + k] + A.eps);
        a[k] = invstd * (A.weight ? A.weight[c0 + k] : 1.f);
        b[k] = A.bias ? A.bias[c0 + k] : 0.f;
      }
    } else {
      float n = 0.f, m2[N];
#pragma unroll
      for (int k = 0; k < N; ++k) mean[k] = m2[k] = 0.f;
      if (P.world > 1) {
        for (int p = 0; p < P.world; ++p) {  // rank order: every rank computes identical statistics
          const float* o = xr + p * kXrStride;
          welford_merge<N>(n, mean, m2, o[2 * cw], o + cv * N, o + cw + cv * N);
        }
      } else {
        n = ld_cg_f32(A.stats_local + 2 * C);
#pragma unroll
        for (int k = 0; k < N; ++k) {
          mean[k] = ld_cg_f32(A.stats_local + c0 + k);
          m2[k] = ld_cg_f32(A.stats_local + C + c0 + k);
        }
      }
      const float inv_n = n > 0.f ? 1.0f / n : 0.f;
      const float inv_nm1 = n > 1.f ? 1.0f / (n - 1.f) : 0.f;
      const bool writer = blockIdx.x == 0;
#pragma unroll
      for (int k = 0; k < N; ++k) {
        const float var = m2[k] * inv_n;
        const float invstd = 1.0f / sqrtf(var + A.eps);
        a[k] = invstd * (A.weight ? A.weight[c0 + k] : 1.f);
        b[k] = A.bias ? A.bias[c0 + k] : 0.f;
        if (writer) {
          A.save_mean[c0 + k] = mean[k];
          A.save_invstd[c0 + k] = invstd;
          if (A.running_mean)
            A.running_mean[c0 + k] =
                (1.f - A.momentum) * A.running_mean[c0 + k] + A.momentum * mean[k];
          if (A.running_var)
            A.running_var[c0 + k] =
                (1.f - A.momentum) * A.running_var[c0 + k] + A.momentum * (m2[k] * inv_nm1);
        }
      }
      if (writer && tid == 0 && blockIdx.y == 0) A.save_count[0] = n;
    }
#pragma unroll
    for (int k = 0; k < N; ++k) mean[k] = -mean[k];
    sts_vec<N>(&coef[0][cv * N], mean);
    sts_vec<N>(&coef[1][cv * N], a);
    sts_vec<N>(&coef[2][cv * N], b);
  }
  __syncthreads();
  float nmean[N], a[N], b[N];
  ldg_vec<N>(&coef[0][cv * N], nmean);  // (generic loads from shared memory)
  ldg_vec<N>(&coef[1][cv * N], a);
  ldg_vec<N>(&coef[2][cv * N], b);

  auto proc = [&](int k, const uint4& vx, const uint4& vr, const uint4&) {
    float f[N];
    if constexpr (is_bf16<T>::value) {
#pragma unroll
      for (int j = 0; j < N; ++j) {
        float o = fmaf(add_bf16_f32(vec_half(vx, j), nmean[j]), a[j], b[j]);
        if (RES) o = add_bf16_f32(vec_half(vr, j), o);
        f[j] = RELU ? fmaxf(o, 0.f) : o;
      }
    } else {
      float g[N];
      Vec<T>::unpack(vx, f);
      if (RES) Vec<T>::unpack(vr, g);
#pragma unroll
      for (int j = 0; j < N; ++j) {
        float o = fmaf(f[j] + nmean[j], a[j], b[j]);
        if (RES) o += g[j];
        f[j] = RELU ? fmaxf(o, 0.f) : o;
      }
    }
    // y is read again by the next convolution: leave it in L2
    const uint4 pk = Vec<T>::pack(f);
    *reinterpret_cast<uint4*>(yb + off0 + k * step) = pk;
    if constexpr (RELU && is_bf16<T>::value) {
      if (A.maskw) {  // post-ReLU halves are >= +0: positive <=> non-zero
        const uint32_t w[4] = {pk.x, pk.y, pk.z, pk.w};
        uint32_t bits = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          bits |= ((w[j] & 0xffffu) != 0u ? 1u : 0u) << (2 * j);
          bits |= ((w[j] >> 16) != 0u ? 1u : 0u) << (2 * j + 1);
        }
        A.maskw[(off0 + k * step) >> 3] = bits;  // element offset / 8 = vector index
      }
    }
  };
  ring.run(nrows, ptrs, proc);
}

// ============================= backward: reduce ==============================
// RELU: 0 none, 1 mask recomputed from x (no residual was fused), 2 mask from y,
// 3 (bf16) mask from the forward's bit words (one uint32 per 16-byte vector; `yo` points at them).
template <typename T, int RELU>
__global__ void __cluster_dims__(kCluster, 1, 1) __launch_bounds__(kT, kOccFwd)
bn_bwd_reduce_kernel(const T* __restrict__ dy, const T* __restrict__ x, const T* __restrict__ yo,
                     int64_t M, int C, const float* save_mean, const float* save_invstd,
                     const float* weight, const float* bias, void* scratch, float* sums_local,
                     float* grad_weight, float* grad_bias, const __grid_constant__ b200dp_peers P, uint32_t seq, int rev) {
  constexpr int N = Vec<T>::N;
  using Ring = RowRing<RELU >= 2 ? 3 : 4, RELU >= 2 ? 3 : 2, false, RELU == 3>;  // 32 or 36 KiB (static smem limit 48 KiB)
  constexpr int kTreeBytes = 2 * N * kT * 4;
  __shared__ __align__(16) unsigned char smem[Ring::kBytes > kTreeBytes ? Ring::kBytes : kTreeBytes];
  float* sA = reinterpret_cast<float*>(smem);
  float* sB = sA + N * kT;
  const int tid = threadIdx.x, r = tid / kCV, cv = tid % kCV;
  const int c0 = blockIdx.y * (kCV * N) + cv * N;
  const bool chan_ok = c0 < C;
  auto started = cg::this_cluster().barrier_arrive();  // waited on just before the DSMEM push

  float s1[N], s2[N];
#pragma unroll
  for (int k = 0; k < N; ++k) s1[k] = s2[k] = 0.f;
  pdl_wait();  // no-op unless launched as a programmatic dependent of the kernel that writes dy
  if (chan_ok) {
    const RowRange rr = my_rows(M, r);
    const int64_t step = rev ? -rr.step * C : rr.step * C;
    const int64_t off0 = rr.first * C + c0 + (rev ? (int64_t)(rr.count - 1) * rr.step * C : 0);
    auto ptrs = [&](int k, const void*& p0, const void*& p1, const void*& p2) {
      p0 = dy + off0 + k * step;
      p1 = x + off0 + k * step;
      p2 = RELU == 2   ? (const void*)(yo + off0 + k * step)
           : RELU == 3 ? (const void*)(reinterpret_cast<const uint32_t*>(yo) + ((off0 + k * step) >> 3))
                       : nullptr;
    };
    Ring ring;
    ring.init(smem);
    ring.prime(rr.count, ptrs);  // in flight while the coefficients are fetched
    float mean[N], nmean[N], a[N], b[N];
    ldg_vec<N>(save_mean + c0, mean);
#pragma unroll
    for (int k = 0; k < N; ++k) nmean[k] = -mean[k];
    if (RELU == 1) {
      ldg_vec<N>(save_invstd + c0, a);
#pragma unroll
      for (int k = 0; k < N; ++k) b[k] = 0.f;
      if (weight) {
        float w[N];
        ldg_vec<N>(weight + c0, w);
#pragma unroll
        for (int k = 0; k < N; ++k) a[k] *= w[k];
      }
      if (bias) ldg_vec<N>(bias + c0, b);
    }
    auto proc = [&](int, const uint4& vg, const uint4& vx, const uint4& vy) {
      if constexpr (is_bf16<T>::value) {
        // s2 accumulates sum g*x (both bf16: one FHFMA, exact product); the mean is taken
        // out once at the end: sum g*(x-mean) = sum g*x - mean * sum g.
        uint4 g = vg;
        if (RELU == 2) {  // g *= (y > 0), two elements per instruction, exact
          g.x = bf16x2_mul(vg.x, bf16x2_gt0_mask(vy.x));
          g.y = bf16x2_mul(vg.y, bf16x2_gt0_mask(vy.y));
          g.z = bf16x2_mul(vg.z, bf16x2_gt0_mask(vy.z));
          g.w = bf16x2_mul(vg.w, bf16x2_gt0_mask(vy.w));
        }
        if (RELU == 3) g = apply_mask_word(vg, vy.x);
#pragma unroll
        for (int k = 0; k < N; ++k) {
          const uint16_t gh = vec_half(g, k), xh = vec_half(vx, k);
          if (RELU == 1) {
            const bool on = fmaf(add_bf16_f32(xh, nmean[k]), a[k], b[k]) > 0.f;
            s1[k] = on ? add_bf16_f32(gh, s1[k]) : s1[k];
            s2[k] = on ? fma_bf16_f32(gh, xh, s2[k]) : s2[k];
          } else {
            s1[k] = add_bf16_f32(gh, s1[k]);
            s2[k] = fma_bf16_f32(gh, xh, s2[k]);
          }
        }
      } else {
        float g[N], f[N], o[N];
        Vec<T>::unpack(vg, g);
        Vec<T>::unpack(vx, f);
        if (RELU == 2) Vec<T>::unpack(vy, o);
#pragma unroll
        for (int k = 0; k < N; ++k) {
          const float xm = f[k] + nmean[k];
          float gk = g[k];
          if (RELU == 1) gk = fmaf(xm, a[k], b[k]) > 0.f ? gk : 0.f;
          if (RELU == 2) gk = o[k] > 0.f ? gk : 0.f;
          s1[k] += gk;
          s2[k] = fmaf(gk, xm, s2[k]);
        }
      }
    };
    ring.run(rr.count, ptrs, proc);
    if constexpr (is_bf16<T>::value) {
#pragma unroll
      for (int k = 0; k < N; ++k) s2[k] = fmaf(nmean[k], s1[k], s2[k]);
    }
  }
  pdl_launch_dependents();
  __syncthreads();  // ring drained everywhere before the tree reuses the memory
  float dummy = 0.f;
  cta_tree<N, false>(sA, sB, nullptr, dummy, s1, s2);
  if (!grid_merge<N, false>(scratch, C, c0, chan_ok, sA, sB, nullptr, dummy, s1, s2, std::move(started))) return;

  if (r == 0 && chan_ok) {
#pragma unroll
    for (int k = 0; k < N; ++k) {
      sums_local[c0 + k] = s1[k];
      sums_local[C + c0 + k] = s2[k];
      if (grad_bias) grad_bias[c0 + k] = s1[k];
      if (grad_weight) grad_weight[c0 + k] = s2[k] * save_invstd[c0 + k];
    }
  }
  if (P.world > 1) {
    constexpr int cw = kCV * N;
    __syncthreads();
    if (r == 0) {
#pragma unroll
      for (int k = 0; k < N; ++k) {
        sA[cv * N + k] = s1[k];
        sA[cw + cv * N + k] = s2[k];
      }
      if (tid == 0) sA[2 * cw] = 0.f;
    }
    __syncthreads();
    push_chunk(P, seq, C, blockIdx.y * cw, cw, sA);
  }
}

// ============================== backward: apply ==============================
struct BwdArgs {
  const void* dy;
  const void* x;
  const void* yo;
  void* dx;
  void* dres;
  int64_t M;
  int C;
  const float* save_mean;
  const float* save_invstd;
  const float* weight;
  const float* bias;
  const float* sums_local;
  const float* save_count;
  int rev;
};

template <typename T, int RELU, bool DRES>
__global__ void __launch_bounds__(kT, kOccBwd) bn_bwd_apply_kernel(const __grid_constant__ BwdArgs A, const __grid_constant__ b200dp_peers P, uint32_t seq) {
  constexpr int N = Vec<T>::N;
  using Ring = RowRing<RELU >= 2 ? 3 : 4, RELU >= 2 ? 3 : 2, true, RELU == 3>;
  __shared__ __align__(16) unsigned char smem[Ring::kBytes];
  __shared__ __align__(16) float coef[5][kChunkMax];  // -mean, a, b, -mean_dy, -k
  const int tid = threadIdx.x, r = tid / kCV, cv = tid % kCV;
  const int C = A.C;
  const int c0 = blockIdx.y * (kCV * N) + cv * N;
  const bool chan_ok = c0 < C;

  const T* dyb = static_cast<const T*>(A.dy);
  const T* xb = static_cast<const T*>(A.x);
  const T* yob = static_cast<const T*>(A.yo);
  T* dxb = static_cast<T*>(A.dx);
  T* dresb = static_cast<T*>(A.dres);
  const RowRange rr = my_rows(A.M, r);
  const int64_t step = A.rev ? -rr.step * C : rr.step * C;
  const int64_t off0 = rr.first * C + c0 + (A.rev ? (int64_t)(rr.count - 1) * rr.step * C : 0);
  auto ptrs = [&](int k, const void*& p0, const void*& p1, const void*& p2) {
    p0 = dyb + off0 + k * step;
    p1 = xb + off0 + k * step;
    p2 = RELU == 2   ? (const void*)(yob + off0 + k * step)
         : RELU == 3 ? (const void*)(reinterpret_cast<const uint32_t*>(yob) + ((off0 + k * step) >> 3))
                     : nullptr;
  };
  Ring ring;
  ring.init(smem);
  const int nrows = chan_ok ? rr.count : 0;
  ring.prime(nrows, ptrs);  // in flight during the flag wait and the prologue
  pdl_wait();

  __shared__ __align__(16) float xr[B200DP_MAX_WORLD * kXrStride];
  constexpr int cw = kCV * N;
  if (P.world > 1) wait_and_gather(P, seq, C, blockIdx.y * cw, cw, xr);
  if (tid < kCV && chan_ok) {
    float s1[N], s2[N];
#pragma unroll
    for (int k = 0; k < N; ++k) s1[k] = s2[k] = 0.f;
    if (P.world > 1) {
      for (int p = 0; p < P.world; ++p) {
        const float* o = xr + p * kXrStride;
#pragma unroll
        for (int k = 0; k < N; ++k) {
          s1[k] += o[cv * N + k];
          s2[k] += o[cw + cv * N + k];
        }
      }
    } else {
#pragma unroll
      for (int k = 0; k < N; ++k) {
        s1[k] = ld_cg_f32(A.sums_local + c0 + k);
        s2[k] = ld_cg_f32(A.sums_local + C + c0 + k);
      }
    }
    const float cnt = ld_cg_f32(A.save_count);
    const float inv_cnt = cnt > 0.f ? 1.0f / cnt : 0.f;
    float nmean[N], a[N], b[N], nmdy[N], nkx[N];
#pragma unroll
    for (int k = 0; k < N; ++k) {
      const float invstd = A.save_invstd[c0 + k];
      nmean[k] = -A.save_mean[c0 + k];
      a[k] = invstd * (A.weight ? A.weight[c0 + k] : 1.f);
      b[k] = A.bias ? A.bias[c0 + k] : 0.f;
      nmdy[k] = -(s1[k] * inv_cnt);
      nkx[k] = -(s2[k] * inv_cnt * invstd * invstd);
    }
    sts_vec<N>(&coef[0][cv * N], nmean);
    sts_vec<N>(&coef[1][cv * N], a);
    sts_vec<N>(&coef[2][cv * N], b);
    sts_vec<N>(&coef[3][cv * N], nmdy);
    sts_vec<N>(&coef[4][cv * N], nkx);
  }
  __syncthreads();
  float nmean[N], a[N], b[N], nmdy[N], nkx[N];
  ldg_vec<N>(&coef[0][cv * N], nmean);
  ldg_vec<N>(&coef[1][cv * N], a);
  if (RELU == 1) ldg_vec<N>(&coef[2][cv * N], b);
  ldg_vec<N>(&coef[3][cv * N], nmdy);
  ldg_vec<N>(&coef[4][cv * N], nkx);

  // dx = (g - mean_dy - (x-mean)*k) * a   with k = invstd^2 * sum(g*(x-mean)) / count
  auto proc = [&](int k, const uint4& vg, const uint4& vx, const uint4& vy) {
    float f[N];
    if constexpr (is_bf16<T>::value) {
      uint4 g = vg;
      if (RELU == 2) {
        g.x = bf16x2_mul(vg.x, bf16x2_gt0_mask(vy.x));
        g.y = bf16x2_mul(vg.y, bf16x2_gt0_mask(vy.y));
        g.z = bf16x2_mul(vg.z, bf16x2_gt0_mask(vy.z));
        g.w = bf16x2_mul(vg.w, bf16x2_gt0_mask(vy.w));
      }
      if (RELU == 3) g = apply_mask_word(vg, vy.x);
      float gm[N];
#pragma unroll
      for (int j = 0; j < N; ++j) {
        const float xm = add_bf16_f32(vec_half(vx, j), nmean[j]);
        float t = add_bf16_f32(vec_half(g, j), nmdy[j]);
        if (RELU == 1) {
          const bool on = fmaf(xm, a[j], b[j]) > 0.f;
          t = on ? t : nmdy[j];
          if (DRES) gm[j] = on ? add_bf16_f32(vec_half(g, j), 0.f) : 0.f;
        }
        f[j] = fmaf(xm, nkx[j], t) * a[j];
      }
      *reinterpret_cast<uint4*>(dxb + off0 + k * step) = Vec<T>::pack(f);
      if (DRES) {
        if (RELU == 1) *reinterpret_cast<uint4*>(dresb + off0 + k * step) = Vec<T>::pack(gm);
        else *reinterpret_cast<uint4*>(dresb + off0 + k * step) = g;  // already masked, already bf16
      }
    } else {
      float g[N], o[N];
      Vec<T>::unpack(vg, g);
      Vec<T>::unpack(vx, f);
      if (RELU == 2) Vec<T>::unpack(vy, o);
#pragma unroll
      for (int j = 0; j < N; ++j) {
        const float xm = f[j] + nmean[j];
        float gk = g[j];
        if (RELU == 1) gk = fmaf(xm, a[j], b[j]) > 0.f ? gk : 0.f;
        if (RELU == 2) gk = o[j] > 0.f ? gk : 0.f;
        g[j] = gk;
        f[j] = fmaf(xm, nkx[j], gk + nmdy[j]) * a[j];
      }
      *reinterpret_cast<uint4*>(dxb + off0 + k * step) = Vec<T>::pack(f);
      if (DRES) *reinterpret_cast<uint4*>(dresb + off0 + k * step) = Vec<T>::pack(g);
    }
  };
  ring.run(nrows, ptrs, proc);
}

// ==================== one kernel per direction for layers that fit the L2 ====================
// The two-kernel path pays a second launch, a second prologue and a grid-wide hand-over (partials in
// global memory, a ticket, a last CTA) per layer and direction: ~15 us that do not shrink with the layer,
// and 40 of ResNet-50's 53 BN layers are small enough for that to be most of their time.  Here ONE
// thread-block cluster owns a channel chunk for ALL rows: it streams the rows once for the statistics
// (or the backward sums), folds them over its CTAs through distributed shared memory — hardware
// co-schedules the CTAs of a cluster, so that wait cannot starve whatever else runs on the GPU, which a
// grid-wide spin could — exchanges them across ranks, and streams the rows a second time to apply,
// this time out of the L2 they were just read into.
//
// Geometry.  A cluster has at most 16 CTAs and there is one cluster per chunk, so the GPU is filled with
// LARGE CTAs (1024 threads forward, 768 backward: one per SM, with the same 8-9 rows of 16 bytes in
// flight per thread as the two-kernel path, i.e. the same ~100 KB per SM outstanding) and chunks of 2, 4
// or 8 sixteen-byte vectors per row, chosen so that chunks x cluster size comes close to the SM count.
constexpr int kFusedMaxCS = 16;          // CTAs per cluster (sizes above 8 are an opt-in per kernel)
constexpr int kFusedMaxGroups = 128;     // cluster size x vectors per row chunk at most
constexpr int kFwdThreads = 1024;
constexpr int kBwdThreads = 768;

struct FusedGeom {
  int cvw;   // 16-byte vectors per row chunk (2, 4, 8)
  int cs;    // cluster size = row splits of the chunk
  int rev2;  // second pass walks the rows last -> first
};

// `nsrc` (<= 32) partials of channel group cv lie in shared memory at src[(l * cvw + cv) * W], l < nsrc.
// Warp cv (cv < cvw) folds them with a shuffle tree; lane 0 of that warp ends up with the result.
template <int N, bool WELFORD>
__device__ __forceinline__ void tree_fold(int cvw, int nsrc, const float* src, float& n, float* a, float* b) {
  constexpr int W = 2 * N + 1;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (w >= cvw) return;
  n = 0.f;
#pragma unroll
  for (int k = 0; k < N; ++k) a[k] = b[k] = 0.f;
  if (lane < nsrc) {
    const float* s = src + (lane * cvw + w) * W;
#pragma unroll
    for (int k = 0; k < N; ++k) {
      a[k] = s[k];
      b[k] = s[N + k];
    }
    n = s[2 * N];
  }
  warp_fold<N, WELFORD>(1, n, a, b);
}

// Rows of a CTA -> lane 0 of warps 0 .. cvw-1.  wp: shared memory, (blockDim.x / 32) * cvw * W floats.
template <int N, bool WELFORD>
__device__ __forceinline__ void cta_fold_big(int cvw, float* wp, float& n, float* a, float* b) {
  constexpr int W = 2 * N + 1;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  warp_fold<N, WELFORD>(cvw, n, a, b);
  if (lane < cvw) {
    float* d = wp + (w * cvw + lane) * W;
#pragma unroll
    for (int k = 0; k < N; ++k) {
      d[k] = a[k];
      d[N + k] = b[k];
    }
    d[2 * N] = n;
  }
  __syncthreads();
  tree_fold<N, WELFORD>(cvw, blockDim.x >> 5, wp, n, a, b);
}

// Every CTA of the cluster hands its partial (lane 0 of warps < cvw) to every CTA, itself included; after
// the cluster barrier each folds the cs partials with the same tree, so all hold bit-identical results
// (again in lane 0 of warps < cvw).  mine: smem [cvw * W]; xchg: smem [cs * cvw * W], written remotely.
template <int N, bool WELFORD>
__device__ __forceinline__ void cluster_fold(int cvw, int cs, float* mine, float* xchg, float& n, float* a,
                                             float* b, cg::cluster_group::arrival_token&& started) {
  constexpr int W = 2 * N + 1;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned rank = cluster.block_rank();
  if (lane == 0 && w < cvw) {
    float* d = mine + w * W;
#pragma unroll
    for (int k = 0; k < N; ++k) {
      d[k] = a[k];
      d[N + k] = b[k];
    }
    d[2 * N] = n;
  }
  __syncthreads();
  cluster.barrier_wait(std::move(started));  // every CTA of the cluster has started: its shared memory exists
  const int per = cvw * W;
  for (int it = tid; it < cs * per; it += blockDim.x) {
    const int q = it / per, e = it - q * per;
    float* dst = cluster.map_shared_rank(xchg + rank * per + e, q);
    *dst = mine[e];
  }
  cluster.sync();  // release / acquire at cluster scope
  tree_fold<N, WELFORD>(cvw, cs, xchg, n, a, b);
}

struct FusedRows {
  int64_t first, step;
  int count;
};
__device__ __forceinline__ FusedRows fused_rows(int64_t M, int r, int rp, int cs) {
  FusedRows rr;
  rr.step = (int64_t)cs * rp;
  rr.first = (int64_t)blockIdx.x * rp + r;
  rr.count = rr.first < M ? (int)((M - rr.first + rr.step - 1) / rr.step) : 0;
  return rr;
}

extern __shared__ __align__(16) unsigned char fused_dyn_smem[];  // the rings of the one-kernel path

template <typename T, bool RELU, bool RES>
__global__ void __launch_bounds__(kFwdThreads, 1)
bn_fwd_fused_kernel(const __grid_constant__ FwdArgs A, const FusedGeom G, const __grid_constant__ b200dp_peers P,
                    uint32_t seq) {
  constexpr int N = Vec<T>::N;
  constexpr int W = 2 * N + 1;
  constexpr int NT = kFwdThreads;
  using Ring1 = RowRing<8, 1, false, false, NT>;
  using Ring2 = RowRing<RES ? 4 : 8, RES ? 2 : 1, true, false, NT>;
  static_assert(Ring2::kBytes <= Ring1::kBytes, "ring");
  unsigned char* smem = fused_dyn_smem;  // Ring1::kBytes: the rings, and in between the warp partials of the fold
  __shared__ __align__(16) float xchg[kFusedMaxGroups * W];
  __shared__ __align__(16) float mine[2 * kChunkMax + 8];  // partial for the cluster, then the exchange stage
  __shared__ __align__(16) float coef[3][kChunkMax];
  __shared__ __align__(16) float xr[B200DP_MAX_WORLD * kXrStride];
  const int tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
  const int cvw = G.cvw, cs = G.cs, rp = NT / cvw;
  const int cv = tid % cvw, r = tid / cvw;
  const int C = A.C;
  const int cw = cvw * N, cb = blockIdx.y * cw;
  const int c0 = cb + cv * N;
  const bool chan_ok = c0 < C;
  auto started = cg::this_cluster().barrier_arrive();

  const T* xb = static_cast<const T*>(A.x);
  const T* rb = static_cast<const T*>(A.res);
  T* yb = static_cast<T*>(A.y);
  const FusedRows rr = fused_rows(A.M, r, rp, cs);
  const int nrows = chan_ok ? rr.count : 0;
  const int64_t rstep = rr.step * C;
  const int64_t off_first = rr.first * C + c0;

  // ---- pass 1: statistics of this CTA's rows ----
  float n = 0.f, mean[N], m2[N];
#pragma unroll
  for (int k = 0; k < N; ++k) mean[k] = m2[k] = 0.f;
  pdl_wait();
  if (A.batches_tracked && blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) *A.batches_tracked += 1;
  {
    auto ptrs = [&](int k, const void*& p0, const void*& p1, const void*& p2) {
      p0 = xb + off_first + k * rstep;
      p1 = p2 = nullptr;
    };
    auto proc = [&](int, const uint4& v, const uint4&, const uint4&) {
      if constexpr (is_bf16<T>::value) {
#pragma unroll
        for (int k = 0; k < N; ++k) {
          const uint16_t h = vec_half(v, k);
          mean[k] = add_bf16_f32(h, mean[k]);
          m2[k] = fma_bf16_f32(h, h, m2[k]);
        }
      } else {
        float f[N];
        Vec<T>::unpack(v, f);
        n += 1.f;
        const float rn = __frcp_rn(n);
#pragma unroll
        for (int k = 0; k < N; ++k) {
          const float d = f[k] - mean[k];
          mean[k] = fmaf(d, rn, mean[k]);
          m2[k] = fmaf(d, f[k] - mean[k], m2[k]);
        }
      }
    };
    Ring1 ring;
    ring.init(smem);
    ring.prime(nrows, ptrs);
    ring.run(nrows, ptrs, proc);
    if constexpr (is_bf16<T>::value) {
      n = (float)nrows;
      if (nrows > 0) {
        const float rn = 1.0f / n;
#pragma unroll
        for (int k = 0; k < N; ++k) {
          const float s = mean[k];
          mean[k] = s * rn;
          m2[k] = fmaxf(fmaf(-s, mean[k], m2[k]), 0.f);
        }
      }
    }
  }
  __syncthreads();  // rings drained: the fold may use the memory
  cta_fold_big<N, true>(cvw, reinterpret_cast<float*>(smem), n, mean, m2);
  cluster_fold<N, true>(cvw, cs, mine, xchg, n, mean, m2, std::move(started));
  // lane 0 of warps < cvw of every CTA now holds this rank's (n, mean, M2) of channel group `wrp`
  const bool holder = lane == 0 && wrp < cvw;
  const int hc0 = cb + wrp * N;  // the holder's first channel

  // second pass: start the rows flying before the coefficients are worked out
  const bool rev = G.rev2 != 0;
  const int64_t off2 = off_first + (rev ? (int64_t)(rr.count - 1) * rstep : 0);
  const int64_t step2 = rev ? -rstep : rstep;
  auto ptrs2 = [&](int k, const void*& p0, const void*& p1, const void*& p2) {
    p0 = xb + off2 + k * step2;
    p1 = RES ? (const void*)(rb + off2 + k * step2) : nullptr;
    p2 = nullptr;
  };
  __syncthreads();  // the fold is done with the ring memory
  Ring2 ring2;
  ring2.init(smem);
  ring2.prime(nrows, ptrs2);

  const bool leader = blockIdx.x == 0;
  if (leader && holder && hc0 < C) {
    float* so = const_cast<float*>(A.stats_local);  // an output here (the apply kernel only reads it)
#pragma unroll
    for (int k = 0; k < N; ++k) {
      so[hc0 + k] = mean[k];
      so[C + hc0 + k] = m2[k];
    }
    if (wrp == 0 && blockIdx.y == 0) so[2 * C] = n;
  }
  if (P.world > 1) {
    if (leader) {
      if (holder) {
#pragma unroll
        for (int k = 0; k < N; ++k) {
          mine[wrp * N + k] = mean[k];
          mine[cw + wrp * N + k] = m2[k];
        }
        if (wrp == 0) mine[2 * cw] = n;
      }
      __syncthreads();
      push_chunk(P, seq, C, cb, cw, mine);
    }
    wait_and_gather(P, seq, C, cb, cw, xr);
  }
  if (holder && hc0 < C) {
    if (P.world > 1) {
      n = 0.f;
#pragma unroll
      for (int k = 0; k < N; ++k) mean[k] = m2[k] = 0.f;
      for (int p = 0; p < P.world; ++p) {
        const float* o = xr + p * kXrStride;
        welford_merge<N>(n, mean, m2, o[2 * cw], o + wrp * N, o + cw + wrp * N);
      }
    }
    const float inv_n = n > 0.f ? 1.0f / n : 0.f;
    const float inv_nm1 = n > 1.f ? 1.0f / (n - 1.f) : 0.f;
    float a[N], b[N];
#pragma unroll
    for (int k = 0; k < N; ++k) {
      const float var = m2[k] * inv_n;
      const float invstd = 1.0f / sqrtf(var + A.eps);
      a[k] = invstd * (A.weight ? A.weight[hc0 + k] : 1.f);
      b[k] = A.bias ? A.bias[hc0 + k] : 0.f;
      if (leader) {
        A.save_mean[hc0 + k] = mean[k];
        A.save_invstd[hc0 + k] = invstd;
        if (A.running_mean)
          A.running_mean[hc0 + k] = (1.f - A.momentum) * A.running_mean[hc0 + k] + A.momentum * mean[k];
        if (A.running_var)
          A.running_var[hc0 + k] = (1.f - A.momentum) * A.running_var[hc0 + k] + A.momentum * (m2[k] * inv_nm1);
      }
      mean[k] = -mean[k];
    }
    if (leader && wrp == 0 && blockIdx.y == 0) A.save_count[0] = n;
    sts_vec<N>(&coef[0][wrp * N], mean);
    sts_vec<N>(&coef[1][wrp * N], a);
    sts_vec<N>(&coef[2][wrp * N], b);
  }
  __syncthreads();
  float nmean[N], a[N], b[N];
  ldg_vec<N>(&coef[0][cv * N], nmean);
  ldg_vec<N>(&coef[1][cv * N], a);
  ldg_vec<N>(&coef[2][cv * N], b);

  // ---- pass 2: normalise (+ residual) (+ ReLU) ----
  auto proc2 = [&](int k, const uint4& vx, const uint4& vr, const uint4&) {
    float f[N];
    if constexpr (is_bf16<T>::value) {
#pragma unroll
      for (int j = 0; j < N; ++j) {
        float o = fmaf(add_bf16_f32(vec_half(vx, j), nmean[j]), a[j], b[j]);
        if (RES) o = add_bf16_f32(vec_half(vr, j), o);
        f[j] = RELU ? fmaxf(o, 0.f) : o;
      }
    } else {
      float g[N];
      Vec<T>::unpack(vx, f);
      if (RES) Vec<T>::unpack(vr, g);
#pragma unroll
      for (int j = 0; j < N; ++j) {
        float o = fmaf(f[j] + nmean[j], a[j], b[j]);
        if (RES) o += g[j];
        f[j] = RELU ? fmaxf(o, 0.f) : o;
      }
    }
    const uint4 pk = Vec<T>::pack(f);
    *reinterpret_cast<uint4*>(yb + off2 + k * step2) = pk;
    if constexpr (RELU && is_bf16<T>::value) {
      if (A.maskw) {
        const uint32_t w[4] = {pk.x, pk.y, pk.z, pk.w};
        uint32_t bits = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          bits |= ((w[j] & 0xffffu) != 0u ? 1u : 0u) << (2 * j);
          bits |= ((w[j] >> 16) != 0u ? 1u : 0u) << (2 * j + 1);
        }
        A.maskw[(off2 + k * step2) >> 3] = bits;
      }
    }
  };
  ring2.run(nrows, ptrs2, proc2);
}

// Backward: sums over the rows, cluster fold, cross-rank exchange, then dx (and d_residual).
struct BwdFusedArgs {
  BwdArgs b;
  float* sums_local;   // out: [2C] this rank's sums (the two-kernel path's contract)
  float* grad_weight;  // out or NULL
  float* grad_bias;    // out or NULL
};

template <typename T, int RELU, bool DRES>
__global__ void __launch_bounds__(kBwdThreads, 1)
bn_bwd_fused_kernel(const __grid_constant__ BwdFusedArgs F, const FusedGeom G,
                    const __grid_constant__ b200dp_peers P, uint32_t seq) {
  constexpr int N = Vec<T>::N;
  constexpr int W = 2 * N + 1;
  constexpr int NT = kBwdThreads;
  using Ring1 = RowRing<RELU >= 2 ? 3 : 4, RELU >= 2 ? 3 : 2, false, RELU == 3, NT>;
  using Ring2 = RowRing<RELU >= 2 ? 3 : 4, RELU >= 2 ? 3 : 2, true, RELU == 3, NT>;
  unsigned char* smem = fused_dyn_smem;  // Ring1::kBytes
  __shared__ __align__(16) float xchg[kFusedMaxGroups * W];
  __shared__ __align__(16) float mine[2 * kChunkMax + 8];
  __shared__ __align__(16) float coef[2][kChunkMax];
  __shared__ __align__(16) float xr[B200DP_MAX_WORLD * kXrStride];
  const BwdArgs& A = F.b;
  const int tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
  const int cvw = G.cvw, cs = G.cs, rp = NT / cvw;
  const int cv = tid % cvw, r = tid / cvw;
  const int C = A.C;
  const int cw = cvw * N, cb = blockIdx.y * cw;
  const int c0 = cb + cv * N;
  const bool chan_ok = c0 < C;
  auto started = cg::this_cluster().barrier_arrive();

  const T* dyb = static_cast<const T*>(A.dy);
  const T* xb = static_cast<const T*>(A.x);
  const T* yob = static_cast<const T*>(A.yo);
  T* dxb = static_cast<T*>(A.dx);
  T* dresb = static_cast<T*>(A.dres);
  const FusedRows rr = fused_rows(A.M, r, rp, cs);
  const int nrows = chan_ok ? rr.count : 0;
  const int64_t rstep = rr.step * C;
  const int64_t off_first = rr.first * C + c0;

  float s1[N], s2[N];
#pragma unroll
  for (int k = 0; k < N; ++k) s1[k] = s2[k] = 0.f;
  float nmean[N], a[N], b[N];
#pragma unroll
  for (int k = 0; k < N; ++k) { nmean[k] = 0.f; a[k] = 1.f; b[k] = 0.f; }
  pdl_wait();
  {
    auto ptrs = [&](int k, const void*& p0, const void*& p1, const void*& p2) {
      p0 = dyb + off_first + k * rstep;
      p1 = xb + off_first + k * rstep;
      p2 = RELU == 2   ? (const void*)(yob + off_first + k * rstep)
           : RELU == 3 ? (const void*)(reinterpret_cast<const uint32_t*>(yob) + ((off_first + k * rstep) >> 3))
                       : nullptr;
    };
    Ring1 ring;
    ring.init(smem);
    ring.prime(nrows, ptrs);
    if (chan_ok) {
      float mean[N];
      ldg_vec<N>(A.save_mean + c0, mean);
#pragma unroll
      for (int k = 0; k < N; ++k) nmean[k] = -mean[k];
      ldg_vec<N>(A.save_invstd + c0, a);
      if (A.weight) {
        float w[N];
        ldg_vec<N>(A.weight + c0, w);
#pragma unroll
        for (int k = 0; k < N; ++k) a[k] *= w[k];
      }
      if (A.bias) ldg_vec<N>(A.bias + c0, b);
    }
    auto proc = [&](int, const uint4& vg, const uint4& vx, const uint4& vy) {
      if constexpr (is_bf16<T>::value) {
        uint4 g = vg;
        if (RELU == 2) {
          g.x = bf16x2_mul(vg.x, bf16x2_gt0_mask(vy.x));
          g.y = bf16x2_mul(vg.y, bf16x2_gt0_mask(vy.y));
          g.z = bf16x2_mul(vg.z, bf16x2_gt0_mask(vy.z));
          g.w = bf16x2_mul(vg.w, bf16x2_gt0_mask(vy.w));
        }
        if (RELU == 3) g = apply_mask_word(vg, vy.x);
#pragma unroll
        for (int k = 0; k < N; ++k) {
          const uint16_t gh = vec_half(g, k), xh = vec_half(vx, k);
          if (RELU == 1) {
            const bool on = fmaf(add_bf16_f32(xh, nmean[k]), a[k], b[k]) > 0.f;
            s1[k] = on ? add_bf16_f32(gh, s1[k]) : s1[k];
            s2[k] = on ? fma_bf16_f32(gh, xh, s2[k]) : s2[k];
          } else {
            s1[k] = add_bf16_f32(gh, s1[k]);
            s2[k] = fma_bf16_f32(gh, xh, s2[k]);
          }
        }
      } else {
        float g[N], f[N], o[N];
        Vec<T>::unpack(vg, g);
        Vec<T>::unpack(vx, f);
        if (RELU == 2) Vec<T>::unpack(vy, o);
#pragma unroll
        for (int k = 0; k < N; ++k) {
          const float xm = f[k] + nmean[k];
          float gk = g[k];
          if (RELU == 1) gk = fmaf(xm, a[k], b[k]) > 0.f ? gk : 0.f;
          if (RELU == 2) gk = o[k] > 0.f ? gk : 0.f;
          s1[k] += gk;
          s2[k] = fmaf(gk, xm, s2[k]);
        }
      }
    };
    ring.run(nrows, ptrs, proc);
    if constexpr (is_bf16<T>::value) {
#pragma unroll
      for (int k = 0; k < N; ++k) s2[k] = fmaf(nmean[k], s1[k], s2[k]);
    }
  }
  __syncthreads();
  float dummy = 0.f;
  cta_fold_big<N, false>(cvw, reinterpret_cast<float*>(smem), dummy, s1, s2);
  cluster_fold<N, false>(cvw, cs, mine, xchg, dummy, s1, s2, std::move(started));
  const bool holder = lane == 0 && wrp < cvw;
  const int hc0 = cb + wrp * N;

  const bool rev = G.rev2 != 0;
  const int64_t off2 = off_first + (rev ? (int64_t)(rr.count - 1) * rstep : 0);
  const int64_t step2 = rev ? -rstep : rstep;
  auto ptrs2 = [&](int k, const void*& p0, const void*& p1, const void*& p2) {
    p0 = dyb + off2 + k * step2;
    p1 = xb + off2 + k * step2;
    p2 = RELU == 2   ? (const void*)(yob + off2 + k * step2)
         : RELU == 3 ? (const void*)(reinterpret_cast<const uint32_t*>(yob) + ((off2 + k * step2) >> 3))
                     : nullptr;
  };
  __syncthreads();
  Ring2 ring2;
  ring2.init(smem);
  ring2.prime(nrows, ptrs2);

  const bool leader = blockIdx.x == 0;
  if (leader && holder && hc0 < C) {
#pragma unroll
    for (int k = 0; k < N; ++k) {
      F.sums_local[hc0 + k] = s1[k];
      F.sums_local[C + hc0 + k] = s2[k];
      if (F.grad_bias) F.grad_bias[hc0 + k] = s1[k];
      if (F.grad_weight) F.grad_weight[hc0 + k] = s2[k] * A.save_invstd[hc0 + k];
    }
  }
  if (P.world > 1) {
    if (leader) {
      if (holder) {
#pragma unroll
        for (int k = 0; k < N; ++k) {
          mine[wrp * N + k] = s1[k];
          mine[cw + wrp * N + k] = s2[k];
        }
        if (wrp == 0) mine[2 * cw] = 0.f;
      }
      __syncthreads();
      push_chunk(P, seq, C, cb, cw, mine);
    }
    wait_and_gather(P, seq, C, cb, cw, xr);
  }
  if (holder && hc0 < C) {
    if (P.world > 1) {
#pragma unroll
      for (int k = 0; k < N; ++k) s1[k] = s2[k] = 0.f;
      for (int p = 0; p < P.world; ++p) {
        const float* o = xr + p * kXrStride;
#pragma unroll
        for (int k = 0; k < N; ++k) {
          s1[k] += o[wrp * N + k];
          s2[k] += o[cw + wrp * N + k];
        }
      }
    }
    const float cnt = ld_cg_f32(A.save_count);
    const float inv_cnt = cnt > 0.f ? 1.0f / cnt : 0.f;
    float nmdy[N], nkx[N];
#pragma unroll
    for (int k = 0; k < N; ++k) {
      const float invstd = A.save_invstd[hc0 + k];
      nmdy[k] = -(s1[k] * inv_cnt);
      nkx[k] = -(s2[k] * inv_cnt * invstd * invstd);
    }
    sts_vec<N>(&coef[0][wrp * N], nmdy);
    sts_vec<N>(&coef[1][wrp * N], nkx);
  }
  __syncthreads();
  float nmdy[N], nkx[N];
  ldg_vec<N>(&coef[0][cv * N], nmdy);
  ldg_vec<N>(&coef[1][cv * N], nkx);

  auto proc2 = [&](int k, const uint4& vg, const uint4& vx, const uint4& vy) {
    float f[N];
    if constexpr (is_bf16<T>::value) {
      uint4 g = vg;
      if (RELU == 2) {
        g.x = bf16x2_mul(vg.x, bf16x2_gt0_mask(vy.x));
        g.y = bf16x2_mul(vg.y, bf16x2_gt0_mask(vy.y));
        g.z = bf16x2_mul(vg.z, bf16x2_gt0_mask(vy.z));
        g.w = bf16x2_mul(vg.w, bf16x2_gt0_mask(vy.w));
      }
      if (RELU == 3) g = apply_mask_word(vg, vy.x);
      float gm[N];
#pragma unroll
      for (int j = 0; j < N; ++j) {
        const float xm = add_bf16_f32(vec_half(vx, j), nmean[j]);
        float t = add_bf16_f32(vec_half(g, j), nmdy[j]);
        if (RELU == 1) {
          const bool on = fmaf(xm, a[j], b[j]) > 0.f;
          t = on ? t : nmdy[j];
          if (DRES) gm[j] = on ? add_bf16_f32(vec_half(g, j), 0.f) : 0.f;
        }
        f[j] = fmaf(xm, nkx[j], t) * a[j];
      }
      *reinterpret_cast<uint4*>(dxb + off2 + k * step2) = Vec<T>::pack(f);
      if (DRES) {
        if (RELU == 1) *reinterpret_cast<uint4*>(dresb + off2 + k * step2) = Vec<T>::pack(gm);
        else *reinterpret_cast<uint4*>(dresb + off2 + k * step2) = g;
      }
    } else {
      float g[N], o[N];
      Vec<T>::unpack(vg, g);
      Vec<T>::unpack(vx, f);
      if (RELU == 2) Vec<T>::unpack(vy, o);
#pragma unroll
      for (int j = 0; j < N; ++j) {
        const float xm = f[j] + nmean[j];
        float gk = g[j];
        if (RELU == 1) gk = fmaf(xm, a[j], b[j]) > 0.f ? gk : 0.f;
        if (RELU == 2) gk = o[j] > 0.f ? gk : 0.f;
        g[j] = gk;
        f[j] = fmaf(xm, nkx[j], gk + nmdy[j]) * a[j];
      }
      *reinterpret_cast<uint4*>(dxb + off2 + k * step2) = Vec<T>::pack(f);
      if (DRES) *reinterpret_cast<uint4*>(dresb + off2 + k * step2) = Vec<T>::pack(g);
    }
  };
  ring2.run(nrows, ptrs2, proc2);
}

// Row traversal order of the four training kernels: bit 0 statistics, bit 1 forward apply,
// bit 2 backward reduce, bit 3 backward apply; a set bit walks the rows from the last to the
// first.  A kernel that runs against the direction of the one before it starts on the rows
// that kernel touched last, which are the ones still in L2 (126 MB of it on B200).
// B200DP_BN_ORDER overrides the default (read once).
int bn_order() {
  static const int order = [] {
    const char* e = getenv("B200DP_BN_ORDER");
    return e && e[0] ? atoi(e) : kDefaultBnOrder;
  }();
  return order;
}

// The reduction kernels are launched as programmatic dependents of whatever precedes them in the
// stream (a convolution, usually): when that kernel signals early, their launch latency and set-up
// overlap its tail; they block in griddepcontrol.wait before the first load.  B200DP_BN_PDL_FIRST=0
// turns it off.
bool pdl_first() {
  static const bool on = [] { const char* e = getenv("B200DP_BN_PDL_FIRST"); return !(e && e[0] == '0'); }();
  return on;
}

// Launch with (or without) the programmatic-stream-serialization attribute.
thread_local bool t_pdl_next = false;  // set by b200dp_bn_forward / _backward around their second launch

template <typename... KArgs, typename... Args>
cudaError_t launch_ex(void (*kernel)(KArgs...), dim3 grid, cudaStream_t s, bool pdl, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(kT, 1, 1);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// ---- which layers take the one-kernel path, and with what geometry ----
// B200DP_BN_FUSED=0 turns it off; B200DP_BN_FUSED_MB / B200DP_BN_FUSED_BWD_MB bound the tensor size
// (MiB of x; default: what stays in the 126 MB L2 next to the tensors written meanwhile);
// B200DP_BN_FUSED_CS=8 keeps to the portable cluster size.
struct FusedCfg {
  bool on;
  int64_t fwd_bytes, bwd_bytes;
  int cs;
  bool any_geometry;  // take every shape under the size limits, whatever geometry fits (tests, measurements)
};
FusedCfg& fused_cfg() {  // read from the environment once; b200dp_bn_set_one_kernel_limits() may change it later
  static FusedCfg cfg = [] {
    FusedCfg c;
    const char* e = getenv("B200DP_BN_FUSED");
    c.on = !(e && e[0] == '0');
    e = getenv("B200DP_BN_FUSED_MB");
    c.fwd_bytes = (int64_t)(e && e[0] ? atoi(e) : 56) << 20;
    e = getenv("B200DP_BN_FUSED_BWD_MB");
    c.bwd_bytes = (int64_t)(e && e[0] ? atoi(e) : 56) << 20;
    e = getenv("B200DP_BN_FUSED_CS");
    c.cs = (e && atoi(e) >= 1 && atoi(e) <= kFusedMaxCS) ? atoi(e) : kFusedMaxCS;
    e = getenv("B200DP_BN_FUSED_ANY");
    c.any_geometry = e && e[0] == '1';
    return c;
  }();
  return cfg;
}

struct FusedPlan {
  bool ok;
  FusedGeom g;
  int chunks;
};

// Geometry of the one-kernel path, and whether it is worth taking.  Measured on B200 (profiles/
// r2_bn_warm8_*.txt, back-to-back launches): it beats the two kernels only where a chunk is a whole
// 128-byte line (narrower chunks cost one L1 wavefront per 32 or 64 bytes and lose more than the second
// launch costs), the clusters fill most of the SMs, and the clusters are small or the tensor is tiny
// (16-CTA clusters pay more at the cluster barrier than a ~5 us launch).  That is: C >= 1024 bf16
// channels (cluster size <= 4 ... 8) or tensors up to 16 MiB.  With `any_geometry` (tests, measurements)
// every shape under the size limit goes through it, with whatever geometry fits.
template <typename T, typename KernelT>
FusedPlan plan_fused(KernelT kernel, int threads, int dyn_smem, int64_t M, int C, int64_t limit_bytes) {
  FusedPlan pl;
  pl.ok = false;
  const FusedCfg& cfg = fused_cfg();
  const int64_t bytes = M * C * (int64_t)sizeof(T);
  if (!cfg.on || limit_bytes <= 0 || bytes > limit_bytes) return pl;
  const bool forced = cfg.any_geometry;
  const int nvec = C / Vec<T>::N;
  int best_ctas = 0;
  for (int cvw = 8; cvw >= 2; cvw >>= 1) {
    const int chunks = (nvec + cvw - 1) / cvw;
    int cs = 0;
    for (int c = kFusedMaxCS; c >= 1; c >>= 1) {
      if (c > cfg.cs) continue;
      if (max_active_clusters(kernel, c, threads, dyn_smem) >= chunks) { cs = c; break; }
    }
    if (cs == 0) continue;  // more chunks than SMs even with single-CTA clusters: try a wider chunk (none: two kernels)
    const int ctas = chunks * cs;
    if (ctas > best_ctas) {
      best_ctas = ctas;
      pl.g.cvw = cvw;
      pl.g.cs = cs;
      pl.chunks = chunks;
    }
    if (ctas >= 96 || !forced) break;  // unforced: whole lines (cvw = 8) or nothing
  }
  if (best_ctas == 0) return pl;
  if (!forced && (best_ctas < 96 || (pl.g.cs > 4 && bytes > ((int64_t)16 << 20)))) return pl;
  pl.g.rev2 = 1;
  pl.ok = true;
  return pl;
}

template <typename... KArgs, typename... Args>
cudaError_t launch_cluster(void (*kernel)(KArgs...), int cs, int chunks, int threads, int dyn_smem, cudaStream_t s,
                           bool pdl, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(cs, chunks, 1);
  cfg.blockDim = dim3(threads, 1, 1);
  cfg.dynamicSmemBytes = dyn_smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = cs;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 2 : 1;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

constexpr int kNotFused = -1000;  // internal: the layer takes the two-kernel path

template <typename T>
int launch_fwd_fused(const FwdArgs& A, bool relu, const b200dp_peers& P, uint32_t seq, cudaStream_t s) {
  const bool res = A.res != nullptr;
  cudaError_t e;
#define B200DP_FWD_FUSED(RELU, RES)                                                              \
  do {                                                                                           \
    constexpr int smem = 8 * kFwdThreads * 16; /* the deeper of the two rings */                 \
    const FusedPlan pl = plan_fused<T>(bn_fwd_fused_kernel<T, RELU, RES>, kFwdThreads, smem, A.M, A.C, fused_cfg().fwd_bytes); \
    if (!pl.ok) return kNotFused;                                                                \
    e = launch_cluster(bn_fwd_fused_kernel<T, RELU, RES>, pl.g.cs, pl.chunks, kFwdThreads, smem, s, pdl_first(), A, pl.g, P, seq); \
  } while (0)
  if (relu && res) B200DP_FWD_FUSED(true, true);
  else if (relu) B200DP_FWD_FUSED(true, false);
  else if (res) B200DP_FWD_FUSED(false, true);
  else B200DP_FWD_FUSED(false, false);
#undef B200DP_FWD_FUSED
  if (e != cudaSuccess) return (int)e;
  B200DP_LAUNCH_CHECK();
  return B200DP_OK;
}

template <typename T>
int launch_bwd_fused(const BwdFusedArgs& F, int relu, const b200dp_peers& P, uint32_t seq, cudaStream_t s) {
  const bool dres = F.b.dres != nullptr;
  cudaError_t e = cudaSuccess;
#define B200DP_BWD_FUSED(R, D)                                                                   \
  do {                                                                                           \
    constexpr int smem = (R >= 2 ? 9 : 8) * kBwdThreads * 16;                                    \
    const FusedPlan pl = plan_fused<T>(bn_bwd_fused_kernel<T, R, D>, kBwdThreads, smem, F.b.M, F.b.C, fused_cfg().bwd_bytes); \
    if (!pl.ok) return kNotFused;                                                                \
    e = launch_cluster(bn_bwd_fused_kernel<T, R, D>, pl.g.cs, pl.chunks, kBwdThreads, smem, s, pdl_first(), F, pl.g, P, seq); \
  } while (0)
#define B200DP_BWD_FUSED_R(R)            \
  do {                                   \
    if (dres) B200DP_BWD_FUSED(R, true); \
    else B200DP_BWD_FUSED(R, false);     \
  } while (0)
  if (relu == 0) B200DP_BWD_FUSED_R(0);
  else if (relu == 1) B200DP_BWD_FUSED_R(1);
  else if (relu == 2) B200DP_BWD_FUSED_R(2);
  else {
    if constexpr (is_bf16<T>::value) B200DP_BWD_FUSED_R(3);
    else return B200DP_EUNSUPPORTED;
  }
#undef B200DP_BWD_FUSED_R
#undef B200DP_BWD_FUSED
  if (e != cudaSuccess) return (int)e;
  B200DP_LAUNCH_CHECK();
  return B200DP_OK;
}

// ================================ host side ==================================
int check_common(const void* p0, const void* p1, const void* p2, int dtype, int64_t M, int C) {
  if (M < 0 || C <= 0) return B200DP_EINVAL;
  if (dtype != B200DP_F32 && dtype != B200DP_BF16) return B200DP_EUNSUPPORTED;
  const int vec = dtype == B200DP_BF16 ? 8 : 4;
  if (C % vec != 0 || C > B200DP_BN_MAX_CHANNELS) return B200DP_EUNSUPPORTED;
  if (M > 0 && !p0) return B200DP_EINVAL;
  if (!aligned16(p0) || !aligned16(p1) || !aligned16(p2)) return B200DP_EALIGN;
  return B200DP_OK;
}

int check_comm(const b200dp_comm* comm, uint32_t seq) {
  if (comm && comm->world > 1) {
    if (!comm->peers_open) return B200DP_ESTATE;
    if (seq == 0) return B200DP_EINVAL;
  }
  return B200DP_OK;
}

}  // namespace

extern "C" size_t b200dp_bn_scratch_bytes(int C) {
  (void)C;
  return kScratchBytes;
}

extern "C" int b200dp_bn_fwd_stats(const void* x, int dtype, int64_t M, int C, void* scratch,
                                   float* stats_local, b200dp_comm* comm, uint32_t seq,
                                   b200dp_stream_t stream) {
  int rc = check_common(x, nullptr, nullptr, dtype, M, C);
  if (rc) return rc;
  if (!scratch || !stats_local) return B200DP_EINVAL;
  if ((rc = check_comm(comm, seq))) return rc;
  const b200dp_peers P = b200dp_make_peers(comm);
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == B200DP_BF16) {
    const Grid2 g = plan_grid_clustered<__nv_bfloat16>(M, C, max_active_clusters(bn_stats_kernel<__nv_bfloat16>));
    B200DP_CUDA_TRY(launch_ex(bn_stats_kernel<__nv_bfloat16>, dim3(g.gx, g.gy), s, pdl_first(),
                              static_cast<const __nv_bfloat16*>(x), M, C, scratch, stats_local, P, seq, bn_order() & 1));
  } else {
    const Grid2 g = plan_grid_clustered<float>(M, C, max_active_clusters(bn_stats_kernel<float>));
    B200DP_CUDA_TRY(launch_ex(bn_stats_kernel<float>, dim3(g.gx, g.gy), s, pdl_first(), static_cast<const float*>(x), M, C,
                              scratch, stats_local, P, seq, bn_order() & 1));
  }
  B200DP_LAUNCH_CHECK();
  return B200DP_OK;
}

// ---- statistics without any allocation ---------------------------------------
// Some hosts may not let a library allocate device memory: a serving process or
// another framework's arena that has handed out every byte of HBM to its own pool
// and passes us nothing but the input and output buffers, or a launch path that must
// never block (cudaMalloc takes a device-wide lock and can synchronise the GPU).
// For them the reduction scratch of the statistics kernel lives in this fixed
// array, part of the library image: every GPU that loads libb200dp.so holds these
// kBnStaticScratchBytes bytes until the process exits.
namespace {
constexpr size_t kBnStaticScratchBytes = kScratchBytes;  // 307,712 bytes
static __device__ __align__(256) unsigned char g_bn_static_scratch[kBnStaticScratchBytes];

constexpr int kMaxDevices = 64;
void* g_bn_static_scratch_addr[kMaxDevices];  // device address of the array, per device ordinal
std::mutex g_bn_static_mutex;

// In host code `g_bn_static_scratch` is only a handle for the symbol, not an address a
// kernel may dereference: ask the runtime for the device address, once per device, with
// that device current.
int bn_static_scratch(void** out) {
  int dev = -1;
  B200DP_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return B200DP_EUNSUPPORTED;
  std::lock_guard<std::mutex> lock(g_bn_static_mutex);
  if (!g_bn_static_scratch_addr[dev]) {
    void* p = nullptr;
    B200DP_CUDA_TRY(cudaGetSymbolAddress(&p, g_bn_static_scratch));
    if (!p) return B200DP_ESTATE;
    g_bn_static_scratch_addr[dev] = p;
  }
  *out = g_bn_static_scratch_addr[dev];
  return B200DP_OK;
}
}  // namespace

extern "C" int b200dp_bn_fwd_stats_noalloc(const void* x, int dtype, int64_t M, int C,
                                           float* stats_local, b200dp_stream_t stream) {
  if (C > B200DP_BN_NOALLOC_MAX_CHANNELS) return B200DP_ETOOBIG;  // the fixed scratch is sized for this
  int rc = check_common(x, nullptr, nullptr, dtype, M, C);
  if (rc) return rc;
  if (!stats_local) return B200DP_EINVAL;
  void* scratch = nullptr;
  if ((rc = bn_static_scratch(&scratch))) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  const b200dp_peers P = b200dp_make_peers(nullptr);
  Grid2 g;
  if (dtype == B200DP_BF16) {
    g = plan_grid_clustered<__nv_bfloat16>(M, C, max_active_clusters(bn_stats_kernel<__nv_bfloat16>));
    if ((size_t)g.gx * g.gy * kCV * Vec<__nv_bfloat16>::N > (size_t)kMaxPartials) return B200DP_ETOOBIG;
  } else {
    g = plan_grid_clustered<float>(M, C, max_active_clusters(bn_stats_kernel<float>));
    if ((size_t)g.gx * g.gy * kCV * Vec<float>::N > (size_t)kMaxPartials) return B200DP_ETOOBIG;
  }
  // The array is shared by every call and keeps what the last one left in it.  Only the
  // "last CTA" tickets have to start at zero; partials and counts are written before they
  // are read.  Clearing them here also recovers from a call that was cut short.
  B200DP_CUDA_TRY(cudaMemsetAsync((char*)scratch + kScratchTickets, 0, kScratchCounts - kScratchTickets, s));
  if (dtype == B200DP_BF16) {
    bn_stats_kernel<__nv_bfloat16><<<dim3(g.gx, g.gy), kT, 0, s>>>(
        static_cast<const __nv_bfloat16*>(x), M, C, scratch, stats_local, P, 0, 0);
  } else {
    bn_stats_kernel<float><<<dim3(g.gx, g.gy), kT, 0, s>>>(static_cast<const float*>(x), M, C,
                                                           scratch, stats_local, P, 0, 0);
  }
  B200DP_LAUNCH_CHECK();
  return B200DP_OK;
}

namespace {
template <typename T>
int launch_fwd_apply(const FwdArgs& A, bool relu, const b200dp_peers& P, uint32_t seq,
                     cudaStream_t s) {
  const bool res = A.res != nullptr;
  const Grid2 g = plan_grid<T>(A.M, A.C, kOccFwd);
  const dim3 grid(g.gx, g.gy);
  const bool pdl = t_pdl_next;
  cudaError_t e;
  if (relu && res) e = launch_ex(bn_fwd_apply_kernel<T, true, true>, grid, s, pdl, A, P, seq);
  else if (relu) e = launch_ex(bn_fwd_apply_kernel<T, true, false>, grid, s, pdl, A, P, seq);
  else if (res) e = launch_ex(bn_fwd_apply_kernel<T, false, true>, grid, s, pdl, A, P, seq);
  else e = launch_ex(bn_fwd_apply_kernel<T, false, false>, grid, s, pdl, A, P, seq);
  if (e != cudaSuccess) return (int)e;
  B200DP_LAUNCH_CHECK();
  return B200DP_OK;
}
}  // namespace

static int fwd_apply_impl(const void* x, const void* residual, void* y, int dtype,
                          int64_t M, int C, const float* stats_local,
                          const float* weight, const float* bias, float eps,
                          float momentum, float* running_mean, float* running_var,
                          float* save_mean, float* save_invstd, float* save_count,
                          int relu, b200dp_comm* comm, uint32_t seq,
                          b200dp_stream_t stream, uint32_t* relu_mask, int64_t* batches_tracked) {
  int rc = check_common(x, residual, y, dtype, M, C);
  if (rc) return rc;
  if (relu_mask && (dtype != B200DP_BF16 || !relu)) return B200DP_EUNSUPPORTED;
  if (!stats_local || !save_mean || !save_invstd || !save_count) return B200DP_EINVAL;
  if (M > 0 && !y) return B200DP_EINVAL;
  if ((rc = check_comm(comm, seq))) return rc;
  FwdArgs A{x, residual, y, M, C, stats_local, weight, bias, eps, momentum, running_mean,
            running_var, save_mean, save_invstd, save_count, nullptr, nullptr, relu_mask, (bn_order() >> 1) & 1,
            reinterpret_cast<long long*>(batches_tracked)};
  const b200dp_peers P = b200dp_make_peers(comm);
  cudaStream_t s = (cudaStream_t)stream;
  return dtype == B200DP_BF16 ? launch_fwd_apply<__nv_bfloat16>(A, relu != 0, P, seq, s)
                              : launch_fwd_apply<float>(A, relu != 0, P, seq, s);
}

extern "C" int b200dp_bn_fwd_apply(const void* x, const void* residual, void* y, int dtype,
                                   int64_t M, int C, const float* stats_local,
                                   const float* weight, const float* bias, float eps,
                                   float momentum, float* running_mean, float* running_var,
                                   float* save_mean, float* save_invstd, float* save_count,
                                   int relu, b200dp_comm* comm, uint32_t seq,
                                   b200dp_stream_t stream, uint32_t* relu_mask) {
  return fwd_apply_impl(x, residual, y, dtype, M, C, stats_local, weight, bias, eps, momentum, running_mean,
                        running_var, save_mean, save_invstd, save_count, relu, comm, seq, stream, relu_mask,
                        nullptr);
}

extern "C" int b200dp_bn_inference(const void* x, const void* residual, void* y, int dtype,
                                   int64_t M, int C, const float* running_mean,
                                   const float* running_var, const float* weight,
                                   const float* bias, float eps, int relu,
                                   b200dp_stream_t stream) {
  int rc = check_common(x, residual, y, dtype, M, C);
  if (rc) return rc;
  if (!running_mean || !running_var) return B200DP_EINVAL;
  if (M == 0) return B200DP_OK;
  if (!y) return B200DP_EINVAL;
  FwdArgs A{x, residual, y, M, C, nullptr, weight, bias, eps, 0.f, nullptr, nullptr, nullptr,
            nullptr, nullptr, running_mean, running_var, nullptr, 0, nullptr};
  const b200dp_peers P = b200dp_make_peers(nullptr);
  cudaStream_t s = (cudaStream_t)stream;
  return dtype == B200DP_BF16 ? launch_fwd_apply<__nv_bfloat16>(A, relu != 0, P, 0, s)
                              : launch_fwd_apply<float>(A, relu != 0, P, 0, s);
}

namespace {
template <typename T>
int launch_bwd_reduce(const void* dy, const void* x, const void* yo, int64_t M, int C,
                      const float* save_mean, const float* save_invstd, const float* weight,
                      const float* bias, int relu, void* scratch, float* sums_local,
                      float* grad_weight, float* grad_bias, const b200dp_peers& P, uint32_t seq,
                      cudaStream_t s) {
  const int maxc = relu == 0   ? max_active_clusters(bn_bwd_reduce_kernel<T, 0>)
                   : relu == 1 ? max_active_clusters(bn_bwd_reduce_kernel<T, 1>)
                   : relu == 2 ? max_active_clusters(bn_bwd_reduce_kernel<T, 2>)
                               : max_active_clusters(bn_bwd_reduce_kernel<T, 3>);
  const Grid2 g = plan_grid_clustered<T>(M, C, maxc);
  const dim3 grid(g.gx, g.gy);
  const T* pdy = static_cast<const T*>(dy);
  const T* px = static_cast<const T*>(x);
  const T* py = static_cast<const T*>(yo);
#define B200DP_BWD_REDUCE(R)                                                                   \
  B200DP_CUDA_TRY(launch_ex(bn_bwd_reduce_kernel<T, R>, grid, s, pdl_first(), pdy, px, py, M, C, save_mean,   \
                            save_invstd, weight, bias, scratch, sums_local, grad_weight, grad_bias, P, seq,  \
                            (bn_order() >> 2) & 1))
  if (relu == 0) B200DP_BWD_REDUCE(0);
  else if (relu == 1) B200DP_BWD_REDUCE(1);
  else if (relu == 2) B200DP_BWD_REDUCE(2);
  else B200DP_BWD_REDUCE(3);
#undef B200DP_BWD_REDUCE
  B200DP_LAUNCH_CHECK();
  return B200DP_OK;
}

template <typename T>
int launch_bwd_apply(const BwdArgs& A, int relu, const b200dp_peers& P, uint32_t seq,
                     cudaStream_t s) {
  const Grid2 g = plan_grid<T>(A.M, A.C, kOccBwd);
  const dim3 grid(g.gx, g.gy);
  const bool dres = A.dres != nullptr;
  const bool pdl = t_pdl_next;
  cudaError_t e = cudaSuccess;
#define B200DP_BWD_APPLY(R)                                                     \
  do {                                                                          \
    if (dres) e = launch_ex(bn_bwd_apply_kernel<T, R, true>, grid, s, pdl, A, P, seq);   \
    else e = launch_ex(bn_bwd_apply_kernel<T, R, false>, grid, s, pdl, A, P, seq);       \
  } while (0)
  if (relu == 0) B200DP_BWD_APPLY(0);
  else if (relu == 1) B200DP_BWD_APPLY(1);
  else if (relu == 2) B200DP_BWD_APPLY(2);
  else B200DP_BWD_APPLY(3);
#undef B200DP_BWD_APPLY
  if (e != cudaSuccess) return (int)e;
  B200DP_LAUNCH_CHECK();
  return B200DP_OK;
}
}  // namespace

extern "C" int b200dp_bn_bwd_reduce(const void* dy, const void* x, const void* y_out, int dtype,
                                    int64_t M, int C, const float* save_mean,
                                    const float* save_invstd, const float* weight,
                                    const float* bias, int relu, void* scratch,
                                    float* sums_local, float* grad_weight, float* grad_bias,
                                    b200dp_comm* comm, uint32_t seq, b200dp_stream_t stream) {
  int rc = check_common(dy, x, y_out, dtype, M, C);
  if (rc) return rc;
  if (relu < 0 || relu > 3) return B200DP_EINVAL;
  if (relu == 3 && dtype != B200DP_BF16) return B200DP_EUNSUPPORTED;
  if (M > 0 && (!x || (relu >= 2 && !y_out))) return B200DP_EINVAL;
  if (!save_mean || !save_invstd || !scratch || !sums_local) return B200DP_EINVAL;
  if ((rc = check_comm(comm, seq))) return rc;
  const b200dp_peers P = b200dp_make_peers(comm);
  cudaStream_t s = (cudaStream_t)stream;
  return dtype == B200DP_BF16
             ? launch_bwd_reduce<__nv_bfloat16>(dy, x, y_out, M, C, save_mean, save_invstd,
                                                weight, bias, relu, scratch, sums_local,
                                                grad_weight, grad_bias, P, seq, s)
             : launch_bwd_reduce<float>(dy, x, y_out, M, C, save_mean, save_invstd, weight, bias,
                                        relu, scratch, sums_local, grad_weight, grad_bias, P,
                                        seq, s);
}

extern "C" int b200dp_bn_bwd_apply(const void* dy, const void* x, const void* y_out, void* dx,
                                   void* d_residual, int dtype, int64_t M, int C,
                                   const float* save_mean, const float* save_invstd,
                                   const float* weight, const float* bias,
                                   const float* sums_local, const float* save_count, int relu,
                                   b200dp_comm* comm, uint32_t seq, b200dp_stream_t stream) {
  int rc = check_common(dy, x, dx, dtype, M, C);
  if (rc) return rc;
  if (!aligned16(y_out) || !aligned16(d_residual)) return B200DP_EALIGN;
  if (relu < 0 || relu > 3) return B200DP_EINVAL;
  if (relu == 3 && dtype != B200DP_BF16) return B200DP_EUNSUPPORTED;
  if (M > 0 && (!x || !dx || (relu >= 2 && !y_out))) return B200DP_EINVAL;
  if (!save_mean || !save_invstd || !sums_local || !save_count) return B200DP_EINVAL;
  if ((rc = check_comm(comm, seq))) return rc;
  BwdArgs A{dy, x, y_out, dx, d_residual, M, C, save_mean, save_invstd, weight, bias,
            sums_local, save_count, (bn_order() >> 3) & 1};
  const b200dp_peers P = b200dp_make_peers(comm);
  cudaStream_t s = (cudaStream_t)stream;
  return dtype == B200DP_BF16 ? launch_bwd_apply<__nv_bfloat16>(A, relu, P, seq, s)
                              : launch_bwd_apply<float>(A, relu, P, seq, s);
}

// ---- one call per direction -------------------------------------------------------
// The Python host pays several microseconds per foreign call; for the small layers of
// CIFAR / 64x64 models that is as long as the kernels themselves.  These two entry points
// issue the same two launches as the four above with a single crossing of the boundary.
extern "C" int b200dp_bn_forward(const void* x, const void* residual, void* y, int dtype, int64_t M,
                                 int C, void* scratch, float* stats_local, const float* weight,
                                 const float* bias, float eps, float momentum, float* running_mean,
                                 float* running_var, float* save, int relu, b200dp_comm* comm,
                                 uint32_t seq, b200dp_stream_t stream, uint32_t* relu_mask,
                                 int64_t* num_batches_tracked) {
  if (!save) return B200DP_EINVAL;
  if ((uintptr_t)num_batches_tracked & 7) return B200DP_EALIGN;
  int rc;
  if (fused_cfg().on) {  // small layer: statistics and normalise in one cluster kernel
    rc = check_common(x, residual, y, dtype, M, C);
    if (rc) return rc;
    if (relu_mask && (dtype != B200DP_BF16 || !relu)) return B200DP_EUNSUPPORTED;
    if (!stats_local || !scratch) return B200DP_EINVAL;
    if (M > 0 && !y) return B200DP_EINVAL;
    if ((rc = check_comm(comm, seq))) return rc;
    FwdArgs A{x, residual, y, M, C, stats_local, weight, bias, eps, momentum, running_mean,
              running_var, save, save + C, save + 2 * C, nullptr, nullptr, relu_mask, 0,
              reinterpret_cast<long long*>(num_batches_tracked)};
    const b200dp_peers P = b200dp_make_peers(comm);
    rc = dtype == B200DP_BF16 ? launch_fwd_fused<__nv_bfloat16>(A, relu != 0, P, seq, (cudaStream_t)stream)
                              : launch_fwd_fused<float>(A, relu != 0, P, seq, (cudaStream_t)stream);
    if (rc != kNotFused) return rc;
  }
  rc = b200dp_bn_fwd_stats(x, dtype, M, C, scratch, stats_local, comm, seq, stream);
  if (rc) return rc;
  t_pdl_next = true;  // the normalise kernel may start while the statistics kernel finishes its tail
  rc = fwd_apply_impl(x, residual, y, dtype, M, C, stats_local, weight, bias, eps, momentum,
                      running_mean, running_var, save, save + C, save + 2 * C, relu, comm, seq,
                      stream, relu_mask, num_batches_tracked);
  t_pdl_next = false;
  return rc;
}

extern "C" int b200dp_bn_backward(const void* dy, const void* x, const void* y_out, void* dx,
                                  void* d_residual, int dtype, int64_t M, int C, const float* save,
                                  const float* weight, const float* bias, int relu, void* scratch,
                                  float* sums_local, float* grad_weight, float* grad_bias,
                                  b200dp_comm* comm, uint32_t seq, b200dp_stream_t stream) {
  if (!save) return B200DP_EINVAL;
  int rc;
  if (fused_cfg().on) {  // small layer: sums and apply in one cluster kernel
    rc = check_common(dy, x, dx, dtype, M, C);
    if (rc) return rc;
    if (!aligned16(y_out) || !aligned16(d_residual)) return B200DP_EALIGN;
    if (relu < 0 || relu > 3) return B200DP_EINVAL;
    if (relu == 3 && dtype != B200DP_BF16) return B200DP_EUNSUPPORTED;
    if (M > 0 && (!x || !dx || (relu >= 2 && !y_out))) return B200DP_EINVAL;
    if (!scratch || !sums_local) return B200DP_EINVAL;
    if ((rc = check_comm(comm, seq))) return rc;
    BwdFusedArgs F{{dy, x, y_out, dx, d_residual, M, C, save, save + C, weight, bias, sums_local, save + 2 * C, 0},
                   sums_local, grad_weight, grad_bias};
    const b200dp_peers P = b200dp_make_peers(comm);
    rc = dtype == B200DP_BF16 ? launch_bwd_fused<__nv_bfloat16>(F, relu, P, seq, (cudaStream_t)stream)
                              : launch_bwd_fused<float>(F, relu, P, seq, (cudaStream_t)stream);
    if (rc != kNotFused) return rc;
  }
  rc = b200dp_bn_bwd_reduce(dy, x, y_out, dtype, M, C, save, save + C, weight, bias, relu, scratch,
                            sums_local, grad_weight, grad_bias, comm, seq, stream);
  if (rc) return rc;
  t_pdl_next = true;
  rc = b200dp_bn_bwd_apply(dy, x, y_out, dx, d_residual, dtype, M, C, save, save + C, weight, bias,
                           sums_local, save + 2 * C, relu, comm, seq, stream);
  t_pdl_next = false;
  return rc;
}

// 1 when b200dp_bn_forward (backward != 0: b200dp_bn_backward) runs this layer as ONE kernel (a thread-block
// cluster per channel chunk, tensors that fit the L2), 0 when it issues the statistics / reduce kernel and
// the apply kernel.  For callers that time or count launches; the results are the same either way.
// Limits of the one-kernel path at run time (tests and A/B measurements; not thread-safe against calls in
// flight on other host threads): tensor size bounds in MiB for the forward / backward (0 = never, < 0 =
// leave as it is), the largest cluster size (1..16, anything else = leave as it is), and whether every shape
// under the limits takes the path whatever geometry fits (1) or only the shapes it was measured to win on
// (0, the default; < 0 = leave as it is).
extern "C" int b200dp_bn_set_one_kernel_limits(int fwd_mb, int bwd_mb, int cluster_size, int any_geometry) {
  FusedCfg& cfg = fused_cfg();
  if (any_geometry >= 0) cfg.any_geometry = any_geometry != 0;
  if (fwd_mb >= 0) cfg.fwd_bytes = (int64_t)fwd_mb << 20;
  if (bwd_mb >= 0) cfg.bwd_bytes = (int64_t)bwd_mb << 20;
  if (cluster_size >= 1 && cluster_size <= kFusedMaxCS) cfg.cs = cluster_size;  // an upper bound: powers of two up to it are tried
  cfg.on = cfg.fwd_bytes > 0 || cfg.bwd_bytes > 0;
  return B200DP_OK;
}

extern "C" int b200dp_bn_one_kernel(int dtype, int64_t M, int C, int backward) {
  const FusedCfg& cfg = fused_cfg();
  if (!cfg.on || (dtype != B200DP_F32 && dtype != B200DP_BF16) || C <= 0 || M < 0) return 0;
  // the planner's answer does not depend on the epilogue variant: ask with the plain one
  if (dtype == B200DP_BF16) {
    if (backward)
      return plan_fused<__nv_bfloat16>(bn_bwd_fused_kernel<__nv_bfloat16, 0, false>, kBwdThreads, 8 * kBwdThreads * 16, M, C, cfg.bwd_bytes).ok;
    return plan_fused<__nv_bfloat16>(bn_fwd_fused_kernel<__nv_bfloat16, false, false>, kFwdThreads, 8 * kFwdThreads * 16, M, C, cfg.fwd_bytes).ok;
  }
  if (backward) return plan_fused<float>(bn_bwd_fused_kernel<float, 0, false>, kBwdThreads, 8 * kBwdThreads * 16, M, C, cfg.bwd_bytes).ok;
  return plan_fused<float>(bn_fwd_fused_kernel<float, false, false>, kFwdThreads, 8 * kFwdThreads * 16, M, C, cfg.fwd_bytes).ok;
}
