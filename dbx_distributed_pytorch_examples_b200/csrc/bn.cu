// bn.cu — NHWC SyncBatchNorm for sm_100a: Welford statistics, cross-rank
// exchange by peer stores, fused BN(+residual)+ReLU forward and backward.
//
// Data layout.  An activation is [M, C] row-major (channels_last), M = N*H*W.
// A CTA of 256 threads is 32 rows x 8 sixteen-byte vectors: it owns one
// 128-byte channel chunk (64 bf16 / 32 fp32 channels) and walks down the rows,
// so every warp request is four full 128-byte lines and every thread keeps the
// same channels for its whole life (their statistics / coefficients stay in
// registers).  grid = (row splits, channel chunks), one resident wave: 4 CTAs
// per SM for the forward kernels (<= 64 registers), 3 for the backward ones.
//
// Reductions (stats, backward sums): thread -> CTA tree in shared memory ->
// per-CTA partial in `scratch` -> the last CTA of each chunk (ticket counter)
// merges the partials, writes the rank-local result and, when world > 1,
// stores it into every peer's exchange slot and raises that chunk's flag
// (st.release.sys).  The element-wise kernels that follow start by waiting on
// the flags of their own chunk, merge the ranks' vectors in rank order (so all
// ranks get bit-identical statistics) and stream.
//
// Algorithmic traffic per element (bf16): stats 2 B; fwd apply 4 B (+2 with a
// residual); bwd reduce 4 B; bwd apply 6 B (+2 for d_residual); the ReLU mask
// is recomputed from x (mode 1) or read from y (mode 2, +2 B per bwd pass).
#include <mutex>

#include "common.cuh"

namespace {

using namespace b200dp;

constexpr int kT = 256;           // threads per CTA
constexpr int kCV = 8;            // 16-byte vectors per chunk row (128 B)
constexpr int kRows = kT / kCV;   // 32 rows per CTA pass
constexpr int kOccFwd = 4;        // resident CTAs per SM, forward kernels
constexpr int kOccBwd = 3;        // resident CTAs per SM, backward kernels
constexpr int kMaxCtas = kOccFwd * B200DP_NUM_SMS;
constexpr int kMaxPartials = kMaxCtas * 64;  // (row split, channel) pairs
constexpr int kUnroll = 4;

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

// Tree over the 32 row-slots of a CTA.  sm is laid out [k][tid] (conflict free).
template <int N, bool WELFORD>
__device__ __forceinline__ void cta_tree(float* sA, float* sB, float* sN, float& n, float* a,
                                         float* b) {
  const int tid = threadIdx.x;
  const int r = tid / kCV;
#pragma unroll
  for (int k = 0; k < N; ++k) {
    sA[k * kT + tid] = a[k];
    sB[k * kT + tid] = b[k];
  }
  if (WELFORD) sN[tid] = n;
  __syncthreads();
  for (int s = kRows / 2; s > 0; s >>= 1) {
    if (r < s) {
      const int o = tid + s * kCV;
      float oa[N], ob[N];
#pragma unroll
      for (int k = 0; k < N; ++k) {
        oa[k] = sA[k * kT + o];
        ob[k] = sB[k * kT + o];
      }
      if (WELFORD) {
        welford_merge<N>(n, a, b, sN[o], oa, ob);
        sN[tid] = n;
      } else {
#pragma unroll
        for (int k = 0; k < N; ++k) {
          a[k] += oa[k];
          b[k] += ob[k];
        }
      }
#pragma unroll
      for (int k = 0; k < N; ++k) {
        sA[k * kT + tid] = a[k];
        sB[k * kT + tid] = b[k];
      }
    }
    __syncthreads();
  }
}

// After the CTA tree, threads with r == 0 hold the CTA's (n, a[], b[]) for their
// channels.  Publish the partial, elect the last CTA of this chunk, and let it
// merge all partials.  Returns true in the electing CTA, where r == 0 threads
// then hold the chunk's final values.
template <int N, bool WELFORD>
__device__ __forceinline__ bool grid_merge(void* scratch, int C, int chbase, bool chan_ok,
                                           float* sA, float* sB, float* sN, float& n, float* a,
                                           float* b) {
  __shared__ bool is_last;
  unsigned int* tickets = reinterpret_cast<unsigned int*>((char*)scratch + kScratchTickets);
  float* counts = reinterpret_cast<float*>((char*)scratch + kScratchCounts);
  float* pa = reinterpret_cast<float*>((char*)scratch + kScratchA);
  float* pb = reinterpret_cast<float*>((char*)scratch + kScratchB);
  const int tid = threadIdx.x, r = tid / kCV;
  const int bx = blockIdx.x, by = blockIdx.y, gx = gridDim.x;
  // partial index: (by * gx + bx) * chunk + local channel  (< kMaxPartials by construction)
  const int chunk = kCV * N;
  const int lc = (tid % kCV) * N;
  if (gx == 1) return true;  // single row split: the CTA result is final

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

// Store this rank's chunk of an exchange vector into every rank's slot and
// raise the chunk flag.  Called by the whole electing CTA; `a`/`b`/`cnt` are
// valid in threads with r == 0.
template <int N>
__device__ __forceinline__ void push_chunk(const b200dp_peers& P, uint32_t seq, int C, int c0,
                                           bool chan_ok, const float* a, const float* b,
                                           float cnt) {
  const int tid = threadIdx.x, r = tid / kCV;
  const int slot = seq % B200DP_XCHG_SLOTS;
  const int by = blockIdx.y;
  if (r == 0 && chan_ok) {
    for (int p = 0; p < P.world; ++p) {
      float* d = P.sync[p]->bn_data[slot][P.rank];
#pragma unroll
      for (int k = 0; k < N; ++k) {
        d[c0 + k] = a[k];
        d[C + c0 + k] = b[k];
      }
      if (tid == 0) d[2 * B200DP_BN_MAX_CHANNELS + by] = cnt;
    }
    __threadfence_system();
  }
  __syncthreads();
  if (tid < P.world) st_release_sys(&P.sync[tid]->bn_flag[slot][P.rank][by], seq);
}

// Wait until every rank's chunk `by` of exchange `seq` has landed here.
__device__ __forceinline__ void wait_chunk(const b200dp_peers& P, uint32_t seq) {
  const int slot = seq % B200DP_XCHG_SLOTS;
  if (threadIdx.x < P.world) {
    if (!wait_flag(&P.sync[P.rank]->bn_flag[slot][threadIdx.x][blockIdx.y], seq))
      atomicExch(&P.sync[P.rank]->error, 1u);
  }
  __syncthreads();
}

// =============================== forward: stats ===============================
template <typename T>
__global__ void __launch_bounds__(kT, kOccFwd)
bn_stats_kernel(const T* __restrict__ x, int64_t M, int C, void* scratch, float* stats_local,
                b200dp_peers P, uint32_t seq) {
  constexpr int N = Vec<T>::N;
  __shared__ float sA[N * kT], sB[N * kT], sN[kT];
  const int tid = threadIdx.x, r = tid / kCV, cv = tid % kCV;
  const int c0 = blockIdx.y * (kCV * N) + cv * N;
  const bool chan_ok = c0 < C;

  float n = 0.f, mean[N], m2[N];
#pragma unroll
  for (int k = 0; k < N; ++k) mean[k] = m2[k] = 0.f;

  if (chan_ok) {
    const int64_t stride = (int64_t)gridDim.x * kRows;
    for (int64_t row0 = (int64_t)blockIdx.x * kRows + r; row0 < M; row0 += kUnroll * stride) {
      uint4 v[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int64_t row = row0 + u * stride;
        if (row < M) v[u] = ld_keep(x + row * C + c0);
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int64_t row = row0 + u * stride;
        if (row < M) {
          float f[N];
          Vec<T>::unpack(v[u], f);
          n += 1.f;
          const float rn = __frcp_rn(n);
#pragma unroll
          for (int k = 0; k < N; ++k) {
            const float d = f[k] - mean[k];
            mean[k] = fmaf(d, rn, mean[k]);
            m2[k] = fmaf(d, f[k] - mean[k], m2[k]);
          }
        }
      }
    }
  }
  cta_tree<N, true>(sA, sB, sN, n, mean, m2);
  if (!grid_merge<N, true>(scratch, C, c0, chan_ok, sA, sB, sN, n, mean, m2)) return;

  if (r == 0 && chan_ok) {
#pragma unroll
    for (int k = 0; k < N; ++k) {
      stats_local[c0 + k] = mean[k];
      stats_local[C + c0 + k] = m2[k];
    }
    if (tid == 0 && blockIdx.y == 0) stats_local[2 * C] = n;
  }
  if (P.world > 1) push_chunk<N>(P, seq, C, c0, chan_ok, mean, m2, n);
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
};

template <typename T, bool RELU, bool RES>
__global__ void __launch_bounds__(kT, RES ? kOccBwd : kOccFwd) bn_fwd_apply_kernel(FwdArgs A, b200dp_peers P, uint32_t seq) {
  constexpr int N = Vec<T>::N;
  const int tid = threadIdx.x, r = tid / kCV, cv = tid % kCV;
  const int C = A.C;
  const int c0 = blockIdx.y * (kCV * N) + cv * N;
  const bool chan_ok = c0 < C;

  float mean[N], a[N], b[N];
  if (A.given_mean) {
    if (chan_ok) {
#pragma unroll
      for (int k = 0; k < N; ++k) {
        mean[k] = A.given_mean[c0 + k];
        const float invstd = 1.0f / sqrtf(A.given_var[c0 + k] + A.eps);
        a[k] = invstd * (A.weight ? A.weight[c0 + k] : 1.f);
        b[k] = A.bias ? A.bias[c0 + k] : 0.f;
      }
    }
  } else {
    float n = 0.f, m2[N];
#pragma unroll
    for (int k = 0; k < N; ++k) mean[k] = m2[k] = 0.f;
    if (P.world > 1) {
      wait_chunk(P, seq);
      if (chan_ok) {
        const int slot = seq % B200DP_XCHG_SLOTS;
        for (int p = 0; p < P.world; ++p) {
          const float* d = P.sync[P.rank]->bn_data[slot][p];
          float om[N], oq[N];
#pragma unroll
          for (int k = 0; k < N; ++k) {
            om[k] = ld_cg_f32(d + c0 + k);
            oq[k] = ld_cg_f32(d + C + c0 + k);
          }
          const float on = ld_cg_f32(d + 2 * B200DP_BN_MAX_CHANNELS + blockIdx.y);
          welford_merge<N>(n, mean, m2, on, om, oq);
        }
      }
    } else if (chan_ok) {
      n = ld_cg_f32(A.stats_local + 2 * C);
#pragma unroll
      for (int k = 0; k < N; ++k) {
        mean[k] = ld_cg_f32(A.stats_local + c0 + k);
        m2[k] = ld_cg_f32(A.stats_local + C + c0 + k);
      }
    }
    if (chan_ok) {
      const float inv_n = n > 0.f ? 1.0f / n : 0.f;
      const float inv_nm1 = n > 1.f ? 1.0f / (n - 1.f) : 0.f;
      const bool writer = (blockIdx.x == 0) && (r == 0);
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
  }
  if (!chan_ok) return;

  const T* x = static_cast<const T*>(A.x);
  const T* res = static_cast<const T*>(A.res);
  T* y = static_cast<T*>(A.y);
  const int64_t M = A.M;
  const int64_t stride = (int64_t)gridDim.x * kRows;
  for (int64_t row0 = (int64_t)blockIdx.x * kRows + r; row0 < M; row0 += kUnroll * stride) {
    uint4 v[kUnroll], w[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int64_t row = row0 + u * stride;
      if (row < M) {
        v[u] = ld_stream(x + row * C + c0);
        if (RES) w[u] = ld_stream(res + row * C + c0);
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int64_t row = row0 + u * stride;
      if (row < M) {
        float f[N], g[N];
        Vec<T>::unpack(v[u], f);
        if (RES) Vec<T>::unpack(w[u], g);
#pragma unroll
        for (int k = 0; k < N; ++k) {
          float o = fmaf(f[k] - mean[k], a[k], b[k]);
          if (RES) o += g[k];
          if (RELU) o = fmaxf(o, 0.f);
          f[k] = o;
        }
        // y is read again by the next convolution: leave it in L2
        *reinterpret_cast<uint4*>(y + row * C + c0) = Vec<T>::pack(f);
      }
    }
  }
}

// ============================= backward: reduce ==============================
// RELU: 0 none, 1 mask recomputed from x (no residual was fused), 2 mask from y.
template <typename T, int RELU>
__global__ void __launch_bounds__(kT, kOccBwd)
bn_bwd_reduce_kernel(const T* __restrict__ dy, const T* __restrict__ x, const T* __restrict__ yo,
                     int64_t M, int C, const float* save_mean, const float* save_invstd,
                     const float* weight, const float* bias, void* scratch, float* sums_local,
                     float* grad_weight, float* grad_bias, b200dp_peers P, uint32_t seq) {
  constexpr int N = Vec<T>::N;
  __shared__ float sA[N * kT], sB[N * kT];
  const int tid = threadIdx.x, r = tid / kCV, cv = tid % kCV;
  const int c0 = blockIdx.y * (kCV * N) + cv * N;
  const bool chan_ok = c0 < C;

  float mean[N], a[N], b[N], s1[N], s2[N];
#pragma unroll
  for (int k = 0; k < N; ++k) s1[k] = s2[k] = 0.f;
  if (chan_ok) {
#pragma unroll
    for (int k = 0; k < N; ++k) {
      mean[k] = save_mean[c0 + k];
      if (RELU == 1) {
        a[k] = save_invstd[c0 + k] * (weight ? weight[c0 + k] : 1.f);
        b[k] = bias ? bias[c0 + k] : 0.f;
      }
    }
    const int64_t stride = (int64_t)gridDim.x * kRows;
    for (int64_t row0 = (int64_t)blockIdx.x * kRows + r; row0 < M; row0 += kUnroll * stride) {
      uint4 vg[kUnroll], vx[kUnroll], vy[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int64_t row = row0 + u * stride;
        if (row < M) {
          vg[u] = ld_keep(dy + row * C + c0);
          vx[u] = ld_keep(x + row * C + c0);
          if (RELU == 2) vy[u] = ld_keep(yo + row * C + c0);
        }
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int64_t row = row0 + u * stride;
        if (row < M) {
          float g[N], f[N], o[N];
          Vec<T>::unpack(vg[u], g);
          Vec<T>::unpack(vx[u], f);
          if (RELU == 2) Vec<T>::unpack(vy[u], o);
#pragma unroll
          for (int k = 0; k < N; ++k) {
            const float xm = f[k] - mean[k];
            float gk = g[k];
            if (RELU == 1) gk = fmaf(xm, a[k], b[k]) > 0.f ? gk : 0.f;
            if (RELU == 2) gk = o[k] > 0.f ? gk : 0.f;
            s1[k] += gk;
            s2[k] = fmaf(gk, xm, s2[k]);
          }
        }
      }
    }
  }
  float dummy = 0.f;
  cta_tree<N, false>(sA, sB, nullptr, dummy, s1, s2);
  if (!grid_merge<N, false>(scratch, C, c0, chan_ok, sA, sB, nullptr, dummy, s1, s2)) return;

  if (r == 0 && chan_ok) {
#pragma unroll
    for (int k = 0; k < N; ++k) {
      sums_local[c0 + k] = s1[k];
      sums_local[C + c0 + k] = s2[k];
      if (grad_bias) grad_bias[c0 + k] = s1[k];
      if (grad_weight) grad_weight[c0 + k] = s2[k] * save_invstd[c0 + k];
    }
  }
  if (P.world > 1) push_chunk<N>(P, seq, C, c0, chan_ok, s1, s2, 0.f);
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
};

template <typename T, int RELU, bool DRES>
__global__ void __launch_bounds__(kT, kOccBwd) bn_bwd_apply_kernel(BwdArgs A, b200dp_peers P, uint32_t seq) {
  constexpr int N = Vec<T>::N;
  constexpr int U = N == 8 ? 2 : kUnroll;  // 40 coefficient registers at N = 8: fewer rows in flight
  const int tid = threadIdx.x, r = tid / kCV, cv = tid % kCV;
  const int C = A.C;
  const int c0 = blockIdx.y * (kCV * N) + cv * N;
  const bool chan_ok = c0 < C;

  float s1[N], s2[N];
#pragma unroll
  for (int k = 0; k < N; ++k) s1[k] = s2[k] = 0.f;
  if (P.world > 1) {
    wait_chunk(P, seq);
    if (chan_ok) {
      const int slot = seq % B200DP_XCHG_SLOTS;
      for (int p = 0; p < P.world; ++p) {
        const float* d = P.sync[P.rank]->bn_data[slot][p];
#pragma unroll
        for (int k = 0; k < N; ++k) {
          s1[k] += ld_cg_f32(d + c0 + k);
          s2[k] += ld_cg_f32(d + C + c0 + k);
        }
      }
    }
  } else if (chan_ok) {
#pragma unroll
    for (int k = 0; k < N; ++k) {
      s1[k] = ld_cg_f32(A.sums_local + c0 + k);
      s2[k] = ld_cg_f32(A.sums_local + C + c0 + k);
    }
  }
  if (!chan_ok) return;

  float mean[N], a[N], b[N], mdy[N], kx[N];
  {
    const float cnt = ld_cg_f32(A.save_count);
    const float inv_cnt = cnt > 0.f ? 1.0f / cnt : 0.f;
#pragma unroll
    for (int k = 0; k < N; ++k) {
      const float invstd = A.save_invstd[c0 + k];
      mean[k] = A.save_mean[c0 + k];
      a[k] = invstd * (A.weight ? A.weight[c0 + k] : 1.f);
      b[k] = A.bias ? A.bias[c0 + k] : 0.f;
      mdy[k] = s1[k] * inv_cnt;
      kx[k] = s2[k] * inv_cnt * invstd * invstd;
    }
  }

  const T* dy = static_cast<const T*>(A.dy);
  const T* x = static_cast<const T*>(A.x);
  const T* yo = static_cast<const T*>(A.yo);
  T* dx = static_cast<T*>(A.dx);
  T* dres = static_cast<T*>(A.dres);
  const int64_t M = A.M;
  const int64_t stride = (int64_t)gridDim.x * kRows;
  for (int64_t row0 = (int64_t)blockIdx.x * kRows + r; row0 < M; row0 += U * stride) {
    uint4 vg[U], vx[U], vy[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t row = row0 + u * stride;
      if (row < M) {
        vg[u] = ld_stream(dy + row * C + c0);
        vx[u] = ld_stream(x + row * C + c0);
        if (RELU == 2) vy[u] = ld_stream(yo + row * C + c0);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t row = row0 + u * stride;
      if (row < M) {
        float g[N], f[N], o[N];
        Vec<T>::unpack(vg[u], g);
        Vec<T>::unpack(vx[u], f);
        if (RELU == 2) Vec<T>::unpack(vy[u], o);
#pragma unroll
        for (int k = 0; k < N; ++k) {
          const float xm = f[k] - mean[k];
          float gk = g[k];
          if (RELU == 1) gk = fmaf(xm, a[k], b[k]) > 0.f ? gk : 0.f;
          if (RELU == 2) gk = o[k] > 0.f ? gk : 0.f;
          g[k] = gk;
          f[k] = (gk - mdy[k] - xm * kx[k]) * a[k];
        }
        *reinterpret_cast<uint4*>(dx + row * C + c0) = Vec<T>::pack(f);
        if (DRES) *reinterpret_cast<uint4*>(dres + row * C + c0) = Vec<T>::pack(g);
      }
    }
  }
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
    const Grid2 g = plan_grid<__nv_bfloat16>(M, C, kOccFwd);
    bn_stats_kernel<__nv_bfloat16><<<dim3(g.gx, g.gy), kT, 0, s>>>(
        static_cast<const __nv_bfloat16*>(x), M, C, scratch, stats_local, P, seq);
  } else {
    const Grid2 g = plan_grid<float>(M, C, kOccFwd);
    bn_stats_kernel<float><<<dim3(g.gx, g.gy), kT, 0, s>>>(static_cast<const float*>(x), M, C,
                                                           scratch, stats_local, P, seq);
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
  // The array is shared by every call and keeps what the last one left in it.  Only the
  // "last CTA" tickets have to start at zero; partials and counts are written before they
  // are read.  Clearing them here also recovers from a call that was cut short.
  B200DP_CUDA_TRY(cudaMemsetAsync((char*)scratch + kScratchTickets, 0, kScratchCounts - kScratchTickets, s));
  const b200dp_peers P = b200dp_make_peers(nullptr);
  Grid2 g;
  if (dtype == B200DP_BF16) {
    g = plan_grid<__nv_bfloat16>(M, C, kOccFwd);
    if ((size_t)g.gx * g.gy * kCV * Vec<__nv_bfloat16>::N > (size_t)kMaxPartials) return B200DP_ETOOBIG;
    bn_stats_kernel<__nv_bfloat16><<<dim3(g.gx, g.gy), kT, 0, s>>>(
        static_cast<const __nv_bfloat16*>(x), M, C, scratch, stats_local, P, 0);
  } else {
    g = plan_grid<float>(M, C, kOccFwd);
    if ((size_t)g.gx * g.gy * kCV * Vec<float>::N > (size_t)kMaxPartials) return B200DP_ETOOBIG;
    bn_stats_kernel<float><<<dim3(g.gx, g.gy), kT, 0, s>>>(static_cast<const float*>(x), M, C,
                                                           scratch, stats_local, P, 0);
  }
  B200DP_LAUNCH_CHECK();
  return B200DP_OK;
}

namespace {
template <typename T>
int launch_fwd_apply(const FwdArgs& A, bool relu, const b200dp_peers& P, uint32_t seq,
                     cudaStream_t s) {
  const bool res = A.res != nullptr;
  const Grid2 g = plan_grid<T>(A.M, A.C, res ? kOccBwd : kOccFwd);
  const dim3 grid(g.gx, g.gy);
  if (relu && res) bn_fwd_apply_kernel<T, true, true><<<grid, kT, 0, s>>>(A, P, seq);
  else if (relu) bn_fwd_apply_kernel<T, true, false><<<grid, kT, 0, s>>>(A, P, seq);
  else if (res) bn_fwd_apply_kernel<T, false, true><<<grid, kT, 0, s>>>(A, P, seq);
  else bn_fwd_apply_kernel<T, false, false><<<grid, kT, 0, s>>>(A, P, seq);
  B200DP_LAUNCH_CHECK();
  return B200DP_OK;
}
}  // namespace

extern "C" int b200dp_bn_fwd_apply(const void* x, const void* residual, void* y, int dtype,
                                   int64_t M, int C, const float* stats_local,
                                   const float* weight, const float* bias, float eps,
                                   float momentum, float* running_mean, float* running_var,
                                   float* save_mean, float* save_invstd, float* save_count,
                                   int relu, b200dp_comm* comm, uint32_t seq,
                                   b200dp_stream_t stream) {
  int rc = check_common(x, residual, y, dtype, M, C);
  if (rc) return rc;
  if (!stats_local || !save_mean || !save_invstd || !save_count) return B200DP_EINVAL;
  if (M > 0 && !y) return B200DP_EINVAL;
  if ((rc = check_comm(comm, seq))) return rc;
  FwdArgs A{x, residual, y, M, C, stats_local, weight, bias, eps, momentum, running_mean,
            running_var, save_mean, save_invstd, save_count, nullptr, nullptr};
  const b200dp_peers P = b200dp_make_peers(comm);
  cudaStream_t s = (cudaStream_t)stream;
  return dtype == B200DP_BF16 ? launch_fwd_apply<__nv_bfloat16>(A, relu != 0, P, seq, s)
                              : launch_fwd_apply<float>(A, relu != 0, P, seq, s);
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
            nullptr, nullptr, running_mean, running_var};
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
  const Grid2 g = plan_grid<T>(M, C, kOccBwd);
  const dim3 grid(g.gx, g.gy);
  const T* pdy = static_cast<const T*>(dy);
  const T* px = static_cast<const T*>(x);
  const T* py = static_cast<const T*>(yo);
#define B200DP_BWD_REDUCE(R)                                                                   \
  bn_bwd_reduce_kernel<T, R><<<grid, kT, 0, s>>>(pdy, px, py, M, C, save_mean, save_invstd,    \
                                                 weight, bias, scratch, sums_local,            \
                                                 grad_weight, grad_bias, P, seq)
  if (relu == 0) B200DP_BWD_REDUCE(0);
  else if (relu == 1) B200DP_BWD_REDUCE(1);
  else B200DP_BWD_REDUCE(2);
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
#define B200DP_BWD_APPLY(R)                                                     \
  do {                                                                          \
    if (dres) bn_bwd_apply_kernel<T, R, true><<<grid, kT, 0, s>>>(A, P, seq);   \
    else bn_bwd_apply_kernel<T, R, false><<<grid, kT, 0, s>>>(A, P, seq);       \
  } while (0)
  if (relu == 0) B200DP_BWD_APPLY(0);
  else if (relu == 1) B200DP_BWD_APPLY(1);
  else B200DP_BWD_APPLY(2);
#undef B200DP_BWD_APPLY
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
  if (relu < 0 || relu > 2) return B200DP_EINVAL;
  if (M > 0 && (!x || (relu == 2 && !y_out))) return B200DP_EINVAL;
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
  if (relu < 0 || relu > 2) return B200DP_EINVAL;
  if (M > 0 && (!x || !dx || (relu == 2 && !y_out))) return B200DP_EINVAL;
  if (!save_mean || !save_invstd || !sums_local || !save_count) return B200DP_EINVAL;
  if ((rc = check_comm(comm, seq))) return rc;
  BwdArgs A{dy, x, y_out, dx, d_residual, M, C, save_mean, save_invstd, weight, bias,
            sums_local, save_count};
  const b200dp_peers P = b200dp_make_peers(comm);
  cudaStream_t s = (cudaStream_t)stream;
  return dtype == B200DP_BF16 ? launch_bwd_apply<__nv_bfloat16>(A, relu, P, seq, s)
                              : launch_bwd_apply<float>(A, relu, P, seq, s);
}
