// update.cu — gradient bucket pack, and the all-reduce fused with the 1/N
// average and the optimizer step, reading peer HBM directly over NVLink.
//
// Two-shot (large buckets):  rank r owns the r-th 1/world slice of the bucket.
//   wait until every rank has raised `ready`            (grads are packed)
//   for each 16-byte vector of my slice (one per thread and pass):
//       g  = sum over ranks, in rank order, of that unit read from the rank's
//            bucket (ld.global.cg on the IPC-mapped pointer: an NVLink load)
//       SGD-momentum / Adam on my master weights and state (local HBM, fp32)
//       store the new parameters into EVERY rank's parameter buffer
//   raise `done` on every rank, wait for every rank's `done`.
// Bytes over NVLink per rank: (world-1)/world * S read + (world-1)/world * S_p
// written — the reduce-scatter + all-gather minimum — and the optimizer pass
// costs no extra trip through HBM for the gradient.
// One-shot (small buckets): every rank reads all buckets and updates all of its
// own parameters; nothing is pushed, (world-1)*S read.
#include "common.cuh"

namespace {

using namespace b200dp;

constexpr int kT = 512;
constexpr int kPackMax = 112;  // tensors per pack launch (kernel parameter space)

// ------------------------------- bucket pack -------------------------------
struct PackTable {
  const void* src[kPackMax];
  long long numel[kPackMax];
  long long off[kPackMax];
  int n;
};

template <typename T>
__global__ void __launch_bounds__(kT) pack_kernel(const PackTable tab, T* __restrict__ dst,
                                                  float scale) {
  constexpr int N = Vec<T>::N;
  const long long gtid = (long long)blockIdx.x * kT + threadIdx.x;
  const long long gsz = (long long)gridDim.x * kT;
  for (int t = 0; t < tab.n; ++t) {
    const T* s = static_cast<const T*>(tab.src[t]);
    T* d = dst + tab.off[t];
    const long long n = tab.numel[t];
    long long done = 0;
    if (aligned16(s)) {
      const long long nv = n / N;
      for (long long i = gtid; i < nv; i += gsz) {
        float f[N];
        Vec<T>::unpack(ld_stream(s + i * N), f);
#pragma unroll
        for (int k = 0; k < N; ++k) f[k] *= scale;
        *reinterpret_cast<uint4*>(d + i * N) = Vec<T>::pack(f);
      }
      done = nv * N;
    }
    for (long long i = done + gtid; i < n; i += gsz) d[i] = (T)((float)s[i] * scale);
  }
}

// ---------------------------- fused reduce+update ---------------------------
struct FusedArgs {
  const void* grad[B200DP_MAX_WORLD];  // bucket base in every rank's heap
  void* param[B200DP_MAX_WORLD];       // parameter base in every rank's heap
  float* master;                       // local fp32 master, or NULL (PT == float)
  float* s1;                           // momentum buffer / exp_avg
  float* s2;                           // exp_avg_sq
  const void* grad_mc;                 // multicast mapping of the bucket (NVLS path) or NULL
  void* param_mc;                      // multicast mapping of the parameter area or NULL
  long long begin, end;                // element range this rank reduces
  b200dp_opt_args opt;
  int one_shot;
  int write_grads;
  uint32_t seq;
};

// A thread's unit of work is ONE 16-byte vector of the bucket (E = 4 fp32 or 8 bf16 elements), so that a
// warp's request to a rank is 512 contiguous bytes whatever the element type.  (Two adjacent 16-byte
// loads per thread, as an 8-element fp32 unit would need, make every warp request touch each 32-byte
// sector twice; peer memory is not cached in the local L2, so that doubled the NVLink traffic of fp32
// buckets.)  fp32 state for E elements is E/4 vectors.
template <int E>
__device__ __forceinline__ void load_f32(const float* p, float* f) {
#pragma unroll
  for (int k = 0; k < E; k += 4) Vec<float>::unpack(ld_cg(p + k), f + k);
}
template <int E>
__device__ __forceinline__ void store_f32(float* p, const float* f) {
#pragma unroll
  for (int k = 0; k < E; k += 4) *reinterpret_cast<uint4*>(p + k) = Vec<float>::pack(f + k);
}

// One optimizer step on E elements.  g: averaged gradient (in), p/s1/s2 in-out.
// Operation order follows torch's foreach kernels so that fp32 results agree to
// the last bit or two (TORCH/optim/sgd.py:322-380, TORCH/optim/adam.py:347+).
template <int OPT, int E>
__device__ __forceinline__ void opt_step(const b200dp_opt_args& o, float* g, float* p, float* s1,
                                         float* s2) {
#pragma unroll
  for (int k = 0; k < E; ++k) {
    float gk = g[k];
    if (o.weight_decay != 0.f) gk = fmaf(p[k], o.weight_decay, gk);
    if (OPT == B200DP_OPT_SGD) {
      if (o.momentum != 0.f) {
        float buf;
        if (o.first_step) {
          buf = gk;
        } else {
          buf = __fmul_rn(s1[k], o.momentum);
          buf = fmaf(gk, 1.f - o.dampening, buf);
        }
        s1[k] = buf;
        gk = o.nesterov ? fmaf(buf, o.momentum, gk) : buf;
      }
      p[k] = fmaf(gk, -o.lr, p[k]);
    } else if (OPT == B200DP_OPT_ADAM) {
      const float m = fmaf(1.f - o.beta1, gk - s1[k], s1[k]);  // lerp
      float v = __fmul_rn(s2[k], o.beta2);
      v = fmaf(__fmul_rn(1.f - o.beta2, gk), gk, v);           // addcmul
      s1[k] = m;
      s2[k] = v;
      const float denom = __fdiv_rn(sqrtf(v), o.bias_correction2_sqrt) + o.eps;
      p[k] = fmaf(__fdiv_rn(m, denom), -o.step_size, p[k]);
    }
  }
}

template <typename GT, typename PT, int OPT, int WORLD>
__global__ void __launch_bounds__(kT, 1) fused_kernel(const __grid_constant__ FusedArgs A,
                                                      const __grid_constant__ b200dp_peers P) {
  __shared__ bool is_last;
  __shared__ int s_ok;
  constexpr int E = Vec<GT>::N;  // elements per 16-byte vector
  const int tid = threadIdx.x;
  const int slot = A.seq % B200DP_BUCKET_FLAG_SLOTS;
  b200dp_sync_area* me = P.sync[P.rank];
  constexpr int world = WORLD;

  if (world > 1) {
    if (tid == 0) s_ok = 1;
    __syncthreads();
    if (blockIdx.x == 0 && tid < world) st_release_sys(&P.sync[tid]->ready[slot][P.rank], A.seq);
    if (tid < world) {
      if (!wait_flag(&me->ready[slot][tid], A.seq, P.timeout)) {
        atomicExch(&me->error, 1u);
        s_ok = 0;
      }
    }
    __syncthreads();
    // A peer never showed up: its bucket is stale.  Leave gradients, state and parameters alone
    // (the host sees the error word and stops) rather than push a wrong update to everybody.
    if (!s_ok) return;
  }

  const bool push = (world > 1) && !A.one_shot;
  PT* my_param = static_cast<PT*>(A.param[P.rank]);
  const long long gsz = (long long)gridDim.x * kT * E;
  for (long long i = A.begin + ((long long)blockIdx.x * kT + tid) * E; i < A.end; i += gsz) {
    // all ranks' loads are issued before the first add (WORLD is a constant,
    // so this unrolls into WORLD independent NVLink requests per thread)
    uint4 raw[WORLD];
#pragma unroll
    for (int r = 0; r < world; ++r) raw[r] = ld_cg(static_cast<const GT*>(A.grad[r]) + i);
    float g[E];
    Vec<GT>::unpack(raw[0], g);
#pragma unroll
    for (int r = 1; r < world; ++r) {  // rank order: every rank computes the same sum
      float h[E];
      Vec<GT>::unpack(raw[r], h);
#pragma unroll
      for (int k = 0; k < E; ++k) g[k] += h[k];
    }
    if (A.write_grads) {
      const uint4 gv = Vec<GT>::pack(g);
      if (push) {
#pragma unroll
        for (int r = 0; r < world; ++r)
          *reinterpret_cast<uint4*>(static_cast<GT*>(const_cast<void*>(A.grad[r])) + i) = gv;
      } else {
        *reinterpret_cast<uint4*>(static_cast<GT*>(const_cast<void*>(A.grad[P.rank])) + i) = gv;
      }
    }
    if (OPT != B200DP_OPT_NONE) {
      float p[E], s1[E], s2[E];
      if (A.master) {
        load_f32<E>(A.master + i, p);
      } else {
        Vec<PT>::unpack(ld_cg(my_param + i), p);
      }
      if (OPT == B200DP_OPT_ADAM || (A.opt.momentum != 0.f && !A.opt.first_step))
        load_f32<E>(A.s1 + i, s1);
      if (OPT == B200DP_OPT_ADAM) load_f32<E>(A.s2 + i, s2);
      opt_step<OPT, E>(A.opt, g, p, s1, s2);
      if (A.master) store_f32<E>(A.master + i, p);
      if (OPT == B200DP_OPT_ADAM || A.opt.momentum != 0.f) store_f32<E>(A.s1 + i, s1);
      if (OPT == B200DP_OPT_ADAM) store_f32<E>(A.s2 + i, s2);
      const uint4 pv = Vec<PT>::pack(p);
      if (push) {
#pragma unroll
        for (int r = 0; r < world; ++r) *reinterpret_cast<uint4*>(static_cast<PT*>(A.param[r]) + i) = pv;
      } else {
        *reinterpret_cast<uint4*>(my_param + i) = pv;
      }
    }
  }

  if (world > 1) {
    // all of this rank's peer stores are out before `done` is raised
    __threadfence_system();
    __syncthreads();
    if (tid == 0) {
      const unsigned int t = atomicAdd(&me->ticket[slot], 1u);
      is_last = (t == gridDim.x - 1);
      if (is_last) me->ticket[slot] = 0;
    }
    __syncthreads();
    if (is_last && tid < world) {
      __threadfence_system();
      st_release_sys(&P.sync[tid]->done[slot][P.rank], A.seq);
      if (!wait_flag(&me->done[slot][tid], A.seq, P.timeout)) atomicExch(&me->error, 1u);
    }
  }
}

// ---- NVLS variant of the two-shot kernel ------------------------------------------------------------
// With a multicast mapping of the heap the NVSwitch does the cross-rank part: multimem.ld_reduce returns
// the sum of the same 16 bytes on all ranks (each GPU receives 1/world of the bucket instead of reading
// (world-1)/world of it from its peers) and multimem.st stores the new parameters on every rank with one
// store instead of one per peer.  Per direction a GPU moves ~S instead of ~1.75 S at 8 ranks.
// bf16 sums are accumulated in fp32 inside the switch and delivered as bf16 (what an NCCL all-reduce of a
// bf16 bucket — the reference DDP path for bf16 parameters — delivers).
template <typename T>
__device__ __forceinline__ uint4 multimem_ld_reduce(const T* mc);
template <>
__device__ __forceinline__ uint4 multimem_ld_reduce<float>(const float* mc) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(mc)
               : "memory");
  return r;
}
template <>
__device__ __forceinline__ uint4 multimem_ld_reduce<__nv_bfloat16>(const __nv_bfloat16* mc) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(mc)
               : "memory");
  return r;
}
template <typename T>
__device__ __forceinline__ void multimem_st(T* mc, const uint4& v);
template <>
__device__ __forceinline__ void multimem_st<float>(float* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}
template <>
__device__ __forceinline__ void multimem_st<__nv_bfloat16>(__nv_bfloat16* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}

constexpr int kNvlsUnroll = 4;  // independent reductions in flight per thread

template <typename GT, int OPT>
__global__ void __launch_bounds__(kT, 1) fused_nvls_kernel(const __grid_constant__ FusedArgs A,
                                                           const __grid_constant__ b200dp_peers P) {
  __shared__ bool is_last;
  __shared__ int s_ok;
  constexpr int E = Vec<GT>::N;
  const int tid = threadIdx.x;
  const int slot = A.seq % B200DP_BUCKET_FLAG_SLOTS;
  b200dp_sync_area* me = P.sync[P.rank];
  const int world = P.world;

  if (tid == 0) s_ok = 1;
  __syncthreads();
  if (blockIdx.x == 0 && tid < world) st_release_sys(&P.sync[tid]->ready[slot][P.rank], A.seq);
  if (tid < world) {
    if (!wait_flag(&me->ready[slot][tid], A.seq, P.timeout)) {
      atomicExch(&me->error, 1u);
      s_ok = 0;
    }
  }
  __syncthreads();
  if (!s_ok) return;

  const GT* gmc = static_cast<const GT*>(A.grad_mc);
  GT* pmc = static_cast<GT*>(A.param_mc);
  const GT* my_param = static_cast<const GT*>(A.param[P.rank]);
  const long long stride = (long long)gridDim.x * kT * E;
  for (long long i0 = A.begin + ((long long)blockIdx.x * kT + tid) * E; i0 < A.end; i0 += stride * kNvlsUnroll) {
    uint4 raw[kNvlsUnroll];
#pragma unroll
    for (int u = 0; u < kNvlsUnroll; ++u) {
      const long long i = i0 + u * stride;
      if (i < A.end) raw[u] = multimem_ld_reduce<GT>(gmc + i);
    }
#pragma unroll
    for (int u = 0; u < kNvlsUnroll; ++u) {
      const long long i = i0 + u * stride;
      if (i >= A.end) break;
      float g[E];
      Vec<GT>::unpack(raw[u], g);
      if (A.write_grads) multimem_st<GT>(const_cast<GT*>(gmc) + i, raw[u]);
      if (OPT != B200DP_OPT_NONE) {
        float p[E], s1[E], s2[E];
        if (A.master) {
          load_f32<E>(A.master + i, p);
        } else {
          Vec<GT>::unpack(ld_cg(my_param + i), p);
        }
        if (OPT == B200DP_OPT_ADAM || (A.opt.momentum != 0.f && !A.opt.first_step)) load_f32<E>(A.s1 + i, s1);
        if (OPT == B200DP_OPT_ADAM) load_f32<E>(A.s2 + i, s2);
        opt_step<OPT, E>(A.opt, g, p, s1, s2);
        if (A.master) store_f32<E>(A.master + i, p);
        if (OPT == B200DP_OPT_ADAM || A.opt.momentum != 0.f) store_f32<E>(A.s1 + i, s1);
        if (OPT == B200DP_OPT_ADAM) store_f32<E>(A.s2 + i, s2);
        multimem_st<GT>(pmc + i, Vec<GT>::pack(p));
      }
    }
  }

  __threadfence_system();
  __syncthreads();
  if (tid == 0) {
    const unsigned int t = atomicAdd(&me->ticket[slot], 1u);
    is_last = (t == gridDim.x - 1);
    if (is_last) me->ticket[slot] = 0;
  }
  __syncthreads();
  if (is_last && tid < world) {
    __threadfence_system();
    st_release_sys(&P.sync[tid]->done[slot][P.rank], A.seq);
    if (!wait_flag(&me->done[slot][tid], A.seq, P.timeout)) atomicExch(&me->error, 1u);
  }
}

template <typename GT>
int launch_nvls(const FusedArgs& A, const b200dp_peers& P, int grid, cudaStream_t s) {
  switch (A.opt.kind) {
    case B200DP_OPT_NONE: fused_nvls_kernel<GT, B200DP_OPT_NONE><<<grid, kT, 0, s>>>(A, P); break;
    case B200DP_OPT_SGD: fused_nvls_kernel<GT, B200DP_OPT_SGD><<<grid, kT, 0, s>>>(A, P); break;
    case B200DP_OPT_ADAM: fused_nvls_kernel<GT, B200DP_OPT_ADAM><<<grid, kT, 0, s>>>(A, P); break;
    default: return B200DP_EINVAL;
  }
  B200DP_LAUNCH_CHECK();
  return B200DP_OK;
}

// B200DP_NVLS=0 keeps the peer load/store kernel even when the heap has a multicast mapping.
bool nvls_enabled() {
  static const bool on = [] { const char* e = getenv("B200DP_NVLS"); return !(e && e[0] == '0'); }();
  return on;
}

template <typename GT, typename PT, int WORLD>
int launch_fused_w(const FusedArgs& A, const b200dp_peers& P, int grid, cudaStream_t s) {
  switch (A.opt.kind) {
    case B200DP_OPT_NONE:
      fused_kernel<GT, PT, B200DP_OPT_NONE, WORLD><<<grid, kT, 0, s>>>(A, P);
      break;
    case B200DP_OPT_SGD:
      fused_kernel<GT, PT, B200DP_OPT_SGD, WORLD><<<grid, kT, 0, s>>>(A, P);
      break;
    case B200DP_OPT_ADAM:
      fused_kernel<GT, PT, B200DP_OPT_ADAM, WORLD><<<grid, kT, 0, s>>>(A, P);
      break;
    default:
      return B200DP_EINVAL;
  }
  B200DP_LAUNCH_CHECK();
  return B200DP_OK;
}

template <typename GT, typename PT>
int launch_fused(const FusedArgs& A, const b200dp_peers& P, int grid, cudaStream_t s) {
  switch (P.world) {
    case 1: return launch_fused_w<GT, PT, 1>(A, P, grid, s);
    case 2: return launch_fused_w<GT, PT, 2>(A, P, grid, s);
    case 4: return launch_fused_w<GT, PT, 4>(A, P, grid, s);
    case 8: return launch_fused_w<GT, PT, 8>(A, P, grid, s);
    default: return B200DP_EUNSUPPORTED;  // world sizes 1, 2, 4, 8 (one NVSwitch node)
  }
}

}  // namespace

extern "C" int b200dp_bucket_pack(const void* const* srcs, const int64_t* numels,
                                  const int64_t* offsets, int n_tensors, void* dst, int dtype,
                                  float scale, b200dp_stream_t stream) {
  if (n_tensors < 0 || n_tensors > 1024) return B200DP_ETOOBIG;
  if (n_tensors == 0) return B200DP_OK;
  if (!srcs || !numels || !offsets || !dst) return B200DP_EINVAL;
  if (dtype != B200DP_F32 && dtype != B200DP_BF16) return B200DP_EUNSUPPORTED;
  if (!b200dp::aligned16(dst)) return B200DP_EALIGN;
  cudaStream_t s = (cudaStream_t)stream;
  for (int base = 0; base < n_tensors; base += kPackMax) {
    PackTable tab;
    tab.n = n_tensors - base < kPackMax ? n_tensors - base : kPackMax;
    long long total = 0;
    for (int t = 0; t < tab.n; ++t) {
      if (numels[base + t] < 0 || offsets[base + t] < 0 || (offsets[base + t] & 7))
        return B200DP_EINVAL;
      if (numels[base + t] > 0 && !srcs[base + t]) return B200DP_EINVAL;
      tab.src[t] = srcs[base + t];
      tab.numel[t] = numels[base + t];
      tab.off[t] = offsets[base + t];
      total += numels[base + t];
    }
    long long want = (total / 8 + kT - 1) / kT;
    int grid = (int)(want < 1 ? 1 : (want > 2 * B200DP_NUM_SMS ? 2 * B200DP_NUM_SMS : want));
    if (dtype == B200DP_BF16)
      pack_kernel<__nv_bfloat16><<<grid, kT, 0, s>>>(tab, static_cast<__nv_bfloat16*>(dst), scale);
    else
      pack_kernel<float><<<grid, kT, 0, s>>>(tab, static_cast<float*>(dst), scale);
    B200DP_LAUNCH_CHECK();
  }
  return B200DP_OK;
}

extern "C" int b200dp_allreduce_update(b200dp_comm* comm, size_t grad_heap_off, int grad_dtype,
                                       size_t param_heap_off, int param_dtype,
                                       int64_t elem_offset, int64_t numel, float* master,
                                       float* state1, float* state2, const b200dp_opt_args* opt,
                                       int one_shot, int write_grads, int max_ctas, uint32_t seq,
                                       b200dp_stream_t stream) {
  if (!comm || !opt) return B200DP_EINVAL;
  if (!comm->peers_open) return B200DP_ESTATE;
  const int world = comm->world;
  if (numel < 0 || elem_offset < 0 || (elem_offset & 7)) return B200DP_EINVAL;
  if (numel % (8 * world) != 0) return B200DP_EINVAL;
  if ((grad_dtype != B200DP_F32 && grad_dtype != B200DP_BF16) ||
      (param_dtype != B200DP_F32 && param_dtype != B200DP_BF16))
    return B200DP_EUNSUPPORTED;
  if (grad_dtype != param_dtype) return B200DP_EUNSUPPORTED;  // one element type per arena
  if ((grad_heap_off & 15) || (param_heap_off & 15)) return B200DP_EALIGN;
  const size_t gsz = grad_dtype == B200DP_BF16 ? 2 : 4, psz = param_dtype == B200DP_BF16 ? 2 : 4;
  if (grad_heap_off + (size_t)(elem_offset + numel) * gsz > comm->heap_bytes) return B200DP_EINVAL;
  if (world > 1 && seq == 0) return B200DP_EINVAL;
  if (one_shot && write_grads && world > 1) return B200DP_EINVAL;
  const int kind = opt->kind;
  if (kind == B200DP_OPT_NONE) {
    if (!write_grads) return B200DP_EINVAL;
  } else {
    if (param_heap_off + (size_t)(elem_offset + numel) * psz > comm->heap_bytes)
      return B200DP_EINVAL;
    if (param_dtype == B200DP_BF16 && !master) return B200DP_EINVAL;
    if (kind == B200DP_OPT_ADAM && (!state1 || !state2)) return B200DP_EINVAL;
    if (kind == B200DP_OPT_SGD && opt->momentum != 0.f && !state1) return B200DP_EINVAL;
    if (!b200dp::aligned16(master) || !b200dp::aligned16(state1) || !b200dp::aligned16(state2))
      return B200DP_EALIGN;
  }
  if (numel == 0) return B200DP_OK;

  FusedArgs A;
  for (int r = 0; r < B200DP_MAX_WORLD; ++r) {
    A.grad[r] = nullptr;
    A.param[r] = nullptr;
  }
  for (int r = 0; r < world; ++r) {
    A.grad[r] = (const char*)comm->heap[r] + grad_heap_off + (size_t)elem_offset * gsz;
    A.param[r] = (char*)comm->heap[r] + param_heap_off + (size_t)elem_offset * psz;
  }
  // master/state pointers address the first element of the bucket; the kernel
  // indexes everything with the in-bucket element index.
  A.grad_mc = A.param_mc = nullptr;
  if (comm->heap_mc) {
    A.grad_mc = (const char*)comm->heap_mc + grad_heap_off + (size_t)elem_offset * gsz;
    A.param_mc = (char*)comm->heap_mc + param_heap_off + (size_t)elem_offset * psz;
  }
  A.master = master;
  A.s1 = state1;
  A.s2 = state2;
  if (world > 1 && !one_shot) {
    const long long slice = numel / world;
    A.begin = slice * comm->rank;
    A.end = A.begin + slice;
  } else {
    A.begin = 0;
    A.end = numel;
  }
  A.opt = *opt;
  A.one_shot = one_shot;
  A.write_grads = write_grads;
  A.seq = seq;

  const long long units = (A.end - A.begin) / (grad_dtype == B200DP_BF16 ? 8 : 4);  // 16-byte vectors
  long long want = (units + kT - 1) / kT;
  // library default: a bounded grid for buckets that overlap the backward pass, the whole GPU for large ones
  const size_t bucket_bytes = (size_t)numel * gsz;
  int cap = max_ctas > 0 ? max_ctas : (world > 1 ? (bucket_bytes >= (32u << 20) ? B200DP_NUM_SMS : 64) : B200DP_NUM_SMS);
  if (cap > B200DP_NUM_SMS) cap = B200DP_NUM_SMS;  // every CTA must be resident: they elect a last one
  int grid = (int)(want < 1 ? 1 : (want > cap ? cap : want));
  const b200dp_peers P = b200dp_make_peers(comm);
  cudaStream_t s = (cudaStream_t)stream;
  if (world > 1 && !one_shot && comm->heap_mc && nvls_enabled() && grad_dtype == param_dtype) {
    // each thread keeps kNvlsUnroll reductions in flight, and the switch does the fan-in: 32 CTAs already
    // saturate the links (profiles/r2_allreduce_sweep_nvls_n8.jsonl: 714 GB/s bus at 64 MB with 32 CTAs, 683
    // with 64, 614 with 148), so that is the library default
    long long want4 = (units + (long long)kT * kNvlsUnroll - 1) / ((long long)kT * kNvlsUnroll);
    const int cap4 = max_ctas > 0 ? cap : 32;
    int g4 = (int)(want4 < 1 ? 1 : (want4 > cap4 ? cap4 : want4));
    return grad_dtype == B200DP_F32 ? launch_nvls<float>(A, P, g4, s) : launch_nvls<__nv_bfloat16>(A, P, g4, s);
  }
  if (grad_dtype == B200DP_F32 && param_dtype == B200DP_F32)
    return launch_fused<float, float>(A, P, grid, s);
  if (grad_dtype == B200DP_BF16 && param_dtype == B200DP_BF16)
    return launch_fused<__nv_bfloat16, __nv_bfloat16>(A, P, grid, s);
  return B200DP_EUNSUPPORTED;  // gradient and parameter element types must match
}
