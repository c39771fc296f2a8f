// common.cuh — shared device helpers for libb200dp (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../include/b200dp.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libb200dp is written for sm_100a (Blackwell B200) only"
#endif

#define B200DP_NUM_SMS 148

#define B200DP_CUDA_TRY(expr)                 \
  do {                                        \
    cudaError_t _e = (expr);                  \
    if (_e != cudaSuccess) return (int)_e;    \
  } while (0)

// Kernels this library has launched in this process (comm.cu owns the counter; b200dp_kernel_launches()).
extern "C" unsigned long long b200dp_count_launch(void);

// After every kernel launch: a launch error must not stay sticky in the runtime's last-error slot; a
// successful launch is counted.
#define B200DP_LAUNCH_CHECK()                 \
  do {                                        \
    cudaError_t _e = cudaGetLastError();      \
    if (_e != cudaSuccess) return (int)_e;    \
    b200dp_count_launch();                    \
  } while (0)

namespace b200dp {

// ---- 128-bit streaming loads / stores --------------------------------------
// Activations are read once per pass: keep them out of L1 and mark them
// evict-first so they do not push the small per-channel tables out of L2.
__device__ __forceinline__ uint64_t evict_first_policy() {
  uint64_t pol;
  asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ uint4 ld_stream(const void* p) {
  uint4 r;
  const uint64_t pol = evict_first_policy();  // not volatile: hoisted out of loops
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p), "l"(pol));
  return r;
}
// Same, but leaves the line in L2 with normal priority: for a first pass whose
// data the next kernel reads again (statistics before normalise, reduce before
// apply); layers smaller than the 126 MB L2 then take their second pass from L2.
__device__ __forceinline__ uint4 ld_keep(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
// Plain coherent 128-bit load (data another kernel or peer may have written).
__device__ __forceinline__ uint4 ld_cg(const void* p) {
  uint4 r;
  asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}
__device__ __forceinline__ float ld_cg_f32(const float* p) {
  float r;
  asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(r) : "l"(p));
  return r;
}

// ---- system-scope flags (cross-GPU signalling over NVLink) ------------------
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_ll(uint2* p, float v, uint32_t seq) {
  asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(__float_as_uint(v)), "r"(seq) : "memory");
}
// Two adjacent LL words with one 16-byte store (p 16-byte aligned).  Each 8-byte half is a complete
// {value, seq} word of its own, so a reader that sees only one half updated simply keeps polling the other.
__device__ __forceinline__ void st_ll2(uint2* p, float v0, float v1, uint32_t seq) {
  const unsigned long long hi = (unsigned long long)seq << 32;
  asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(hi | __float_as_uint(v0)),
               "l"(hi | __float_as_uint(v1))
               : "memory");
}
// Default bound of every wait on a peer, in SM cycles: about 10 minutes at 2 GHz, the order of NCCL's
// watchdog.  B200DP_TIMEOUT_S overrides it (read once by b200dp_make_peers).
#define B200DP_DEFAULT_TIMEOUT_CYCLES (600LL * 2000000000LL)
// Poll an LL word until it carries `seq`; false once `timeout` cycles have passed.
__device__ __forceinline__ bool ld_ll(const uint2* p, uint32_t seq, float* out, long long timeout) {
  uint32_t a, b;
  long long t0 = 0;
  while (true) {
    asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(a), "=r"(b) : "l"(p) : "memory");
    if (b == seq) break;
    if (t0 == 0) t0 = clock64();
    if (clock64() - t0 > timeout) { *out = 0.f; return false; }
  }
  *out = __uint_as_float(a);
  return true;
}
// Same for two adjacent words read with one 16-byte load.
__device__ __forceinline__ bool ld_ll2(const uint2* p, uint32_t seq, float* out0, float* out1, long long timeout) {
  unsigned long long a, b;
  long long t0 = 0;
  while (true) {
    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
    if ((uint32_t)(a >> 32) == seq && (uint32_t)(b >> 32) == seq) break;
    if (t0 == 0) t0 = clock64();
    if (clock64() - t0 > timeout) { *out0 = *out1 = 0.f; return false; }
  }
  *out0 = __uint_as_float((uint32_t)a);
  *out1 = __uint_as_float((uint32_t)b);
  return true;
}
// Spin until *p has reached `seq` (wrap-safe).  Returns false on timeout so a
// dead peer turns into an error flag instead of a hung GPU.
__device__ __forceinline__ bool wait_flag(const uint32_t* p, uint32_t seq, long long timeout) {
  long long t0 = clock64();
  while ((int32_t)(ld_acquire_sys(p) - seq) < 0) {
    __nanosleep(20);
    if (clock64() - t0 > timeout) return false;
  }
  return true;
}

// ---- bf16 <-> fp32 on packed words -----------------------------------------
__device__ __forceinline__ float bf16lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

// ---- sm_100 mixed-precision FP32 arithmetic on bf16 operands ----------------
// `add/fma.rn.f32.bf16` take bf16 sources straight from a register half (SASS FHADD.BF16 /
// FHFMA.BF16 with .H0/.H1 selectors) and accumulate in fp32: no unpack instruction, exact
// products.  They halve the instruction count of the bf16 reduction kernels, which are
// otherwise issue-bound before they are HBM-bound.
__device__ __forceinline__ float add_bf16_f32(uint16_t a, float c) {  // float(a) + c
  float d;
  asm("add.rn.f32.bf16 %0, %1, %2;" : "=f"(d) : "h"(a), "f"(c));
  return d;
}
__device__ __forceinline__ float fma_bf16_f32(uint16_t a, uint16_t b, float c) {  // float(a)*float(b) + c
  float d;
  asm("fma.rn.f32.bf16 %0, %1, %2, %3;" : "=f"(d) : "h"(a), "h"(b), "f"(c));
  return d;
}
__device__ __forceinline__ uint16_t half_lo(uint32_t w) { return (uint16_t)(w & 0xffffu); }
__device__ __forceinline__ uint16_t half_hi(uint32_t w) { return (uint16_t)(w >> 16); }
// 16-bit half j (0..7) of a 16-byte vector
__device__ __forceinline__ uint16_t vec_half(const uint4& v, int j) {
  const uint32_t w = (j >> 1) == 0 ? v.x : (j >> 1) == 1 ? v.y : (j >> 1) == 2 ? v.z : v.w;
  return (j & 1) ? half_hi(w) : half_lo(w);
}
// packed bf16x2: (a > 0) ? 1.0 : 0.0 per half, and a * b per half (both exact)
__device__ __forceinline__ uint32_t bf16x2_gt0_mask(uint32_t a) {
  uint32_t d;
  asm("set.gt.bf16x2.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(0u));
  return d;
}
__device__ __forceinline__ uint32_t bf16x2_mul(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}

// A 16-byte vector of T seen as floats: 4 (fp32) or 8 (bf16) lanes.
template <typename T>
struct Vec;
template <>
struct Vec<float> {
  static constexpr int N = 4;
  __device__ __forceinline__ static void unpack(const uint4& r, float* f) {
    f[0] = __uint_as_float(r.x);
    f[1] = __uint_as_float(r.y);
    f[2] = __uint_as_float(r.z);
    f[3] = __uint_as_float(r.w);
  }
  __device__ __forceinline__ static uint4 pack(const float* f) {
    return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]),
                      __float_as_uint(f[3]));
  }
};
template <>
struct Vec<__nv_bfloat16> {
  static constexpr int N = 8;
  __device__ __forceinline__ static void unpack(const uint4& r, float* f) {
    f[0] = bf16lo(r.x);
    f[1] = bf16hi(r.x);
    f[2] = bf16lo(r.y);
    f[3] = bf16hi(r.y);
    f[4] = bf16lo(r.z);
    f[5] = bf16hi(r.z);
    f[6] = bf16lo(r.w);
    f[7] = bf16hi(r.w);
  }
  __device__ __forceinline__ static uint4 pack(const float* f) {
    return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]),
                      pack_bf16(f[6], f[7]));
  }
};

__host__ __device__ __forceinline__ bool aligned16(const void* p) {
  return (reinterpret_cast<uintptr_t>(p) & 15u) == 0;
}

}  // namespace b200dp

// ---- communicator layout (comm.cu owns it; bn.cu / update.cu read it) -------
// The "sync area" of every rank is one cudaMalloc'd block, IPC-mapped into all
// peers:  flags first, then the SyncBN exchange slots.
#define B200DP_XCHG_SLOTS 4
#define B200DP_BN_MAX_CHUNKS 128                       /* 4096 ch / 32 (fp32 chunk) */
#define B200DP_XCHG_FLOATS (2 * B200DP_BN_MAX_CHANNELS + B200DP_BN_MAX_CHUNKS)
#define B200DP_BUCKET_FLAG_SLOTS 64

struct b200dp_sync_area {
  // Gradient buckets: ready = "my bucket is packed", done = "I finished
  // reading/writing everybody's slice".
  uint32_t ready[B200DP_BUCKET_FLAG_SLOTS][B200DP_MAX_WORLD];
  uint32_t done[B200DP_BUCKET_FLAG_SLOTS][B200DP_MAX_WORLD];
  uint32_t barrier[B200DP_MAX_WORLD];
  uint32_t error;  // set by a kernel whose flag wait timed out
  uint32_t pad_[7];
  uint32_t ticket[B200DP_BUCKET_FLAG_SLOTS];  // local "last CTA" elections (self-resetting)
  // SyncBN exchange, "LL" style: every 8-byte word is {float bits, seq}.  A word is valid for
  // exchange `seq` once its second half equals seq; 8-byte stores are atomic, so no fence and
  // no separate flag are needed and the exchange costs one NVLink traversal.
  uint2 bn_ll[B200DP_XCHG_SLOTS][B200DP_MAX_WORLD][B200DP_XCHG_FLOATS];
};

struct b200dp_comm {
  int rank;
  int world;
  int device;
  size_t heap_bytes;
  void* heap[B200DP_MAX_WORLD];               // heap[rank] is the local allocation
  b200dp_sync_area* sync[B200DP_MAX_WORLD];   // sync[rank] is the local one
  bool peers_open;
  uint32_t seq;
  // b200dp_comm_create_external: the heap belongs to the caller (all ranks' mappings handed in), and
  // heap_mc, when not NULL, is a multicast (NVLS) mapping of the same memory on all ranks
  bool heap_external;
  void* heap_mc;
};

// Passed by value to kernels (as a __grid_constant__ parameter wherever sync[] is indexed at run time,
// so that the index goes to the constant bank instead of a local-memory copy of the struct).
struct b200dp_peers {
  int rank;
  int world;
  long long timeout;  // cycles a kernel waits for a peer before it raises the error word and gives up
  b200dp_sync_area* sync[B200DP_MAX_WORLD];
};

// Wait bound in cycles: B200DP_TIMEOUT_S seconds (default 600) at a nominal 2 GHz, read once.
static inline long long b200dp_timeout_cycles() {
  static const long long cycles = [] {
    const char* e = getenv("B200DP_TIMEOUT_S");
    const double s = e && e[0] ? atof(e) : 0.0;
    return s > 0.0 ? (long long)(s * 2.0e9) : B200DP_DEFAULT_TIMEOUT_CYCLES;
  }();
  return cycles;
}

static inline b200dp_peers b200dp_make_peers(const b200dp_comm* c) {
  b200dp_peers p;
  p.rank = c ? c->rank : 0;
  p.world = c ? c->world : 1;
  p.timeout = b200dp_timeout_cycles();
  for (int i = 0; i < B200DP_MAX_WORLD; ++i) p.sync[i] = c ? c->sync[i] : nullptr;
  return p;
}
