// pool.cu — NHWC max pooling (window <= 3x3) forward and backward.
//
// The pooling layer that follows the stem's BN + ReLU in the reference model
// (setup/resnet18.py:37,58: nn.MaxPool2d(kernel_size=3, stride=2, padding=1) applied to
// relu(bn1(conv1(x)))) and in torchvision's resnet18/50.  It is the last piece of that
// BN -> ReLU -> pool epilogue and, on ResNet-50 224^2, torch's NHWC kernels spend 2.5 ms
// of a 27 ms step in it (0.76 ms forward + 1.73 ms backward for 411 MB in / 103 MB out).
// Arithmetic = torch's max_pool2d: scan the window rows then columns, take a value when
// `(val > max) || isnan(val)` (so the FIRST maximum wins a tie, which after a ReLU happens
// all the time), gradient routed to that one position.
//
// Forward: one thread = one output pixel x one 16-byte channel vector; it also writes which
// tap won (0..k*k-1, one byte per output element), so that the backward is a pure gather:
// one thread = one INPUT pixel x one channel vector, it looks at the <= 4 windows that cover
// it and sums the gradients of those whose recorded tap is this pixel (fp32 accumulation).
// Both are HBM-bound: forward reads H*W*C*es and writes Ho*Wo*C*(es+1) bytes, backward
// reads Ho*Wo*C*(es+1) and writes H*W*C*es.
#include <stdlib.h>

#include "common.cuh"

namespace {

using namespace b200dp;

struct PoolArgs {
  const void* in;    // forward: x; backward: dy
  void* out;         // forward: y; backward: dx
  uint8_t* idx;      // winning tap per output element
  int N, H, W, C, Ho, Wo, k, stride, pad;
  int vecs;          // 16-byte vectors per pixel
  long long items;   // forward: N*Ho*Wo*vecs; backward: N*H*W*vecs
};

template <typename T> struct PV;
template <> struct PV<float> {
  static constexpr int N = 4;
  static __device__ __forceinline__ void unpack(const uint4& v, float* f) {
    f[0] = __uint_as_float(v.x); f[1] = __uint_as_float(v.y); f[2] = __uint_as_float(v.z); f[3] = __uint_as_float(v.w);
  }
  static __device__ __forceinline__ uint4 pack(const float* f) {
    return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]), __float_as_uint(f[3]));
  }
};
template <> struct PV<__nv_bfloat16> {
  static constexpr int N = 8;
  static __device__ __forceinline__ void unpack(const uint4& v, float* f) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      f[2 * i] = __uint_as_float(w[i] << 16);
      f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
  static __device__ __forceinline__ uint4 pack(const float* f) {  // round to nearest even
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __nv_bfloat162 p = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
      w[i] = *reinterpret_cast<const uint32_t*>(&p);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
};

__device__ __forceinline__ uint4 ldg_nc16(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

template <typename T>
__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const PoolArgs a) {
  constexpr int NV = PV<T>::N;
  const T* x = static_cast<const T*>(a.in);
  T* y = static_cast<T*>(a.out);
  const long long stride_items = (long long)gridDim.x * 256;
  for (long long it = (long long)blockIdx.x * 256 + threadIdx.x; it < a.items; it += stride_items) {
    const int v = (int)(it % a.vecs);
    long long pix = it / a.vecs;
    const int ow = (int)(pix % a.Wo);
    pix /= a.Wo;
    const int oh = (int)(pix % a.Ho);
    const int n = (int)(pix / a.Ho);
    const int h0 = oh * a.stride - a.pad, w0 = ow * a.stride - a.pad;
    float m[NV];
    int best[NV];
    // torch starts from -inf at the first in-range position of the window
    const int first_tap = (h0 < 0 ? -h0 : 0) * a.k + (w0 < 0 ? -w0 : 0);
#pragma unroll
    for (int c = 0; c < NV; ++c) { m[c] = -INFINITY; best[c] = first_tap; }
    for (int i = 0; i < a.k; ++i) {
      const int h = h0 + i;
      if (h < 0 || h >= a.H) continue;
      for (int j = 0; j < a.k; ++j) {
        const int w = w0 + j;
        if (w < 0 || w >= a.W) continue;
        float f[NV];
        PV<T>::unpack(ldg_nc16(x + (((size_t)n * a.H + h) * a.W + w) * a.C + v * NV), f);
        const int tap = i * a.k + j;
#pragma unroll
        for (int c = 0; c < NV; ++c) {
          if (f[c] > m[c] || f[c] != f[c]) { m[c] = f[c]; best[c] = tap; }
        }
      }
    }
    const size_t o = (((size_t)n * a.Ho + oh) * a.Wo + ow) * a.C + v * NV;
    *reinterpret_cast<uint4*>(y + o) = PV<T>::pack(m);   // values are inputs: packing them back is exact
    if (NV == 8) {
      uint2 b;
      b.x = best[0] | (best[1] << 8) | (best[2] << 16) | (best[3] << 24);
      b.y = best[4] | (best[5] << 8) | (best[6] << 16) | (best[7] << 24);
      *reinterpret_cast<uint2*>(a.idx + o) = b;
    } else {
      *reinterpret_cast<uint32_t*>(a.idx + o) = best[0] | (best[1] << 8) | (best[2] << 16) | (best[3] << 24);
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const PoolArgs a) {
  constexpr int NV = PV<T>::N;
  const T* dy = static_cast<const T*>(a.in);
  T* dx = static_cast<T*>(a.out);
  const long long stride_items = (long long)gridDim.x * 256;
  for (long long it = (long long)blockIdx.x * 256 + threadIdx.x; it < a.items; it += stride_items) {
    const int v = (int)(it % a.vecs);
    long long pix = it / a.vecs;
    const int w = (int)(pix % a.W);
    pix /= a.W;
    const int h = (int)(pix % a.H);
    const int n = (int)(pix / a.H);
    // windows (oh, ow) with oh*stride - pad <= h <= oh*stride - pad + k - 1
    const int hp = h + a.pad, wp = w + a.pad;
    const int oh_lo = hp < a.k ? 0 : (hp - a.k) / a.stride + 1, oh_hi = min(hp / a.stride, a.Ho - 1);
    const int ow_lo = wp < a.k ? 0 : (wp - a.k) / a.stride + 1, ow_hi = min(wp / a.stride, a.Wo - 1);
    float g[NV];
#pragma unroll
    for (int c = 0; c < NV; ++c) g[c] = 0.f;
    for (int oh = oh_lo; oh <= oh_hi; ++oh) {
      for (int ow = ow_lo; ow <= ow_hi; ++ow) {
        const uint32_t mine = (uint32_t)((hp - oh * a.stride) * a.k + (wp - ow * a.stride));
        const size_t o = (((size_t)n * a.Ho + oh) * a.Wo + ow) * a.C + v * NV;
        uint32_t ib[2];
        if (NV == 8) {
          const uint2 b = *reinterpret_cast<const uint2*>(a.idx + o);
          ib[0] = b.x; ib[1] = b.y;
        } else {
          ib[0] = *reinterpret_cast<const uint32_t*>(a.idx + o); ib[1] = 0;
        }
        float f[NV];
        PV<T>::unpack(ldg_nc16(dy + o), f);
#pragma unroll
        for (int c = 0; c < NV; ++c)
          if (((ib[c >> 2] >> (8 * (c & 3))) & 255u) == mine) g[c] += f[c];
      }
    }
    *reinterpret_cast<uint4*>(dx + (((size_t)n * a.H + h) * a.W + w) * a.C + v * NV) = PV<T>::pack(g);
  }
}

// ---- bf16, 3x3 window, stride 2, padding 1 (the ResNet stem): packed bf16x2 arithmetic ------------
// Forward: the maximum of the 9 taps with HMNMX2 (NaN-propagating), then the FIRST tap equal to
// it, found by walking the taps from the last to the first with packed compares (== torch's
// `val > max` scan: the first maximum wins).  A window that holds a NaN is redone by the scalar
// rule (`|| isnan(val)`: the last NaN wins).  One block = one output row.
__device__ __forceinline__ uint32_t bf2_max_nan(uint32_t a, uint32_t b) {
  const __nv_bfloat162 r = __hmax2_nan(*reinterpret_cast<const __nv_bfloat162*>(&a), *reinterpret_cast<const __nv_bfloat162*>(&b));
  return *reinterpret_cast<const uint32_t*>(&r);
}
__device__ __forceinline__ uint32_t bf2_eq_mask(uint32_t a, uint32_t b) {
  return __heq2_mask(*reinterpret_cast<const __nv_bfloat162*>(&a), *reinterpret_cast<const __nv_bfloat162*>(&b));
}
__device__ __forceinline__ uint32_t bf2_nan_mask(uint32_t a) {  // 0xffff per lane that is NaN
  return __hneu2_mask(*reinterpret_cast<const __nv_bfloat162*>(&a), *reinterpret_cast<const __nv_bfloat162*>(&a));
}

__global__ void __launch_bounds__(256) maxpool_fwd_3x3s2_bf16_kernel(const PoolArgs a) {
  const __nv_bfloat16* x = static_cast<const __nv_bfloat16*>(a.in);
  __nv_bfloat16* y = static_cast<__nv_bfloat16*>(a.out);
  const int n = blockIdx.x / a.Ho, oh = blockIdx.x - n * a.Ho;
  const int h0 = oh * 2 - 1;
  const int row_items = a.Wo * a.vecs;
  constexpr uint32_t kNegInf2 = 0xff80ff80u;
  for (int it = threadIdx.x; it < row_items; it += 256) {
    const int ow = it / a.vecs, v = it - ow * a.vecs;
    const int w0 = ow * 2 - 1;
    uint4 t[9];
    bool ok[9];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        const int h = h0 + i, w = w0 + j;
        ok[i * 3 + j] = h >= 0 && h < a.H && w >= 0 && w < a.W;
        t[i * 3 + j] = ok[i * 3 + j] ? ldg_nc16(x + (((size_t)n * a.H + h) * a.W + w) * a.C + v * 8)
                                     : make_uint4(kNegInf2, kNegInf2, kNegInf2, kNegInf2);
      }
    }
    uint32_t m[4] = {t[0].x, t[0].y, t[0].z, t[0].w};
#pragma unroll
    for (int k = 1; k < 9; ++k) {
      m[0] = bf2_max_nan(m[0], t[k].x); m[1] = bf2_max_nan(m[1], t[k].y);
      m[2] = bf2_max_nan(m[2], t[k].z); m[3] = bf2_max_nan(m[3], t[k].w);
    }
    const uint32_t first_tap = (uint32_t)((h0 < 0 ? 3 : 0) + (w0 < 0 ? 1 : 0));
    uint32_t best[4];  // two 16-bit lanes per word
#pragma unroll
    for (int q = 0; q < 4; ++q) best[q] = first_tap * 0x00010001u;
#pragma unroll
    for (int k = 8; k >= 0; --k) {
      if (!ok[k]) continue;
      const uint32_t tw[4] = {t[k].x, t[k].y, t[k].z, t[k].w};
      const uint32_t kk = (uint32_t)k * 0x00010001u;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint32_t e = bf2_eq_mask(tw[q], m[q]);
        best[q] = (kk & e) | (best[q] & ~e);
      }
    }
    if ((bf2_nan_mask(m[0]) | bf2_nan_mask(m[1]) | bf2_nan_mask(m[2]) | bf2_nan_mask(m[3])) != 0u) {
      // torch's rule with NaNs: a NaN is always taken, so the last NaN of the scan wins
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        for (int lane = 0; lane < 2; ++lane) {
          float mv = -INFINITY;
          uint32_t bi = first_tap;
          for (int k = 0; k < 9; ++k) {
            if (!ok[k]) continue;
            const uint32_t tw = q == 0 ? t[k].x : q == 1 ? t[k].y : q == 2 ? t[k].z : t[k].w;
            const float f = __uint_as_float(lane ? (tw & 0xffff0000u) : (tw << 16));
            if (f > mv || f != f) { mv = f; bi = (uint32_t)k; }
          }
          const uint32_t sh = lane * 16;
          best[q] = (best[q] & ~(0xffffu << sh)) | (bi << sh);
          m[q] = (m[q] & ~(0xffffu << sh)) | ((__float_as_uint(mv) >> 16) << sh);
        }
      }
    }
    const size_t o = (((size_t)n * a.Ho + oh) * a.Wo + ow) * a.C + v * 8;
    *reinterpret_cast<uint4*>(y + o) = make_uint4(m[0], m[1], m[2], m[3]);
    uint2 b;
    b.x = __byte_perm(best[0], best[1], 0x6420);  // low byte of each 16-bit lane
    b.y = __byte_perm(best[2], best[3], 0x6420);
    *reinterpret_cast<uint2*>(a.idx + o) = b;
  }
}

// Backward: one thread = a 2 x 2 block of input pixels (rows 2a, 2a+1; columns 2b, 2b+1) x one channel
// vector.  Exactly four windows overlap such a block — (a, b), (a, b+1), (a+1, b), (a+1, b+1) — and
// each of its pixels is a fixed tap of each of them, so every window is loaded once per thread
// (a per-pixel gather visits 9 window/pixel pairs with 9 loads; here 9 pairs, 4 loads, all issued
// before any is used).  Four recorded taps are compared at a time (`__vcmpeq4`), the selected bf16
// gradients are accumulated in fp32 in ascending window order.
__device__ __forceinline__ void pool_acc(float* g, const uint2& ib, const uint4& d, uint32_t tap) {
  const uint32_t mine = tap * 0x01010101u;
  const uint32_t e0 = __vcmpeq4(ib.x, mine), e1 = __vcmpeq4(ib.y, mine);  // 0xff per matching channel
  const uint32_t dw[4] = {d.x, d.y, d.z, d.w};
  const uint32_t mk[4] = {__byte_perm(e0, 0, 0x1100), __byte_perm(e0, 0, 0x3322), __byte_perm(e1, 0, 0x1100),
                          __byte_perm(e1, 0, 0x3322)};
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const uint32_t sel = dw[q] & mk[q];
    g[2 * q] += __uint_as_float(sel << 16);
    g[2 * q + 1] += __uint_as_float(sel & 0xffff0000u);
  }
}

__global__ void __launch_bounds__(256) maxpool_bwd_3x3s2_bf16_kernel(const PoolArgs a) {
  const __nv_bfloat16* dy = static_cast<const __nv_bfloat16*>(a.in);
  __nv_bfloat16* dx = static_cast<__nv_bfloat16*>(a.out);
  const int hpairs = (a.H + 1) >> 1, wpairs = (a.W + 1) >> 1;
  const int pair_items = wpairs * a.vecs;
  const int chunks = (pair_items + 255) >> 8;
  const int rp = blockIdx.x / chunks;
  const int it = (blockIdx.x - rp * chunks) * 256 + threadIdx.x;
  if (it >= pair_items) return;
  const int n = rp / hpairs, pa = rp - n * hpairs;
  const int pb = it / a.vecs, v = it - pb * a.vecs;
  uint2 ib[4];
  uint4 d[4];
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    const int oh = pa + (s >> 1), ow = pb + (s & 1);
    const bool ok = oh < a.Ho && ow < a.Wo;
    const size_t o = (((size_t)n * a.Ho + (ok ? oh : 0)) * a.Wo + (ok ? ow : 0)) * a.C + v * 8;
    ib[s] = ok ? *reinterpret_cast<const uint2*>(a.idx + o) : make_uint2(0xffffffffu, 0xffffffffu);  // tap 255: no match
    d[s] = ok ? ldg_nc16(dy + o) : make_uint4(0u, 0u, 0u, 0u);
  }
  float g[4][8];  // [0] (2a, 2b), [1] (2a, 2b+1), [2] (2a+1, 2b), [3] (2a+1, 2b+1)
#pragma unroll
  for (int p = 0; p < 4; ++p)
#pragma unroll
    for (int c = 0; c < 8; ++c) g[p][c] = 0.f;
  // window (a, b): rows 2a-1..2a+1, columns 2b-1..2b+1
  pool_acc(g[0], ib[0], d[0], 4); pool_acc(g[1], ib[0], d[0], 5); pool_acc(g[2], ib[0], d[0], 7); pool_acc(g[3], ib[0], d[0], 8);
  // window (a, b+1): columns 2b+1..2b+3
  pool_acc(g[1], ib[1], d[1], 3); pool_acc(g[3], ib[1], d[1], 6);
  // window (a+1, b): rows 2a+1..2a+3
  pool_acc(g[2], ib[2], d[2], 1); pool_acc(g[3], ib[2], d[2], 2);
  // window (a+1, b+1)
  pool_acc(g[3], ib[3], d[3], 0);
  const int h = 2 * pa, w = 2 * pb;
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    const int hh = h + (p >> 1), ww = w + (p & 1);
    if (hh < a.H && ww < a.W)
      *reinterpret_cast<uint4*>(dx + (((size_t)n * a.H + hh) * a.W + ww) * a.C + v * 8) = PV<__nv_bfloat16>::pack(g[p]);
  }
}

int fill_args(PoolArgs& a, const void* in, void* out, const uint8_t* idx, int dtype, int N, int H, int W, int C,
              int k, int stride, int pad, bool forward) {
  if (N < 0 || H <= 0 || W <= 0 || C <= 0 || k < 1 || stride < 1 || pad < 0) return B200DP_EINVAL;
  if (dtype != B200DP_F32 && dtype != B200DP_BF16) return B200DP_EUNSUPPORTED;
  if (k > 3) return B200DP_EUNSUPPORTED;          // taps are recorded in one byte and looped, 3x3 is what the models use
  if (2 * pad > k) return B200DP_EINVAL;           // torch: pad should be at most half of the kernel size
  const int nv = dtype == B200DP_BF16 ? 8 : 4;
  if (C % nv) return B200DP_EUNSUPPORTED;
  if (H + 2 * pad < k || W + 2 * pad < k) return B200DP_EINVAL;
  a.N = N; a.H = H; a.W = W; a.C = C; a.k = k; a.stride = stride; a.pad = pad;
  a.Ho = (H + 2 * pad - k) / stride + 1;
  a.Wo = (W + 2 * pad - k) / stride + 1;
  a.vecs = C / nv;
  a.items = (long long)N * (forward ? (long long)a.Ho * a.Wo : (long long)H * W) * a.vecs;
  if (N == 0) return B200DP_OK;
  if (!in || !out || !idx) return B200DP_EINVAL;
  if (!aligned16(in) || !aligned16(out) || ((uintptr_t)idx & 7)) return B200DP_EALIGN;
  a.in = in; a.out = out; a.idx = const_cast<uint8_t*>(idx);
  return B200DP_OK;
}

// the packed kernels: bf16, 3x3 / 2 / 1, one block per (pair of) row(s);
// B200DP_POOL_GENERIC=1 keeps the generic kernels reachable for tests
bool stem_geometry(const PoolArgs& a, int dtype) {
  static const bool generic_only = [] { const char* e = getenv("B200DP_POOL_GENERIC"); return e && e[0] == '1'; }();
  return !generic_only && dtype == B200DP_BF16 && a.k == 3 && a.stride == 2 && a.pad == 1 &&
         (long long)a.N * a.H * ((a.W * a.vecs + 255) / 256) < 0x7fffffffLL;
}

int grid_for(long long items) {
  const long long want = (items + 255) / 256;
  const long long cap = (long long)B200DP_NUM_SMS * 8 * 4;  // a few waves of 8 CTAs per SM
  return (int)(want < cap ? (want < 1 ? 1 : want) : cap);
}

}  // namespace

extern "C" int b200dp_maxpool_fwd(const void* x, void* y, uint8_t* idx, int dtype, int N, int H, int W, int C,
                                  int k, int stride, int pad, b200dp_stream_t stream) {
  PoolArgs a;
  const int rc = fill_args(a, x, y, idx, dtype, N, H, W, C, k, stride, pad, true);
  if (rc || N == 0) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  if (stem_geometry(a, dtype)) maxpool_fwd_3x3s2_bf16_kernel<<<a.N * a.Ho, 256, 0, s>>>(a);
  else if (dtype == B200DP_BF16) maxpool_fwd_kernel<__nv_bfloat16><<<grid_for(a.items), 256, 0, s>>>(a);
  else maxpool_fwd_kernel<float><<<grid_for(a.items), 256, 0, s>>>(a);
  B200DP_LAUNCH_CHECK();
  return B200DP_OK;
}

extern "C" int b200dp_maxpool_bwd(const void* dy, const uint8_t* idx, void* dx, int dtype, int N, int H, int W,
                                  int C, int k, int stride, int pad, b200dp_stream_t stream) {
  PoolArgs a;
  const int rc = fill_args(a, dy, dx, idx, dtype, N, H, W, C, k, stride, pad, false);
  if (rc || N == 0) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  if (stem_geometry(a, dtype)) {
    const int chunks = (((a.W + 1) / 2) * a.vecs + 255) / 256;
    maxpool_bwd_3x3s2_bf16_kernel<<<a.N * ((a.H + 1) / 2) * chunks, 256, 0, s>>>(a);
  } else if (dtype == B200DP_BF16) maxpool_bwd_kernel<__nv_bfloat16><<<grid_for(a.items), 256, 0, s>>>(a);
  else maxpool_bwd_kernel<float><<<grid_for(a.items), 256, 0, s>>>(a);
  B200DP_LAUNCH_CHECK();
  return B200DP_OK;
}
