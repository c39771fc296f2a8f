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
  if (dtype == B200DP_BF16) maxpool_fwd_kernel<__nv_bfloat16><<<grid_for(a.items), 256, 0, s>>>(a);
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
  if (dtype == B200DP_BF16) maxpool_bwd_kernel<__nv_bfloat16><<<grid_for(a.items), 256, 0, s>>>(a);
  else maxpool_bwd_kernel<float><<<grid_for(a.items), 256, 0, s>>>(a);
  B200DP_LAUNCH_CHECK();
  return B200DP_OK;
}
