// resample.cu — uint8 NHWC -> crop box -> Pillow-exact bilinear resize -> h-flip -> normalize -> bf16 NHWC.
//
// What the reference does per sample on the CPU with PIL images: transforms.Resize((S, S))
// (utils/hf_dataset_utilities.py:64-66) and transforms.RandomResizedCrop(224)
// (02_deepspeed/03_1k_imagenet_deepspeed_resnet.py:45-53), i.e. img.crop + img.resize(BILINEAR):
// Pillow's two-pass resampler with 22-bit fixed-point coefficients and an 8-bit
// intermediate image (Pillow 12.2 src/libImaging/Resample.c).  The result has to be
// bit-exact, so this is integer work laid out the way Pillow does it:
//   * coefficients: the bilinear weights are evaluated in IEEE double with explicitly
//     rounded operations (no FMA contraction), normalised, and rounded to int32 << 22;
//   * horizontal pass first, over exactly the source rows the vertical pass of this tile
//     touches, rounded and clipped to uint8 into shared memory (the flip is applied here,
//     as a mirrored column index);
//   * vertical pass from shared memory, 8 bytes of a row per thread, then the
//     ((v/255) - mean) / std normalisation as a 3x256 lookup of bf16 results computed with
//     the same fp32 operations torchvision performs, one 16-byte store per thread.
// One CTA = one tile of output rows of one image.  Algorithmic traffic per image:
// h*w*3 bytes of the crop box read + H*W*3*2 bytes written.
#include <math.h>

#include "common.cuh"

namespace {

using namespace b200dp;

constexpr int kThreads = 256;
constexpr int kPrecisionBits = 32 - 8 - 2;
constexpr int kTileRows = 32;
constexpr int kSmemBudget = 160 * 1024;

struct ResampleArgs {
  const uint8_t* src;
  const int32_t* params;  // [B,5] = i, j, h, w, flip
  uint16_t* dst;
  int Hs, Ws, H, W;
  int tile_rows;   // T: output rows per CTA
  int temp_rows;   // R: rows of the 8-bit intermediate a tile can need
  int pitch;       // bytes per intermediate row (W*3 rounded up to 16)
  int kh, kv;      // coefficient strides (max taps) of the two passes
  float mean[3], std[3];
};

// Resample.c precompute_coeffs + normalize_coeffs_8bpc for output index xx: taps into k[0..kstride),
// returns (xmin, xmax) through the pointers.  in0 = 0, in1 = in_size (the crop is resized whole).
__device__ __forceinline__ void coeffs_for(int xx, int in_size, int out_size, int kstride, int32_t* k,
                                           int* xmin_out, int* xmax_out) {
  const double scale = __ddiv_rn((double)(float)in_size, (double)out_size);
  const double filterscale = scale < 1.0 ? 1.0 : scale;
  const double support = filterscale;  // bilinear: 1.0 * filterscale
  const double ss = __ddiv_rn(1.0, filterscale);
  const double center = __dmul_rn((double)xx + 0.5, scale);  // in0 + ..., in0 = 0
  int xmin = (int)__dadd_rn(__dsub_rn(center, support), 0.5);
  if (xmin < 0) xmin = 0;
  int xmax = (int)__dadd_rn(__dadd_rn(center, support), 0.5);
  if (xmax > in_size) xmax = in_size;
  xmax -= xmin;
  if (xmax > kstride) xmax = kstride;  // cannot happen for a box inside the slab
  double ww = 0.0;
  for (int x = 0; x < xmax; ++x) {
    double a = __dmul_rn(__dadd_rn(__dsub_rn((double)(x + xmin), center), 0.5), ss);
    if (a < 0.0) a = -a;
    ww = __dadd_rn(ww, a < 1.0 ? __dsub_rn(1.0, a) : 0.0);
  }
  for (int x = 0; x < kstride; ++x) {
    double w = 0.0;
    if (x < xmax) {
      double a = __dmul_rn(__dadd_rn(__dsub_rn((double)(x + xmin), center), 0.5), ss);
      if (a < 0.0) a = -a;
      w = a < 1.0 ? __dsub_rn(1.0, a) : 0.0;
      if (ww != 0.0) w = __ddiv_rn(w, ww);
    }
    // bilinear weights are never negative: the +0.5 branch of normalize_coeffs_8bpc
    k[x] = (int)__dadd_rn(0.5, __dmul_rn(w, (double)(1 << kPrecisionBits)));
  }
  *xmin_out = xmin;
  *xmax_out = xmax;
}

__device__ __forceinline__ int clip8(int acc) {
  return min(max(acc >> kPrecisionBits, 0), 255);
}

// Horizontal pass of one output column over the tile's source rows r0, r0+rstep, …, with the
// column's TAPS coefficients and byte offsets in registers.  Coefficients past xmax are zero
// (coeffs_for pads them) and read the column's first pixel again, which is always inside the box.
// No clip: the weights are non-negative and sum to 2^22 (+ at most a few units of rounding), so
// 2^21 + sum(p*k) < 256 * 2^22 and the shift alone yields 0..255, as Pillow's clip8 would.
template <int TAPS>
__device__ __forceinline__ void hpass_column(const uint8_t* __restrict__ img, int row_stride, int nrows, int r0,
                                             int rstep, const int32_t* k, int xmin, int xmax, uint8_t* tcol,
                                             int pitch) {
  int kk[TAPS];
  int off[TAPS];
#pragma unroll
  for (int x = 0; x < TAPS; ++x) {
    kk[x] = k[x];
    off[x] = (xmin + (x < xmax ? x : 0)) * 3;
  }
  const uint8_t* row = img + (size_t)r0 * row_stride;
  uint8_t* t = tcol + r0 * pitch;
  for (int r = r0; r < nrows; r += rstep) {
    int s0 = 1 << (kPrecisionBits - 1), s1 = s0, s2 = s0;
#pragma unroll
    for (int x = 0; x < TAPS; ++x) {
      const uint8_t* px = row + off[x];
      s0 += (int)__ldg(px) * kk[x];
      s1 += (int)__ldg(px + 1) * kk[x];
      s2 += (int)__ldg(px + 2) * kk[x];
    }
    t[0] = (uint8_t)(s0 >> kPrecisionBits);
    t[1] = (uint8_t)(s1 >> kPrecisionBits);
    t[2] = (uint8_t)(s2 >> kPrecisionBits);
    row += (size_t)rstep * row_stride;
    t += rstep * pitch;
  }
}

// Any number of taps, coefficients read from shared memory (down-scaling by more than 2.5x).
__device__ __forceinline__ void hpass_column_any(const uint8_t* __restrict__ img, int row_stride, int nrows, int r0,
                                                 int rstep, const int32_t* k, int xmin, int xmax, uint8_t* tcol,
                                                 int pitch) {
  for (int r = r0; r < nrows; r += rstep) {
    const uint8_t* px = img + (size_t)r * row_stride + xmin * 3;
    int s0 = 1 << (kPrecisionBits - 1), s1 = s0, s2 = s0;
    for (int x = 0; x < xmax; ++x) {
      const int kx = k[x];
      s0 += (int)__ldg(px + 3 * x) * kx;
      s1 += (int)__ldg(px + 3 * x + 1) * kx;
      s2 += (int)__ldg(px + 3 * x + 2) * kx;
    }
    uint8_t* t = tcol + r * pitch;
    t[0] = (uint8_t)clip8(s0);
    t[1] = (uint8_t)clip8(s1);
    t[2] = (uint8_t)clip8(s2);
  }
}

template <bool VEC>
__global__ void __launch_bounds__(kThreads, 5) resample_kernel(const ResampleArgs a) {
  extern __shared__ __align__(16) uint8_t smem[];
  // layout: temp [R][pitch] | kh coef [W][kh] | kv coef [T][kv] | bounds h [W][2] | bounds v [T][2] | lut [3][256]
  uint8_t* temp = smem;
  int32_t* coef_h = reinterpret_cast<int32_t*>(temp + (size_t)a.temp_rows * a.pitch);
  int32_t* coef_v = coef_h + a.W * a.kh;
  int* bnd_h = coef_v + a.tile_rows * a.kv;
  int* bnd_v = bnd_h + 2 * a.W;
  uint16_t* lut = reinterpret_cast<uint16_t*>(bnd_v + 2 * a.tile_rows);
  __shared__ int s_taps_h;

  const int b = blockIdx.y;
  const int y0 = blockIdx.x * a.tile_rows;
  const int rows = min(a.tile_rows, a.H - y0);
  const int tid = threadIdx.x;
  if (tid == 0) s_taps_h = 0;

  // the crop box, forced inside the slab (torchvision never produces anything else)
  const int32_t* p = a.params + (size_t)b * 5;
  int bh = min(max(p[2], 1), a.Hs), bw = min(max(p[3], 1), a.Ws);
  const int bi = min(max(p[0], 0), a.Hs - bh), bj = min(max(p[1], 0), a.Ws - bw);
  const bool flip = p[4] != 0;
  __syncthreads();

  int my_taps = 0;
  for (int c = tid; c < a.W; c += kThreads) {
    coeffs_for(c, bw, a.W, a.kh, coef_h + c * a.kh, &bnd_h[2 * c], &bnd_h[2 * c + 1]);
    my_taps = max(my_taps, bnd_h[2 * c + 1]);
  }
  for (int y = tid; y < rows; y += kThreads)
    coeffs_for(y0 + y, bh, a.H, a.kv, coef_v + y * a.kv, &bnd_v[2 * y], &bnd_v[2 * y + 1]);
  for (int e = tid; e < 768; e += kThreads) {
    const int c = e >> 8;
    const float x = __fdiv_rn((float)(e & 255), 255.0f);                 // ToTensor: .div(255)
    const float y = __fdiv_rn(__fsub_rn(x, a.mean[c]), a.std[c]);         // Normalize: sub_(mean).div_(std)
    lut[e] = __bfloat16_as_ushort(__float2bfloat16_rn(y));
  }
  my_taps = __reduce_max_sync(0xffffffffu, my_taps);
  if ((tid & 31) == 0 && my_taps) atomicMax(&s_taps_h, my_taps);
  __syncthreads();

  // ---- horizontal pass over the source rows this tile's vertical pass touches --------------
  // A thread owns one output column (and every rgroups-th row when the image is narrow).
  const int first = bnd_v[0];
  int nrows = bnd_v[2 * (rows - 1)] + bnd_v[2 * (rows - 1) + 1] - first;
  nrows = min(nrows, a.temp_rows);
  const uint8_t* img = a.src + ((size_t)b * a.Hs + (bi + first)) * a.Ws * 3 + (size_t)bj * 3;
  {
    const int cw = min((a.W + 31) & ~31, kThreads);  // columns per pass, whole warps
    const int rgroups = kThreads / cw, rg = tid / cw;
    const int taps = s_taps_h, row_stride = a.Ws * 3;
    if (rg < rgroups) {
      for (int c = tid - rg * cw; c < a.W; c += cw) {
        const int xmin = bnd_h[2 * c], xmax = bnd_h[2 * c + 1];
        const int32_t* k = coef_h + c * a.kh;
        uint8_t* tcol = temp + (flip ? a.W - 1 - c : c) * 3;
        if (taps <= 3) hpass_column<3>(img, row_stride, nrows, rg, rgroups, k, xmin, xmax, tcol, a.pitch);
        else if (taps == 4) hpass_column<4>(img, row_stride, nrows, rg, rgroups, k, xmin, xmax, tcol, a.pitch);
        else if (taps == 5) hpass_column<5>(img, row_stride, nrows, rg, rgroups, k, xmin, xmax, tcol, a.pitch);
        else hpass_column_any(img, row_stride, nrows, rg, rgroups, k, xmin, xmax, tcol, a.pitch);
      }
    }
  }
  __syncthreads();

  // ---- vertical pass + normalisation ---------------------------------------------------------
  // A thread owns one group of 8 consecutive bytes of the row (and every vslots-th output row).
  const int row_elems = a.W * 3;
  const int groups = (row_elems + 7) >> 3;
  const int gw = min(groups, kThreads);
  const int vslots = kThreads / gw, vs = tid / gw;
  if (vs < vslots) {
    for (int g = tid - vs * gw; g < groups; g += gw) {
      // channel of element e of this group is (2g + e) % 3: three table bases per thread
      const int ch0 = (2 * g) % 3;
      const uint16_t* lt0 = lut + ch0 * 256;
      const uint16_t* lt1 = lut + (ch0 == 2 ? 0 : ch0 + 1) * 256;
      const uint16_t* lt2 = lut + (ch0 == 0 ? 2 : ch0 - 1) * 256;
      for (int y = vs; y < rows; y += vslots) {
        const int ymin = bnd_v[2 * y] - first, ymax = bnd_v[2 * y + 1];
        const int32_t* k = coef_v + y * a.kv;
        int acc[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = 1 << (kPrecisionBits - 1);
        const uint8_t* col = temp + ymin * a.pitch + g * 8;
        for (int x = 0; x < ymax; ++x) {
          const uint2 v = *reinterpret_cast<const uint2*>(col + x * a.pitch);
          const int kx = k[x];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            acc[e] += (int)__byte_perm(v.x, 0, 0x4440 | e) * kx;
            acc[4 + e] += (int)__byte_perm(v.y, 0, 0x4440 | e) * kx;
          }
        }
        uint32_t o[8];
        // no clip, as in the horizontal pass: the shift alone yields 0..255
        o[0] = lt0[acc[0] >> kPrecisionBits]; o[1] = lt1[acc[1] >> kPrecisionBits]; o[2] = lt2[acc[2] >> kPrecisionBits];
        o[3] = lt0[acc[3] >> kPrecisionBits]; o[4] = lt1[acc[4] >> kPrecisionBits]; o[5] = lt2[acc[5] >> kPrecisionBits];
        o[6] = lt0[acc[6] >> kPrecisionBits]; o[7] = lt1[acc[7] >> kPrecisionBits];
        uint16_t* out = a.dst + ((size_t)b * a.H + (y0 + y)) * row_elems + g * 8;
        if (VEC) {
          uint4 w;
          w.x = o[0] | (o[1] << 16);
          w.y = o[2] | (o[3] << 16);
          w.z = o[4] | (o[5] << 16);
          w.w = o[6] | (o[7] << 16);
          *reinterpret_cast<uint4*>(out) = w;
        } else {
#pragma unroll
          for (int e = 0; e < 8; ++e)
            if (g * 8 + e < row_elems) out[e] = (uint16_t)o[e];
        }
      }
    }
  }
}

}  // namespace

// Largest number of taps precompute_coeffs can produce for in_size -> out_size.
static int max_taps(int in_size, int out_size) {
  double scale = (double)in_size / (double)out_size;
  if (scale < 1.0) scale = 1.0;
  return (int)ceil(scale) * 2 + 1;
}

extern "C" int b200dp_resized_crop_u8_bf16(const uint8_t* src, int B, int Hs, int Ws,
                                           const int32_t* params, int H, int W,
                                           const float* mean, const float* std, uint16_t* dst,
                                           b200dp_stream_t stream) {
  if (B < 0 || Hs <= 0 || Ws <= 0 || H <= 0 || W <= 0) return B200DP_EINVAL;
  if (B == 0) return B200DP_OK;
  if (!src || !params || !dst || !mean || !std) return B200DP_EINVAL;
  if (!b200dp::aligned16(dst)) return B200DP_EALIGN;
  if (B > 65535) return B200DP_ETOOBIG;  // gridDim.y
  if ((long long)Hs * Ws * 3 > 0x7fffffffLL || (long long)H * W * 3 > 0x7fffffffLL) return B200DP_ETOOBIG;
  for (int c = 0; c < 3; ++c)
    if (!(std[c] != 0.0f)) return B200DP_EINVAL;

  ResampleArgs a;
  a.src = src; a.params = params; a.dst = dst;
  a.Hs = Hs; a.Ws = Ws; a.H = H; a.W = W;
  for (int c = 0; c < 3; ++c) { a.mean[c] = mean[c]; a.std[c] = std[c]; }
  a.pitch = ((W * 3 + 15) / 16) * 16;
  a.kh = max_taps(Ws, W);  // a box is never wider / taller than the slab
  a.kv = max_taps(Hs, H);
  // A tile of T output rows reads source rows [xmin(first), xmin(last) + xmax(last)), at most
  // T*scale + 2*support + 2 of them with scale <= Hs/H: shrink T until the 8-bit intermediate
  // fits the shared-memory budget.
  const double sv = (double)Hs / (double)H;
  const double support = sv < 1.0 ? 1.0 : sv;
  size_t smem = 0;
  int T = kTileRows < H ? kTileRows : H;
  for (;; T = (T + 1) / 2) {
    a.tile_rows = T;
    a.temp_rows = (int)floor(T * sv + 2.0 * support) + 3;
    if (a.temp_rows > Hs) a.temp_rows = Hs;
    smem = (size_t)a.temp_rows * a.pitch + sizeof(int32_t) * ((size_t)W * a.kh + (size_t)T * a.kv) +
           sizeof(int) * 2 * ((size_t)W + T) + 768 * sizeof(uint16_t);
    if (smem <= (size_t)kSmemBudget) break;
    if (T == 1) return B200DP_EUNSUPPORTED;  // a single output row needs more than the budget
  }
  const bool vec = (W * 3) % 8 == 0;
  auto kern = vec ? resample_kernel<true> : resample_kernel<false>;
  B200DP_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBudget));
  B200DP_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
  dim3 grid((H + T - 1) / T, B);
  kern<<<grid, kThreads, smem, (cudaStream_t)stream>>>(a);
  B200DP_LAUNCH_CHECK();
  return B200DP_OK;
}
