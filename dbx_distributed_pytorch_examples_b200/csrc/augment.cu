// augment.cu — uint8 NHWC -> random-crop / h-flip / normalize -> bf16 NHWC.
//
// One persistent kernel over (image, row-tile) tiles:
//   * the source rows of a tile are staged into shared memory with 16-byte
//     cp.async from the aligned-down address (double buffered, so the next
//     tile's loads are in flight while this tile is being written out);
//   * the normalisation is a lookup: there are only 3*256 possible results of
//     ((v/255) - mean[c]) / std[c], so they are computed once per CTA with the
//     same two IEEE fp32 operations torchvision performs and replicated 32x in
//     shared memory, one copy per bank, which makes every lookup conflict-free
//     whatever the pixel values are;
//   * every thread produces 8 consecutive bf16 outputs and stores them with one
//     128-bit store.  Crop offset, flip and zero padding are index arithmetic
//     on the staged bytes.
// Algorithmic traffic: 1 B read + 2 B written per output element.
#include "common.cuh"

namespace {

using namespace b200dp;

constexpr int kThreads = 1024;              // one CTA per SM
constexpr int kLutFloats = 3 * 256 * 32;       // 96 KiB
constexpr int kMaxRows = 64;                   // rows per tile upper bound
constexpr int kTileBytes = 48 * 1024;          // staged source bytes per buffer

struct AugArgs {
  const uint8_t* src;
  const int32_t* params;
  uint16_t* dst;
  int B, Hs, Ws, pad, H, W;
  int rows_per_tile;      // R
  int tiles_per_img;      // ceil(H / R)
  int pitch;              // bytes per staged row (multiple of 16)
  int chunks_per_row;     // W*3/8 (vector path) or W*3 (scalar path)
  uint32_t cpr_magic;     // floor(2^32 / chunks_per_row) + 1
  long long src_bytes;    // B*Hs*Ws*3
  float mean[3], std[3];
};

__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void* gptr, int src_size) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_addr), "l"(gptr),
               "r"(src_size)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

struct TileInfo {
  int b, y0, rows;   // image, first output row, number of rows
  int i, j, flip;    // crop params of the image
};

__device__ __forceinline__ TileInfo tile_info(const AugArgs& a, int tile) {
  TileInfo t;
  t.b = tile / a.tiles_per_img;
  int rt = tile - t.b * a.tiles_per_img;
  t.y0 = rt * a.rows_per_tile;
  t.rows = min(a.rows_per_tile, a.H - t.y0);
  // clamped, so that bad parameters cannot send the staging out of bounds
  t.i = min(max(__ldg(a.params + 3 * t.b + 0), 0), a.Hs + 2 * a.pad - a.H);
  t.j = min(max(__ldg(a.params + 3 * t.b + 1), 0), a.Ws + 2 * a.pad - a.W);
  t.flip = __ldg(a.params + 3 * t.b + 2) != 0;
  return t;
}

// Stage the source bytes of one tile.  rowshift[r] = offset to add to a byte
// index inside the SOURCE ROW (column*3 + channel) to get the byte's offset in
// the staged row, or INT_MIN when the whole row is padding.
__device__ __forceinline__ void stage_tile(const AugArgs& a, const TileInfo& t, uint8_t* buf,
                                           int* rowshift) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int sx0 = t.j - a.pad;
  const int lo = max(sx0, 0) * 3;
  const int hi = min(sx0 + a.W, a.Ws) * 3;
  for (int r = warp; r < t.rows; r += kThreads / 32) {
    const int sy = t.i + t.y0 + r - a.pad;
    const bool valid = (sy >= 0) && (sy < a.Hs) && (hi > lo);
    if (!valid) {
      if (lane == 0) rowshift[r] = INT_MIN;
      continue;
    }
    const long long row_off = ((long long)t.b * a.Hs + sy) * (long long)a.Ws * 3;
    const long long g0 = row_off + lo;        // first needed byte (offset from src)
    const long long g0a = g0 & ~15LL;         // aligned down (src itself is 16B aligned)
    const int nchunks = (int)((row_off + hi - g0a + 15) >> 4);
    if (lane == 0) rowshift[r] = (int)(g0 - g0a) - lo;
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(buf + r * a.pitch);
    for (int k = lane; k < nchunks; k += 32) {
      const long long ga = g0a + 16LL * k;
      const long long left = a.src_bytes - ga;
      cp_async16(sbase + 16 * k, a.src + ga, left >= 16 ? 16 : (int)left);
    }
  }
}

template <bool PAD>
__device__ __forceinline__ void emit_tile_vec8(const AugArgs& a, const TileInfo& t,
                                               const uint8_t* buf, const int* rowshift,
                                               const float* lut_lane) {
  const int cpr = a.chunks_per_row;
  const int nchunks = t.rows * cpr;
  const int sx0 = t.j - a.pad;
  uint16_t* out_img = a.dst + ((long long)t.b * a.H + t.y0) * (long long)a.W * 3;
  for (int idx = threadIdx.x; idx < nchunks; idx += kThreads) {
    const int r = (int)__umulhi((uint32_t)idx, a.cpr_magic);
    const int cc = idx - r * cpr;
    const int e0 = cc * 8;
    const int x0 = e0 / 3;
    const int c0 = e0 - 3 * x0;
    const int shift = rowshift[r];
    const bool rowok = (shift != INT_MIN);
    // source pixel of output pixel x0, and the per-pixel step
    const int sxs = sx0 + (t.flip ? (a.W - 1 - x0) : x0);
    const int step = t.flip ? -1 : 1;
    const uint8_t* rowp = buf + r * a.pitch + (rowok ? shift : 0);
    // LUT bases of the three channels in the order this chunk meets them
    const float* lp[3];
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      int c = c0 + q;
      c = c >= 3 ? c - 3 : c;
      lp[q] = lut_lane + c * (256 * 32);
    }
    float f[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int pos = c0 + k;                 // 0..9
      const int dxp = (pos * 11) >> 5;        // pos / 3 for pos < 32
      const int c = pos - 3 * dxp;            // absolute channel is (c0+k)%3 == c
      const int sx = sxs + step * dxp;
      int v;
      if (PAD) {
        const bool ok = rowok && (sx >= 0) && (sx < a.Ws);
        v = ok ? (int)rowp[sx * 3 + c] : 0;
      } else {
        v = (int)rowp[sx * 3 + c];
      }
      f[k] = lp[k % 3][v * 32];
    }
    uint4 o = make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]),
                         pack_bf16(f[6], f[7]));
    st_stream(out_img + (long long)r * a.W * 3 + e0, o);
  }
}

// Scalar path for widths where a row is not a whole number of 8-element chunks.
template <bool PAD>
__device__ __forceinline__ void emit_tile_scalar(const AugArgs& a, const TileInfo& t,
                                                 const uint8_t* buf, const int* rowshift,
                                                 const float* lut_lane) {
  const int epr = a.chunks_per_row;  // elements per row
  const int n = t.rows * epr;
  const int sx0 = t.j - a.pad;
  uint16_t* out_img = a.dst + ((long long)t.b * a.H + t.y0) * (long long)a.W * 3;
  for (int idx = threadIdx.x; idx < n; idx += kThreads) {
    const int r = (int)__umulhi((uint32_t)idx, a.cpr_magic);
    const int e = idx - r * epr;
    const int x = e / 3, c = e - 3 * x;
    const int shift = rowshift[r];
    const int sx = sx0 + (t.flip ? (a.W - 1 - x) : x);
    const bool ok = (shift != INT_MIN) && (sx >= 0) && (sx < a.Ws);
    const int v = ok ? (int)buf[r * a.pitch + shift + sx * 3 + c] : 0;
    __nv_bfloat16 h = __float2bfloat16_rn(lut_lane[c * (256 * 32) + v * 32]);
    out_img[(long long)r * epr + e] = *reinterpret_cast<uint16_t*>(&h);
  }
}

template <bool VEC8, bool PAD>
__global__ void __launch_bounds__(kThreads, 1) augment_kernel(const AugArgs a, int total_tiles) {
  extern __shared__ __align__(16) uint8_t smem[];
  float* lut = reinterpret_cast<float*>(smem);                       // [3][256][32]
  uint8_t* tilebuf = smem + kLutFloats * sizeof(float);              // [2][kTileBytes]
  int* rowshift = reinterpret_cast<int*>(tilebuf + 2 * kTileBytes);  // [2][kMaxRows]
  float* lut768 = reinterpret_cast<float*>(rowshift + 2 * kMaxRows); // [768]

  int tile = blockIdx.x;
  if (tile >= total_tiles) return;
  TileInfo cur = tile_info(a, tile);
  stage_tile(a, cur, tilebuf, rowshift);
  cp_async_commit();

  // Build the table while the first loads are in flight: torchvision's
  // ToTensor (v / 255) and Normalize ((x - mean) / std), both IEEE fp32.
  for (int e = threadIdx.x; e < 768; e += kThreads) {
    const int c = e >> 8, v = e & 255;
    const float x = __fdiv_rn((float)v, 255.0f);
    lut768[e] = __fdiv_rn(__fsub_rn(x, a.mean[c]), a.std[c]);
  }
  __syncthreads();
  for (int e = threadIdx.x; e < kLutFloats; e += kThreads) lut[e] = lut768[e >> 5];
  const float* lut_lane = lut + (threadIdx.x & 31);

  int p = 0;
  while (true) {
    const int next = tile + gridDim.x;
    TileInfo nxt;
    if (next < total_tiles) {
      nxt = tile_info(a, next);
      stage_tile(a, nxt, tilebuf + (p ^ 1) * kTileBytes, rowshift + (p ^ 1) * kMaxRows);
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();  // staged bytes, rowshift (and, first time round, the LUT) visible
    if (VEC8)
      emit_tile_vec8<PAD>(a, cur, tilebuf + p * kTileBytes, rowshift + p * kMaxRows, lut_lane);
    else
      emit_tile_scalar<PAD>(a, cur, tilebuf + p * kTileBytes, rowshift + p * kMaxRows, lut_lane);
    if (next >= total_tiles) break;
    __syncthreads();  // everyone is done with buffer p before it is refilled
    tile = next;
    cur = nxt;
    p ^= 1;
  }
}

constexpr size_t kSmemBytes =
    kLutFloats * sizeof(float) + 2 * kTileBytes + 2 * kMaxRows * sizeof(int) + 768 * sizeof(float);

template <bool VEC8, bool PAD>
int launch(const AugArgs& a, int total_tiles, cudaStream_t s) {
  B200DP_CUDA_TRY(cudaFuncSetAttribute(augment_kernel<VEC8, PAD>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)kSmemBytes));
  int grid = total_tiles < B200DP_NUM_SMS ? total_tiles : B200DP_NUM_SMS;
  augment_kernel<VEC8, PAD><<<grid, kThreads, kSmemBytes, s>>>(a, total_tiles);
  B200DP_LAUNCH_CHECK();
  return B200DP_OK;
}

}  // namespace

extern "C" int b200dp_augment_u8_bf16(const uint8_t* src, int B, int Hs, int Ws,
                                      const int32_t* params, int pad, int H, int W,
                                      const float* mean, const float* std, uint16_t* dst,
                                      b200dp_stream_t stream) {
  if (B < 0 || Hs <= 0 || Ws <= 0 || H <= 0 || W <= 0 || pad < 0) return B200DP_EINVAL;
  if (B == 0) return B200DP_OK;
  if (!src || !params || !dst || !mean || !std) return B200DP_EINVAL;
  if (H > Hs + 2 * pad || W > Ws + 2 * pad) return B200DP_EINVAL;
  if (!b200dp::aligned16(src) || !b200dp::aligned16(dst)) return B200DP_EALIGN;
  for (int c = 0; c < 3; ++c)
    if (!(std[c] != 0.0f)) return B200DP_EINVAL;

  AugArgs a;
  a.src = src;
  a.params = params;
  a.dst = dst;
  a.B = B; a.Hs = Hs; a.Ws = Ws; a.pad = pad; a.H = H; a.W = W;
  a.src_bytes = (long long)B * Hs * Ws * 3;
  for (int c = 0; c < 3; ++c) { a.mean[c] = mean[c]; a.std[c] = std[c]; }
  // a staged row holds at most the W*3 needed bytes plus 15 of alignment slack
  a.pitch = ((W * 3 + 15 + 15) / 16) * 16;
  if (a.pitch > kTileBytes) return B200DP_EUNSUPPORTED;  // rows wider than ~4000 pixels
  int R = kTileBytes / a.pitch;
  if (R > kMaxRows) R = kMaxRows;
  if (R > H) R = H;
  a.tiles_per_img = (H + R - 1) / R;
  a.rows_per_tile = (H + a.tiles_per_img - 1) / a.tiles_per_img;  // equal-sized tiles
  const bool vec8 = ((W * 3) % 8) == 0;
  a.chunks_per_row = vec8 ? (W * 3) / 8 : W * 3;
  a.cpr_magic = (uint32_t)(0x100000000ULL / (unsigned)a.chunks_per_row) + 1u;
  const long long total = (long long)B * a.tiles_per_img;
  if (total > 0x7fffffffLL) return B200DP_ETOOBIG;
  cudaStream_t s = (cudaStream_t)stream;
  if (vec8) return pad ? launch<true, true>(a, (int)total, s) : launch<true, false>(a, (int)total, s);
  return pad ? launch<false, true>(a, (int)total, s) : launch<false, false>(a, (int)total, s);
}
