/*
 * b200dp.h — C ABI of libb200dp.so, the sm_100a data-parallel hot path.
 *
 * Every entry point takes plain pointers and sizes (no torch types), enqueues
 * its work on the CUDA stream it is given and returns 0 on success, a negative
 * B200DP_E* code for a contract violation, or a positive cudaError_t value
 * when the CUDA runtime reports one.  Nothing throws across this boundary and
 * nothing in here has a CPU fallback: without a Blackwell GPU the calls fail.
 *
 * The reference (alexxx-db/dbx-distributed-pytorch-examples) has no FFI layer
 * of its own; each declaration below cites the reference call site (relative
 * to the reference root) or the torch routine (TORCH/ = site-packages/torch,
 * TV/ = site-packages/torchvision) whose work it takes over.  See
 * INTEGRATION.md for the ctypes binding a maintainer adds on the Python side.
 *
 * Pointers are DEVICE pointers unless stated otherwise.  `stream` is a
 * cudaStream_t passed as void*.  dtype codes: B200DP_F32 / B200DP_BF16.
 */
#ifndef B200DP_H_
#define B200DP_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200DP_VERSION 100

/* element types of activation / gradient buffers */
#define B200DP_F32 0
#define B200DP_BF16 1

/* optimizer kinds for the fused update */
#define B200DP_OPT_NONE 0 /* only average the gradients */
#define B200DP_OPT_SGD 1  /* TORCH/optim/sgd.py:322-380  */
#define B200DP_OPT_ADAM 2 /* TORCH/optim/adam.py:347+ (coupled L2 weight decay) */

/* error codes (negative); positive values are cudaError_t */
#define B200DP_OK 0
#define B200DP_EINVAL (-1)      /* bad argument (null pointer, negative size, …)   */
#define B200DP_EUNSUPPORTED (-2) /* shape / dtype outside what the kernels handle   */
#define B200DP_EALIGN (-3)      /* pointer not aligned as the entry point requires */
#define B200DP_ESTATE (-4)      /* comm not ready (peers not opened, bad rank, …)  */
#define B200DP_ETOOBIG (-5)     /* size beyond a documented fixed limit            */

#define B200DP_MAX_WORLD 8
#define B200DP_BN_MAX_CHANNELS 4096

typedef void* b200dp_stream_t;
typedef struct b200dp_comm b200dp_comm; /* opaque: IPC-mapped peer buffers + flags */

int b200dp_version(void);
/* Number of CUDA kernels this library has launched in this process so far (all entry points, all
 * threads).  bench.py reports the difference over its timed region as `gpu_launches`. */
unsigned long long b200dp_kernel_launches(void);
/* Static string for a code returned by any entry point (never NULL). */
const char* b200dp_error_string(int code);

/* ------------------------------------------------------------------------
 * (3) Input transform: uint8 NHWC -> random crop / h-flip / normalize -> bf16 NHWC
 *
 * Replaces, for a whole batch in one launch, the per-sample CPU pipeline of
 *   utils/hf_dataset_utilities.py:46-53,64-77 (CustomDataset.__getitem__ +
 *   default_image_transforms), the CIFAR recipe
 *   04_accelerate/01_cifar_accelerate.ipynb:295-298 (RandomCrop(32,padding=4),
 *   RandomHorizontalFlip, ToTensor, Normalize) and the crop/flip/normalize tail
 *   of 02_deepspeed/03_1k_imagenet_deepspeed_resnet.py:45-53, followed by the
 *   H2D copy at 01_torch_distributor/01_basic_torch_distributor.py:225.
 * Arithmetic per pixel is torchvision's: x/255 (TV/transforms/functional.py:
 * 174-176) then (x-mean)/std in fp32 (TV/transforms/_functional_tensor.py:
 * 921-928), rounded once to bf16.
 *
 *  src     uint8 [B, Hs, Ws, 3]; device memory OR pinned (mapped) host memory
 *  params  int32 [B, 3] = (top i, left j, flip) per sample, in the coordinates
 *          of the zero-padded image (torchvision RandomCrop.get_params,
 *          TV/transforms/transforms.py:653-654,676-677); device or pinned host
 *  pad     zero padding on each side before the crop (0 for none)
 *  dst     bf16 [B, H, W, 3] (the memory of a [B,3,H,W] channels_last tensor)
 *  mean,std  HOST pointers to 3 floats each
 * Requires 0 <= i <= Hs+2*pad-H and 0 <= j <= Ws+2*pad-W (not checked on the
 * device; out-of-range sources read as padding).
 * ---------------------------------------------------------------------- */
int b200dp_augment_u8_bf16(const uint8_t* src, int B, int Hs, int Ws,
                           const int32_t* params, int pad, int H, int W,
                           const float* mean, const float* std, uint16_t* dst,
                           b200dp_stream_t stream);

/* (3b) Input transform with a resize: uint8 NHWC -> crop box -> bilinear resize ->
 *      h-flip -> normalize -> bf16 NHWC, one launch per batch.
 *
 * Replaces the per-sample PIL path of transforms.Resize((S, S)) + RandomHorizontalFlip +
 * ToTensor + Normalize (utils/hf_dataset_utilities.py:64-79) and of
 * transforms.RandomResizedCrop(224) + ... (02_deepspeed/03_1k_imagenet_deepspeed_resnet.py:
 * 45-53).  For PIL images torchvision's resized_crop is img.crop + img.resize(BILINEAR)
 * (TV/transforms/functional.py resized_crop -> TV/transforms/_functional_pil.py resize), so
 * the resize is Pillow's two-pass fixed-point resampler (Pillow src/libImaging/Resample.c,
 * 22-bit coefficients, uint8 intermediate): the uint8 result is reproduced bit for bit and
 * then normalised exactly as in (3).
 *
 *  src     uint8 [B, Hs, Ws, 3], device memory (or mapped pinned host memory)
 *  params  int32 [B, 5] = (top i, left j, height h, width w, flip) per sample: the box
 *          RandomResizedCrop.get_params returns (TV/transforms/transforms.py:929-970) or
 *          (0, 0, Hs, Ws, flip) for a plain Resize; device or pinned host.  A box outside
 *          the slab is moved / shrunk into it on the device (torchvision never makes one).
 *  dst     bf16 [B, H, W, 3], 16-byte aligned
 *  mean,std  HOST pointers to 3 floats each
 * B200DP_EUNSUPPORTED when one output row's source span does not fit shared memory
 * (down-scaling by more than ~200x); B200DP_ETOOBIG for B > 65535. */
int b200dp_resized_crop_u8_bf16(const uint8_t* src, int B, int Hs, int Ws,
                                const int32_t* params, int H, int W,
                                const float* mean, const float* std, uint16_t* dst,
                                b200dp_stream_t stream);

/* ------------------------------------------------------------------------
 * (2b) NHWC max pooling, the layer that follows the stem's BN + ReLU.
 *
 * Replaces nn.MaxPool2d(kernel_size=3, stride=2, padding=1) as the reference model applies
 * it to relu(bn1(conv1(x))) (setup/resnet18.py:37,58) and as torchvision's resnet18/50 do
 * (01_torch_distributor/02_cifar_torch_distributor_resnet.py:146,
 * 03_tiny_imagenet_torch_distributor_resnet.py:130).  Arithmetic is torch's max_pool2d
 * (ATen max_pool2d_with_indices: the window is scanned rows then columns and a value is
 * taken when `val > max || isnan(val)`, so the first maximum wins a tie); the backward
 * routes each output gradient to that one input position and sums in fp32.
 *
 *  x / dx   [N, H, W, C] (channels_last), float32 or bfloat16, 16-byte aligned
 *  y / dy   [N, Ho, Wo, C], Ho = (H + 2*pad - k) / stride + 1 (floor), same for Wo
 *  idx      uint8 [N, Ho, Wo, C]: which tap of the k*k window won (row * k + column,
 *           relative to the unclipped window), written by fwd, read by bwd; 8-byte aligned
 *  k <= 3, pad <= k/2, C a multiple of 8 (bf16) / 4 (fp32); dilation 1, ceil_mode off.
 * ---------------------------------------------------------------------- */
int b200dp_maxpool_fwd(const void* x, void* y, uint8_t* idx, int dtype, int N, int H, int W, int C,
                       int k, int stride, int pad, b200dp_stream_t stream);
int b200dp_maxpool_bwd(const void* dy, const uint8_t* idx, void* dx, int dtype, int N, int H, int W,
                       int C, int k, int stride, int pad, b200dp_stream_t stream);

/* ------------------------------------------------------------------------
 * (2) SyncBatchNorm, NHWC ([M, C] row-major, M = N*H*W), training mode.
 *
 * Together these four take over TORCH/nn/modules/_functions.py:7-209
 * (SyncBatchNorm.forward/backward: batch_norm_stats, all_gather,
 * batch_norm_gather_stats_with_counts, batch_norm_elemt,
 * batch_norm_backward_reduce, all_reduce, batch_norm_backward_elemt) for the
 * nn.BatchNorm2d layers of setup/resnet18.py:8,12,19,35 and torchvision's
 * resnet18/50, plus the ReLU / residual-add epilogues of
 * setup/resnet18.py:23-26,57.
 *
 * x/y/dy/dx/residual are `dtype` (F32 or BF16); weight, bias, statistics and
 * sums are fp32.  C must be a multiple of 8 (BF16) or 4 (F32) and
 * <= B200DP_BN_MAX_CHANNELS; activation pointers must be 16-byte aligned.
 * `comm` is NULL (or a world-size-1 comm) for single-GPU batch norm; with a
 * comm of world size > 1 the statistics are exchanged across ranks inside the
 * kernels (peer stores + flags) and `seq` must be the value returned by
 * b200dp_comm_next_seq() for this exchange, called in the same order on every
 * rank.  `scratch` is device memory of b200dp_bn_scratch_bytes(C) bytes that
 * was zero-filled once when it was allocated; calls that share a scratch must
 * be ordered on one stream.
 * ---------------------------------------------------------------------- */
size_t b200dp_bn_scratch_bytes(int C);

/* Per-GPU Welford mean / M2 over the M rows of x (batch_norm_stats).
 * stats_local: fp32 [2*C+1] = mean[C], M2[C] (sum of squared deviations),
 * count.  With world > 1 the same vector is also stored into every peer's
 * exchange slot.  M may be 0 (the rank still takes part in the exchange). */
int b200dp_bn_fwd_stats(const void* x, int dtype, int64_t M, int C,
                        void* scratch, float* stats_local, b200dp_comm* comm,
                        uint32_t seq, b200dp_stream_t stream);

/* b200dp_bn_fwd_stats for callers that must not allocate device memory.
 * No counterpart in the reference: this is an addition of this library for hosts
 * that give a library no way to allocate — a serving process or another
 * framework's arena that owns all of HBM through its own pool and hands over only
 * the input and output buffers, or a launch path that may never block on the
 * allocator's device-wide lock.  The reduction scratch is a fixed array inside the
 * library image (307,712 bytes, held by every GPU that loads the library until
 * the process exits); the call performs no cudaMalloc / cudaFree and uses no
 * allocating library routine.  Single-GPU statistics only (no communicator).
 *   - C larger than B200DP_BN_NOALLOC_MAX_CHANNELS returns B200DP_ETOOBIG.
 *   - The first call on a device looks up the array's device address
 *     (cudaGetSymbolAddress) with that device current; make that first call
 *     outside any stream capture.
 *   - All calls on a device share the one array, which keeps the previous call's
 *     contents: the entry point clears the part that must start at zero on the
 *     given stream before the kernel.  Calls that overlap on the GPU (different
 *     streams without ordering between them) race on it — serialise them.
 * Output as b200dp_bn_fwd_stats: stats_local fp32 [2*C+1] = mean, M2, count. */
#define B200DP_BN_NOALLOC_MAX_CHANNELS 4096
int b200dp_bn_fwd_stats_noalloc(const void* x, int dtype, int64_t M, int C,
                                float* stats_local, b200dp_stream_t stream);

/* Merge the ranks' statistics by count (batch_norm_gather_stats_with_counts),
 * then y = act((x-mean)*invstd*weight + bias [+ residual]), act = ReLU when
 * `relu` != 0 (batch_norm_elemt + the epilogues of setup/resnet18.py:23-26).
 * Writes save_mean[C], save_invstd[C], save_count[1] (global count, fp32) for
 * the backward, and when running_mean/var are non-NULL updates them with
 * factor `momentum` (running_var uses the unbiased variance; TORCH/nn/modules/
 * batchnorm.py:760-767 picks the factor).  residual may be NULL. */
int b200dp_bn_fwd_apply(const void* x, const void* residual, void* y, int dtype,
                        int64_t M, int C, const float* stats_local,
                        const float* weight, const float* bias, float eps,
                        float momentum, float* running_mean, float* running_var,
                        float* save_mean, float* save_invstd, float* save_count,
                        int relu, b200dp_comm* comm, uint32_t seq,
                        b200dp_stream_t stream, uint32_t* relu_mask);
/* relu_mask (may be NULL; BF16 with relu != 0 only): M*C/8 words, one per 16-byte vector
 * of y, bit j set when y[j] > 0.  The backward entry points read it in mode relu == 3
 * (pass it as y_out) instead of re-reading y: 0.5 B instead of 2 B per element and pass. */

/* Eval-mode batch norm from running statistics (F.batch_norm(training=False),
 * TORCH/nn/modules/batchnorm.py:818-828) with the same fused epilogues. */
int b200dp_bn_inference(const void* x, const void* residual, void* y, int dtype,
                        int64_t M, int C, const float* running_mean,
                        const float* running_var, const float* weight,
                        const float* bias, float eps, int relu,
                        b200dp_stream_t stream);

/* batch_norm_backward_reduce.  g = dy masked by the ReLU of the forward:
 *   relu == 0  no activation was fused, g = dy;
 *   relu == 1  the mask is recomputed from x as (x-mean)*invstd*weight+bias > 0
 *              (only valid when no residual was fused; costs no extra bytes);
 *   relu == 2  the mask is y_out > 0, y_out being the forward OUTPUT (read only
 *              in this mode; may be NULL otherwise);
 *   relu == 3  (BF16) y_out points at the relu_mask words written by the forward.
 * sums_local[2*C] = sum g, sum g*(x-mean); grad_weight = sum g*(x-mean)*invstd
 * and grad_bias = sum g from the LOCAL sums (DDP averages them later); either
 * may be NULL.  weight / bias may be NULL (affine=False: 1 and 0).  With
 * world > 1 the sums are also stored into every peer's exchange slot. */
int b200dp_bn_bwd_reduce(const void* dy, const void* x, const void* y_out,
                         int dtype, int64_t M, int C, const float* save_mean,
                         const float* save_invstd, const float* weight,
                         const float* bias, int relu, void* scratch,
                         float* sums_local, float* grad_weight,
                         float* grad_bias, b200dp_comm* comm, uint32_t seq,
                         b200dp_stream_t stream);

/* batch_norm_backward_elemt: sums the ranks' vectors in rank order, then
 * dx = (g - sum_g/Mtot - (x-mean)*invstd^2*sum_gxmu/Mtot) * invstd*weight.
 * d_residual (may be NULL) receives g, the gradient of the fused residual
 * input.  save_count is the device scalar written by b200dp_bn_fwd_apply;
 * `relu` as in b200dp_bn_bwd_reduce. */
int b200dp_bn_bwd_apply(const void* dy, const void* x, const void* y_out,
                        void* dx, void* d_residual, int dtype, int64_t M, int C,
                        const float* save_mean, const float* save_invstd,
                        const float* weight, const float* bias,
                        const float* sums_local, const float* save_count,
                        int relu, b200dp_comm* comm, uint32_t seq,
                        b200dp_stream_t stream);

/* Forward = b200dp_bn_fwd_stats + b200dp_bn_fwd_apply, backward = b200dp_bn_bwd_reduce +
 * b200dp_bn_bwd_apply, each pair issued by ONE call (same kernels, same arguments; `save`
 * is fp32 [2*C+1] = save_mean | save_invstd | save_count).  They stand for the whole of
 * SyncBatchNorm.forward / .backward (TORCH/nn/modules/_functions.py:7-125 and :127-209) and
 * exist because a foreign call from Python costs about as much as a small layer's kernel. */
int b200dp_bn_forward(const void* x, const void* residual, void* y, int dtype,
                      int64_t M, int C, void* scratch, float* stats_local,
                      const float* weight, const float* bias, float eps,
                      float momentum, float* running_mean, float* running_var,
                      float* save, int relu, b200dp_comm* comm, uint32_t seq,
                      b200dp_stream_t stream, uint32_t* relu_mask,
                      int64_t* num_batches_tracked /* optional: the layer's int64 counter on the device,
                         incremented by one (TORCH/nn/modules/batchnorm.py:760-763 does it with a kernel
                         of its own per layer and step); NULL to leave it to the caller */);
int b200dp_bn_backward(const void* dy, const void* x, const void* y_out, void* dx,
                       void* d_residual, int dtype, int64_t M, int C,
                       const float* save, const float* weight, const float* bias,
                       int relu, void* scratch, float* sums_local,
                       float* grad_weight, float* grad_bias, b200dp_comm* comm,
                       uint32_t seq, b200dp_stream_t stream);

/* 1 when b200dp_bn_forward (backward != 0: b200dp_bn_backward) runs a layer of this shape as ONE
 * kernel, 0 when it issues two.  Small layers (tensor within B200DP_BN_FUSED_MB / B200DP_BN_FUSED_BWD_MB
 * MiB — default 56: what stays in the 126 MB L2 — and a shape the path was measured to win on, see
 * below; B200DP_BN_FUSED=0 turns the path off) are handled by one
 * thread-block cluster per channel chunk: rows streamed once for the statistics (sums), folded over the
 * cluster through distributed shared memory, exchanged across ranks, and streamed again out of the L2
 * for the normalise (dx) pass.  Same arguments, same results, same exchange words as the two-kernel
 * path (ranks with different local batch sizes may take different paths in the same step).
 * For callers that time or count launches. */
int b200dp_bn_one_kernel(int dtype, int64_t M, int C, int backward);
/* The same limits at run time (tests, A/B measurements): MiB bounds for the forward / backward
 * (0 = never take the one-kernel path, < 0 = leave unchanged), the largest cluster size (1..16;
 * anything else = leave unchanged), and any_geometry: 1 = every shape under the bounds takes the path
 * with whatever geometry fits, 0 = only the shapes it was measured to win on (rows of >= 2 KB, i.e.
 * C >= 1024 bf16 channels, or tensors up to 16 MiB; the default), < 0 = leave unchanged.
 * Environment equivalents: B200DP_BN_FUSED_MB, B200DP_BN_FUSED_BWD_MB, B200DP_BN_FUSED_CS,
 * B200DP_BN_FUSED_ANY.  Not synchronised with calls in flight on other host threads. */
int b200dp_bn_set_one_kernel_limits(int fwd_mb, int bwd_mb, int cluster_size, int any_geometry);

/* ------------------------------------------------------------------------
 * Communicator: library-owned symmetric buffers mapped into every rank.
 *
 * Stands where dist.init_process_group("nccl") + ProcessGroupNCCL stand for
 * the gradient path (01_torch_distributor/01_basic_torch_distributor.py:269,
 * 291).  One process per GPU; rank r allocates `heap_bytes` of device memory
 * plus a small flag/exchange area, exports both as CUDA IPC handles, the host
 * side all-gathers the handles (through the existing gloo/NCCL group or a
 * TCPStore) and every rank opens its peers' handles.  Kernels then load and
 * store peer memory directly over NVLink.
 * ---------------------------------------------------------------------- */
#define B200DP_IPC_HANDLE_BYTES 64
#define B200DP_COMM_HANDLE_BYTES (2 * B200DP_IPC_HANDLE_BYTES)

/* Allocates on the CURRENT device.  world==1 is valid (no peers to open). */
int b200dp_comm_create(int rank, int world, size_t heap_bytes,
                       b200dp_comm** out);
/* The same over a heap the CALLER owns and has mapped on every rank: heap_ptrs[r] (HOST array of `world`
 * device pointers, 16-byte aligned) is rank r's heap as seen from this process, heap_multicast (may be
 * NULL) a multicast mapping of the same bytes on all ranks (NVLink SHARP / NVLS: cuMulticastCreate +
 * cuMulticastBindMem, which torch.distributed._symmetric_memory does for its allocations).  With a
 * multicast mapping the two-shot bucket kernel reduces in the switch (multimem.ld_reduce) and pushes the
 * new parameters with one multicast store instead of one store per peer.  Only the flag / exchange area
 * is allocated (and exported / opened through the two calls below) by the library; the heap is never
 * freed by b200dp_comm_destroy. */
int b200dp_comm_create_external(int rank, int world, size_t heap_bytes,
                                void* const* heap_ptrs, void* heap_multicast,
                                b200dp_comm** out);
/* 1 when the communicator's heap has a multicast mapping. */
int b200dp_comm_has_multicast(const b200dp_comm* comm);
/* Copies this rank's B200DP_COMM_HANDLE_BYTES export blob to `out` (host). */
int b200dp_comm_export(b200dp_comm* comm, void* out);
/* all_handles: host, world * B200DP_COMM_HANDLE_BYTES, indexed by rank. */
int b200dp_comm_open_peers(b200dp_comm* comm, const void* all_handles);
int b200dp_comm_destroy(b200dp_comm* comm);
int b200dp_comm_rank(const b200dp_comm* comm);
int b200dp_comm_world(const b200dp_comm* comm);
/* Device address of the symmetric heap as seen from THIS process:
 * peer == own rank gives the local allocation. */
void* b200dp_comm_heap(b200dp_comm* comm, int peer);
size_t b200dp_comm_heap_bytes(const b200dp_comm* comm);
/* Next exchange sequence number (1, 2, 3, …).  Every collective entry point
 * below and every SyncBN exchange consumes one, in the same order on every
 * rank. */
uint32_t b200dp_comm_next_seq(b200dp_comm* comm);
/* All-gather of k 32-bit words (float32 or int32, bit-exact) per rank: out[r*k + j] = rank r's in[j].
 * One small kernel over the peer-mapped exchange slots; stands for the per-epoch metric gathers of the
 * reference (04_accelerate/01_cifar_accelerate.ipynb:638-640: three accelerator.gather calls on
 * one-element tensors, then .sum().item()).  Collective: every rank calls it in the same order
 * relative to its other collective calls.  k <= 2*B200DP_BN_MAX_CHANNELS + 128. */
int b200dp_comm_all_gather32(b200dp_comm* comm, const void* in, int k, void* out,
                             b200dp_stream_t stream);
/* Device-side barrier across the ranks on `stream` (a kernel that signals
 * every peer and waits for every peer); used at set-up and in tests. */
int b200dp_comm_barrier(b200dp_comm* comm, b200dp_stream_t stream);
/* Synchronous read of the rank's error word: non-zero once a kernel on this
 * rank gave up waiting for a peer (flag or exchange word).  The bound of every
 * such wait is B200DP_TIMEOUT_S seconds (environment, default 600: the order of
 * the NCCL watchdog's 10 minutes); a kernel whose wait ran out raises the word,
 * skips its update / push and returns, so the replicas are no longer in step
 * and the host must stop.  Takes the place of the NCCL watchdog /
 * TORCH_DISTRIBUTED_DEBUG checks (setup/00_setup.py:66-67). */
int b200dp_comm_error(b200dp_comm* comm);
/* The same word copied into `host_word` (PINNED host memory, 4 bytes) on
 * `stream` without blocking the host: enqueue it every few steps and look at
 * the value the previous copy delivered. */
int b200dp_comm_error_async(b200dp_comm* comm, uint32_t* host_word,
                            b200dp_stream_t stream);

/* ------------------------------------------------------------------------
 * (1) Gradient bucket: pack -> all-reduce fused with 1/N and the optimizer.
 * ---------------------------------------------------------------------- */

/* Gathers n_tensors gradient tensors into one flat bucket, scaling by `scale`
 * (1/world: DDP divides BEFORE it sums, TORCH/distributed/algorithms/
 * ddp_comm_hooks/default_hooks.py:18-32; Reducer div_factor_).  Takes the
 * place of the Reducer's grad -> bucket copy (TORCH/include/torch/csrc/
 * distributed/c10d/reducer.hpp:280-286).
 *  srcs     HOST array of n_tensors device pointers (dtype elements)
 *  numels   HOST array of element counts
 *  offsets  HOST array of element offsets into dst (each a multiple of 8)
 * At most 1024 tensors per call. */
int b200dp_bucket_pack(const void* const* srcs, const int64_t* numels,
                       const int64_t* offsets, int n_tensors, void* dst,
                       int dtype, float scale, b200dp_stream_t stream);

typedef struct b200dp_opt_args {
  int kind;           /* B200DP_OPT_*                                         */
  float lr;
  float momentum;     /* SGD                                                  */
  float dampening;    /* SGD                                                  */
  int nesterov;       /* SGD                                                  */
  int first_step;     /* SGD: momentum buffer := grad (sgd.py:343-352)        */
  float weight_decay; /* coupled L2: g += wd * p (both optimizers)            */
  float beta1, beta2, eps;   /* Adam                                          */
  float step_size;           /* Adam: lr / (1 - beta1^t), and                 */
  float bias_correction2_sqrt; /*     sqrt(1 - beta2^t); both computed on the  */
                             /*       host in double, as adam.py does         */
} b200dp_opt_args;

/* One gradient bucket of `numel` elements starting `elem_offset` elements into
 * every rank's symmetric heap region at byte offset `grad_heap_off`:
 *   g = sum over ranks of the (pre-scaled) bucket, read from peer memory
 *       (DDP all-reduce, 01_basic_torch_distributor.py:229 under :291);
 *   optimizer step on `master`/`state1`/`state2` (fp32, LOCAL pointers to the
 *       first element of the bucket; optimizer.step(), :230);
 *   the new parameter values are stored, as `param_dtype`, into every rank's
 *       heap at byte offset `param_heap_off` (+ elem_offset elements), which is
 *       where the module's parameters live.
 * `master` may be NULL when param_dtype is F32: the heap copy is then the
 * master copy.
 * Two-shot: rank r reduces and updates the r-th 1/world slice of the bucket
 * and pushes it to all peers; optimizer state for the other slices is not
 * touched on this rank.  One-shot (`one_shot` != 0, meant for small buckets):
 * every rank reduces and updates the whole bucket itself, nothing is pushed.
 * With `write_grads` != 0 (two-shot only) the averaged gradient is also stored
 * back into the grad bucket of every rank (what DDP leaves in .grad).
 * `max_ctas` bounds the grid (0 = library default), so that the kernel can
 * share the GPU with the rest of the backward pass.
 * numel must be a multiple of 8*world; offsets keep 16-byte alignment.
 * world == 1 degenerates to the plain optimizer step on local memory.
 * kind == B200DP_OPT_NONE only averages (requires write_grads). */
int b200dp_allreduce_update(b200dp_comm* comm, size_t grad_heap_off,
                            int grad_dtype, size_t param_heap_off,
                            int param_dtype, int64_t elem_offset, int64_t numel,
                            float* master, float* state1, float* state2,
                            const b200dp_opt_args* opt, int one_shot,
                            int write_grads, int max_ctas, uint32_t seq,
                            b200dp_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* B200DP_H_ */
