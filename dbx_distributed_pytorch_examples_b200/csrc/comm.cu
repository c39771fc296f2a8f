// comm.cu — the communicator: library-owned buffers that every rank maps into
// its address space with CUDA IPC, so that kernels reach peer HBM with plain
// loads/stores over NVLink 5 / NVSwitch.  One process per GPU.
//
// Each rank owns two cudaMalloc'd blocks:
//   heap  (heap_bytes)        gradient buckets and module parameters
//   sync  (b200dp_sync_area)  flags + SyncBN exchange slots, zero-filled once
// Flags only ever grow (they hold exchange sequence numbers), so nothing is
// reset between steps and a late reader can never see a "cleared" flag.
#include <atomic>
#include <new>
#include <string.h>

#include "common.cuh"

namespace {

using namespace b200dp;

__global__ void barrier_kernel(const __grid_constant__ b200dp_peers P, uint32_t seq) {
  const int t = threadIdx.x;
  if (t < P.world) {
    __threadfence_system();
    st_release_sys(&P.sync[t]->barrier[P.rank], seq);
    if (!wait_flag(&P.sync[P.rank]->barrier[t], seq, P.timeout)) atomicExch(&P.sync[P.rank]->error, 1u);
  }
}

// All-gather of k 32-bit words per rank through the SyncBN exchange slots ("LL" words: {payload, seq}
// written with one 8-byte store into every peer's slot, polled by the reader): one kernel, one NVLink
// traversal, no separate flag.  Thread idx handles (rank idx / k, word idx % k) in both phases.
__global__ void gather32_kernel(const __grid_constant__ b200dp_peers P, uint32_t seq, const uint32_t* __restrict__ in, int k,
                                uint32_t* __restrict__ out) {
  const int slot = seq % B200DP_XCHG_SLOTS;
  const int total = k * P.world;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    const int peer = idx / k, j = idx - peer * k;
    st_ll(&P.sync[peer]->bn_ll[slot][P.rank][j], __uint_as_float(in[j]), seq);
  }
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    const int src = idx / k, j = idx - src * k;
    float v;
    if (!ld_ll(&P.sync[P.rank]->bn_ll[slot][src][j], seq, &v, P.timeout)) atomicExch(&P.sync[P.rank]->error, 1u);
    out[idx] = __float_as_uint(v);
  }
}

struct ExportBlob {
  cudaIpcMemHandle_t heap;
  cudaIpcMemHandle_t sync;
};
static_assert(sizeof(cudaIpcMemHandle_t) == B200DP_IPC_HANDLE_BYTES, "IPC handle size");
static_assert(sizeof(ExportBlob) == B200DP_COMM_HANDLE_BYTES, "export blob size");

}  // namespace

extern "C" int b200dp_version(void) { return B200DP_VERSION; }

namespace {
std::atomic<unsigned long long> g_launches{0};
}
extern "C" unsigned long long b200dp_count_launch(void) { return ++g_launches; }
extern "C" unsigned long long b200dp_kernel_launches(void) { return g_launches.load(); }

extern "C" const char* b200dp_error_string(int code) {
  switch (code) {
    case B200DP_OK: return "ok";
    case B200DP_EINVAL: return "b200dp: invalid argument";
    case B200DP_EUNSUPPORTED: return "b200dp: unsupported shape or dtype";
    case B200DP_EALIGN: return "b200dp: pointer is not 16-byte aligned";
    case B200DP_ESTATE: return "b200dp: communicator is not ready (peers not opened?)";
    case B200DP_ETOOBIG: return "b200dp: size exceeds a fixed limit of this entry point";
    default: break;
  }
  if (code > 0) return cudaGetErrorString((cudaError_t)code);
  return "b200dp: unknown error";
}

extern "C" int b200dp_comm_create(int rank, int world, size_t heap_bytes, b200dp_comm** out) {
  if (!out || world < 1 || world > B200DP_MAX_WORLD || rank < 0 || rank >= world)
    return B200DP_EINVAL;
  b200dp_comm* c = new (std::nothrow) b200dp_comm();
  if (!c) return (int)cudaErrorMemoryAllocation;
  memset(c, 0, sizeof(*c));
  c->rank = rank;
  c->world = world;
  // a 2 MiB granule keeps every bucket offset aligned however it is carved up
  c->heap_bytes = ((heap_bytes + (2u << 20) - 1) >> 21) << 21;
  if (c->heap_bytes == 0) c->heap_bytes = 2u << 20;
  cudaError_t e = cudaGetDevice(&c->device);
  void* heap = nullptr;
  void* sync = nullptr;
  if (e == cudaSuccess) e = cudaMalloc(&heap, c->heap_bytes);
  if (e == cudaSuccess) e = cudaMalloc(&sync, sizeof(b200dp_sync_area));
  if (e == cudaSuccess) e = cudaMemset(sync, 0, sizeof(b200dp_sync_area));
  if (e == cudaSuccess) e = cudaMemset(heap, 0, c->heap_bytes);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    if (heap) cudaFree(heap);
    if (sync) cudaFree(sync);
    delete c;
    return (int)e;
  }
  c->heap[rank] = heap;
  c->sync[rank] = static_cast<b200dp_sync_area*>(sync);
  c->peers_open = (world == 1);
  c->seq = 0;
  *out = c;
  return B200DP_OK;
}

// A communicator over a heap the CALLER allocated and mapped on every rank (e.g. torch's symmetric
// memory: cuMemCreate + fd exchange + cuMulticastBindMem), so that the bucket kernel can use the
// switch's multicast address.  Only the small flag / exchange area is still allocated and IPC-exported
// here.  heap_ptrs[r] = rank r's heap as mapped in THIS process; heap_multicast may be NULL.
extern "C" int b200dp_comm_create_external(int rank, int world, size_t heap_bytes, void* const* heap_ptrs,
                                           void* heap_multicast, b200dp_comm** out) {
  if (!out || !heap_ptrs || world < 1 || world > B200DP_MAX_WORLD || rank < 0 || rank >= world) return B200DP_EINVAL;
  for (int r = 0; r < world; ++r)
    if (!heap_ptrs[r] || ((uintptr_t)heap_ptrs[r] & 15)) return B200DP_EALIGN;
  if ((uintptr_t)heap_multicast & 15) return B200DP_EALIGN;
  b200dp_comm* c = new (std::nothrow) b200dp_comm();
  if (!c) return (int)cudaErrorMemoryAllocation;
  memset(c, 0, sizeof(*c));
  c->rank = rank;
  c->world = world;
  c->heap_bytes = heap_bytes;
  c->heap_external = true;
  c->heap_mc = heap_multicast;
  cudaError_t e = cudaGetDevice(&c->device);
  void* sync = nullptr;
  if (e == cudaSuccess) e = cudaMalloc(&sync, sizeof(b200dp_sync_area));
  if (e == cudaSuccess) e = cudaMemset(sync, 0, sizeof(b200dp_sync_area));
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    if (sync) cudaFree(sync);
    delete c;
    return (int)e;
  }
  for (int r = 0; r < world; ++r) c->heap[r] = heap_ptrs[r];
  c->sync[rank] = static_cast<b200dp_sync_area*>(sync);
  c->peers_open = (world == 1);
  *out = c;
  return B200DP_OK;
}

extern "C" int b200dp_comm_has_multicast(const b200dp_comm* comm) { return comm && comm->heap_mc ? 1 : 0; }

extern "C" int b200dp_comm_export(b200dp_comm* comm, void* out) {
  if (!comm || !out) return B200DP_EINVAL;
  ExportBlob blob;
  memset(&blob, 0, sizeof(blob));
  if (comm->world > 1) {
    if (!comm->heap_external) B200DP_CUDA_TRY(cudaIpcGetMemHandle(&blob.heap, comm->heap[comm->rank]));
    B200DP_CUDA_TRY(cudaIpcGetMemHandle(&blob.sync, comm->sync[comm->rank]));
  }
  memcpy(out, &blob, sizeof(blob));
  return B200DP_OK;
}

extern "C" int b200dp_comm_open_peers(b200dp_comm* comm, const void* all_handles) {
  if (!comm || !all_handles) return B200DP_EINVAL;
  if (comm->peers_open) return B200DP_OK;
  const ExportBlob* blobs = static_cast<const ExportBlob*>(all_handles);
  for (int p = 0; p < comm->world; ++p) {
    if (p == comm->rank) continue;
    void* heap = nullptr;
    void* sync = nullptr;
    if (!comm->heap_external) {
      B200DP_CUDA_TRY(cudaIpcOpenMemHandle(&heap, blobs[p].heap, cudaIpcMemLazyEnablePeerAccess));
      comm->heap[p] = heap;
    }
    B200DP_CUDA_TRY(cudaIpcOpenMemHandle(&sync, blobs[p].sync, cudaIpcMemLazyEnablePeerAccess));
    comm->sync[p] = static_cast<b200dp_sync_area*>(sync);
  }
  comm->peers_open = true;
  return B200DP_OK;
}

extern "C" int b200dp_comm_destroy(b200dp_comm* comm) {
  if (!comm) return B200DP_OK;
  cudaDeviceSynchronize();
  for (int p = 0; p < comm->world; ++p) {
    if (p == comm->rank) continue;
    if (comm->heap[p] && !comm->heap_external) cudaIpcCloseMemHandle(comm->heap[p]);
    if (comm->sync[p]) cudaIpcCloseMemHandle(comm->sync[p]);
  }
  if (comm->heap[comm->rank] && !comm->heap_external) cudaFree(comm->heap[comm->rank]);
  if (comm->sync[comm->rank]) cudaFree(comm->sync[comm->rank]);
  delete comm;
  return B200DP_OK;
}

extern "C" int b200dp_comm_rank(const b200dp_comm* comm) { return comm ? comm->rank : -1; }
extern "C" int b200dp_comm_world(const b200dp_comm* comm) { return comm ? comm->world : -1; }
extern "C" void* b200dp_comm_heap(b200dp_comm* comm, int peer) {
  if (!comm || peer < 0 || peer >= comm->world) return nullptr;
  return comm->heap[peer];
}
extern "C" size_t b200dp_comm_heap_bytes(const b200dp_comm* comm) {
  return comm ? comm->heap_bytes : 0;
}
extern "C" uint32_t b200dp_comm_next_seq(b200dp_comm* comm) { return comm ? ++comm->seq : 0; }

extern "C" int b200dp_comm_barrier(b200dp_comm* comm, b200dp_stream_t stream) {
  if (!comm) return B200DP_EINVAL;
  if (!comm->peers_open) return B200DP_ESTATE;
  if (comm->world == 1) return B200DP_OK;
  const uint32_t seq = ++comm->seq;
  barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(b200dp_make_peers(comm), seq);
  B200DP_LAUNCH_CHECK();
  return B200DP_OK;
}

extern "C" int b200dp_comm_all_gather32(b200dp_comm* comm, const void* in, int k, void* out,
                                        b200dp_stream_t stream) {
  if (!comm || k < 0) return B200DP_EINVAL;
  if (k == 0) return B200DP_OK;
  if (!in || !out) return B200DP_EINVAL;
  if (((uintptr_t)in | (uintptr_t)out) & 3) return B200DP_EALIGN;
  if (k > B200DP_XCHG_FLOATS) return B200DP_ETOOBIG;
  cudaStream_t s = (cudaStream_t)stream;
  if (comm->world == 1) {
    B200DP_CUDA_TRY(cudaMemcpyAsync(out, in, (size_t)k * 4, cudaMemcpyDeviceToDevice, s));
    return B200DP_OK;
  }
  if (!comm->peers_open) return B200DP_ESTATE;
  const uint32_t seq = ++comm->seq;
  const int total = k * comm->world;
  const int blocks = total <= 256 ? 1 : (total + 255) / 256 > 32 ? 32 : (total + 255) / 256;
  gather32_kernel<<<blocks, 256, 0, s>>>(b200dp_make_peers(comm), seq, static_cast<const uint32_t*>(in), k,
                                         static_cast<uint32_t*>(out));
  B200DP_LAUNCH_CHECK();
  return B200DP_OK;
}

// The error word copied to pinned host memory on `stream`, without blocking the host: the training loop
// enqueues this every few steps and looks at what the previous copy brought back.
extern "C" int b200dp_comm_error_async(b200dp_comm* comm, uint32_t* host_word, b200dp_stream_t stream) {
  if (!comm || !host_word) return B200DP_EINVAL;
  B200DP_CUDA_TRY(cudaMemcpyAsync(host_word, &comm->sync[comm->rank]->error, sizeof(uint32_t),
                                  cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  return B200DP_OK;
}

// Non-zero once any kernel on this rank gave up waiting for a peer.
extern "C" int b200dp_comm_error(b200dp_comm* comm) {
  if (!comm) return B200DP_EINVAL;
  uint32_t v = 0;
  B200DP_CUDA_TRY(cudaMemcpy(&v, &comm->sync[comm->rank]->error, sizeof(v), cudaMemcpyDeviceToHost));
  return (int)v;
}
