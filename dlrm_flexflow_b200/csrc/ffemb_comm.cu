// ffemb_comm.cu -- peer-memory plumbing for the table-wise sharded path.
//
// The reference moves embedding outputs and their gradients between GPUs
// implicitly: Embedding maps its output to zero-copy host memory
// (src/ops/embedding.cu:280-284,343-347) and Concat re-partitions it
// (src/ops/concat.cu:85-93), Legion DMA doing the copies through the host.
// Here every rank (one process per GPU) exposes its [B/N][T*D] output matrix
// and its gradient staging to its peers by CUDA IPC; the forward kernel then
// stores pooled rows straight into the consuming rank's matrix over NVLink,
// push_columns_kernel stores gradient blocks straight into the owning rank's
// staging (or, "pull", the update kernel loads gradient rows from the
// producing rank's matrix), and peer_barrier_kernel orders the ranks.
#include <string.h>

#include "ffemb_common.cuh"

using namespace ffemb;

extern "C" int ffemb_device_malloc(void **ptr, size_t bytes) {
  if (!ptr) return fail(FFEMB_ERR_INVALID_ARGUMENT, "device_malloc: NULL argument");
  FFEMB_CUDA_OK(cudaMalloc(ptr, bytes ? bytes : 1));
  return FFEMB_OK;
}

extern "C" int ffemb_device_free(void *ptr) {
  if (ptr) FFEMB_CUDA_OK(cudaFree(ptr));
  return FFEMB_OK;
}

extern "C" int ffemb_ipc_get_handle(void *device_ptr, uint8_t handle[64]) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
  if (!device_ptr || !handle) return fail(FFEMB_ERR_INVALID_ARGUMENT, "ipc_get_handle: NULL argument");
  cudaIpcMemHandle_t h;
  FFEMB_CUDA_OK(cudaIpcGetMemHandle(&h, device_ptr));
  memcpy(handle, &h, 64);
  return FFEMB_OK;
}

extern "C" int ffemb_ipc_open_handle(const uint8_t handle[64], void **device_ptr) {
  if (!device_ptr || !handle) return fail(FFEMB_ERR_INVALID_ARGUMENT, "ipc_open_handle: NULL argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  FFEMB_CUDA_OK(cudaIpcOpenMemHandle(device_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return FFEMB_OK;
}

extern "C" int ffemb_ipc_close_handle(void *device_ptr) {
  if (device_ptr) FFEMB_CUDA_OK(cudaIpcCloseMemHandle(device_ptr));
  return FFEMB_OK;
}

// ---------------------------------------------------------------------------
// Stream-ordered barrier among the ranks of one box, over peer memory.
// Every rank owns 32 32-bit words, zeroed once: word q < world
// is the flag written by rank q, word 16 is this rank's barrier count.  The
// kernel bumps the count to `epoch` (kept on the device, so a captured launch
// can be replayed), thread q stores `epoch` into peer q's flag[rank] (an
// NVLink store) and then waits until its own flag[q] has reached `epoch`.
// Everything a rank launched before the barrier on this stream has completed
// when its barrier kernel starts, so after the barrier every peer's earlier
// kernels -- and their stores into this rank's memory -- are done.  About two
// NVLink latencies instead of a NCCL all-reduce.  The ranks run on different
// GPUs, so the kernels are co-resident by construction; a peer that never
// arrives trips the timeout and raises the error flag instead of hanging.
// ---------------------------------------------------------------------------
namespace ffemb {

constexpr int kEpochWord = 16;

struct BarrierParams {
  uint32_t *flags[kMaxParts];  // flags[q] = base of rank q's words
  int rank, world;
  int *err;
  unsigned long long timeout_ns;
};

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__global__ void __launch_bounds__(32) peer_barrier_kernel(const __grid_constant__ BarrierParams p) {
  const int q = threadIdx.x;
  uint32_t epoch = 0;
  if (q == 0) {
    epoch = p.flags[p.rank][kEpochWord] + 1;
    p.flags[p.rank][kEpochWord] = epoch;
  }
  epoch = __shfl_sync(0xffffffffu, epoch, 0);
  if (q >= p.world) return;
  __threadfence_system();
  // release store into the peer's memory
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p.flags[q] + p.rank), "r"(epoch) : "memory");
  const uint32_t *mine = p.flags[p.rank] + q;
  const unsigned long long t0 = globaltimer_ns();
  for (;;) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
    if ((int32_t)(v - epoch) >= 0) break;  // counts only grow
    if (globaltimer_ns() - t0 > p.timeout_ns) {
      if (p.err) *p.err = FFEMB_ERR_COMM;
      break;
    }
    __nanosleep(64);
  }
}

}  // namespace ffemb

extern "C" int ffemb_peer_barrier(uint32_t *const *peer_flags, int rank, int world, int *err_flag,
                                  void *stream) {
  if (!peer_flags || world < 1 || world > kMaxParts || rank < 0 || rank >= world)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "peer_barrier: bad arguments");
  BarrierParams p;
  for (int q = 0; q < kMaxParts; ++q) p.flags[q] = q < world ? peer_flags[q] : nullptr;
  for (int q = 0; q < world; ++q)
    if (!p.flags[q]) return fail(FFEMB_ERR_INVALID_ARGUMENT, "peer_barrier: NULL flag pointer");
  p.rank = rank;
  p.world = world;
  p.err = err_flag;
  p.timeout_ns = 5ull * 1000 * 1000 * 1000;
  peer_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(p);
  FFEMB_LAUNCH_OK("peer_barrier");
  return FFEMB_OK;
}

// ---------------------------------------------------------------------------
// Backward exchange as a push: every rank holds the upstream gradient of its
// own samples for ALL tables, [B/N][T*D]; table t's owner needs the gradient
// of all samples.  One kernel copies each table's column block straight into
// its owner's staging matrix [B][T_owner*D] (rows rank*B/N ..) through the peer
// mapping -- remote stores are fire-and-forget, unlike remote loads they do
// not tie up the SM for a round trip -- so the update kernel then reads local
// memory only.  Stands in for the gradient half of the reference's implicit
// repartition at Concat (src/ops/concat.cu:263-320 add_with_stride + Legion
// copies through zero-copy memory, src/ops/embedding.cu:343-347).
// ---------------------------------------------------------------------------
namespace ffemb {

struct PushItem {
  float *dst;
  int64_t dst_ld;
  int64_t src_col;
};

struct PushParams {
  PushItem it[kMaxTables];
  const float *src;
  int64_t src_ld;
  int rows;
  int width;
};

template <bool VEC>
__global__ void __launch_bounds__(256) push_columns_kernel(const __grid_constant__ PushParams p) {
  const PushItem &it = p.it[blockIdx.y];
  const int per_row = VEC ? p.width / 4 : p.width;  // elements of this thread type per row
  const int64_t total = (int64_t)p.rows * per_row;
  const int64_t stride = (int64_t)gridDim.x * 256;
  constexpr int U = 4;
  for (int64_t i0 = (int64_t)blockIdx.x * 256 + threadIdx.x; i0 < total; i0 += stride * U) {
    if (VEC) {
      float4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t i = i0 + u * stride;
        if (i < total) {
          const int64_t r = i / per_row, c = (i - r * per_row) * 4;
          v[u] = ld_row_nc(p.src + r * p.src_ld + it.src_col + c);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t i = i0 + u * stride;
        if (i < total) {
          const int64_t r = i / per_row, c = (i - r * per_row) * 4;
          st_row(it.dst + r * it.dst_ld + c, v[u]);
        }
      }
    } else {
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t i = i0 + u * stride;
        if (i < total) {
          const int64_t r = i / per_row, c = i - r * per_row;
          it.dst[r * it.dst_ld + c] = __ldg(p.src + r * p.src_ld + it.src_col + c);
        }
      }
    }
  }
}

}  // namespace ffemb

extern "C" int ffemb_push_columns(const ffemb_push_t *items, int num_items, int rows, int width,
                                  const float *src, int64_t src_ld, void *stream) {
  if (!items || num_items < 0 || rows < 0 || width <= 0 || !src)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "push_columns: bad arguments");
  if (num_items == 0 || rows == 0) return FFEMB_OK;
  bool vec = width % 4 == 0 && src_ld % 4 == 0 && ((uintptr_t)src & 15) == 0;
  for (int i = 0; i < num_items; ++i) {
    if (!items[i].dst) return fail(FFEMB_ERR_INVALID_ARGUMENT, "push_columns: NULL destination");
    vec = vec && ((uintptr_t)items[i].dst & 15) == 0 && items[i].dst_ld % 4 == 0 && items[i].src_col % 4 == 0;
  }
  PushParams p;
  p.src = src;
  p.src_ld = src_ld;
  p.rows = rows;
  p.width = width;
  for (int i0 = 0; i0 < num_items; i0 += kMaxTables) {
    const int n = num_items - i0 < kMaxTables ? num_items - i0 : kMaxTables;
    for (int i = 0; i < n; ++i) p.it[i] = PushItem{items[i0 + i].dst, items[i0 + i].dst_ld, items[i0 + i].src_col};
    const int64_t work = (int64_t)rows * (vec ? width / 4 : width);
    int64_t bx = (work + 256 * 4 - 1) / (256 * 4);
    // the copy is NVLink-bound: a few CTAs per SM are plenty (148 SMs)
    if (bx > 148 * 4) bx = 148 * 4;
    dim3 grid((unsigned)bx, n);
    if (vec)
      push_columns_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    else
      push_columns_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    FFEMB_LAUNCH_OK("push_columns");
  }
  return FFEMB_OK;
}

// ---------------------------------------------------------------------------
// Replicated small tables: sum of the gradient replicas fused with the SGD
// step, over peer memory.  The reference sums the N gradient replicas of a
// data-parallel embedding into replica 0 and runs the dense sgd_update
// (src/runtime/optimizer_kernel.cu:96-106, src/runtime/model.cc:697-724).
// Here every rank r owns the r-th slice of the replicated tables' flat
// [count] array: it loads that slice of every rank's dense gradient through
// the peer mappings (N loads in flight per lane), adds them in rank order,
// applies W = fmaf(-lr, sum, W) once and stores the new slice into EVERY
// rank's copy of the tables -- a reduce-scatter, the update and an all-gather
// in one kernel, without a collective library and without a second pass over
// the tables.  One rank computes each element, so the replicas stay bit
// identical by construction.  The caller brackets it with two barriers: all
// gradients complete before, all copies written after.
// ---------------------------------------------------------------------------
namespace ffemb {

struct ReplicaParams {
  float *W[kMaxParts];
  const float *G[kMaxParts];
  int64_t count4;  // float4 elements
  int rank, world;
  float lr;
};

__global__ void __launch_bounds__(256) replica_reduce_update_kernel(const __grid_constant__ ReplicaParams p) {
  const int64_t per = (p.count4 + p.world - 1) / p.world;
  const int64_t lo = per * p.rank;
  const int64_t hi = lo + per < p.count4 ? lo + per : p.count4;
  const int64_t stride = (int64_t)gridDim.x * 256;
  for (int64_t i = lo + (int64_t)blockIdx.x * 256 + threadIdx.x; i < hi; i += stride) {
    float4 g[kMaxParts];
#pragma unroll
    for (int q = 0; q < kMaxParts; ++q)
      if (q < p.world) g[q] = __ldcg((const float4 *)p.G[q] + i);
    float4 sum = g[0];  // replica 0 += replica q, q = 1.. (optimizer_kernel.cu:96-101)
#pragma unroll
    for (int q = 1; q < kMaxParts; ++q)
      if (q < p.world) add4(sum, g[q]);
    const float4 w = __ldcg((const float4 *)p.W[p.rank] + i);
    const float4 nw = make_float4(fmaf(-p.lr, sum.x, w.x), fmaf(-p.lr, sum.y, w.y),
                                  fmaf(-p.lr, sum.z, w.z), fmaf(-p.lr, sum.w, w.w));
#pragma unroll
    for (int q = 0; q < kMaxParts; ++q)
      if (q < p.world) ((float4 *)p.W[q])[i] = nw;
  }
}

}  // namespace ffemb

extern "C" int ffemb_replica_reduce_update(float *const *weights, const float *const *grads,
                                           int64_t count, int rank, int world, float lr,
                                           void *stream) {
  if (!weights || !grads || count < 0 || world < 1 || world > kMaxParts || rank < 0 || rank >= world)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "replica_reduce_update: bad arguments");
  if (count % 4)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "replica_reduce_update: count must be a multiple of 4 floats");
  if (count == 0) return FFEMB_OK;
  ReplicaParams p;
  for (int q = 0; q < kMaxParts; ++q) {
    p.W[q] = q < world ? weights[q] : nullptr;
    p.G[q] = q < world ? grads[q] : nullptr;
    if (q < world && (!p.W[q] || !p.G[q] || ((uintptr_t)p.W[q] & 15) || ((uintptr_t)p.G[q] & 15)))
      return fail(FFEMB_ERR_INVALID_ARGUMENT, "replica_reduce_update: NULL or unaligned replica pointer");
  }
  p.count4 = count / 4;
  p.rank = rank;
  p.world = world;
  p.lr = lr;
  const int64_t per = (p.count4 + world - 1) / world;
  int64_t blocks = (per + 255) / 256;
  if (blocks > 148 * 4) blocks = 148 * 4;
  replica_reduce_update_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(p);
  FFEMB_LAUNCH_OK("replica_reduce_update");
  return FFEMB_OK;
}
