// ffemb_backward.cu -- deterministic backward fused with sparse SGD (sm_100a).
//
// Replaces, for weight_decay == momentum == 0 (DLRM's optimiser,
// examples/cpp/DLRM/dlrm.cc:134), three dense reference passes per step:
//   zero of the [N][D] gradient       src/runtime/initializer_kernel.cu:232-262
//   atomicAdd scatter embed_backward  src/ops/embedding.cu:199-224
//   dense sgd_update over N*D         src/runtime/optimizer_kernel.cu:23-41
// by work proportional to the lookups of the step:
//   1. make_keys   key = table_row_base + row, val = table*B + bag
//   2. stable LSD radix sort of (key, val) on the significant key bits
//   3. reduce_kernel  one sub-warp per chunk of 32 sorted positions: gradient
//                  rows added in sorted (= original (bag, position)) order; a
//                  run inside the chunk is applied at once,
//                  W[row,:] = fmaf(-lr, gsum, W[row,:]); runs cut by chunk
//                  boundaries are chained in shared memory inside the block;
//                  only runs that leave the block reach memory as partials,
//                  and the block in which such a run ENDS looks back at the
//                  earlier blocks' published partials, adds them in order and
//                  applies the row -- all in the one launch.
// No floating-point atomics; the result depends only on the inputs (bitwise
// repeatable).  Rows whose lookups all fall in one chunk get the exact
// sequential sum of the reference's CPU backward (src/ops/embedding.cc:54-72);
// longer runs are summed piecewise in a fixed order, within 1e-5 of it.
#include <cub/device/device_radix_sort.cuh>

#include "ffemb_common.cuh"
#include "ffemb_csort.cuh"
#include "ffemb_sort.cuh"

namespace ffemb {

// 0: CUB DeviceRadixSort (library); 1: the segmented per-table sort of ffemb_sort.cuh
// (three launches per pass, through global memory); 2: the one-launch on-chip
// cluster sort of ffemb_csort.cuh where every table's segment fits (<= 65536
// lookups per table), else 0 -- for segments of half a million lookups (C4) the
// library sort's 7 launches hide better under the gather than the 10 of the own
// global sort (forward + sort 0.51 vs 0.61 ms).  Launch groups with 2^32-1 rows
// or more: always 0.
#ifndef FFEMB_DEFAULT_SORT
#define FFEMB_DEFAULT_SORT 2
#endif
std::atomic<int> g_sort_impl{FFEMB_DEFAULT_SORT};

constexpr int kChunk = 32;  // sorted positions per sub-warp in reduce_kernel
constexpr int kBwdThreads = 256;
#ifndef FFEMB_BU
#define FFEMB_BU 8
#endif
constexpr int kBU = FFEMB_BU;     // positions whose rows are in flight together
#ifndef FFEMB_L1_MINBLOCKS
#define FFEMB_L1_MINBLOCKS 2
#endif

struct KeyTable {  // make_keys view
  const int64_t *idx;
  const int64_t *off;
  int64_t N;
  int64_t row_base;
  int64_t pos_base;
  int64_t n;
  int L;
};

struct KeyParams {
  KeyTable t[kMaxTables];
  void *keys;
  uint32_t *vals;
  int B;
  int *err;
};

struct BwdTable {
  float *W;
  const int64_t *off;
  int64_t row_base;
  int64_t col;
  int L;
};

struct BwdParams {
  BwdTable t[kMaxTables];
  Matrix grad;
  const void *keys;
  const uint32_t *vals;
  float *part0;      // [blocks][D] partial of the run that enters a block
  float *part1;      // [blocks][D] partial of the run that starts in a block and leaves it
  uint32_t *status;  // [2][blocks] 0 -> 1 once part0[b] / part1[b] is published
  int *err;
  int n;
  int T;
  int B;
  int D;
  int avg;
  float lr;
};

template <typename KeyT>
struct KeyTraits;
template <>
struct KeyTraits<uint32_t> {
  static constexpr uint32_t kInvalid = 0xFFFFFFFFu;
};
template <>
struct KeyTraits<uint64_t> {
  static constexpr uint64_t kInvalid = ~0ull;
};

// ---------------------------------------------------------------------------
// 1. keys.  Fixed bag size: one thread per lookup.  CSR bags: one thread per
// bag.  An out-of-range index gets the invalid key (sorts last, is skipped)
// and raises the error flag.
// ---------------------------------------------------------------------------
template <typename KeyT>
__global__ void __launch_bounds__(256) make_keys_kernel(const __grid_constant__ KeyParams p) {
  const KeyTable &tb = p.t[blockIdx.y];
  KeyT *keys = (KeyT *)p.keys;
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  bool bad = false;
  if (!tb.off) {
    if (i < tb.n) {
      const int64_t row = tb.idx[i];
      const uint32_t bag = tb.L == 1 ? (uint32_t)i : (uint32_t)(i / tb.L);
      KeyT k = KeyTraits<KeyT>::kInvalid;
      if ((uint64_t)row < (uint64_t)tb.N)
        k = (KeyT)(tb.row_base + row);
      else
        bad = true;
      keys[tb.pos_base + i] = k;
      p.vals[tb.pos_base + i] = blockIdx.y * (uint32_t)p.B + bag;
    }
  } else if (i < p.B) {
    const int64_t s = tb.off[i], e = tb.off[i + 1];
    for (int64_t j = s; j < e; ++j) {
      const int64_t row = tb.idx[j];
      KeyT k = KeyTraits<KeyT>::kInvalid;
      if ((uint64_t)row < (uint64_t)tb.N)
        k = (KeyT)(tb.row_base + row);
      else
        bad = true;
      keys[tb.pos_base + j] = k;
      p.vals[tb.pos_base + j] = blockIdx.y * (uint32_t)p.B + (uint32_t)i;
    }
  }
  if (bad && p.err) *p.err = FFEMB_ERR_INDEX_OUT_OF_RANGE;
}

// ---------------------------------------------------------------------------
// Column access of one sub-warp lane.  VEC: lane gl owns the float4 at column
// 4*gl (D == 4*G).  Scalar: lane owns columns d0+lane+{0,32,64,96} < D.
// ---------------------------------------------------------------------------
template <bool VEC>
struct Cols {
  int gl, d0, D;
  __device__ __forceinline__ float4 ld_nc(const float *row) const {
    if (VEC) return ld_row_nc(row + gl * 4);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    const int d = d0 + gl;
    if (d < D) v.x = __ldg(row + d);
    if (d + 32 < D) v.y = __ldg(row + d + 32);
    if (d + 64 < D) v.z = __ldg(row + d + 64);
    if (d + 96 < D) v.w = __ldg(row + d + 96);
    return v;
  }
  __device__ __forceinline__ float4 ld(const float *row) const {
    if (VEC) return ld_row(row + gl * 4);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    const int d = d0 + gl;
    if (d < D) v.x = row[d];
    if (d + 32 < D) v.y = row[d + 32];
    if (d + 64 < D) v.z = row[d + 64];
    if (d + 96 < D) v.w = row[d + 96];
    return v;
  }
  // partials published by other blocks: read at L2, never from this SM's L1
  __device__ __forceinline__ float4 ld_cg(const float *row) const {
    if (VEC) return __ldcg((const float4 *)(row + gl * 4));
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    const int d = d0 + gl;
    if (d < D) v.x = __ldcg(row + d);
    if (d + 32 < D) v.y = __ldcg(row + d + 32);
    if (d + 64 < D) v.z = __ldcg(row + d + 64);
    if (d + 96 < D) v.w = __ldcg(row + d + 96);
    return v;
  }
  __device__ __forceinline__ void st(float *row, const float4 &v) const {
    if (VEC) {
      st_row(row + gl * 4, v);
      return;
    }
    const int d = d0 + gl;
    if (d < D) row[d] = v.x;
    if (d + 32 < D) row[d + 32] = v.y;
    if (d + 64 < D) row[d + 64] = v.z;
    if (d + 96 < D) row[d + 96] = v.w;
  }
};

__device__ __forceinline__ float4 sgd_row(float lr, const float4 &g, const float4 &w) {
  // W -= lr*g as one FMA (reference: nvcc contracts optimizer_kernel.cu:38)
  return make_float4(fmaf(-lr, g.x, w.x), fmaf(-lr, g.y, w.y), fmaf(-lr, g.z, w.z),
                     fmaf(-lr, g.w, w.w));
}

template <typename P>
__device__ __forceinline__ P *shfl_ptr(unsigned mask, P *ptr, int src, int width) {
  return (P *)__shfl_sync(mask, (unsigned long long)ptr, src, width);
}

// Largest t with row_base[t] <= key (T <= 64: six halvings).
template <typename KeyT>
__device__ __forceinline__ int table_of(const BwdParams &p, KeyT key) {
  int lo = 0, hi = p.T;
#pragma unroll
  for (int it = 0; it < 6; ++it) {
    const int mid = (lo + hi) >> 1;
    if (hi - lo > 1) {
      if (p.t[mid].row_base <= (int64_t)key)
        lo = mid;
      else
        hi = mid;
    }
  }
  return lo;
}

// ---------------------------------------------------------------------------
// 3. segmented reduce + SGD, one launch.
//
// Per chunk of 32 sorted positions, each lane of the sub-warp first works out
// its own positions' bookkeeping (key, neighbours' keys, table, gradient-row
// and table-row addresses, head/tail of a run); ballots turn head/tail/valid
// into bit masks.  The walk itself then only shuffles the two addresses of a
// position round the sub-warp: 8 gradient rows and the table rows of the runs
// ending among them are in flight together and are consumed strictly in
// order.  Runs cut by a chunk boundary leave a partial in shared memory;
// after one barrier the sub-warp in whose chunk such a run starts adds the
// following chunks' partials in order, so only runs that leave the BLOCK
// (256/G chunks) reach global memory as partials: part0[b] for the run that
// enters block b, part1[b] for the run that starts in b and leaves it.
//
// Runs that cross block boundaries ("chains") are completed by the block in
// which they END, decoupled-look-back style: the sub-warp that stores a
// partial then raises its status word (fence, then a relaxed store); the end
// block finds the chain's first block by galloping + bisecting over the
// (sorted) first keys of the blocks, waits for the status words of the blocks
// before it, and adds part1[first], part0[first+1 .. end] in that order.
// Waits only ever point at lower block indices and cannot cycle.
// ---------------------------------------------------------------------------
constexpr int kHasLead = 1;   // first run of the chunk started before the chunk
constexpr int kLeadCont = 2;  // ... and also continues past the end of the chunk
constexpr int kHasTrail = 4;  // last run started in the chunk and continues past it

__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_gpu(uint32_t *p, uint32_t v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long bwd_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

template <typename KeyT, int G, bool VEC>
__global__ void __launch_bounds__(kBwdThreads, FFEMB_L1_MINBLOCKS)
reduce_kernel(const __grid_constant__ BwdParams p) {
  constexpr int U = kBU;
  constexpr int S = kBwdThreads / G;  // chunks (sub-warps) per block
  constexpr int KPL = kChunk / G;     // positions per lane
  constexpr int kSpan = S * kChunk;   // positions per block
  constexpr KeyT kInv = KeyTraits<KeyT>::kInvalid;
  static_assert(G >= 4 && G <= 32, "sub-warp of 4..32 lanes");
  __shared__ float4 s_lead[S][G];
  __shared__ float4 s_trail[S][G];
  __shared__ int s_flags[S];
  const KeyT *__restrict__ keys = (const KeyT *)p.keys;
  const int s = threadIdx.x / G, gl = threadIdx.x % G;
  const int subshift = (threadIdx.x & 31) / G * G;
  const unsigned gmask = G == 32 ? 0xffffffffu : ((1u << G) - 1u);
  const unsigned submask = gmask << subshift;
  const int n = p.n;
  // Blocks are dispatched in index order, so every block a waiter looks back
  // at (lower index) is resident or finished; the wait below is bounded anyway.
  const int bid = blockIdx.x;
  uint32_t *const status0 = p.status;              // part0[b] published
  uint32_t *const status1 = p.status + gridDim.x;  // part1[b] published
  const int i0 = (bid * S + s) * kChunk;
  const int cnt = n - i0 < kChunk ? n - i0 : kChunk;  // positions in this chunk (<= 0: none)
  const int D = p.D;

  // ---- bookkeeping, one position per lane and register --------------------
  const float *gptr[KPL];
  float *wptr[KPL];
  float glen[KPL];
  unsigned headbits = 0, tailbits = 0, validbits = 0;
#pragma unroll
  for (int j = 0; j < KPL; ++j) {
    const int q = j * G + gl, pos = i0 + q;
    KeyT k = kInv, kp = kInv, kn = kInv;
    uint32_t v = 0;
    if (q < cnt) {
      k = keys[pos];
      v = p.vals[pos];
      if (pos > 0) kp = keys[pos - 1];
      if (pos + 1 < n) kn = keys[pos + 1];
    }
    const bool valid = k != kInv;
    const bool head = valid && (pos == 0 || kp != k);
    const bool tail = valid && kn != k;
    gptr[j] = nullptr;
    wptr[j] = nullptr;
    glen[j] = 1.f;
    if (valid) {
      const int t = table_of(p, k);
      const BwdTable &tb = p.t[t];
      const int b = (int)(v - (uint32_t)t * (uint32_t)p.B);
      gptr[j] = matrix_row(p.grad, b) + tb.col;
      wptr[j] = tb.W + ((int64_t)k - tb.row_base) * D;
      if (p.avg) {
        // per-position gradient is g / (float)len (reference embedding.cu:213-218).
        // For len a power of two, g * (1/len) is the same correctly rounded
        // value and costs one multiply: stored positive.  Otherwise keep the
        // IEEE division: stored as -len.
        const float len = tb.off ? (float)(int)(tb.off[b + 1] - tb.off[b]) : (float)tb.L;
        glen[j] = (__float_as_uint(len) & 0x7FFFFFu) == 0 ? 1.0f / len : -len;
      }
    }
    validbits |= ((__ballot_sync(0xffffffffu, valid) >> subshift) & gmask) << (j * G);
    headbits |= ((__ballot_sync(0xffffffffu, head) >> subshift) & gmask) << (j * G);
    tailbits |= ((__ballot_sync(0xffffffffu, tail) >> subshift) & gmask) << (j * G);
  }

  Cols<VEC> cols{gl, 0, D};
  for (int d0 = 0; d0 < (VEC ? 1 : D); d0 += 128) {
    cols.d0 = d0;
    int flags = 0;
    bool head_in_chunk = false;
    float *wlast = nullptr;  // table row of the run that leaves the chunk
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int base = 0; base < kChunk; base += U) {
      if (base < cnt) {
        float4 gv[U], wv[U];
        float *wp[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int q = base + u;
          const float *gp = shfl_ptr(submask, gptr[q / G], q % G, G);
          wp[u] = shfl_ptr(submask, wptr[q / G], q % G, G);
          gv[u] = make_float4(0.f, 0.f, 0.f, 0.f);
          wv[u] = gv[u];
          if ((validbits >> q) & 1u) {
            gv[u] = cols.ld_nc(gp);
            if ((tailbits >> q) & 1u) wv[u] = cols.ld(wp[u]);
          }
        }
        if (p.avg) {
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const int q = base + u;
            const float sc = __shfl_sync(submask, glen[q / G], q % G, G);
            if (sc > 0.f) {
              gv[u].x *= sc;
              gv[u].y *= sc;
              gv[u].z *= sc;
              gv[u].w *= sc;
            } else {
              gv[u].x /= -sc;
              gv[u].y /= -sc;
              gv[u].z /= -sc;
              gv[u].w /= -sc;
            }
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int q = base + u;
          if ((validbits >> q) & 1u) {
            if ((headbits >> q) & 1u) {
              acc = make_float4(0.f, 0.f, 0.f, 0.f);
              head_in_chunk = true;
            }
            add4(acc, gv[u]);
            if ((tailbits >> q) & 1u) {  // true end of the run
              if (head_in_chunk) {
                cols.st(wp[u], sgd_row(p.lr, acc, wv[u]));
              } else {
                s_lead[s][gl] = acc;
                flags |= kHasLead;
              }
            } else if (q == cnt - 1) {  // run continues in the next chunk
              if (head_in_chunk) {
                s_trail[s][gl] = acc;
                flags |= kHasTrail;
              } else {
                s_lead[s][gl] = acc;
                flags |= kHasLead | kLeadCont;
              }
              wlast = wp[u];
            }
          }
        }
      }
    }
    if (gl == 0) s_flags[s] = flags;
    __syncthreads();
    // Chains inside the block.  The run entering the block (chunk 0's lead)
    // and every run that starts in a chunk and leaves it are each summed by
    // one sub-warp, following chunks in order.
    if ((flags & kHasLead) && s == 0) {
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
      add4(a, s_lead[0][gl]);
      bool cont = flags & kLeadCont;
      for (int s2 = 1; cont && s2 < S; ++s2) {
        add4(a, s_lead[s2][gl]);
        cont = s_flags[s2] & kLeadCont;
      }
      cols.st(p.part0 + (int64_t)bid * D, a);
      if (VEC || d0 + 128 >= D) {  // whole row written: publish it
        __threadfence();
        __syncwarp(submask);
        if (gl == 0) st_relaxed_gpu(status0 + bid, 1u);
      }
    }
    if (flags & kHasTrail) {
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
      add4(a, s_trail[s][gl]);
      bool cont = true;
      for (int s2 = s + 1; cont && s2 < S; ++s2) {
        add4(a, s_lead[s2][gl]);
        cont = s_flags[s2] & kLeadCont;
      }
      if (cont) {  // leaves the block: the block in which it ends completes it
        cols.st(p.part1 + (int64_t)bid * D, a);
        if (VEC || d0 + 128 >= D) {
          __threadfence();
          __syncwarp(submask);
          if (gl == 0) st_relaxed_gpu(status1 + bid, 1u);
        }
      } else {
        cols.st(wlast, sgd_row(p.lr, a, cols.ld(wlast)));
      }
    }
    if (!VEC) __syncthreads();  // shared partials are reused by the next column tile
  }

  // ---- complete the chain that ends in this block ---------------------------
  if (s != 0 || !(s_flags[0] & kHasLead)) return;
  for (int s2 = 0; s2 < S; ++s2) {
    if (!(s_flags[s2] & kLeadCont)) break;
    if (s2 == S - 1) return;  // the entering run also leaves the block: a later block ends it
  }
  const KeyT key = keys[(int64_t)bid * kSpan];  // the entering run's key
  // first block whose first key is `key` (the blocks' first keys are sorted)
  int hi = bid, lo = -1;
  for (int step = 1;; step *= 2) {
    const int j = hi - step;
    if (j < 0) break;
    if (keys[(int64_t)j * kSpan] != key) {
      lo = j;
      break;
    }
    hi = j;
  }
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (keys[(int64_t)mid * kSpan] == key)
      hi = mid;
    else
      lo = mid;
  }
  // the run starts in block hi at its first position, or inside block hi - 1
  const int c = (hi > 0 && keys[(int64_t)hi * kSpan - 1] == key) ? hi - 1 : hi;
  {
    const unsigned long long t0 = bwd_timer_ns();
    for (int j = c + gl; j < bid; j += G) {
      const uint32_t *flag = j == c ? status1 + j : status0 + j;
      while (ld_acquire_gpu(flag) == 0u) {
        if (bwd_timer_ns() - t0 > 2000000000ull) {  // 2 s: a lower ticket never published
          if (p.err) *p.err = FFEMB_ERR_CUDA;
          break;
        }
        __nanosleep(32);
      }
    }
    __syncwarp(submask);
  }
  const int t = table_of(p, key);
  float *wrow = p.t[t].W + ((int64_t)key - p.t[t].row_base) * D;
  for (int d0 = 0; d0 < (VEC ? 1 : D); d0 += 128) {
    cols.d0 = d0;
    const float4 w = cols.ld(wrow);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    add4(acc, cols.ld_cg(p.part1 + (int64_t)c * D));
    for (int b = c + 1; b <= bid; b += 8) {
      float4 v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (b + j <= bid) v[j] = cols.ld_cg(p.part0 + (int64_t)(b + j) * D);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (b + j <= bid) add4(acc, v[j]);
    }
    cols.st(wrow, sgd_row(p.lr, acc, w));
  }
}

template <typename KeyT, int G, bool VEC>
static void launch_reduce(const BwdParams &p, cudaStream_t s) {
  constexpr int per_block = kBwdThreads / G;
  const int64_t chunks = ((int64_t)p.n + kChunk - 1) / kChunk;
  const int64_t blocks = (chunks + per_block - 1) / per_block;
  reduce_kernel<KeyT, G, VEC><<<(unsigned)blocks, kBwdThreads, 0, s>>>(p);
  g_launches.fetch_add(1, std::memory_order_relaxed);
}

template <typename KeyT>
static void launch_reduce_g(const BwdParams &p, bool vec, cudaStream_t s) {
  if (!vec) return launch_reduce<KeyT, 32, false>(p, s);
  switch (p.D / 4) {
    case 4: return launch_reduce<KeyT, 4, true>(p, s);
    case 8: return launch_reduce<KeyT, 8, true>(p, s);
    case 16: return launch_reduce<KeyT, 16, true>(p, s);
    case 32: return launch_reduce<KeyT, 32, true>(p, s);
    default: return launch_reduce<KeyT, 32, false>(p, s);  // D = 4, 8: scalar columns
  }
}

// ---- workspace layout --------------------------------------------------------
static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

struct Workspace {
  size_t key_bytes;  // per key
  void *keys_a, *keys_b;
  uint32_t *vals_a, *vals_b;
  void *sort_tmp;
  size_t sort_tmp_bytes;
  float *part0, *part1;
  uint32_t *status;  // [2][blocks]: "part0 published", "part1 published"
  size_t status_bytes;
  size_t total;
};

static bool wide_keys(int64_t total_entries) { return total_entries >= 0xFFFFFFFFll; }

static int key_bits(int64_t total_entries) {
  int b = 1;
  while (b < 64 && (total_entries >> b) != 0) ++b;
  return b;  // floor(log2)+1: every valid key < 2^b - 1, the invalid key sorts last
}

static size_t sort_tmp_bytes(int64_t n, bool wide) {
  size_t bytes = 0;
  if (wide)
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint64_t *)nullptr, (uint64_t *)nullptr,
                                    (const uint32_t *)nullptr, (uint32_t *)nullptr, n);
  else
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint32_t *)nullptr, (uint32_t *)nullptr,
                                    (const uint32_t *)nullptr, (uint32_t *)nullptr, n);
  return bytes;
}

static Workspace carve(void *base, int64_t n, int64_t total_entries, int D) {
  Workspace w;
  const bool wide = wide_keys(total_entries);
  w.key_bytes = wide ? 8 : 4;
  // one partial pair per level-1 block; the smallest block span is 256 positions
  const int64_t blocks = (n + 255) / 256 + 1;
  char *p = (char *)base;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void *q = p ? p + off : nullptr;
    off += align256(bytes);
    return q;
  };
  // keys_a and vals_a are adjacent: the on-chip sort uses the pair as one
  // array of n 8-byte (row, bag) records between its passes
  w.keys_a = take(n * w.key_bytes);
  w.vals_a = (uint32_t *)take(n * 4);
  w.keys_b = take(n * w.key_bytes);
  w.vals_b = (uint32_t *)take(n * 4);
  w.sort_tmp_bytes = sort_tmp_bytes(n, wide);
  if (!wide && own_sort_hist_bytes(n) > w.sort_tmp_bytes) w.sort_tmp_bytes = own_sort_hist_bytes(n);
  w.sort_tmp = take(w.sort_tmp_bytes);
  w.part0 = (float *)take((size_t)blocks * D * 4);
  w.part1 = (float *)take((size_t)blocks * D * 4);
  w.status_bytes = (size_t)blocks * 2 * sizeof(uint32_t);
  w.status = (uint32_t *)take(w.status_bytes);
  w.total = off;
  return w;
}

static bool aligned16(const void *p) { return ((uintptr_t)p & 15) == 0; }

struct Group {  // one batched launch group (<= kMaxTables tables)
  int64_t n;              // lookups in the group
  int64_t total_entries;  // rows in the group
};

int check_lookups(const char *who, const ffemb_lookup_t *lookups, int num_tables, int batch,
                  int out_dim) {
  if (!lookups || num_tables < 0)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, std::string(who) + ": lookups is NULL");
  if (batch < 0 || out_dim <= 0)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, std::string(who) + ": bad batch/out_dim");
  for (int t = 0; t < num_tables; ++t) {
    const ffemb_lookup_t &l = lookups[t];
    if (!l.weight || (!l.indices && l.num_indices > 0) || l.num_entries <= 0 || l.num_indices < 0)
      return fail(FFEMB_ERR_INVALID_ARGUMENT,
                  std::string(who) + ": table " + std::to_string(t) + " is malformed");
    if (!l.offsets && (l.bag_size < 0 || l.num_indices != (int64_t)batch * l.bag_size))
      return fail(FFEMB_ERR_INVALID_ARGUMENT,
                  std::string(who) + ": num_indices != batch*bag_size for table " + std::to_string(t));
  }
  if ((int64_t)batch * (num_tables < kMaxTables ? num_tables : kMaxTables) >= 0xFFFFFFFFll)
    return fail(FFEMB_ERR_UNSUPPORTED, std::string(who) + ": batch * tables per launch >= 2^32");
  return FFEMB_OK;
}

static Group group_of(const ffemb_lookup_t *lookups, int T) {
  Group g{0, 0};
  for (int t = 0; t < T; ++t) {
    g.n += lookups[t].num_indices;
    g.total_entries += lookups[t].num_entries;
  }
  return g;
}

template <typename KeyT>
static int sort_group(const ffemb_lookup_t *lookups, int T, int batch, const Workspace &w,
                      const Group &g, int *err, cudaStream_t s) {
  const int impl = g_sort_impl.load(std::memory_order_relaxed);
  if (sizeof(KeyT) == 4 && impl == 2) {
    bool done = false;
    if (int rc = cluster_sort_group(lookups, T, batch, (uint32_t *)w.keys_b, w.vals_b, (uint2 *)w.keys_a,
                                    err, s, &done))
      return rc;
    if (done) return FFEMB_OK;
  }
  if (sizeof(KeyT) == 4 && impl == 1)
    return own_sort_group(lookups, T, batch, (uint32_t *)w.keys_a, (uint32_t *)w.keys_b, w.vals_a,
                          w.vals_b, (uint32_t *)w.sort_tmp, err, s);
  KeyParams kp;
  int64_t rows = 0, pos = 0, max_n = 0;
  bool any_csr = false;
  for (int t = 0; t < T; ++t) {
    const ffemb_lookup_t &l = lookups[t];
    kp.t[t] = KeyTable{l.indices, l.offsets, l.num_entries, rows, pos, l.num_indices, l.bag_size};
    rows += l.num_entries;
    pos += l.num_indices;
    if (l.offsets) any_csr = true;
    if (l.num_indices > max_n) max_n = l.num_indices;
  }
  kp.keys = w.keys_a;
  kp.vals = w.vals_a;
  kp.B = batch;
  kp.err = err;
  int64_t work = max_n;
  if (any_csr && batch > work) work = batch;
  if (work == 0) return FFEMB_OK;
  dim3 grid((unsigned)((work + 255) / 256), T);
  make_keys_kernel<KeyT><<<grid, 256, 0, s>>>(kp);
  FFEMB_LAUNCH_OK("make_keys");
  size_t tmp = w.sort_tmp_bytes;
  FFEMB_CUDA_OK(cub::DeviceRadixSort::SortPairs(w.sort_tmp, tmp, (const KeyT *)w.keys_a,
                                                (KeyT *)w.keys_b, (const uint32_t *)w.vals_a,
                                                w.vals_b, g.n, 0, key_bits(g.total_entries), s));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return FFEMB_OK;
}

static int update_group(const ffemb_lookup_t *lookups, int T, int batch, int out_dim, int avg,
                        const Matrix &grad, float lr, const Workspace &w, const Group &g, int *err,
                        cudaStream_t s) {
  if (g.n == 0) return FFEMB_OK;
  BwdParams bp;
  int64_t rows = 0;
  const int gq = out_dim / 4;
  bool vec = out_dim % 4 == 0 && gq <= 32 && (gq & (gq - 1)) == 0 && grad.ld % 4 == 0;
  for (int r = 0; r < grad.num_parts; ++r) vec = vec && aligned16(grad.base[r]);
  for (int t = 0; t < T; ++t) {
    const ffemb_lookup_t &l = lookups[t];
    bp.t[t] = BwdTable{l.weight, l.offsets, rows, l.col, l.bag_size};
    rows += l.num_entries;
    vec = vec && aligned16(l.weight) && l.col % 4 == 0;
  }
  bp.grad = grad;
  bp.keys = w.keys_b;
  bp.vals = w.vals_b;
  bp.part0 = w.part0;
  bp.part1 = w.part1;
  bp.status = w.status;
  bp.err = err;
  bp.n = (int)g.n;
  bp.T = T;
  bp.B = batch;
  bp.D = out_dim;
  bp.avg = avg;
  bp.lr = lr;
  // every block's "published" words start from zero
  FFEMB_CUDA_OK(cudaMemsetAsync(w.status, 0, w.status_bytes, s));
  if (wide_keys(g.total_entries))
    launch_reduce_g<uint64_t>(bp, vec, s);
  else
    launch_reduce_g<uint32_t>(bp, vec, s);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess)
    return fail(FFEMB_ERR_CUDA, std::string("launch sgd_update: ") + cudaGetErrorString(e));
  return FFEMB_OK;
}

enum Phase { kSort = 1, kUpdate = 2 };

static int backward_impl(int phases, const ffemb_lookup_t *lookups, int num_tables, int batch,
                         int out_dim, int avg, const ffemb_matrix_t *grad, float lr,
                         void *workspace, size_t workspace_bytes, int *err, void *stream) {
  const char *who = phases == kSort ? "backward_sort" : phases == kUpdate ? "sgd_update" : "backward_sgd";
  if (int rc = check_lookups(who, lookups, num_tables, batch, out_dim)) return rc;
  if (num_tables == 0 || batch == 0) return FFEMB_OK;
  Matrix gm;
  if (phases & kUpdate)
    if (int rc = to_matrix(grad, batch, &gm)) return rc;
  if (!workspace || ((uintptr_t)workspace & 255))
    return fail(FFEMB_ERR_INVALID_ARGUMENT, std::string(who) + ": workspace NULL or not 256-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  // With more than kMaxTables tables the groups are processed one after the
  // other and each needs its own region (sort results must survive until the
  // update phase).
  size_t off = 0;
  for (int t0 = 0; t0 < num_tables; t0 += kMaxTables) {
    const int T = num_tables - t0 < kMaxTables ? num_tables - t0 : kMaxTables;
    const Group g = group_of(lookups + t0, T);
    if (g.n >= 0x7FFFFF00ll)
      return fail(FFEMB_ERR_UNSUPPORTED, std::string(who) + ": more than 2^31 lookups in one launch group");
    Workspace w = carve((char *)workspace + off, g.n, g.total_entries, out_dim);
    if (off + w.total > workspace_bytes)
      return fail(FFEMB_ERR_WORKSPACE_TOO_SMALL,
                  std::string(who) + ": workspace has " + std::to_string(workspace_bytes) +
                      " bytes, needs at least " + std::to_string(off + w.total));
    off += w.total;
    if (phases & kSort) {
      int rc = wide_keys(g.total_entries)
                   ? sort_group<uint64_t>(lookups + t0, T, batch, w, g, err, s)
                   : sort_group<uint32_t>(lookups + t0, T, batch, w, g, err, s);
      if (rc) return rc;
    }
    if (phases & kUpdate)
      if (int rc = update_group(lookups + t0, T, batch, out_dim, avg, gm, lr, w, g, err, s)) return rc;
  }
  return FFEMB_OK;
}

}  // namespace ffemb

using namespace ffemb;

extern "C" int ffemb_set_sort_impl(int impl) {
  if (impl < 0 || impl > 2)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "set_sort_impl: 0 (library), 1 (own, global passes) or 2 (own, on-chip)");
  g_sort_impl.store(impl);
  return FFEMB_OK;
}
extern "C" int ffemb_get_sort_impl(void) { return g_sort_impl.load(); }

extern "C" size_t ffemb_embed_backward_workspace_bytes(int64_t total_indices,
                                                       int64_t total_entries, int out_dim) {
  if (total_indices < 0 || total_entries < 0 || out_dim <= 0) return 0;
  // One group's layout for the whole problem, plus per-group slack for
  // callers with more than kMaxTables tables.
  Workspace w = carve(nullptr, total_indices, total_entries, out_dim);
  return w.total + (1u << 20);
}

extern "C" int ffemb_embed_backward_sgd(const ffemb_lookup_t *lookups, int num_tables, int batch,
                                        int out_dim, enum AggrMode aggr,
                                        const ffemb_matrix_t *grad, float lr, void *workspace,
                                        size_t workspace_bytes, int *err_flag, void *stream) {
  if (aggr != AGGR_MODE_SUM && aggr != AGGR_MODE_AVG)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "backward_sgd: aggr must be SUM or AVG");
  return backward_impl(kSort | kUpdate, lookups, num_tables, batch, out_dim,
                       aggr == AGGR_MODE_AVG, grad, lr, workspace, workspace_bytes, err_flag,
                       stream);
}

extern "C" int ffemb_embed_backward_sort(const ffemb_lookup_t *lookups, int num_tables, int batch,
                                         int out_dim, void *workspace, size_t workspace_bytes,
                                         int *err_flag, void *stream) {
  return backward_impl(kSort, lookups, num_tables, batch, out_dim, 0, nullptr, 0.f, workspace,
                       workspace_bytes, err_flag, stream);
}

extern "C" int ffemb_embed_sgd_update(const ffemb_lookup_t *lookups, int num_tables, int batch,
                                      int out_dim, enum AggrMode aggr, const ffemb_matrix_t *grad,
                                      float lr, void *workspace, size_t workspace_bytes,
                                      void *stream) {
  if (aggr != AGGR_MODE_SUM && aggr != AGGR_MODE_AVG)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "sgd_update: aggr must be SUM or AVG");
  return backward_impl(kUpdate, lookups, num_tables, batch, out_dim, aggr == AGGR_MODE_AVG, grad,
                       lr, workspace, workspace_bytes, nullptr, stream);
}

// ---------------------------------------------------------------------------
// ffemb_embed_backward_sgd_static -- the same backward + sparse SGD, for
// callers that may not allocate device memory.
//
// Who that is: code running where cudaMalloc is forbidden or unsafe -- a
// plugin called from inside a framework that owns the whole memory pool and
// hands out no scratch, a real-time / serving loop that must not hit the
// allocator's device-wide lock, a caller between other libraries' captures.
// They cannot provide the workspace ffemb_embed_backward_sgd asks for, so
// this variant brings its own: a fixed array that is part of the library
// image.  Every GPU on which the library is loaded holds those
// FFEMB_STATIC_SCRATCH_BYTES (8 MiB) until the process exits, used or not.
//
// Consequences the caller has to know:
//   * a batch whose scratch need (ffemb_embed_backward_workspace_bytes) is
//     larger than the array is refused with FFEMB_ERR_WORKSPACE_TOO_SMALL --
//     at out_dim 128 that is roughly 250k lookups per call;
//   * the first call on each device looks the array's device address up
//     (cudaGetSymbolAddress, once per device, with that device current);
//     make that first call outside any stream capture;
//   * all calls on one device share the one array.  It still holds whatever
//     the previous call left there; the only part that has to start from
//     zero -- the blocks' "published" words of the reduce kernel -- is
//     cleared by every call on its stream before use.
//     Two calls that overlap on the GPU (different streams, or different host
//     threads) would trample each other's keys and partial sums: serialise
//     them.
// Nothing here calls cudaMalloc, cudaFree or a library routine that does.
// ---------------------------------------------------------------------------
#define FFEMB_STATIC_SCRATCH_BYTES (8u << 20)

// Lives in device memory; in host code this name is only a handle for
// cudaGetSymbolAddress, never a pointer a kernel could be given.
static __device__ __align__(256) unsigned char g_static_scratch[FFEMB_STATIC_SCRATCH_BYTES];

namespace ffemb {

constexpr int kMaxDevices = 64;
// device address of the array, per device; nullptr = not looked up yet.  Two host
// threads making their first call at once both store the same value.
static std::atomic<void *> g_static_scratch_ptr[kMaxDevices];

// Device address of g_static_scratch on the current device (looked up once).
static int static_scratch(void **ptr) {
  int dev = -1;
  FFEMB_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices)
    return fail(FFEMB_ERR_UNSUPPORTED, "backward_sgd_static: device ordinal out of range");
  void *p = g_static_scratch_ptr[dev].load(std::memory_order_acquire);
  if (!p) {
    FFEMB_CUDA_OK(cudaGetSymbolAddress(&p, g_static_scratch));
    if (!p) return fail(FFEMB_ERR_CUDA, "backward_sgd_static: the scratch array has no device address");
    g_static_scratch_ptr[dev].store(p, std::memory_order_release);
  }
  *ptr = p;
  return FFEMB_OK;
}

}  // namespace ffemb

extern "C" size_t ffemb_static_scratch_bytes(void) { return FFEMB_STATIC_SCRATCH_BYTES; }

extern "C" int ffemb_embed_backward_sgd_static(const ffemb_lookup_t *lookups, int num_tables,
                                               int batch, int out_dim, enum AggrMode aggr,
                                               const ffemb_matrix_t *grad, float lr, int *err_flag,
                                               void *stream) {
  if (aggr != AGGR_MODE_SUM && aggr != AGGR_MODE_AVG)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "backward_sgd_static: aggr must be SUM or AVG");
  if (int rc = check_lookups("backward_sgd_static", lookups, num_tables, batch, out_dim)) return rc;
  if (num_tables > kMaxTables)
    return fail(FFEMB_ERR_UNSUPPORTED, "backward_sgd_static: at most " + std::to_string(kMaxTables) +
                                           " tables per call");
  const Group g = group_of(lookups, num_tables);
  const size_t need = carve(nullptr, g.n, g.total_entries, out_dim).total;
  if (need > FFEMB_STATIC_SCRATCH_BYTES)
    return fail(FFEMB_ERR_WORKSPACE_TOO_SMALL,
                "backward_sgd_static: this batch (" + std::to_string(g.n) + " lookups, out_dim " +
                    std::to_string(out_dim) + ") needs " + std::to_string(need) +
                    " bytes of scratch, the built-in array has " +
                    std::to_string(FFEMB_STATIC_SCRATCH_BYTES) +
                    "; split the batch or use ffemb_embed_backward_sgd with your own workspace");
  void *scratch = nullptr;
  if (int rc = static_scratch(&scratch)) return rc;
  return backward_impl(kSort | kUpdate, lookups, num_tables, batch, out_dim, aggr == AGGR_MODE_AVG,
                       grad, lr, scratch, FFEMB_STATIC_SCRATCH_BYTES, err_flag, stream);
}
