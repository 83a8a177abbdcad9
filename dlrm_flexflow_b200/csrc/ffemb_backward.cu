// ffemb_backward.cu -- deterministic backward fused with sparse SGD (sm_100a).
//
// Replaces, for weight_decay == momentum == 0 (DLRM's optimiser,
// examples/cpp/DLRM/dlrm.cc:134), three dense reference passes per step:
//   zero of the [N][D] gradient    src/runtime/initializer_kernel.cu:232-262
//   atomicAdd scatter embed_backward  src/ops/embedding.cu:199-224
//   dense sgd_update over N*D      src/runtime/optimizer_kernel.cu:23-41
// by work proportional to the lookups of the step:
//   1. make_keys   key = table_row_base + row, val = table*B + bag
//   2. stable LSD radix sort of (key, val) on the significant key bits
//   3. reduce_l1   one sub-warp per chunk of 32 sorted positions: gradient
//                  rows added in sorted (= original (bag, position)) order;
//                  a run that starts and ends inside the chunk is applied at
//                  once: W[row,:] = fmaf(-lr, gsum, W[row,:]); a run cut by a
//                  chunk boundary leaves a partial sum
//   4. reduce_l2   the chunk where such a run starts adds the partials in
//                  chunk order and applies the update.
// No atomics; the result depends only on the inputs (bitwise repeatable).
// Rows whose lookups all fall in one chunk get the exact sequential sum of
// the reference's CPU backward (src/ops/embedding.cc:54-72); longer runs are
// summed chunk-wise (still a fixed order), within 1e-5 of it.
#include <cub/device/device_radix_sort.cuh>

#include "ffemb_common.cuh"

namespace ffemb {

constexpr int kChunk = 32;  // sorted positions per sub-warp in reduce_l1
constexpr int kBwdThreads = 256;
constexpr int kBU = 8;  // positions whose rows are in flight together

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
  float *part0;  // [chunks][D] partial of the run that enters the chunk
  float *part1;  // [chunks][D] partial of the run that leaves the chunk
  int64_t n;
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

// ---------------------------------------------------------------------------
// 3. level-1 segmented reduce + SGD on complete runs.
// ---------------------------------------------------------------------------
template <typename KeyT, int G, bool VEC>
__global__ void __launch_bounds__(kBwdThreads, 2)
reduce_l1_kernel(const __grid_constant__ BwdParams p) {
  constexpr int U = kBU;
  constexpr KeyT kInv = KeyTraits<KeyT>::kInvalid;
  const KeyT *__restrict__ keys = (const KeyT *)p.keys;
  const int64_t sub = (int64_t)blockIdx.x * (kBwdThreads / G) + threadIdx.x / G;
  const int64_t i0 = sub * kChunk;
  if (i0 >= p.n) return;
  const int64_t i1 = i0 + kChunk < p.n ? i0 + kChunk : p.n;
  const int D = p.D;
  Cols<VEC> cols{(int)(threadIdx.x % G), 0, D};

  for (int d0 = 0; d0 < (VEC ? 1 : D); d0 += 128) {
    cols.d0 = d0;
    KeyT kprev = i0 > 0 ? keys[i0 - 1] : kInv;
    bool head_in_chunk = (i0 == 0) || (kprev != keys[i0]);
    const bool first_is_head = head_in_chunk;
    int t = 0;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int64_t base = i0; base < i1; base += U) {
      KeyT k[U + 1];
      uint32_t gr[U];
#pragma unroll
      for (int u = 0; u <= U; ++u) k[u] = (base + u < p.n) ? keys[base + u] : kInv;
#pragma unroll
      for (int u = 0; u < U; ++u) gr[u] = (base + u < i1) ? p.vals[base + u] : 0u;
      float4 gv[U], wv[U];
      float *wrow[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        gv[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        wv[u] = gv[u];
        wrow[u] = nullptr;
        if (base + u < i1 && k[u] != kInv) {
          while (t + 1 < p.T && (int64_t)k[u] >= p.t[t + 1].row_base) ++t;
          const BwdTable &tb = p.t[t];
          const int b = (int)(gr[u] - (uint32_t)t * (uint32_t)p.B);
          gv[u] = cols.ld_nc(matrix_row(p.grad, b) + tb.col);
          if (p.avg) {
            const float len = tb.off ? (float)(int)(tb.off[b + 1] - tb.off[b]) : (float)tb.L;
            gv[u].x /= len;
            gv[u].y /= len;
            gv[u].z /= len;
            gv[u].w /= len;
          }
          if (k[u] != k[u + 1]) {
            wrow[u] = tb.W + ((int64_t)k[u] - tb.row_base) * D;
            wv[u] = cols.ld(wrow[u]);
          }
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (base + u < i1 && k[u] != kInv) {
          const bool is_head =
              (base + u == i0) ? first_is_head : (k[u] != (u == 0 ? kprev : k[u - 1]));
          if (is_head) {
            acc = make_float4(0.f, 0.f, 0.f, 0.f);
            head_in_chunk = true;
          }
          add4(acc, gv[u]);
          if (k[u] != k[u + 1]) {  // true end of the run
            if (head_in_chunk)
              cols.st(wrow[u], sgd_row(p.lr, acc, wv[u]));
            else
              cols.st(p.part0 + sub * D, acc);
          } else if (base + u == i1 - 1) {  // run continues in the next chunk
            cols.st((head_in_chunk ? p.part1 : p.part0) + sub * D, acc);
          }
        }
      }
      kprev = k[U - 1];
    }
  }
}

// ---------------------------------------------------------------------------
// 4. level-2: the chunk in which a boundary-crossing run starts owns it.
// ---------------------------------------------------------------------------
template <typename KeyT, int G, bool VEC>
__global__ void __launch_bounds__(kBwdThreads)
reduce_l2_kernel(const __grid_constant__ BwdParams p) {
  constexpr KeyT kInv = KeyTraits<KeyT>::kInvalid;
  const KeyT *__restrict__ keys = (const KeyT *)p.keys;
  const int64_t c = (int64_t)blockIdx.x * (kBwdThreads / G) + threadIdx.x / G;
  const int64_t i0 = c * kChunk;
  const int64_t i1 = i0 + kChunk;  // first position of the next chunk
  if (i1 >= p.n) return;           // no next chunk: nothing can leave this one
  const KeyT key = keys[i1 - 1];
  if (key == kInv || keys[i1] != key) return;  // last run ends here
  // the leaving run starts in this chunk?
  if (!(keys[i0] != key || i0 == 0 || keys[i0 - 1] != key)) return;
  int t = 0;
  while (t + 1 < p.T && (int64_t)key >= p.t[t + 1].row_base) ++t;
  const int D = p.D;
  float *wrow = p.t[t].W + ((int64_t)key - p.t[t].row_base) * D;
  Cols<VEC> cols{(int)(threadIdx.x % G), 0, D};
  for (int d0 = 0; d0 < (VEC ? 1 : D); d0 += 128) {
    cols.d0 = d0;
    const float4 w = cols.ld(wrow);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    add4(acc, cols.ld(p.part1 + c * D));
    int64_t cc = c + 1;
    bool more = true;
    while (more) {
      // up to 4 following chunks at a time; chunk cc+j takes part while the
      // run reaches it, i.e. chunks cc..cc+j-1 are wholly inside the run.
      float4 v[4];
      bool in[4], whole[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int64_t s = (cc + j) * kChunk;  // first position of chunk cc+j
        const int64_t e = s + kChunk;
        in[j] = s < p.n && keys[s] == key;
        whole[j] = in[j] && e < p.n && keys[e - 1] == key && keys[e] == key;
        v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (in[j]) v[j] = cols.ld(p.part0 + (cc + j) * D);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (more) {
          if (in[j]) add4(acc, v[j]);
          more = whole[j];
        }
      }
      cc += 4;
    }
    cols.st(wrow, sgd_row(p.lr, acc, w));
  }
}

template <typename KeyT, int G, bool VEC>
static void launch_reduce(const BwdParams &p, cudaStream_t s) {
  const int64_t chunks = (p.n + kChunk - 1) / kChunk;
  const int per_block = kBwdThreads / G;
  const unsigned grid = (unsigned)((chunks + per_block - 1) / per_block);
  reduce_l1_kernel<KeyT, G, VEC><<<grid, kBwdThreads, 0, s>>>(p);
  g_launches.fetch_add(1, std::memory_order_relaxed);
  if (chunks > 1) {
    reduce_l2_kernel<KeyT, G, VEC><<<grid, kBwdThreads, 0, s>>>(p);
    g_launches.fetch_add(1, std::memory_order_relaxed);
  }
}

template <typename KeyT>
static void launch_reduce_g(const BwdParams &p, bool vec, cudaStream_t s) {
  if (!vec) return launch_reduce<KeyT, 32, false>(p, s);
  switch (p.D / 4) {
    case 1: return launch_reduce<KeyT, 1, true>(p, s);
    case 2: return launch_reduce<KeyT, 2, true>(p, s);
    case 4: return launch_reduce<KeyT, 4, true>(p, s);
    case 8: return launch_reduce<KeyT, 8, true>(p, s);
    case 16: return launch_reduce<KeyT, 16, true>(p, s);
    default: return launch_reduce<KeyT, 32, true>(p, s);
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
  const int64_t chunks = (n + kChunk - 1) / kChunk;
  char *p = (char *)base;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void *q = p ? p + off : nullptr;
    off += align256(bytes);
    return q;
  };
  w.keys_a = take(n * w.key_bytes);
  w.keys_b = take(n * w.key_bytes);
  w.vals_a = (uint32_t *)take(n * 4);
  w.vals_b = (uint32_t *)take(n * 4);
  w.sort_tmp_bytes = sort_tmp_bytes(n, wide);
  w.sort_tmp = take(w.sort_tmp_bytes);
  w.part0 = (float *)take((size_t)chunks * D * 4);
  w.part1 = (float *)take((size_t)chunks * D * 4);
  w.total = off;
  return w;
}

static bool aligned16(const void *p) { return ((uintptr_t)p & 15) == 0; }

struct Group {  // one batched launch group (<= kMaxTables tables)
  int64_t n;              // lookups in the group
  int64_t total_entries;  // rows in the group
};

static int check_lookups(const char *who, const ffemb_lookup_t *lookups, int num_tables,
                         int batch, int out_dim) {
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
                        const Matrix &grad, float lr, const Workspace &w, const Group &g,
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
  bp.n = g.n;
  bp.T = T;
  bp.B = batch;
  bp.D = out_dim;
  bp.avg = avg;
  bp.lr = lr;
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
      if (int rc = update_group(lookups + t0, T, batch, out_dim, avg, gm, lr, w, g, s)) return rc;
  }
  return FFEMB_OK;
}

}  // namespace ffemb

using namespace ffemb;

extern "C" size_t ffemb_embed_backward_workspace_bytes(int64_t total_indices,
                                                       int64_t total_entries, int out_dim) {
  if (total_indices < 0 || total_entries < 0 || out_dim <= 0) return 0;
  // One group's layout for the whole problem, plus per-group alignment slack
  // for callers with more than kMaxTables tables.
  Workspace w = carve(nullptr, total_indices, total_entries, out_dim);
  return w.total + 16 * 256 * 8;
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
