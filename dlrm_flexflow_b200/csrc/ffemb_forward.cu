// ffemb_forward.cu -- forward gather-and-pool kernels (sm_100a).
//
// Stands in for embed_forward<<<>>> (reference src/ops/embedding.cu:173-197)
// with the arithmetic of the reference's CPU path
// (src/ops/embedding_avx2.cc:5-296): per bag, acc = +0, rows added one by one
// in bag order, AVG = acc * (1.0f / len).  That order is the only one that is
// bit-exact against the CPU path, so the bag is walked sequentially and the
// parallelism is over bags (warps / sub-warps) and over the row's columns
// (lanes, one float4 each); there is no cross-lane reduction.
//
// HBM-bound random-row gather: no contraction, no tensor cores.  What matters
// is (1) 128-bit row loads with every lane of a sub-warp on one row (full
// 32-byte sectors), (2) many independent row loads in flight per lane (8),
// (3) indices fetched once per warp, coalesced, and handed round by shuffle.
#include "ffemb_common.cuh"

namespace ffemb {

struct FwdTable {
  const float *W;
  const int64_t *idx;
  const int64_t *off;
  int64_t N;
  int64_t col;
  int L;
};

struct FwdParams {
  FwdTable t[kMaxTables];
  Matrix out;
  int B;
  int D;
  int avg;
  int *err;
};

constexpr int kFwdThreads = 256;
#ifndef FFEMB_FWD_UNROLL
#define FFEMB_FWD_UNROLL 8
#endif
constexpr int kUnroll = FFEMB_FWD_UNROLL;  // independent row loads in flight per lane

// ---------------------------------------------------------------------------
// One-hot (bag size 1): a warp owns 32 consecutive bags of one table.  Lane i
// loads index b0+i (one coalesced 256-byte request), then the warp makes G
// passes; in each pass its 32/G sub-warps fetch 32/G rows, the row number
// arriving by shuffle from the lane that loaded it.
// ---------------------------------------------------------------------------
template <int G>  // lanes per row; D == 4*G
__global__ void __launch_bounds__(kFwdThreads)
fwd_onehot_kernel(const __grid_constant__ FwdParams p) {
  constexpr int BPP = 32 / G;  // bags per pass
  constexpr int U = G < kUnroll ? G : kUnroll;
  const FwdTable &tb = p.t[blockIdx.y];
  const int lane = threadIdx.x & 31;
  const int b0 = (blockIdx.x * (kFwdThreads / 32) + (threadIdx.x >> 5)) * 32;
  if (b0 >= p.B) return;
  int64_t row = -1;
  if (b0 + lane < p.B) {
    row = tb.idx[b0 + lane];
    if ((uint64_t)row >= (uint64_t)tb.N) {
      row = -1;
      if (p.err) *p.err = FFEMB_ERR_INDEX_OUT_OF_RANGE;
    }
  }
  const int gl = lane % G, sub = lane / G;
  const float *Wc = tb.W + gl * 4;
  const int D = 4 * G;
#pragma unroll
  for (int p0 = 0; p0 < G; p0 += U) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int bag = (p0 + u) * BPP + sub;
      const int64_t r = __shfl_sync(0xffffffffu, row, bag);
      v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (r >= 0) v[u] = ld_row_nc(Wc + r * D);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int b = b0 + (p0 + u) * BPP + sub;
      if (b < p.B) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        add4(acc, v[u]);  // +0 + w, as the CPU path does (turns -0 into +0)
        st_row(matrix_row(p.out, b) + tb.col + gl * 4, acc);
      }
    }
  }
}

// ---------------------------------------------------------------------------
// Multi-hot / ragged bags: one sub-warp of G lanes per bag.  The bag's indices
// are read as broadcast loads (all lanes of the sub-warp, same address) one
// batch of 8 ahead of the row loads they feed; rows are loaded 8 at a time and
// added strictly in bag order.
// ---------------------------------------------------------------------------
template <int G>
__global__ void __launch_bounds__(kFwdThreads)
fwd_multihot_kernel(const __grid_constant__ FwdParams p) {
  constexpr int U = kUnroll;
  const FwdTable &tb = p.t[blockIdx.y];
  const int gl = threadIdx.x % G;
  const int b = blockIdx.x * (kFwdThreads / G) + threadIdx.x / G;
  if (b >= p.B) return;
  int64_t start, end;
  if (tb.off) {
    start = tb.off[b];
    end = tb.off[b + 1];
  } else {
    start = (int64_t)b * tb.L;
    end = start + tb.L;
  }
  const int D = 4 * G;
  const float *Wc = tb.W + gl * 4;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  bool bad = false;
  int64_t r[U];
#pragma unroll
  for (int u = 0; u < U; ++u) r[u] = (start + u < end) ? tb.idx[start + u] : 0;
  for (int64_t base = start; base < end; base += U) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (base + u < end) {
        if ((uint64_t)r[u] < (uint64_t)tb.N)
          v[u] = ld_row_nc(Wc + r[u] * D);
        else
          bad = true;
      }
    }
    const int64_t nb = base + U;
#pragma unroll
    for (int u = 0; u < U; ++u) r[u] = (nb + u < end) ? tb.idx[nb + u] : 0;
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (base + u < end) add4(acc, v[u]);
  }
  if (bad && p.err) *p.err = FFEMB_ERR_INDEX_OUT_OF_RANGE;
  if (p.avg && end > start) {
    const float inv = 1.0f / (float)(int)(end - start);
    acc.x *= inv;
    acc.y *= inv;
    acc.z *= inv;
    acc.w *= inv;
  }
  st_row(matrix_row(p.out, b) + tb.col + gl * 4, acc);
}

// ---------------------------------------------------------------------------
// Any out_dim / alignment: one warp per bag, scalar columns.  Correctness
// path for shapes outside {4,8,16,32,64,128}; not tuned.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kFwdThreads)
fwd_generic_kernel(const __grid_constant__ FwdParams p) {
  constexpr int K = 4;
  const FwdTable &tb = p.t[blockIdx.y];
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.x * (kFwdThreads / 32) + (threadIdx.x >> 5);
  if (b >= p.B) return;
  int64_t start, end;
  if (tb.off) {
    start = tb.off[b];
    end = tb.off[b + 1];
  } else {
    start = (int64_t)b * tb.L;
    end = start + tb.L;
  }
  const int D = p.D;
  float *orow = matrix_row(p.out, b) + tb.col;
  bool bad = false;
  for (int d0 = 0; d0 < D; d0 += 32 * K) {
    float acc[K];
#pragma unroll
    for (int k = 0; k < K; ++k) acc[k] = 0.f;
    for (int64_t i = start; i < end; ++i) {
      const int64_t r = tb.idx[i];
      if ((uint64_t)r >= (uint64_t)tb.N) {
        bad = true;
        continue;
      }
      const float *wr = tb.W + r * D;
#pragma unroll
      for (int k = 0; k < K; ++k) {
        const int d = d0 + k * 32 + lane;
        if (d < D) acc[k] += __ldg(wr + d);
      }
    }
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const int d = d0 + k * 32 + lane;
      if (d < D) {
        float a = acc[k];
        if (p.avg && end > start) a *= 1.0f / (float)(int)(end - start);
        orow[d] = a;
      }
    }
  }
  if (bad && p.err) *p.err = FFEMB_ERR_INDEX_OUT_OF_RANGE;
}

template <int G>
static void launch_vec(const FwdParams &p, int T, bool onehot, cudaStream_t s) {
  if (onehot) {
    const int bags_per_block = (kFwdThreads / 32) * 32;
    dim3 grid((p.B + bags_per_block - 1) / bags_per_block, T);
    fwd_onehot_kernel<G><<<grid, kFwdThreads, 0, s>>>(p);
  } else {
    const int bags_per_block = kFwdThreads / G;
    dim3 grid((p.B + bags_per_block - 1) / bags_per_block, T);
    fwd_multihot_kernel<G><<<grid, kFwdThreads, 0, s>>>(p);
  }
}

static bool aligned16(const void *p) { return ((uintptr_t)p & 15) == 0; }

}  // namespace ffemb

using namespace ffemb;

extern "C" int ffemb_embed_forward(const ffemb_lookup_t *lookups, int num_tables,
                                   int batch, int out_dim, enum AggrMode aggr,
                                   const ffemb_matrix_t *out, int *err_flag,
                                   void *stream) {
  if (int rc = check_lookups("embed_forward", lookups, num_tables, batch, out_dim)) return rc;
  if (aggr != AGGR_MODE_SUM && aggr != AGGR_MODE_AVG)
    return fail(FFEMB_ERR_INVALID_ARGUMENT,
                "embed_forward: aggr must be AGGR_MODE_SUM or AGGR_MODE_AVG");
  if (num_tables == 0 || batch == 0) return FFEMB_OK;
  FwdParams p;
  if (int rc = to_matrix(out, batch, &p.out)) return rc;
  p.B = batch;
  p.D = out_dim;
  p.avg = aggr == AGGR_MODE_AVG;
  p.err = err_flag;
  cudaStream_t s = (cudaStream_t)stream;

  const int g = out_dim / 4;
  bool vec = out_dim % 4 == 0 && g <= 32 && (g & (g - 1)) == 0 && p.out.ld % 4 == 0;
  for (int r = 0; r < p.out.num_parts; ++r) vec = vec && aligned16(p.out.base[r]);
  for (int t = 0; t < num_tables; ++t) {
    const ffemb_lookup_t &l = lookups[t];
    vec = vec && aligned16(l.weight) && l.col % 4 == 0;
  }

  for (int t0 = 0; t0 < num_tables; t0 += kMaxTables) {
    const int T = num_tables - t0 < kMaxTables ? num_tables - t0 : kMaxTables;
    bool onehot = true;
    for (int t = 0; t < T; ++t) {
      const ffemb_lookup_t &l = lookups[t0 + t];
      p.t[t] = FwdTable{l.weight, l.indices, l.offsets, l.num_entries, l.col, l.bag_size};
      onehot = onehot && !l.offsets && l.bag_size == 1;
    }
    if (!vec) {
      dim3 grid((batch + kFwdThreads / 32 - 1) / (kFwdThreads / 32), T);
      fwd_generic_kernel<<<grid, kFwdThreads, 0, s>>>(p);
    } else {
      switch (g) {
        case 1: launch_vec<1>(p, T, onehot, s); break;
        case 2: launch_vec<2>(p, T, onehot, s); break;
        case 4: launch_vec<4>(p, T, onehot, s); break;
        case 8: launch_vec<8>(p, T, onehot, s); break;
        case 16: launch_vec<16>(p, T, onehot, s); break;
        default: launch_vec<32>(p, T, onehot, s); break;
      }
    }
    FFEMB_LAUNCH_OK("embed_forward");
  }
  return FFEMB_OK;
}
