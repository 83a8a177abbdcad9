// ffemb_sort.cuh -- hand-written key sort for the backward (sm_100a).
//
// What has to be sorted is special: the lookups of one step arrive grouped by
// table (table t's positions are contiguous) and only have to be ordered by
// row INSIDE each table's segment; a table with N_t rows needs only
// floor(log2 N_t)+1 key bits.  So this is a segmented, stable LSD radix sort
// over per-table row numbers in which every table takes only the passes its
// own size needs: with 9-bit digits the MLPerf-Terabyte tables take 3 passes
// for the eight 40M-row tables, 2 for the mid-sized ones and 1 for the eight
// tables of at most 511 rows, instead of 4 passes over everything.  The last
// pass of a table rewrites the row number into the global key
// (row_base + row) that the reduce kernels expect.
//
// Per pass three small kernels over the tables still active:
//   rs_hist     per 2048-key tile: digit histogram in shared memory
//   rs_scan     per table: exclusive scan of its counts in digit-major order
//   rs_scatter  per tile: stable rank of every key among equal digits
//               (__match_any_sync inside the warp, per-warp counters across
//               rounds, prefix over the 8 warps), then scatter.
// Stability: a tile's keys are taken in the order warp, round, lane, which is
// their order in memory; tiles and digits are ordered by the scan.
//
// Buffers ping-pong per table; a table that needs an odd number of passes
// starts in buffer 0, an even number in buffer 1, so all of them finish in
// buffer 1.  An out-of-range index is carried as row 0xFFFFFFFF: all-ones in
// every digit, so it ends behind the table's valid keys and is written out as
// the invalid key.  32-bit keys only (fewer than 2^32-1 rows in the launch
// group); wider keys use the library sort.
#pragma once

#include "ffemb_common.cuh"

namespace ffemb {

constexpr int kSortThreads = 256;
constexpr int kSortItems = 8;
constexpr int kSortTile = kSortThreads * kSortItems;  // 2048 keys per tile
constexpr int kMaxRadixBits = 11;
constexpr uint32_t kRowInvalid = 0xFFFFFFFFu;

struct SortTable {      // one active table in one pass
  uint32_t pos_base;    // first position of the table's segment
  uint32_t n;           // lookups of the table
  uint32_t row_base;    // global key of its row 0
  uint32_t tile_base;   // first tile of this table among the pass's tiles
  uint32_t hist_base;   // first of its ntiles << radix_bits histogram entries
  uint32_t ntiles;
  uint32_t src;         // buffer (0/1) this pass reads for the table
  uint32_t last;        // the table's last pass: write global keys
};

struct SortParams {
  SortTable t[kMaxTables];
  uint32_t *keys[2];
  uint32_t *vals[2];
  uint32_t *hist;
  int num_tables;
  int shift;
  int radix_bits;
};

struct RowTable {       // make_rows view of one table
  const int64_t *idx;
  const int64_t *off;
  int64_t N;
  uint32_t pos_base;
  uint32_t n;
  int L;
  int start;            // buffer the table's rows are written to
};

struct RowParams {
  RowTable t[kMaxTables];
  uint32_t *keys[2];
  uint32_t *vals[2];
  int B;
  int *err;
};

// rows (not yet global keys) + values, into each table's start buffer
__global__ void __launch_bounds__(256) make_rows_kernel(const __grid_constant__ RowParams p) {
  const RowTable &tb = p.t[blockIdx.y];
  uint32_t *keys = p.keys[tb.start] + tb.pos_base;
  uint32_t *vals = p.vals[tb.start] + tb.pos_base;
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  bool bad = false;
  if (!tb.off) {
    if (i < tb.n) {
      const int64_t row = tb.idx[i];
      const uint32_t bag = tb.L == 1 ? (uint32_t)i : (uint32_t)(i / tb.L);
      uint32_t k = kRowInvalid;
      if ((uint64_t)row < (uint64_t)tb.N)
        k = (uint32_t)row;
      else
        bad = true;
      keys[i] = k;
      vals[i] = blockIdx.y * (uint32_t)p.B + bag;
    }
  } else if (i < p.B) {
    const int64_t s = tb.off[i], e = tb.off[i + 1];
    for (int64_t j = s; j < e; ++j) {
      const int64_t row = tb.idx[j];
      uint32_t k = kRowInvalid;
      if ((uint64_t)row < (uint64_t)tb.N)
        k = (uint32_t)row;
      else
        bad = true;
      keys[j] = k;
      vals[j] = blockIdx.y * (uint32_t)p.B + (uint32_t)i;
    }
  }
  if (bad && p.err) *p.err = FFEMB_ERR_INDEX_OUT_OF_RANGE;
}

__device__ __forceinline__ int sort_table_of(const SortParams &p, uint32_t tile) {
  int t = 0;
  while (t + 1 < p.num_tables && tile >= p.t[t + 1].tile_base) ++t;
  return t;
}

__global__ void __launch_bounds__(kSortThreads) rs_hist_kernel(const __grid_constant__ SortParams p) {
  __shared__ uint32_t h[1 << kMaxRadixBits];
  const int radix = 1 << p.radix_bits;
  const SortTable &tb = p.t[sort_table_of(p, blockIdx.x)];
  const uint32_t tile = blockIdx.x - tb.tile_base;
  for (int d = threadIdx.x; d < radix; d += kSortThreads) h[d] = 0;
  __syncthreads();
  const uint32_t *src = p.keys[tb.src] + tb.pos_base;
  const uint32_t e0 = tile * kSortTile + threadIdx.x;
#pragma unroll
  for (int k = 0; k < kSortItems; ++k) {
    const uint32_t e = e0 + k * kSortThreads;
    if (e < tb.n) atomicAdd(&h[(src[e] >> p.shift) & (radix - 1)], 1u);
  }
  __syncthreads();
  for (int d = threadIdx.x; d < radix; d += kSortThreads)
    p.hist[tb.hist_base + tile * (uint32_t)radix + d] = h[d];
}

// One block per table.  The histogram is stored [tile][digit] (coalesced for
// every kernel); the exclusive scan runs in digit-major order -- all tiles of
// digit 0, then of digit 1, ... -- which is the order of the sorted segment.
// Thread tid owns digits tid, tid+256, ...: it totals a digit over the tiles,
// the block scans the totals, then the thread hands each tile its offset.
__global__ void __launch_bounds__(kSortThreads) rs_scan_kernel(const __grid_constant__ SortParams p) {
  __shared__ uint32_t part[kSortThreads];
  const SortTable &tb = p.t[blockIdx.x];
  const uint32_t radix = 1u << p.radix_bits;
  uint32_t *h = p.hist + tb.hist_base;
  uint32_t carry = 0;  // keys with a digit below the current block of 256 digits
  for (uint32_t d0 = 0; d0 < radix; d0 += kSortThreads) {
    const uint32_t d = d0 + threadIdx.x;
    uint32_t tot = 0;
    if (d < radix) {
#pragma unroll 8
      for (uint32_t tile = 0; tile < tb.ntiles; ++tile) tot += h[tile * radix + d];
    }
    part[threadIdx.x] = tot;
    __syncthreads();
    for (int off = 1; off < kSortThreads; off <<= 1) {
      const uint32_t v = (int)threadIdx.x >= off ? part[threadIdx.x - off] : 0u;
      __syncthreads();
      part[threadIdx.x] += v;
      __syncthreads();
    }
    uint32_t run = carry + part[threadIdx.x] - tot;  // exclusive over digits
    carry += part[kSortThreads - 1];
    if (d < radix) {
      for (uint32_t tile = 0; tile < tb.ntiles; ++tile) {
        const uint32_t v = h[tile * radix + d];
        h[tile * radix + d] = run;
        run += v;
      }
    }
    __syncthreads();  // part[] is reused by the next block of digits
  }
}

__global__ void __launch_bounds__(kSortThreads) rs_scatter_kernel(const __grid_constant__ SortParams p) {
  __shared__ uint16_t cnt[kSortThreads / 32][1 << kMaxRadixBits];  // per-warp digit counters
  __shared__ uint32_t base[1 << kMaxRadixBits];                    // tile's first slot per digit
  const int radix = 1 << p.radix_bits;
  const uint32_t mask = (uint32_t)radix - 1u;
  const SortTable &tb = p.t[sort_table_of(p, blockIdx.x)];
  const uint32_t tile = blockIdx.x - tb.tile_base;
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int d = threadIdx.x; d < radix; d += kSortThreads) {
    base[d] = p.hist[tb.hist_base + tile * (uint32_t)radix + d];
#pragma unroll
    for (int ww = 0; ww < kSortThreads / 32; ++ww) cnt[ww][d] = 0;
  }
  __syncthreads();
  const uint32_t *ksrc = p.keys[tb.src] + tb.pos_base;
  const uint32_t *vsrc = p.vals[tb.src] + tb.pos_base;
  // memory order inside the tile: warp, round, lane
  const uint32_t e0 = tile * kSortTile + (uint32_t)w * (kSortItems * 32) + lane;
  uint32_t row[kSortItems];
  uint32_t rk[kSortItems];
#pragma unroll
  for (int r = 0; r < kSortItems; ++r) {
    const uint32_t e = e0 + r * 32;
    // slots past the end of the segment come last in this order, so giving
    // them the all-ones digit cannot disturb the rank of a real key
    row[r] = e < tb.n ? ksrc[e] : kRowInvalid;
    const uint32_t d = (row[r] >> p.shift) & mask;
    const unsigned peers = __match_any_sync(0xffffffffu, d);
    const unsigned lt = peers & ((1u << lane) - 1u);
    uint32_t b = 0;
    if (lt == 0) {  // lowest lane of its digit group
      b = cnt[w][d];
      cnt[w][d] = (uint16_t)(b + __popc(peers));
    }
    b = __shfl_sync(0xffffffffu, b, __ffs(peers) - 1);
    rk[r] = b + __popc(lt);
    __syncwarp();
  }
  __syncthreads();
  for (int d = threadIdx.x; d < radix; d += kSortThreads) {
    uint32_t run = 0;
#pragma unroll
    for (int ww = 0; ww < kSortThreads / 32; ++ww) {
      const uint32_t c = cnt[ww][d];
      cnt[ww][d] = (uint16_t)run;
      run += c;
    }
  }
  __syncthreads();
  uint32_t *kdst = p.keys[tb.src ^ 1u] + tb.pos_base;
  uint32_t *vdst = p.vals[tb.src ^ 1u] + tb.pos_base;
#pragma unroll
  for (int r = 0; r < kSortItems; ++r) {
    const uint32_t e = e0 + r * 32;
    if (e < tb.n) {
      const uint32_t d = (row[r] >> p.shift) & mask;
      const uint32_t dst = base[d] + cnt[w][d] + rk[r];
      uint32_t k = row[r];
      if (tb.last) k = k == kRowInvalid ? 0xFFFFFFFFu : tb.row_base + k;
      kdst[dst] = k;
      vdst[dst] = vsrc[e];
    }
  }
}

inline size_t own_sort_hist_bytes(int64_t n) {
  return (size_t)((n + kSortTile - 1) / kSortTile + kMaxTables) * (size_t)(1 << kMaxRadixBits) * sizeof(uint32_t);
}

inline int bits_for(int64_t n_rows) {  // floor(log2 n)+1: every row < 2^bits - 1
  int b = 1;
  while (b < 63 && (n_rows >> b) != 0) ++b;
  return b;
}

// keys_b / vals_b (buffer 1) hold the result.  `hist` has own_sort_hist_bytes(n).
inline int own_sort_group(const ffemb_lookup_t *lookups, int T, int batch, uint32_t *keys_a,
                          uint32_t *keys_b, uint32_t *vals_a, uint32_t *vals_b, uint32_t *hist,
                          int *err, cudaStream_t s) {
  int bits[kMaxTables], passes_of[kMaxTables];
  int maxbits = 1;
  for (int t = 0; t < T; ++t) {
    bits[t] = bits_for(lookups[t].num_entries);
    if (bits[t] > 32) return fail(FFEMB_ERR_UNSUPPORTED, "own sort: table with 2^32 rows or more");
    if (lookups[t].num_indices > 0 && bits[t] > maxbits) maxbits = bits[t];
  }
  const int passes = (maxbits + kMaxRadixBits - 1) / kMaxRadixBits;
  int rb = (maxbits + passes - 1) / passes;
  if (rb < 4) rb = 4;
  RowParams rp;
  int64_t pos = 0, max_work = 0;
  bool any_csr = false;
  for (int t = 0; t < T; ++t) {
    const ffemb_lookup_t &l = lookups[t];
    passes_of[t] = (bits[t] + rb - 1) / rb;
    rp.t[t] = RowTable{l.indices, l.offsets, l.num_entries, (uint32_t)pos, (uint32_t)l.num_indices,
                       l.bag_size, (passes_of[t] & 1) ? 0 : 1};
    pos += l.num_indices;
    if (l.offsets) any_csr = true;
    if (l.num_indices > max_work) max_work = l.num_indices;
  }
  if (any_csr && batch > max_work) max_work = batch;
  if (max_work == 0) return FFEMB_OK;
  rp.keys[0] = keys_a;
  rp.keys[1] = keys_b;
  rp.vals[0] = vals_a;
  rp.vals[1] = vals_b;
  rp.B = batch;
  rp.err = err;
  make_rows_kernel<<<dim3((unsigned)((max_work + 255) / 256), T), 256, 0, s>>>(rp);
  FFEMB_LAUNCH_OK("make_rows");
  for (int pass = 0; pass < passes; ++pass) {
    SortParams sp;
    sp.keys[0] = keys_a;
    sp.keys[1] = keys_b;
    sp.vals[0] = vals_a;
    sp.vals[1] = vals_b;
    sp.hist = hist;
    sp.shift = pass * rb;
    sp.radix_bits = rb;
    int na = 0;
    uint32_t tiles = 0, hist_entries = 0;
    int64_t rows = 0, p0 = 0;
    for (int t = 0; t < T; ++t) {
      const ffemb_lookup_t &l = lookups[t];
      if (passes_of[t] > pass && l.num_indices > 0) {
        SortTable &st = sp.t[na++];
        st.pos_base = (uint32_t)p0;
        st.n = (uint32_t)l.num_indices;
        st.row_base = (uint32_t)rows;
        st.tile_base = tiles;
        st.hist_base = hist_entries;
        st.ntiles = (uint32_t)((l.num_indices + kSortTile - 1) / kSortTile);
        st.src = (uint32_t)((rp.t[t].start + pass) & 1);
        st.last = pass == passes_of[t] - 1;
        tiles += st.ntiles;
        hist_entries += st.ntiles << rb;
      }
      rows += l.num_entries;
      p0 += l.num_indices;
    }
    if (na == 0) continue;
    sp.num_tables = na;
    rs_hist_kernel<<<tiles, kSortThreads, 0, s>>>(sp);
    FFEMB_LAUNCH_OK("rs_hist");
    rs_scan_kernel<<<na, kSortThreads, 0, s>>>(sp);
    FFEMB_LAUNCH_OK("rs_scan");
    rs_scatter_kernel<<<tiles, kSortThreads, 0, s>>>(sp);
    FFEMB_LAUNCH_OK("rs_scatter");
  }
  return FFEMB_OK;
}

}  // namespace ffemb
