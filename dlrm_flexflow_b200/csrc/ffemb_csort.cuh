// ffemb_csort.cuh -- the backward's key sort in ONE launch (sm_100a).
//
// The lookups of a step arrive grouped by table and only have to be ordered
// by row inside each table's segment (ffemb_sort.cuh explains why).  When a
// table's segment has at most 65536 lookups -- every DLRM shape with batch <=
// 65536 and one-hot bags, and every shard of a bigger job -- one thread-block
// CLUSTER of up to 8 CTAs sorts the table start to finish in a single launch:
// CTA r owns a slab of up to 16384 consecutive slots of the segment's current
// order and keeps them in shared memory as (row u32, bag u16).
//
//   pass k   stable LSD radix pass on digit k (<= 9 bits; a table takes only
//            ceil(bits(N_t) / 9) passes):
//     load     pass 0 only: the CTA's slab of the int64 index stream, coalesced,
//              with its range check (later passes find their slab in shared
//              memory, where the previous pass left it)
//     count    per-warp digit histograms (packed shared-memory atomics)
//     barrier  cluster-wide; every CTA reads the other CTAs' digit totals
//              through distributed shared memory and derives, per digit, its
//              first slot in the table's sorted order
//     reorder  stable rank among equal digits in the order warp, row, lane
//              (= position order; the lanes of a row that share a digit find
//              each other with one ballot per digit bit), scatter inside the
//              CTA's OWN shared memory so that every digit's keys sit together
//     barrier  cluster-wide: every CTA is done reading its slab
//     copy     each digit's run goes to the slab of the CTA that owns its slots:
//              consecutive lanes hold consecutive slots of a run, so these
//              are coalesced distributed-shared-memory stores (the last pass
//              writes global keys row_base + row and values to memory)
//     barrier  cluster-wide
//
// Elements cross CTAs only as coalesced runs: scattering single 4-byte
// elements into another CTA's shared memory costs a packet each (measured: 170
// us for the MLPerf-Terabyte step done that way, against 80 us now).  Against
// the multi-launch sorts (7-10 dependent launches, latency-bound on 1.7 M
// keys) this is one launch of about 100 CTAs that reads the indices once and
// writes nothing but the result.  A launch with few tables spreads each over
// up to 8 CTAs (the sort is latency-bound per CTA); tiny segments stay in one.
// Tables whose segments do not fit fall back to the library sort.
#pragma once

#include <cooperative_groups.h>

#include "ffemb_sort.cuh"

namespace ffemb {

namespace cg = cooperative_groups;

constexpr int kCsThreads = 512;
constexpr int kCsWarps = kCsThreads / 32;
constexpr int kCsSlab = 16384;  // slots per CTA
constexpr int kCsMaxCluster = 4;
constexpr int kCsMaxBits = 9;
constexpr int kCsRadix = 1 << kCsMaxBits;
constexpr int kCsPerWarp = kCsSlab / kCsWarps;      // 1024 slots per warp
constexpr int kCsPerThread = kCsSlab / kCsThreads;  // 32 slots per thread
#ifndef FFEMB_CS_DSMEM
#define FFEMB_CS_DSMEM 1
#endif
// between passes the digit runs move CTA to CTA through distributed shared
// memory (1) or through L2 (0)
constexpr bool kCsDsmem = FFEMB_CS_DSMEM != 0;
constexpr size_t kCsSmemBytes = (size_t)2 * kCsSlab * 4      // rows: slab + reordered
                                + (size_t)2 * kCsSlab * 2    // bags: slab + reordered
                                + (size_t)kCsWarps * kCsRadix * 2  // per-warp digit counters
                                + (size_t)kCsRadix * 4 * 3   // CTA digit totals, local and global first slots
                                + 64 * 4;                    // scan scratch

struct CsTable {
  const int64_t *idx;
  const int64_t *off;
  int64_t N;
  uint32_t pos_base;  // first position of the table's segment in the output
  uint32_t n;         // lookups of the table
  uint32_t row_base;  // global key of its row 0
  int L;
  int passes;  // radix passes this table takes
  int rb;      // bits per pass
};

struct CsParams {
  CsTable t[kMaxTables];
  uint32_t *keys_out;
  uint32_t *vals_out;
  uint2 *tmp;  // [n] (row, bag) records between passes
  int slab_log2;  // slots per CTA in this launch = 1 << slab_log2 (<= kCsSlab)
  int B;
  int *err;
};

#ifdef FFEMB_CS_PROFILE
__device__ long long g_cs_prof[16];
#define CS_MARK(i)                                                        \
  do {                                                                    \
    if (blockIdx.x == FFEMB_CS_PROFILE && threadIdx.x == 0) {             \
      const long long _t = clock64();                                     \
      atomicAdd((unsigned long long *)&g_cs_prof[i], (unsigned long long)(_t - _t0)); \
      _t0 = _t;                                                           \
    }                                                                     \
  } while (0)
#else
#define CS_MARK(i)
#endif

#ifndef FFEMB_CS_MAXNREG
#define FFEMB_CS_MAXNREG 64
#endif
__global__ void __maxnreg__(FFEMB_CS_MAXNREG) cluster_sort_kernel(const __grid_constant__ CsParams p) {
  extern __shared__ __align__(16) unsigned char cs_smem[];
  uint32_t *s_rows = (uint32_t *)cs_smem;                                    // slab
  uint32_t *s_rows2 = s_rows + kCsSlab;                                      // reordered by digit
  uint16_t *s_bags = (uint16_t *)(cs_smem + (size_t)2 * kCsSlab * 4);
  uint16_t *s_bags2 = s_bags + kCsSlab;
  uint16_t *const s_hist = (uint16_t *)(cs_smem + (size_t)2 * kCsSlab * 6);  // [kCsWarps][kCsRadix]
  uint32_t *const s_cnt = (uint32_t *)(cs_smem + (size_t)2 * kCsSlab * 6 + (size_t)kCsWarps * kCsRadix * 2);
  uint32_t *const s_lbase = s_cnt + kCsRadix;   // first slot of a digit in s_rows2
  uint32_t *const s_gbase = s_lbase + kCsRadix; // this CTA's first slot of a digit in the sorted segment
  uint32_t *const s_scan = s_gbase + kCsRadix;  // [2][kCsWarps]

#ifdef FFEMB_CS_PROFILE
  long long _t0 = clock64();
#endif
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned C = cluster.num_blocks();
  const unsigned r = cluster.block_rank();
  const int t = blockIdx.x / C;
  const CsTable &tb = p.t[t];
  const uint32_t n = tb.n;
  if (n == 0) return;  // the same decision in every CTA of the cluster
  const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
  const uint32_t slab = 1u << p.slab_log2;
  const uint32_t slab0 = r * slab;
  const uint32_t m = n > slab0 ? (n - slab0 < slab ? n - slab0 : slab) : 0u;
  uint2 *const tmp = p.tmp + tb.pos_base;
  bool bad = false;

  if (tb.off) {
    // CSR bags: (row, bag) of every position goes to the between-pass arrays
    // first, one thread per bag; slots no bag covers stay invalid.
    for (uint32_t e = tid; e < m; e += kCsThreads) tmp[slab0 + e] = make_uint2(kRowInvalid, 0u);
    cluster.sync();
    const uint32_t B = (uint32_t)p.B;
    const uint32_t per = (B + C - 1) / C;
    const uint32_t b1 = (r + 1) * per < B ? (r + 1) * per : B;
    for (uint32_t b = r * per + tid; b < b1; b += kCsThreads) {
      const int64_t s = tb.off[b], e = tb.off[b + 1];
      for (int64_t j = s; j < e && j < (int64_t)n; ++j) {
        const int64_t row = tb.idx[j];
        uint32_t k = kRowInvalid;
        if ((uint64_t)row < (uint64_t)tb.N)
          k = (uint32_t)row;
        else
          bad = true;
        tmp[j] = make_uint2(k, b);
      }
    }
  }
  cluster.sync();  // every CTA of the cluster is running; CSR rows are in place
  CS_MARK(0);

  // warp w takes slots [w*Q, (w+1)*Q) of the slab, Q a multiple of 32: all 16
  // warps share a short slab instead of the first few doing 1024 slots each
  const uint32_t Q = ((m + kCsWarps - 1) / kCsWarps + 31u) & ~31u;
  const uint32_t e0 = (uint32_t)w * Q;
  const int nrows = (int)(Q / 32);
  for (int pass = 0; pass < tb.passes; ++pass) {
    const int shift = pass * tb.rb;
    const int rb = tb.rb;
    const uint32_t radix = 1u << rb, mask = radix - 1u;
    const bool last = pass == tb.passes - 1;
    uint16_t *hist = s_hist + w * kCsRadix;

    // ---- load the slab, 16 independent loads in flight per thread -------------
    // (a single-CTA cluster keeps its slab in shared memory from pass to pass)
    if (pass == 0 && !tb.off) {
      const int64_t *src = tb.idx + slab0;
#pragma unroll
      for (int j0 = 0; j0 < kCsPerThread; j0 += 16) {
        if ((uint32_t)j0 * kCsThreads >= m) break;  // uniform: short slabs stop early
        int64_t v[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const uint32_t e = (uint32_t)(j0 + j) * kCsThreads + tid;
          v[j] = e < m ? __ldg(src + e) : 0;
        }
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const uint32_t e = (uint32_t)(j0 + j) * kCsThreads + tid;
          if (e < m) {
            uint32_t k = kRowInvalid;
            if ((uint64_t)v[j] < (uint64_t)tb.N)
              k = (uint32_t)v[j];
            else
              bad = true;
            const uint32_t i = slab0 + e;
            s_rows[e] = k;
            s_bags[e] = (uint16_t)(tb.L == 1 ? i : i / (uint32_t)tb.L);
          }
        }
      }
    } else if (pass == 0 || (C > 1 && !kCsDsmem)) {
#pragma unroll
      for (int j0 = 0; j0 < kCsPerThread; j0 += 16) {
        if ((uint32_t)j0 * kCsThreads >= m) break;  // uniform: short slabs stop early
        uint2 v[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const uint32_t e = (uint32_t)(j0 + j) * kCsThreads + tid;
          v[j] = e < m ? __ldcg(tmp + slab0 + e) : make_uint2(0u, 0u);
        }
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const uint32_t e = (uint32_t)(j0 + j) * kCsThreads + tid;
          if (e < m) {
            s_rows[e] = v[j].x;
            s_bags[e] = (uint16_t)v[j].y;
          }
        }
      }
    }
    for (uint32_t d = lane; d < radix; d += 32) hist[d] = 0;
    __syncthreads();
    CS_MARK(1);

    // ---- count: digit histogram of this warp's slots (order does not matter
    // here: 16-bit counters, two per word, bumped with shared-memory atomics)
    {
      unsigned *hist32 = (unsigned *)hist;
#pragma unroll 8
      for (int k = 0; k < nrows; ++k) {
        const uint32_t e = e0 + k * 32 + lane;
        if (e < m) {
          const uint32_t d = (s_rows[e] >> shift) & mask;
          atomicAdd(hist32 + (d >> 1), (d & 1u) ? 0x10000u : 1u);
        }
      }
    }
    __syncthreads();
    CS_MARK(2);
    // exclusive prefix over the warps of this CTA, CTA total per digit
    if ((uint32_t)tid < radix) {
      uint32_t run = 0;
#pragma unroll
      for (int ww = 0; ww < kCsWarps; ++ww) {
        const uint32_t c = s_hist[ww * kCsRadix + tid];
        s_hist[ww * kCsRadix + tid] = (uint16_t)run;
        run += c;
      }
      s_cnt[tid] = run;
    }
    CS_MARK(3);
    cluster.sync();  // every CTA's digit totals are complete
    CS_MARK(4);

    // ---- first slot of every digit: in the reordered slab, in the sorted segment
    uint32_t tot = 0, pre = 0, own = 0;
    if ((uint32_t)tid < radix) {
      for (unsigned rr = 0; rr < C; ++rr) {
        const uint32_t c = cluster.map_shared_rank(s_cnt, rr)[tid];
        if (rr < r) pre += c;
        if (rr == r) own = c;
        tot += c;
      }
    }
    uint32_t incl = tot, incl_own = own;  // inclusive scans over the 512 threads (= digits)
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
      const uint32_t v2 = __shfl_up_sync(0xffffffffu, incl_own, o);
      if (lane >= o) {
        incl += v;
        incl_own += v2;
      }
    }
    if (lane == 31) {
      s_scan[w] = incl;
      s_scan[kCsWarps + w] = incl_own;
    }
    __syncthreads();
    if (w == 0) {
      uint32_t v = lane < kCsWarps ? s_scan[lane] : 0u;
      uint32_t v2 = lane < kCsWarps ? s_scan[kCsWarps + lane] : 0u;
#pragma unroll
      for (int o = 1; o < kCsWarps; o <<= 1) {
        const uint32_t u = __shfl_up_sync(0xffffffffu, v, o);
        const uint32_t u2 = __shfl_up_sync(0xffffffffu, v2, o);
        if (lane >= o) {
          v += u;
          v2 += u2;
        }
      }
      if (lane < kCsWarps) {
        s_scan[lane] = v;
        s_scan[kCsWarps + lane] = v2;
      }
    }
    __syncthreads();
    if ((uint32_t)tid < radix) {
      s_gbase[tid] = incl - tot + (w ? s_scan[w - 1] : 0u) + pre;
      s_lbase[tid] = incl_own - own + (w ? s_scan[kCsWarps + w - 1] : 0u);
    }
    __syncthreads();
    CS_MARK(5);

    // ---- reorder inside the CTA: stable rank among equal digits ---------------
    for (int k = 0; k < nrows; ++k) {
      const uint32_t e = e0 + k * 32 + lane;
      const bool valid = e < m;
      unsigned peers = __ballot_sync(0xffffffffu, valid);
      if (peers == 0) break;
      const uint32_t row = valid ? s_rows[e] : 0u;
      const uint32_t d = (row >> shift) & mask;
      // all ballots of the row first (they issue back to back), then the combine
      unsigned bal[kCsMaxBits];
#pragma unroll
      for (int bit = 0; bit < kCsMaxBits; ++bit)
        bal[bit] = bit < rb ? __ballot_sync(0xffffffffu, (d >> bit) & 1u) : 0u;
#pragma unroll
      for (int bit = 0; bit < kCsMaxBits; ++bit)
        if (bit < rb) peers &= ((d >> bit) & 1u) ? bal[bit] : ~bal[bit];
      const unsigned lt = peers & ((1u << lane) - 1u);
      uint32_t b = 0;
      if (valid && lt == 0) {  // lowest lane of its digit group
        b = hist[d];
        hist[d] = (uint16_t)(b + __popc(peers));
      }
      b = __shfl_sync(0xffffffffu, b, (__ffs(peers) - 1) & 31);
      if (valid) {
        const uint32_t ls = s_lbase[d] + b + __popc(lt);
        s_rows2[ls] = row;
        s_bags2[ls] = s_bags[e];
      }
      __syncwarp();
    }
    __syncthreads();
    CS_MARK(6);

    if (C == 1 && !last) {  // the reordered slab IS the next pass's input
      uint32_t *tr = s_rows;
      s_rows = s_rows2;
      s_rows2 = tr;
      uint16_t *tg = s_bags;
      s_bags = s_bags2;
      s_bags2 = tg;
      continue;
    }
    // ---- copy every digit's run to its slots, coalesced ------------------------
    if (kCsDsmem && !last) {
      // The runs go straight into the slab (shared memory) of whichever CTA
      // owns their slots: consecutive lanes hold consecutive slots of a run,
      // so these are coalesced distributed-shared-memory stores.  Every CTA
      // of the cluster must be done reading its slab first.
      cluster.sync();
#pragma unroll 8
      for (int j = 0; j < kCsPerThread; ++j) {
        if ((uint32_t)j * kCsThreads >= m) break;
        const uint32_t e = (uint32_t)j * kCsThreads + tid;
        if (e < m) {
          const uint32_t row = s_rows2[e];
          const uint32_t d = (row >> shift) & mask;
          const uint32_t g = s_gbase[d] + (e - s_lbase[d]);
          const unsigned dst = g >> p.slab_log2;
          const uint32_t o = g & (slab - 1u);
          cluster.map_shared_rank(s_rows, dst)[o] = row;
          cluster.map_shared_rank(s_bags, dst)[o] = s_bags2[e];
        }
      }
      CS_MARK(7);
      cluster.sync();  // the next pass's slabs are complete; digit totals may be overwritten
      CS_MARK(8);
      continue;
    }
#pragma unroll 8
    for (int j = 0; j < kCsPerThread; ++j) {
        if ((uint32_t)j * kCsThreads >= m) break;
      const uint32_t e = (uint32_t)j * kCsThreads + tid;
      if (e < m) {
        const uint32_t row = s_rows2[e];
        const uint32_t d = (row >> shift) & mask;
        const uint32_t g = s_gbase[d] + (e - s_lbase[d]);
        if (last) {
          p.keys_out[tb.pos_base + g] = row == kRowInvalid ? 0xFFFFFFFFu : tb.row_base + row;
          p.vals_out[tb.pos_base + g] = (uint32_t)t * (uint32_t)p.B + s_bags2[e];
        } else {
          tmp[g] = make_uint2(row, (uint32_t)s_bags2[e]);
        }
      }
    }
    CS_MARK(7);
    // the pass's output is visible to the cluster; digit totals may be overwritten;
    // after the last pass: nobody reads this CTA's shared memory any more
    cluster.sync();
    CS_MARK(8);
  }
  if (bad && p.err) *p.err = FFEMB_ERR_INDEX_OUT_OF_RANGE;
}

// Sorts the group with cluster_sort_kernel if every table's segment fits on
// chip.  Returns FFEMB_OK and sets *done; *done == false means "not eligible,
// use another sort".
inline int cluster_sort_group(const ffemb_lookup_t *lookups, int T, int batch, uint32_t *keys_out,
                              uint32_t *vals_out, uint2 *tmp, int *err, cudaStream_t s, bool *done) {
  *done = false;
  if (batch > 65536) return FFEMB_OK;  // bags are carried as 16-bit numbers
  int64_t max_n = 0;
  for (int t = 0; t < T; ++t) {
    if (lookups[t].num_indices > (int64_t)kCsSlab * kCsMaxCluster) return FFEMB_OK;
    if (bits_for(lookups[t].num_entries) > 32) return FFEMB_OK;
    if (lookups[t].num_indices > max_n) max_n = lookups[t].num_indices;
  }
  *done = true;
  if (max_n == 0) return FFEMB_OK;
  static std::atomic<int> configured[64];
  int dev = 0;
  FFEMB_CUDA_OK(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !configured[dev].load(std::memory_order_acquire)) {
    FFEMB_CUDA_OK(cudaFuncSetAttribute(cluster_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)kCsSmemBytes));
    configured[dev].store(1, std::memory_order_release);
  }
  CsParams cp;
  int64_t rows = 0, pos = 0;
  for (int t = 0; t < T; ++t) {
    const ffemb_lookup_t &l = lookups[t];
    const int bits = bits_for(l.num_entries);
    const int passes = (bits + kCsMaxBits - 1) / kCsMaxBits;
    int rb = (bits + passes - 1) / passes;
    if (rb < 1) rb = 1;
    cp.t[t] = CsTable{l.indices, l.offsets, l.num_entries, (uint32_t)pos, (uint32_t)l.num_indices,
                      (uint32_t)rows, l.bag_size, passes, rb};
    rows += l.num_entries;
    pos += l.num_indices;
  }
  cp.keys_out = keys_out;
  cp.vals_out = vals_out;
  cp.tmp = tmp;
  cp.B = batch;
  cp.err = err;
  // Cluster size: what the largest segment needs, and beyond that as many CTAs
  // per table (up to the portable 8) as keep the launch within about 64 CTAs --
  // a rank that sorts two tables wants them spread over 16 SMs (the sort is
  // latency-bound per CTA), a GPU that sorts 26 wants no more CTAs than needed.
  int C = 1;
  while ((int64_t)C * kCsSlab < max_n) C *= 2;
  while (C < 8 && T * C * 2 <= 64 && max_n >= (int64_t)C * 2 * 2048) C *= 2;  // slabs of 2048 slots or more
  int slab_log2 = 5;
  while (((int64_t)C << slab_log2) < max_n) ++slab_log2;
  cp.slab_log2 = slab_log2;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(T * C), 1, 1);
  cfg.blockDim = dim3(kCsThreads, 1, 1);
  cfg.dynamicSmemBytes = kCsSmemBytes;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)C;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  FFEMB_CUDA_OK(cudaLaunchKernelEx(&cfg, cluster_sort_kernel, cp));
  FFEMB_LAUNCH_OK("cluster_sort");
  return FFEMB_OK;
}

#ifdef FFEMB_CS_PROFILE
extern "C" int ffemb_debug_cs_prof(long long *out, int reset) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, g_cs_prof, sizeof(long long) * 16);
  if (reset) {
    long long z[16] = {0};
    cudaMemcpyToSymbol(g_cs_prof, z, sizeof z);
  }
  return 0;
}
#endif

}  // namespace ffemb
