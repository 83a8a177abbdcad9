/*
 * ffemb_oracle.c -- CPU restatement of the reference's embedding-bag path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the shipped
 * library: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this file's shared object, and there only as
 * the checker or the timed CPU baseline -- never as the thing being shipped.
 * The product path (dlrm_flexflow_b200/) has no CPU fallback.
 *
 * Parity status: PINNED against the reference's own CPU code, both halves.
 * Forward: the restatement below is checked bit-for-bit against
 * /root/reference/src/ops/embedding_avx2.cc compiled in place into
 * oracle/_ref/libembedding_avx2.so (see oracle/Makefile and
 * tests/test_oracle.py), and against the golden vectors that kernel produced
 * (tests/golden/embedding_avx2_golden.npz, made by tests/golden/make_golden.py).
 * Backward: ffo_embed_backward_dense is checked bit-for-bit, for the case the
 * reference's CPU code covers (one index per sample, SUM, accumulating into
 * the gradient), against embed_backward_generic (src/ops/embedding.cc:54-72)
 * compiled in place into oracle/_ref/libembed_backward_generic.so, and against
 * the golden vectors it produced (tests/golden/embedding_bwd_golden.npz).
 * The reference holds no test or fixture of its own for the embedding
 * operator and no CPU code for bags longer than 1 in backward,
 * for AVG's per-position g/L (CUDA only, src/ops/embedding.cu:213-218) or for
 * the SGD update (CUDA only, src/runtime/optimizer_kernel.cu:23-41): those
 * parts are restatements of the cited lines, cross-checked against
 * torch.nn.EmbeddingBag + torch.optim.SGD at 1e-5 (the reference test-suite's
 * own tolerance, src/ops/tests/test_harness.py:44-60).
 *
 * Build: gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC (oracle/Makefile).
 * -ffp-contract=off keeps every multiply/add exactly as written; the fused
 * operations the reference relies on are spelled fmaf() explicitly.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define FFO_AGGR_SUM 21 /* include/ffconst.h:11-15 */
#define FFO_AGGR_AVG 22

/* ---------------------------------------------------------------------------
 * Forward.  Follows EmbeddingLookup_int64_t_float_float__avx2_fma,
 * src/ops/embedding_avx2.cc:5-296 (generic path :241-295; the D=128/64/32
 * specialisations :17-240 perform the same arithmetic on 8-wide vectors):
 *   per bag: acc[0..D) = +0; for each index in bag order:
 *   acc = fma(wgt = 1.0f, W[idx,:], acc)  ==  acc + W[idx,:]  (exact: x*1 = x)
 *   then, when normalize_by_lengths and len != 0: acc *= (1.0f / len).
 * Index convention (embedding.cu:182-189): table is row-major [N][D], the
 * index stream is flattened [bag][position].
 * `lengths == NULL` means every bag has `fixed_len` entries.
 * `ldo` is the output row stride in floats (>= D).
 * Returns 0, or -1 if an index is out of [0,N) (the reference asserts,
 * embedding_avx2.cc:51-53) -- rows already written stay written.
 * An empty bag yields zeros (generic path behaviour, :247-253).
 * ------------------------------------------------------------------------- */
int ffo_embed_forward(int D, int64_t B, int64_t N, const float *W,
                      const int64_t *indices, const int *lengths,
                      int fixed_len, int normalize_by_lengths, float *out,
                      int64_t ldo)
{
  int64_t pos = 0;
  for (int64_t b = 0; b < B; ++b) {
    float *op = out + b * ldo;
    int len = lengths ? lengths[b] : fixed_len;
    for (int d = 0; d < D; ++d) op[d] = 0.0f;
    for (int j = 0; j < len; ++j, ++pos) {
      int64_t idx = indices[pos];
      if (idx < 0 || idx >= N) return -1;
      const float *ip = W + idx * (int64_t)D;
      for (int d = 0; d < D; ++d) op[d] = fmaf(1.0f, ip[d], op[d]);
    }
    if (normalize_by_lengths && len) {
      float len_inv = 1.0f / (float)len;
      for (int d = 0; d < D; ++d) op[d] = op[d] * len_inv;
    }
  }
  return 0;
}

/* Threaded forward over bags (fixed_len only) -- the "all host cores" CPU
 * baseline of SURVEY.md 8d.  Same arithmetic per bag as ffo_embed_forward. */
int ffo_embed_forward_mt(int D, int64_t B, int64_t N, const float *W,
                         const int64_t *indices, int fixed_len,
                         int normalize_by_lengths, float *out, int64_t ldo,
                         int nthreads)
{
  int bad = 0;
#pragma omp parallel for num_threads(nthreads) schedule(static) reduction(| : bad)
  for (int64_t b = 0; b < B; ++b) {
    float *op = out + b * ldo;
    for (int d = 0; d < D; ++d) op[d] = 0.0f;
    for (int j = 0; j < fixed_len; ++j) {
      int64_t idx = indices[b * fixed_len + j];
      if (idx < 0 || idx >= N) { bad = 1; continue; }
      const float *ip = W + idx * (int64_t)D;
      for (int d = 0; d < D; ++d) op[d] = fmaf(1.0f, ip[d], op[d]);
    }
    if (normalize_by_lengths && fixed_len) {
      float len_inv = 1.0f / (float)fixed_len;
      for (int d = 0; d < D; ++d) op[d] = op[d] * len_inv;
    }
  }
  return bad ? -1 : 0;
}

/* ---------------------------------------------------------------------------
 * Backward into a dense table-shaped gradient.  Follows
 * embed_backward_generic, src/ops/embedding.cc:54-72
 *   for i in 0..B*D: gradW[idx[i/D]*D + i%D] += g[i]
 * generalised from one index per sample to a bag of L (or lengths[b])
 * indices visited in flattened (bag, position) order, with the per-position
 * gradient of the GPU kernel, src/ops/embedding.cu:199-224: g (SUM) or
 * g / (float)L (AVG, true division, :213-218).
 * Accumulates into gradW (caller zeroes it: reference zero_gradients,
 * src/runtime/model.cc:256-280).  `ldg` is the row stride of g in floats.
 * ------------------------------------------------------------------------- */
int ffo_embed_backward_dense(int D, int64_t B, int64_t N,
                             const int64_t *indices, const int *lengths,
                             int fixed_len, int aggr, const float *g,
                             int64_t ldg, float *gradW)
{
  int64_t pos = 0;
  for (int64_t b = 0; b < B; ++b) {
    int len = lengths ? lengths[b] : fixed_len;
    const float *gp = g + b * ldg;
    for (int j = 0; j < len; ++j, ++pos) {
      int64_t idx = indices[pos];
      if (idx < 0 || idx >= N) return -1;
      float *wp = gradW + idx * (int64_t)D;
      if (aggr == FFO_AGGR_AVG) {
        float fl = (float)len;
        for (int d = 0; d < D; ++d) wp[d] += gp[d] / fl;
      } else {
        for (int d = 0; d < D; ++d) wp[d] += gp[d];
      }
    }
  }
  return 0;
}

/* ---------------------------------------------------------------------------
 * Dense SGD.  Follows sgd_update, src/runtime/optimizer_kernel.cu:23-41:
 *   gt = G + wd*W; if (m > 0) { V = V*m + gt; gt = nesterov ? gt + m*V : V; }
 *   W -= lr*gt;
 * nvcc's default -fmad=true contracts `W - lr*gt` into one FFMA, spelled here
 * as fmaf(-lr, gt, W); the same for `G + wd*W`.
 * ------------------------------------------------------------------------- */
void ffo_sgd_update_dense(int64_t count, float lr, float weight_decay,
                          float momentum, int nesterov, const float *G,
                          float *V, float *W)
{
  for (int64_t i = 0; i < count; ++i) {
    float gt = fmaf(weight_decay, W[i], G[i]);
    if (momentum > 0.0f) {
      V[i] = fmaf(V[i], momentum, gt);
      gt = nesterov ? fmaf(momentum, V[i], gt) : V[i];
    }
    W[i] = fmaf(-lr, gt, W[i]);
  }
}

/* ---------------------------------------------------------------------------
 * One reference training step on one table, touched rows only.  Result is
 * identical to  zero(gradW); backward_dense; sgd_update_dense(wd=0, m=0)
 * because an untouched row has G = 0 and fmaf(-lr, 0, w) == w.  `gradW` is a
 * caller-provided [N][D] scratch that must be all zero on entry and is all
 * zero again on exit.  Used by the timed CPU baseline so that the CPU is not
 * charged the reference's dense 4x-table-bytes passes.
 * ------------------------------------------------------------------------- */
int ffo_embed_backward_sgd_touched(int D, int64_t B, int64_t N,
                                   const int64_t *indices, int fixed_len,
                                   int aggr, const float *g, int64_t ldg,
                                   float lr, float *gradW, float *W)
{
  int rc = ffo_embed_backward_dense(D, B, N, indices, NULL, fixed_len, aggr, g,
                                    ldg, gradW);
  if (rc) return rc;
  int64_t n = B * (int64_t)fixed_len;
  for (int64_t p = 0; p < n; ++p) {
    float *gp = gradW + indices[p] * (int64_t)D;
    float *wp = W + indices[p] * (int64_t)D;
    for (int d = 0; d < D; ++d) {
      wp[d] = fmaf(-lr, gp[d], wp[d]);
      gp[d] = 0.0f; /* a later duplicate of this row is then a no-op */
    }
  }
  return 0;
}

/* ---------------------------------------------------------------------------
 * Whole-step CPU baseline over T same-shaped tables, tables spread over
 * `nthreads` OpenMP threads (one table per thread at a time -- tables are
 * independent, which is also how the reference would place one Legion CPU
 * task per table).  Forward writes out[t][B][D]; backward reads g[t][B][D].
 * Returns 0 or -1.
 * ------------------------------------------------------------------------- */
int ffo_step_tables_mt(int T, int D, int64_t B, const int64_t *N,
                       float *const *W, const int64_t *const *indices,
                       int fixed_len, int aggr, float *const *out,
                       const float *const *g, float lr, float *const *gradW,
                       int nthreads)
{
  int bad = 0;
#pragma omp parallel for num_threads(nthreads) schedule(dynamic, 1) reduction(| : bad)
  for (int t = 0; t < T; ++t) {
    int rc = ffo_embed_forward(D, B, N[t], W[t], indices[t], NULL, fixed_len,
                               aggr == FFO_AGGR_AVG, out[t], D);
    if (!rc)
      rc = ffo_embed_backward_sgd_touched(D, B, N[t], indices[t], fixed_len,
                                          aggr, g[t], D, lr, gradW[t], W[t]);
    if (rc) bad = 1;
  }
  return bad ? -1 : 0;
}

/* ---------------------------------------------------------------------------
 * Table initialiser.  The reference draws U(0,1] with cuRAND XORWOW seeded
 * from std::rand() and maps it with scale_kernel  w = (max-min)*u + min
 * (src/runtime/initializer_kernel.cu:24-85, src/runtime/cuda_helper.cu:3-11).
 * XORWOW's stream is not reproducible off-device, so both this oracle and
 * the CUDA library use the same counter-based generator instead: element i
 * of a table gets u = ((mix(seed, i) >> 8) + 1) * 2^-24  in (0,1], then the
 * reference's affine map.  Only the distribution, not the stream, is
 * reference behaviour.
 * ------------------------------------------------------------------------- */
static inline uint64_t ffo_mix64(uint64_t z)
{ /* splitmix64 finaliser */
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}

void ffo_uniform_init(float *W, int64_t count, int seed, float min_val,
                      float max_val)
{
  uint64_t s = ffo_mix64((uint64_t)(uint32_t)seed + 0x9E3779B97F4A7C15ULL);
  for (int64_t i = 0; i < count; ++i) {
    uint64_t r = ffo_mix64(s + (uint64_t)i * 0x9E3779B97F4A7C15ULL);
    float u = (float)((uint32_t)(r >> 40) + 1u) * 5.9604644775390625e-08f;
    W[i] = fmaf(max_val - min_val, u, min_val);
  }
}

int ffo_max_threads(void)
{
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
