// ffemb_common.cuh -- shared device/host helpers of the embedding-bag library.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <string>

#include "ffemb_c.h"

namespace ffemb {

// ---- error plumbing (never abort across the C boundary) --------------------
void set_error(const std::string &msg);
int fail(int status, const std::string &msg);
extern std::atomic<int64_t> g_launches;

#define FFEMB_CUDA_OK(expr)                                                    \
  do {                                                                         \
    cudaError_t _e = (expr);                                                   \
    if (_e != cudaSuccess)                                                     \
      return ::ffemb::fail(FFEMB_ERR_CUDA, std::string(#expr) + ": " +         \
                                               cudaGetErrorString(_e));        \
  } while (0)

#define FFEMB_LAUNCH_OK(name)                                                  \
  do {                                                                         \
    ::ffemb::g_launches.fetch_add(1, std::memory_order_relaxed);               \
    cudaError_t _e = cudaGetLastError();                                       \
    if (_e != cudaSuccess)                                                     \
      return ::ffemb::fail(FFEMB_ERR_CUDA, std::string("launch ") + name +     \
                                               ": " + cudaGetErrorString(_e)); \
  } while (0)

// ---- launch-parameter structs (passed by value in the kernel param bank) ---
constexpr int kMaxTables = 64;  // tables per batched launch
constexpr int kMaxParts = FFEMB_MAX_PARTS;

struct Matrix {  // device view of ffemb_matrix_t
  float *base[kMaxParts];
  int num_parts;
  int rows_per_part;
  int64_t ld;
};

// ---- 128-bit memory ops ------------------------------------------------------
// Table rows in the forward are read exactly once per lookup and never written
// during that kernel: non-coherent path, do not allocate in L1.
__device__ __forceinline__ float4 ld_row_nc(const float *p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p));
  return v;
}
// Rows that the same kernel also writes (SGD update): coherent load.
__device__ __forceinline__ float4 ld_row(const float *p) {
  float4 v;
  asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ void st_row(float *p, const float4 &v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};"
               :
               : "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

__device__ __forceinline__ float *matrix_row(const Matrix &m, int b) {
  if (m.num_parts == 1) return m.base[0] + (int64_t)b * m.ld;
  int r = b / m.rows_per_part;
  int bb = b - r * m.rows_per_part;
  return m.base[r] + (int64_t)bb * m.ld;
}

__device__ __forceinline__ void add4(float4 &a, const float4 &v) {
  a.x += v.x;
  a.y += v.y;
  a.z += v.z;
  a.w += v.w;
}

// Host-side validation shared by the forward and backward entry points: the
// same malformed lookups are refused by both (ffemb_backward.cu).
int check_lookups(const char *who, const ffemb_lookup_t *lookups, int num_tables, int batch,
                  int out_dim);

inline int to_matrix(const ffemb_matrix_t *in, int batch, Matrix *out) {
  if (!in) return fail(FFEMB_ERR_INVALID_ARGUMENT, "matrix is NULL");
  if (in->num_parts < 1 || in->num_parts > kMaxParts)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "matrix num_parts out of range");
  if (in->num_parts > 1 &&
      (in->rows_per_part <= 0 ||
       (int64_t)in->rows_per_part * in->num_parts < batch))
    return fail(FFEMB_ERR_INVALID_ARGUMENT,
                "matrix rows_per_part * num_parts < batch");
  for (int r = 0; r < kMaxParts; ++r)
    out->base[r] = r < in->num_parts ? in->base[r] : nullptr;
  for (int r = 0; r < in->num_parts; ++r)
    if (!in->base[r]) return fail(FFEMB_ERR_INVALID_ARGUMENT, "matrix base is NULL");
  out->num_parts = in->num_parts;
  out->rows_per_part = in->num_parts > 1 ? in->rows_per_part : batch;
  out->ld = in->ld;
  return FFEMB_OK;
}

}  // namespace ffemb
