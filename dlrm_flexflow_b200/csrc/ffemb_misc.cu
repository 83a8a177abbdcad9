// ffemb_misc.cu -- error state, version, table initialiser.
#include "ffemb_common.cuh"

namespace ffemb {

static thread_local std::string t_error;
std::atomic<int64_t> g_launches{0};

void set_error(const std::string &msg) { t_error = msg; }
int fail(int status, const std::string &msg) {
  t_error = msg;
  return status;
}

__device__ __forceinline__ uint64_t mix64(uint64_t z) {  // splitmix64 finaliser
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// Stands in for curandGenerateUniform + scale_kernel
// (reference src/runtime/initializer_kernel.cu:78-81, cuda_helper.cu:3-11):
// u in (0,1], w = (max-min)*u + min as one FMA.  Counter-based, so element i
// does not depend on the launch shape and the stream is reproducible on the
// host (oracle/ffemb_oracle.c:ffo_uniform_init).
__global__ void __launch_bounds__(256)
uniform_init_kernel(float *__restrict__ w, int64_t count, uint64_t s, float min_val, float span) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
    const uint64_t r = mix64(s + (uint64_t)i * 0x9E3779B97F4A7C15ull);
    const float u = (float)((uint32_t)(r >> 40) + 1u) * 5.9604644775390625e-08f;
    w[i] = fmaf(span, u, min_val);
  }
}

static uint64_t host_mix64(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

}  // namespace ffemb

using namespace ffemb;

extern "C" const char *ffemb_last_error(void) { return t_error.c_str(); }
extern "C" const char *ffemb_version(void) { return "0.2.0"; }
extern "C" int ffemb_compiled_sm(void) { return 100; }
extern "C" int64_t ffemb_kernel_launch_count(void) { return g_launches.load(); }

extern "C" int ffemb_uniform_init(float *weight, int64_t count, int seed, float min_val,
                                  float max_val, void *stream) {
  if (!weight || count < 0) return fail(FFEMB_ERR_INVALID_ARGUMENT, "uniform_init: bad arguments");
  if (count == 0) return FFEMB_OK;
  const uint64_t s = host_mix64((uint64_t)(uint32_t)seed + 0x9E3779B97F4A7C15ull);
  int64_t blocks = (count + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;  // 16 resident CTAs of 256 threads per SM
  uniform_init_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(weight, count, s, min_val,
                                                                          max_val - min_val);
  FFEMB_LAUNCH_OK("uniform_init");
  return FFEMB_OK;
}

namespace ffemb {
// Dense SGD for small replicated tables: the reference's sgd_update with
// weight_decay = momentum = 0 (src/runtime/optimizer_kernel.cu:23-41),
// W = fmaf(-lr, G, W), and G is zeroed for the next step in the same pass
// (the reference zeroes it separately, src/runtime/model.cc:256-280).
__global__ void __launch_bounds__(256)
dense_sgd_kernel(float *__restrict__ w, float *__restrict__ g, int64_t count, float lr) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
    w[i] = fmaf(-lr, g[i], w[i]);
    g[i] = 0.f;
  }
}
}  // namespace ffemb

extern "C" int ffemb_dense_sgd(float *weight, float *grad, int64_t count, float lr, void *stream) {
  if (!weight || !grad || count < 0) return fail(FFEMB_ERR_INVALID_ARGUMENT, "dense_sgd: bad arguments");
  if (count == 0) return FFEMB_OK;
  int64_t blocks = (count + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  dense_sgd_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(weight, grad, count, lr);
  FFEMB_LAUNCH_OK("dense_sgd");
  return FFEMB_OK;
}
