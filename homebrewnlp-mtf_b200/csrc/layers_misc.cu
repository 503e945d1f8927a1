// Small kernels behind the registered layers that are not on the mixer hot path (src/model/basic.py:21-30,72-90;
// src/model/frontend.py:39-55): head sums, the sequence/feature swap, element-wise products, rezero's scalar gate and
// inverted dropout.  All are plain bandwidth kernels: grid-stride, coalesced along the contiguous axis.
#include "obst_common.cuh"

namespace obst {

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

static inline unsigned misc_blocks(int64_t n) {
  int64_t b = (n + 255) / 256;
  const int64_t cap = (int64_t)num_sms() * 16;
  if (b > cap) b = cap;
  return (unsigned)(b < 1 ? 1 : b);
}

// out[r, w] = sum_h x[r, h, w]   (reduce_sum over the head dim)
template <typename T>
__global__ void __launch_bounds__(256) sum_heads_kernel(const T* __restrict__ x, T* __restrict__ out, int64_t rows,
                                                         int heads, int width) {
  const int64_t total = rows * width;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / width;
    const int w = (int)(i - r * width);
    float acc = 0.f;
    for (int h = 0; h < heads; ++h) acc += to_f32(x[(r * heads + h) * width + w]);
    out[i] = from_f32<T>(acc);
  }
}

// dx[r, h, w] = dy[r, w]   (its gradient: broadcast back over the heads)
template <typename T>
__global__ void __launch_bounds__(256) broadcast_heads_kernel(const T* __restrict__ dy, T* __restrict__ dx,
                                                               int64_t rows, int heads, int width) {
  const int64_t total = rows * heads * width;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int w = (int)(i % width);
    const int64_t r = i / ((int64_t)heads * width);
    dx[i] = dy[r * width + w];
  }
}

// out[o, j, m, i] = x[o, i, m, j]  for x [outer, a, mid, b] -> out [outer, b, mid, a]
template <typename T>
__global__ void __launch_bounds__(256) swap_axes_kernel(const T* __restrict__ x, T* __restrict__ out, int64_t outer,
                                                         int a, int mid, int b) {
  const int64_t per = (int64_t)a * mid * b;
  const int64_t total = outer * per;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    // idx walks the OUTPUT (coalesced stores): out[o, j, m, i]
    const int64_t o = idx / per;
    int64_t rem = idx - o * per;
    const int i = (int)(rem % a);
    rem /= a;
    const int m = (int)(rem % mid);
    const int j = (int)(rem / mid);
    out[idx] = x[o * per + ((int64_t)i * mid + m) * b + j];
  }
}

template <typename T>
__global__ void __launch_bounds__(256) mul_kernel(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ out,
                                                   int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = from_f32<T>(to_f32(a[i]) * to_f32(b[i]));
}

// out = x * s, s a one-element tensor in device memory (rezero, src/model/basic.py:21-22)
template <typename T>
__global__ void __launch_bounds__(256) scale_kernel(const T* __restrict__ x, const T* __restrict__ s, T* __restrict__ out,
                                                     int64_t n) {
  const float sv = to_f32(s[0]);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = from_f32<T>(to_f32(x[i]) * sv);
}

// out[0] += sum_i a[i] * b[i]   (fp32 accumulate; gradient of the rezero scalar)
template <typename T>
__global__ void __launch_bounds__(256) dot_kernel(const T* __restrict__ a, const T* __restrict__ b,
                                                   float* __restrict__ out, int64_t n) {
  __shared__ float s_part[8];
  float acc = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    acc = fmaf(to_f32(a[i]), to_f32(b[i]), acc);
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < 8; ++w) t += s_part[w];
    atomicAdd(out, t);
  }
}

// inverted dropout (mtf.dropout): out = x * [u < keep] / keep, u = hash(seed, i) -- the same seed gives the same mask,
// which is how the backward pass applies it to dy
template <typename T>
__global__ void __launch_bounds__(256) dropout_kernel(const T* __restrict__ x, T* __restrict__ out, int64_t n,
                                                       float keep, uint64_t seed) {
  const float inv = 1.0f / keep;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float u = (float)(mix64(seed ^ mix64((uint64_t)i)) >> 40) * (1.0f / 16777216.0f);
    out[i] = from_f32<T>(u < keep ? to_f32(x[i]) * inv : 0.f);
  }
}

}  // namespace obst

using namespace obst;

#define OBST_DISPATCH(what, call_bf16, call_f32)          \
  do {                                                    \
    if (dtype == OBST_BF16) {                             \
      call_bf16;                                          \
    } else if (dtype == OBST_F32) {                       \
      call_f32;                                           \
    } else {                                              \
      OBST_REQUIRE(false, what ": bad dtype %d", dtype);  \
    }                                                     \
    count_launch();                                       \
    return check_launch(what);                            \
  } while (0)

extern "C" {

int obst_sum_heads(int dtype, const void* x, void* out, int64_t rows, int32_t heads, int32_t width, void* stream) {
  OBST_REQUIRE(rows >= 0 && heads > 0 && width > 0, "obst_sum_heads: bad shape");
  if (rows == 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned b = misc_blocks(rows * width);
  OBST_DISPATCH("obst_sum_heads",
                (sum_heads_kernel<bf16><<<b, 256, 0, st>>>((const bf16*)x, (bf16*)out, rows, heads, width)),
                (sum_heads_kernel<float><<<b, 256, 0, st>>>((const float*)x, (float*)out, rows, heads, width)));
}

int obst_broadcast_heads(int dtype, const void* dy, void* dx, int64_t rows, int32_t heads, int32_t width,
                         void* stream) {
  OBST_REQUIRE(rows >= 0 && heads > 0 && width > 0, "obst_broadcast_heads: bad shape");
  if (rows == 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned b = misc_blocks(rows * heads * width);
  OBST_DISPATCH("obst_broadcast_heads",
                (broadcast_heads_kernel<bf16><<<b, 256, 0, st>>>((const bf16*)dy, (bf16*)dx, rows, heads, width)),
                (broadcast_heads_kernel<float><<<b, 256, 0, st>>>((const float*)dy, (float*)dx, rows, heads, width)));
}

int obst_swap_axes(int dtype, const void* x, void* out, int64_t outer, int32_t a, int32_t mid, int32_t b,
                   void* stream) {
  OBST_REQUIRE(outer >= 0 && a > 0 && mid > 0 && b > 0, "obst_swap_axes: bad shape");
  OBST_REQUIRE(x != out, "obst_swap_axes: in-place operation is not supported");
  if (outer == 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned blk = misc_blocks(outer * a * mid * b);
  OBST_DISPATCH("obst_swap_axes",
                (swap_axes_kernel<bf16><<<blk, 256, 0, st>>>((const bf16*)x, (bf16*)out, outer, a, mid, b)),
                (swap_axes_kernel<float><<<blk, 256, 0, st>>>((const float*)x, (float*)out, outer, a, mid, b)));
}

int obst_mul(int dtype, const void* a, const void* b, void* out, int64_t n, void* stream) {
  if (n <= 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned blk = misc_blocks(n);
  OBST_DISPATCH("obst_mul", (mul_kernel<bf16><<<blk, 256, 0, st>>>((const bf16*)a, (const bf16*)b, (bf16*)out, n)),
                (mul_kernel<float><<<blk, 256, 0, st>>>((const float*)a, (const float*)b, (float*)out, n)));
}

int obst_scale(int dtype, const void* x, const void* scalar, void* out, int64_t n, void* stream) {
  OBST_REQUIRE(scalar != nullptr, "obst_scale: null scalar");
  if (n <= 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned blk = misc_blocks(n);
  OBST_DISPATCH("obst_scale",
                (scale_kernel<bf16><<<blk, 256, 0, st>>>((const bf16*)x, (const bf16*)scalar, (bf16*)out, n)),
                (scale_kernel<float><<<blk, 256, 0, st>>>((const float*)x, (const float*)scalar, (float*)out, n)));
}

int obst_dot_accumulate(int dtype, const void* a, const void* b, float* out, int64_t n, void* stream) {
  OBST_REQUIRE(out != nullptr, "obst_dot_accumulate: null output");
  if (n <= 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  unsigned blk = misc_blocks(n);
  if (blk > (unsigned)num_sms() * 4) blk = (unsigned)num_sms() * 4;
  OBST_DISPATCH("obst_dot_accumulate",
                (dot_kernel<bf16><<<blk, 256, 0, st>>>((const bf16*)a, (const bf16*)b, out, n)),
                (dot_kernel<float><<<blk, 256, 0, st>>>((const float*)a, (const float*)b, out, n)));
}

int obst_dropout(int dtype, const void* x, void* out, int64_t n, float keep, uint64_t seed, void* stream) {
  OBST_REQUIRE(keep > 0.f && keep <= 1.f, "obst_dropout: keep probability %f outside (0, 1]", (double)keep);
  if (n <= 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned blk = misc_blocks(n);
  OBST_DISPATCH("obst_dropout", (dropout_kernel<bf16><<<blk, 256, 0, st>>>((const bf16*)x, (bf16*)out, n, keep, seed)),
                (dropout_kernel<float><<<blk, 256, 0, st>>>((const float*)x, (float*)out, n, keep, seed)));
}

}  // extern "C"
