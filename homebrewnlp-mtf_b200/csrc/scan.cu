// Running sums along the sequence axis: the `cumsum` / `cummean` layers (src/model/spatial.py:26-39) and their
// gradients.  x is [outer, seq, inner] with `inner` contiguous (for the model: [batch, sequence, heads * features]).
// One thread owns a 16-byte column group of one outer index and walks the sequence with an fp32 running sum, so every
// warp access is a full, contiguous line; UNROLL rows are loaded before the first is consumed.
#include "obst_common.cuh"

namespace obst {

// scale: 0 none | 1 out[s] = sum / (s + 1)  (cummean forward) | 2 in[t] / (t + 1) before summing (cummean backward)
template <typename T, bool REVERSE, int SCALE>
__global__ void __launch_bounds__(256) cumsum_kernel(const T* __restrict__ x, T* __restrict__ out, int64_t outer,
                                                      int seq, int64_t inner) {
  constexpr int N = Vec16<T>::N;
  constexpr int UNROLL = 4;
  const int64_t groups = inner / N;
  const int64_t total = outer * groups;
  for (int64_t id = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; id < total; id += (int64_t)gridDim.x * blockDim.x) {
    const int64_t o = id / groups, gcol = id - o * groups;
    const T* xp = x + o * seq * inner + gcol * N;
    T* op = out + o * seq * inner + gcol * N;
    float run[N];
#pragma unroll
    for (int j = 0; j < N; ++j) run[j] = 0.f;
    for (int s0 = 0; s0 < seq; s0 += UNROLL) {
      Vec16<T> v[UNROLL];
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        const int s = s0 + u;
        if (s < seq) v[u].load(xp + (int64_t)(REVERSE ? seq - 1 - s : s) * inner);
      }
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        const int s = s0 + u;
        if (s >= seq) continue;
        const int pos = REVERSE ? seq - 1 - s : s;
        const float w = SCALE ? 1.0f / (float)(pos + 1) : 1.0f;
        Vec16<T> o16;
#pragma unroll
        for (int j = 0; j < N; ++j) {
          run[j] += SCALE == 2 ? v[u].v[j] * w : v[u].v[j];
          o16.v[j] = SCALE == 1 ? run[j] * w : run[j];
        }
        o16.store(op + (int64_t)pos * inner);
      }
    }
  }
}

template <typename T>
static int cumsum_launch(const void* x, void* out, int64_t outer, int seq, int64_t inner, int reverse, int scale,
                         cudaStream_t st) {
  constexpr int N = Vec16<T>::N;
  OBST_REQUIRE(inner % N == 0, "obst_cumsum: inner size %lld must be a multiple of %d", (long long)inner, N);
  OBST_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0,
               "obst_cumsum: pointers must be 16-byte aligned");
  const int64_t total = outer * (inner / N);
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = (int64_t)num_sms() * 16;
  if (blocks > cap) blocks = cap;
#define OBST_SCAN(R, S) \
  cumsum_kernel<T, R, S><<<(unsigned)blocks, 256, 0, st>>>((const T*)x, (T*)out, outer, seq, inner)
  if (reverse) {
    if (scale == 0) OBST_SCAN(true, 0);
    else if (scale == 1) OBST_SCAN(true, 1);
    else OBST_SCAN(true, 2);
  } else {
    if (scale == 0) OBST_SCAN(false, 0);
    else if (scale == 1) OBST_SCAN(false, 1);
    else OBST_SCAN(false, 2);
  }
#undef OBST_SCAN
  count_launch();
  return check_launch("obst_cumsum");
}

}  // namespace obst

using namespace obst;

extern "C" int obst_cumsum(int dtype, const void* x, void* out, int64_t outer, int32_t seq, int64_t inner,
                           int32_t reverse, int32_t scale, void* stream) {
  OBST_REQUIRE(outer >= 0 && seq >= 0 && inner >= 0 && scale >= 0 && scale <= 2, "obst_cumsum: bad arguments");
  if (outer == 0 || seq == 0 || inner == 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == OBST_BF16) return cumsum_launch<bf16>(x, out, outer, seq, inner, reverse, scale, st);
  if (dtype == OBST_F32) return cumsum_launch<float>(x, out, outer, seq, inner, reverse, scale, st);
  OBST_REQUIRE(false, "obst_cumsum: bad dtype %d", dtype);
}
