// One step of the reference's autoregressive sampling loop (src/run/inference.py:76-95, language branch):
//   token_out = argmax_v( logits + log(-log(U)) * (-temperature) ),  U ~ Uniform(1e-9, 1)
//   token_out = shift(token_out, +1 along the sequence, no wrap)
//   token_x   = token_out at `position`, unchanged elsewhere
// Only sequence index `position` of the shifted argmax is ever used, i.e. the prediction made at position - 1
// (position 0 receives the shift's zero fill), so one warp per batch row scans one logits row.
#include "obst_common.cuh"

namespace obst {

__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

template <typename T>
__global__ void __launch_bounds__(256) sample_kernel(const T* __restrict__ logits, int32_t* __restrict__ tokens,
                                                      const float* __restrict__ uniform, int batch, int seq, int vocab,
                                                      int pos, float temperature, uint64_t seed) {
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (b >= batch) return;
  if (pos == 0) {  // shift(..., wrap=False) fills position 0 with zeros
    if (lane == 0) tokens[(int64_t)b * seq] = 0;
    return;
  }
  const T* row = logits + ((int64_t)b * seq + (pos - 1)) * vocab;
  float best = -INFINITY;
  int best_i = 0x7fffffff;
  for (int v = lane; v < vocab; v += 32) {
    float val = to_f32(row[v]);
    if (temperature != 0.f) {
      float u;
      if (uniform) {
        u = uniform[(int64_t)b * vocab + v];
      } else {  // counter-based: (seed, position, row, vocabulary entry) -> 24 random bits
        const uint64_t h = splitmix64(seed ^ splitmix64(((uint64_t)pos << 40) ^ ((uint64_t)b << 20) ^ (uint64_t)v));
        u = (float)(h >> 40) * (1.0f / 16777216.0f);
      }
      u = fminf(fmaxf(u, 1e-9f), 0.99999994f);  // random_uniform(minval=1e-9, maxval=1)
      val += logf(-logf(u)) * (-temperature);
    }
    if (val > best || (val == best && v < best_i)) {  // argmax: first index among ties
      best = val;
      best_i = v;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, best_i, o);
    if (ov > best || (ov == best && oi < best_i)) {
      best = ov;
      best_i = oi;
    }
  }
  if (lane == 0) tokens[(int64_t)b * seq + pos] = best_i == 0x7fffffff ? 0 : best_i;
}

}  // namespace obst

using namespace obst;

extern "C" int obst_sample_step(int dtype, const void* logits, int32_t* tokens, const float* uniform, int32_t batch,
                                int32_t seq, int32_t vocab, int32_t position, float temperature, uint64_t seed,
                                void* stream) {
  OBST_REQUIRE(batch >= 0 && seq > 0 && vocab > 0, "obst_sample_step: bad shape");
  OBST_REQUIRE(position >= 0 && position < seq, "obst_sample_step: position %d outside [0, %d)", position, seq);
  OBST_REQUIRE(logits && tokens, "obst_sample_step: null argument");
  if (batch == 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned blocks = (unsigned)((batch + 7) / 8);
  if (dtype == OBST_BF16)
    sample_kernel<bf16><<<blocks, 256, 0, st>>>((const bf16*)logits, tokens, uniform, batch, seq, vocab, position,
                                                 temperature, seed);
  else if (dtype == OBST_F32)
    sample_kernel<float><<<blocks, 256, 0, st>>>((const float*)logits, tokens, uniform, batch, seq, vocab, position,
                                                  temperature, seed);
  else
    OBST_REQUIRE(false, "obst_sample_step: bad dtype %d", dtype);
  count_launch();
  return check_launch("obst_sample_step");
}
