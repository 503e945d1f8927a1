// Soft-max stage of the reference's dot-product attention (src/model/spatial.py:42-81) and the small broadcast ops
// around it.  The contractions (q.k^T, p.v and their gradients) are obst_gemm calls; what is left is, per
// (batch, position s, head) row over the key positions t:
//     z_t    = scale * logit_t + bias[h,s,t] * m(s,t)          ('biased_softmax', _masked_map: :19-23,68-69)
//     z_t    = -inf for t > s                                   (compare_range(dim, tmp, less) * 1e38 * -2, :71)
//     soft   = softmax_t(z)                                     (:72-74)
//     p_t    = (soft_t + add[h,s,t] * m(s,t)) * (mul[h,s,t] * m(s,t))   ('biased_attention_map', 'scale_attention_map', :75-78)
// with m(s,t) = [s >= t] when the attention dim is masked, else 1; each map is optional.
// Not on the mixer hot path (its configs use the map-only form, a plain causal GEMM): one warp per row, three passes.
#include "obst_common.cuh"

namespace obst {

template <typename T>
__global__ void __launch_bounds__(256) attn_softmax_fwd_kernel(const T* __restrict__ logits, const T* __restrict__ bias,
                                                                const T* __restrict__ addm, const T* __restrict__ mulm,
                                                                T* __restrict__ soft, T* __restrict__ probs,
                                                                int64_t rows, int S, int H, int Tn, float scale,
                                                                int masked) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int h = (int)(row % H);
  const int s = (int)((row / H) % S);
  const T* l = logits + row * Tn;
  const int64_t moff = ((int64_t)h * S + s) * Tn;
  auto z_of = [&](int t) -> float {
    if (t > s) return -INFINITY;
    float z = scale * to_f32(l[t]);
    if (bias) z += to_f32(bias[moff + t]);  // m(s,t) = 1 for t <= s either way
    return z;
  };
  float mx = -INFINITY;
  for (int t = lane; t < Tn; t += 32) mx = fmaxf(mx, z_of(t));
  mx = warp_max(mx);
  float sum = 0.f;
  for (int t = lane; t < Tn; t += 32) sum += t > s ? 0.f : expf(z_of(t) - mx);
  sum = warp_sum(sum);
  const float inv = 1.0f / sum;
  for (int t = lane; t < Tn; t += 32) {
    const float sf = t > s ? 0.f : expf(z_of(t) - mx) * inv;
    soft[row * Tn + t] = from_f32<T>(sf);
    if (probs) {
      const float m = (masked && t > s) ? 0.f : 1.f;
      float p = sf;
      if (addm) p += to_f32(addm[moff + t]) * m;
      if (mulm) p *= to_f32(mulm[moff + t]) * m;
      probs[row * Tn + t] = from_f32<T>(p);
    }
  }
}

// dP (gradient of p) -> dlogits (in place of dP allowed), map gradients accumulated with fp32 atomics
template <typename T>
__global__ void __launch_bounds__(256) attn_softmax_bwd_kernel(const T* __restrict__ dp, const T* __restrict__ soft,
                                                                const T* __restrict__ addm, const T* __restrict__ mulm,
                                                                T* __restrict__ dlogits, float* __restrict__ dbias,
                                                                float* __restrict__ dadd, float* __restrict__ dmul,
                                                                int64_t rows, int S, int H, int Tn, float scale,
                                                                int masked) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int h = (int)(row % H);
  const int s = (int)((row / H) % S);
  const int64_t moff = ((int64_t)h * S + s) * Tn;
  const T* g = dp + row * Tn;
  const T* sf = soft + row * Tn;
  auto dsoft_of = [&](int t) -> float {  // gradient reaching soft_t
    const float m = (masked && t > s) ? 0.f : 1.f;
    float d = to_f32(g[t]);
    if (mulm) d *= to_f32(mulm[moff + t]) * m;
    return d;
  };
  float dot = 0.f;
  for (int t = lane; t < Tn; t += 32) dot = fmaf(dsoft_of(t), to_f32(sf[t]), dot);
  dot = warp_sum(dot);
  for (int t = lane; t < Tn; t += 32) {
    const float m = (masked && t > s) ? 0.f : 1.f;
    const float sv = to_f32(sf[t]);
    const float ds = dsoft_of(t);
    if (dmul) {
      float base = sv;
      if (addm) base += to_f32(addm[moff + t]) * m;
      const float v = to_f32(g[t]) * base * m;
      if (v != 0.f) atomicAdd(&dmul[moff + t], v);
    }
    if (dadd && ds * m != 0.f) atomicAdd(&dadd[moff + t], ds * m);
    const float dz = t > s ? 0.f : sv * (ds - dot);
    if (dbias && dz != 0.f) atomicAdd(&dbias[moff + t], dz);
    dlogits[row * Tn + t] = from_f32<T>(scale * dz);
  }
}

// out[o, i] = x[o, i] + bias[i]   (a position embedding added to every batch row)
template <typename T>
__global__ void __launch_bounds__(256) bias_add_kernel(const T* __restrict__ x, const T* __restrict__ bias,
                                                        T* __restrict__ out, int64_t outer, int64_t n) {
  const int64_t total = outer * n;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = from_f32<T>(to_f32(x[i]) + to_f32(bias[i % n]));
}

// out[i] += sum_o x[o, i]   (fp32; its gradient)
template <typename T>
__global__ void __launch_bounds__(256) colsum_kernel(const T* __restrict__ x, float* __restrict__ out, int64_t outer,
                                                      int64_t n, int64_t rows_per_block) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int64_t o0 = (int64_t)blockIdx.y * rows_per_block;
  const int64_t o1 = o0 + rows_per_block < outer ? o0 + rows_per_block : outer;
  float acc = 0.f;
  for (int64_t o = o0; o < o1; ++o) acc += to_f32(x[o * n + i]);
  atomicAdd(&out[i], acc);
}

}  // namespace obst

using namespace obst;

extern "C" {

int obst_attn_softmax_fwd(int dtype, const void* logits, const void* bias, const void* add_map, const void* mul_map,
                          void* soft, void* probs, int64_t batch, int32_t seq, int32_t heads, int32_t keys, float scale,
                          int32_t masked, void* stream) {
  OBST_REQUIRE(logits && soft && batch >= 0 && seq > 0 && heads > 0 && keys > 0, "obst_attn_softmax_fwd: bad arguments");
  OBST_REQUIRE(!((add_map || mul_map) && !probs), "obst_attn_softmax_fwd: maps need a separate `probs` output");
  const int64_t rows = batch * seq * heads;
  if (rows == 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned blocks = (unsigned)((rows + 7) / 8);
  if (dtype == OBST_BF16)
    attn_softmax_fwd_kernel<bf16><<<blocks, 256, 0, st>>>((const bf16*)logits, (const bf16*)bias, (const bf16*)add_map,
                                                          (const bf16*)mul_map, (bf16*)soft, (bf16*)probs, rows, seq,
                                                          heads, keys, scale, masked);
  else if (dtype == OBST_F32)
    attn_softmax_fwd_kernel<float><<<blocks, 256, 0, st>>>((const float*)logits, (const float*)bias,
                                                           (const float*)add_map, (const float*)mul_map, (float*)soft,
                                                           (float*)probs, rows, seq, heads, keys, scale, masked);
  else
    OBST_REQUIRE(false, "obst_attn_softmax_fwd: bad dtype %d", dtype);
  count_launch();
  return check_launch("obst_attn_softmax_fwd");
}

int obst_attn_softmax_bwd(int dtype, const void* dprobs, const void* soft, const void* add_map, const void* mul_map,
                          void* dlogits, float* dbias, float* dadd, float* dmul, int64_t batch, int32_t seq,
                          int32_t heads, int32_t keys, float scale, int32_t masked, void* stream) {
  OBST_REQUIRE(dprobs && soft && dlogits && batch >= 0 && seq > 0 && heads > 0 && keys > 0,
               "obst_attn_softmax_bwd: bad arguments");
  const int64_t rows = batch * seq * heads;
  if (rows == 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned blocks = (unsigned)((rows + 7) / 8);
  if (dtype == OBST_BF16)
    attn_softmax_bwd_kernel<bf16><<<blocks, 256, 0, st>>>((const bf16*)dprobs, (const bf16*)soft, (const bf16*)add_map,
                                                          (const bf16*)mul_map, (bf16*)dlogits, dbias, dadd, dmul, rows,
                                                          seq, heads, keys, scale, masked);
  else if (dtype == OBST_F32)
    attn_softmax_bwd_kernel<float><<<blocks, 256, 0, st>>>((const float*)dprobs, (const float*)soft,
                                                           (const float*)add_map, (const float*)mul_map, (float*)dlogits,
                                                           dbias, dadd, dmul, rows, seq, heads, keys, scale, masked);
  else
    OBST_REQUIRE(false, "obst_attn_softmax_bwd: bad dtype %d", dtype);
  count_launch();
  return check_launch("obst_attn_softmax_bwd");
}

int obst_bias_add(int dtype, const void* x, const void* bias, void* out, int64_t outer, int64_t n, void* stream) {
  OBST_REQUIRE(x && bias && out && outer >= 0 && n > 0, "obst_bias_add: bad arguments");
  if (outer == 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  int64_t blocks = (outer * n + 255) / 256;
  if (blocks > (int64_t)num_sms() * 16) blocks = (int64_t)num_sms() * 16;
  if (dtype == OBST_BF16)
    bias_add_kernel<bf16><<<(unsigned)blocks, 256, 0, st>>>((const bf16*)x, (const bf16*)bias, (bf16*)out, outer, n);
  else if (dtype == OBST_F32)
    bias_add_kernel<float><<<(unsigned)blocks, 256, 0, st>>>((const float*)x, (const float*)bias, (float*)out, outer, n);
  else
    OBST_REQUIRE(false, "obst_bias_add: bad dtype %d", dtype);
  count_launch();
  return check_launch("obst_bias_add");
}

int obst_colsum_accumulate(int dtype, const void* x, float* out, int64_t outer, int64_t n, void* stream) {
  OBST_REQUIRE(x && out && outer >= 0 && n > 0, "obst_colsum_accumulate: bad arguments");
  if (outer == 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  int64_t ysplit = (outer + 63) / 64;
  if (ysplit > 1024) ysplit = 1024;
  const int64_t rpb = (outer + ysplit - 1) / ysplit;
  const dim3 grid((unsigned)((n + 255) / 256), (unsigned)((outer + rpb - 1) / rpb));
  if (dtype == OBST_BF16) colsum_kernel<bf16><<<grid, 256, 0, st>>>((const bf16*)x, out, outer, n, rpb);
  else if (dtype == OBST_F32) colsum_kernel<float><<<grid, 256, 0, st>>>((const float*)x, out, outer, n, rpb);
  else OBST_REQUIRE(false, "obst_colsum_accumulate: bad dtype %d", dtype);
  count_launch();
  return check_launch("obst_colsum_accumulate");
}

}  // extern "C"
