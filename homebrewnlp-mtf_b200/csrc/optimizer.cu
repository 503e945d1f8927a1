// Fused multi-tensor optimizer update for
//   adaptive_clip:<c> - sm3 - momentum:<beta>:1:<nesterov> - learning_rate  (+ weight decay, assign_sub)
// src/optimizer/__init__.py:31-66, src/optimizer/optimizers.py:13-14,60-84,114-128.
//
// HBM-bound: one pass computes the two L2 norms per tensor, one pass does everything else.  A tensor
// is viewed as [rows, cols] (cols = last dim).  A warp owns whole rows, lanes own fixed columns, so
//   - the SM3 "max over the other dims" of the last dim lives in registers across rows,
//   - the max for every other dim is one warp reduction per row + one atomicMax,
// and the new accumulators are staged in fp32 (acc_new) because an SM3 accumulator may DEcrease.
#include "optimizer_math.cuh"

namespace obst {

template <typename TS>
__device__ __forceinline__ float ld_state(const void* p, int64_t i) {
  return to_f32(reinterpret_cast<const TS*>(p)[i]);
}

__device__ __forceinline__ float ld_grad(const obst_opt_tensor& T, int64_t i, int grad_bf16) {
  if (T.g_bf16) return to_f32(reinterpret_cast<const bf16*>(T.g)[i]);
  const float g = reinterpret_cast<const float*>(T.g)[i];
  return grad_bf16 ? round_to<bf16>(g) : g;
}

// ------------------------------------------------------------------------------------------------
// pass 1: per-tensor sum g^2 and sum cast_AT(p)^2, in float64 (see optimizer_math.cuh)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) opt_norms_kernel(const obst_opt_tensor* __restrict__ table,
                                                         const int32_t* __restrict__ item_tensor,
                                                         const int32_t* __restrict__ item_row0,
                                                         const int32_t* __restrict__ item_rows, int act_bf16,
                                                         int grad_bf16, double* __restrict__ norms, int carry,
                                                         int skip_g, double* __restrict__ wnext, int carry_first,
                                                         int carry_count, int w_from_shadow) {
  __shared__ double s_part[2][8];
  const int item = blockIdx.x;
  // carry mode: sum cast_AT(p)^2 was accumulated by the previous update (it wrote every p) -- move it into place
  // instead of reading the weights again
  if (carry && blockIdx.x == 0) {
    for (int t = threadIdx.x; t < carry_count; t += blockDim.x) {
      norms[2 * (carry_first + t) + 1] = wnext[carry_first + t];
      wnext[carry_first + t] = 0.0;
    }
  }
  if (carry && skip_g) return;  // nothing to read
  const int ti = item_tensor[item];
  const obst_opt_tensor T = table[ti];
  const int64_t cols = T.dims[T.ndim - 1];
  const int64_t e0 = (int64_t)item_row0[item] * cols;
  const int64_t n = (int64_t)item_rows[item] * cols;
  const float* __restrict__ p = T.p + e0;
  double sg = 0.0, sg2 = 0.0, sw = 0.0;
  const bool g_is_bf16_value = T.g_bf16 || grad_bf16;
  if (!skip_g) {
    if (T.g_bf16) {
      const bf16* __restrict__ g = reinterpret_cast<const bf16*>(T.g) + e0;
      const bool vec_ok = ((e0 & 7) == 0);
      const int64_t nv = vec_ok ? n / 8 : 0;
#pragma unroll 4
      for (int64_t i = threadIdx.x; i < nv; i += blockDim.x) {
        const uint4 q = reinterpret_cast<const uint4*>(g)[i];
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          // the square of a bf16 value is exact in fp32; the squares are summed in float64 (two accumulators
          // shorten the dependent chain)
          const float lo = __uint_as_float(w[j] << 16), hi = __uint_as_float(w[j] & 0xFFFF0000u);
          sg += (double)__fmul_rn(lo, lo);
          sg2 += (double)__fmul_rn(hi, hi);
        }
      }
      for (int64_t i = nv * 8 + threadIdx.x; i < n; i += blockDim.x) sg += square_as_double(to_f32(g[i]), true);
    } else {
      const float* __restrict__ g = reinterpret_cast<const float*>(T.g) + e0;
      for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
        const float gg = grad_bf16 ? round_to<bf16>(g[i]) : g[i];
        sg += square_as_double(gg, g_is_bf16_value);
      }
    }
  }
  if (!carry) {
    const bool from_shadow = w_from_shadow && T.shadow != nullptr;
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
      float ww = act_bf16 ? round_to<bf16>(p[i]) : p[i];
      if (from_shadow) {  // macro-batching chain: see OptHyper::w_from_shadow
        const int64_t e = e0 + i;
        const bool masked = T.tril_s > 0 && (int)(e % T.tril_s) > (int)((e / T.tril_s) % T.tril_s);
        if (!masked)
          ww = act_bf16 ? to_f32(reinterpret_cast<const bf16*>(T.shadow)[e]) : reinterpret_cast<const float*>(T.shadow)[e];
      }
      sw += square_as_double(ww, act_bf16 != 0);
    }
  }
  const double dg = warp_sum_d(sg + sg2), dw = warp_sum_d(sw);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) {
    s_part[0][warp] = dg;
    s_part[1][warp] = dw;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0.0, b = 0.0;
    for (int w = 0; w < 8; ++w) {
      a += s_part[0][w];
      b += s_part[1][w];
    }
    if (!skip_g) atomicAdd(&norms[2 * ti], a);
    if (!carry) atomicAdd(&norms[2 * ti + 1], b);
  }
}

// ------------------------------------------------------------------------------------------------
// gradient hand-over: g (bf16) = round(g_acc fp32), optionally summing the squares of the rounded values
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) grad_finalize_kernel(const obst_opt_tensor* __restrict__ table,
                                                             const int32_t* __restrict__ item_tensor,
                                                             const int32_t* __restrict__ item_row0,
                                                             const int32_t* __restrict__ item_rows,
                                                             double* __restrict__ norms, int with_norms) {
  __shared__ double s_part[8];
  const int item = blockIdx.x;
  const int ti = item_tensor[item];
  const obst_opt_tensor T = table[ti];
  if (!T.g_bf16 || T.g_acc == nullptr) return;
  const int64_t cols = T.dims[T.ndim - 1];
  const int64_t e0 = (int64_t)item_row0[item] * cols;
  const int64_t n = (int64_t)item_rows[item] * cols;
  const float* __restrict__ src = T.g_acc + e0;
  bf16* __restrict__ dst = reinterpret_cast<bf16*>(T.g) + e0;
  double sg = 0.0;
  const bool vec_ok = ((e0 & 7) == 0);
  const int64_t nv = vec_ok ? n / 8 : 0;
  for (int64_t i = threadIdx.x; i < nv; i += blockDim.x) {
    const float4 a = reinterpret_cast<const float4*>(src)[2 * i], b = reinterpret_cast<const float4*>(src)[2 * i + 1];
    uint4 o;
    o.x = pack_bf16x2(a.x, a.y);
    o.y = pack_bf16x2(a.z, a.w);
    o.z = pack_bf16x2(b.x, b.y);
    o.w = pack_bf16x2(b.z, b.w);
    reinterpret_cast<uint4*>(dst)[i] = o;
    if (with_norms) {
      const uint32_t w[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const double lo = bf16_bits_to_double(w[j] << 16), hi = bf16_bits_to_double(w[j] & 0xFFFF0000u);
        sg = fma(lo, lo, sg);
        sg = fma(hi, hi, sg);
      }
    }
  }
  for (int64_t i = nv * 8 + threadIdx.x; i < n; i += blockDim.x) {
    const bf16 v = __float2bfloat16_rn(src[i]);
    dst[i] = v;
    if (with_norms) sg += square_as_double(to_f32(v), true);
  }
  if (!with_norms) return;
  const double dg = warp_sum_d(sg);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) s_part[warp] = dg;
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0.0;
    for (int w = 0; w < 8; ++w) a += s_part[w];
    atomicAdd(&norms[2 * ti], a);
  }
}

// ------------------------------------------------------------------------------------------------
// pass 2, generic kernel: any last dim (VEC = 1) -- the streaming kernel of optimizer_ring.cu takes every launch
// whose last dims are multiples of 4; this one also serves the SM3-only stage of the generic optimizer chains
// ------------------------------------------------------------------------------------------------
// A tensor is viewed as [rows, cols] (cols = last dim).  A warp owns whole rows, lanes own fixed columns, so
//   - the SM3 "max over the other dims" of the last dim lives in registers across rows,
//   - the max for every other dim is one warp reduction per row + one atomicMax,
// and the new accumulators are staged in fp32 (acc_new) because an SM3 accumulator may DEcrease.
// blockIdx.y selects a chunk of 32 * VEC * NI columns, so a lane never owns more than NI vectors of a row.
template <typename TS, int VEC, int NI>
__global__ void __launch_bounds__(256) opt_update_kernel(const obst_opt_tensor* __restrict__ table,
                                                          const int32_t* __restrict__ item_tensor,
                                                          const int32_t* __restrict__ item_row0,
                                                          const int32_t* __restrict__ item_rows,
                                                          const double* __restrict__ norms, const OptHyper hp,
                                                          double* __restrict__ wnext) {
  constexpr int CHUNK = 32 * VEC * NI;
  __shared__ float s_col[CHUNK];     // block-level column maxima of this chunk
  __shared__ obst_opt_tensor sT;     // the table entry (dynamically indexed: keep it out of local memory)
  __shared__ double s_w2[8];
  __shared__ float s_rowmax[OPT_MAX_ITEM_ROWS];
  const int item = blockIdx.x;
  const int ti = item_tensor[item];
  if (item_rows[item] > OPT_MAX_ITEM_ROWS) __trap();  // contract of the work list (include/obst_b200.h)
  if (threadIdx.x == 0) sT = table[ti];
  __syncthreads();
  const obst_opt_tensor& T = sT;
  const int nd = T.ndim;
  const int cols = T.dims[nd - 1];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cbase = (int)blockIdx.y * CHUNK;
  if (cbase >= cols) return;  // whole block: this tensor is narrower than the widest one of the launch
  for (int i = threadIdx.x; i < CHUNK; i += blockDim.x) s_col[i] = 0.f;
  opt_row_minima<TS>(T, item_row0[item], item_rows[item], s_rowmax);
  __syncthreads();
  float* const Tp = T.p;
  void* const Tm = T.m;
  void* const Tshadow = T.shadow;
  const int tril_s = T.tril_s;
  const bool wd_on = T.decay && hp.weight_decay > 0.f;
  const float scale = hp.clip > 0.f ? clip_scale(norms[2 * ti], norms[2 * ti + 1], hp.clip) : 1.f;
  const void* acc_last = T.acc[nd - 1];
  float al[NI * VEC], cm[NI * VEC];
#pragma unroll
  for (int i = 0; i < NI; ++i) {
    const int c = cbase + (lane + 32 * i) * VEC;
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      cm[i * VEC + j] = 0.f;
      al[i * VEC + j] = (c + j < cols) ? ld_state<TS>(acc_last, c + j) : 0.f;
    }
  }
  const int row_begin = item_row0[item];
  const int row_end = row_begin + item_rows[item];
  double w2 = 0.0;  // sum cast_AT(p_new)^2 of this thread: next step's weight norm, for free
  for (int r = row_begin + warp; r < row_end; r += 8) {
    const float rowmin = s_rowmax[r - row_begin];
    float rowmax = 0.f;
    const int64_t rbase = (int64_t)r * cols;
#pragma unroll
    for (int i = 0; i < NI; ++i) {
#pragma unroll
      for (int j = 0; j < VEC; ++j) {
        const int c = cbase + (lane + 32 * i) * VEC + j;
        if (c >= cols) continue;
        const int64_t e = rbase + c;
        const float pv = Tp[e];
        const float g = __fmul_rn(ld_grad(T, e, hp.grad_bf16), scale);
        float w = hp.act_bf16 ? round_to<bf16>(pv) : pv;
        bool masked = false;
        if (tril_s > 0) masked = (int)(e % tril_s) > (int)((e / tril_s) % tril_s);
        // macro-batching chain: the graph reads the previous macro-batch's would-be weights, which live in the
        // shadow (masked entries of a mixing map never move, so there the cast of p is still the value)
        if (hp.w_from_shadow && Tshadow && !masked)
          w = hp.act_bf16 ? to_f32(reinterpret_cast<const bf16*>(Tshadow)[e]) : reinterpret_cast<const float*>(Tshadow)[e];
        const float mv = hp.sm3_only ? 0.f : ld_state<TS>(Tm, e);
        const OptElemOut o = opt_element(g, fminf(rowmin, al[i * VEC + j]), mv, pv, w, hp, wd_on);
        cm[i * VEC + j] = fmaxf(cm[i * VEC + j], o.nu);
        rowmax = fmaxf(rowmax, o.nu);
        if (hp.sm3_only) {
          reinterpret_cast<float*>(T.g)[e] = o.u;
          continue;
        }
        float sh = o.p_new;
        if (hp.preview) {  // subtract(var.value, cast(update)) in the activation dtype
          const float w0 = hp.act_bf16 ? round_to<bf16>(pv) : pv;
          const float us = hp.act_bf16 ? round_to<bf16>(o.step) : o.step;
          sh = __fsub_rn(w0, us);
        }
        if (masked) sh = 0.f;
        if (Tshadow) {
          if (hp.act_bf16) reinterpret_cast<bf16*>(Tshadow)[e] = __float2bfloat16_rn(sh);
          else reinterpret_cast<float*>(Tshadow)[e] = sh;
        }
        reinterpret_cast<TS*>(Tm)[e] = from_f32<TS>(o.m_new);
        if (hp.preview) continue;
        Tp[e] = o.p_new;
        const float wv = hp.act_bf16 ? round_to<bf16>(o.p_new) : o.p_new;
        w2 = add_square(w2, wv, hp.act_bf16 != 0);
      }
    }
    rowmax = warp_max(rowmax);
    if (lane == 0) s_rowmax[r - row_begin] = rowmax;
  }
  __syncthreads();
  opt_row_maxima_commit(T, item_row0[item], item_rows[item], s_rowmax);
#pragma unroll
  for (int i = 0; i < NI; ++i) {
    const int c = (lane + 32 * i) * VEC;
#pragma unroll
    for (int j = 0; j < VEC; ++j)
      if (cbase + c + j < cols && cm[i * VEC + j] > 0.f) atomic_max_nonneg(&s_col[c + j], cm[i * VEC + j]);
  }
  if (wnext) {
    const double dw = warp_sum_d(w2);
    if (lane == 0) s_w2[warp] = dw;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < CHUNK && cbase + i < cols; i += blockDim.x) {
    const float v = s_col[i];
    if (v > 0.f) atomic_max_nonneg(T.acc_new[nd - 1] + cbase + i, v);
  }
  if (wnext && threadIdx.x == 0) {
    double a = 0.0;
    for (int w = 0; w < 8; ++w) a += s_w2[w];
    atomicAdd(&wnext[ti], a);
  }
}

template <typename TS>
__global__ void __launch_bounds__(256) opt_finalize_kernel(const obst_opt_tensor* __restrict__ table) {
  const obst_opt_tensor T = table[blockIdx.x];
  for (int i = 0; i < T.ndim; ++i) {
    TS* acc = reinterpret_cast<TS*>(T.acc[i]);
    float* an = T.acc_new[i];
    for (int k = threadIdx.x; k < T.dims[i]; k += blockDim.x) {
      acc[k] = from_f32<TS>(an[k]);
      an[k] = 0.f;
    }
  }
}

// One update launch over a work list: the streaming kernel when every last dim is a multiple of 4, else the generic one.
static int launch_update(const obst_opt_tensor* table, const int32_t* item_tensor, const int32_t* item_row0,
                         const int32_t* item_rows, int nitems, int max_cols, int vec4, int dtype_state,
                         const double* norms, const OptHyper& hp, double* wnext, cudaStream_t st, const char* what) {
  if (vec4 && !hp.sm3_only && !hp.preview && !hp.w_from_shadow)  // the macro-batching chain takes the generic kernel
    return opt_update_ring_launch(table, item_tensor, item_row0, item_rows, nitems, max_cols, dtype_state, norms, hp,
                                  wnext, st);
  const int vec = vec4 ? 4 : 1;
  const int chunk = 32 * vec * 4;
  const int64_t ychunks = ((int64_t)max_cols + chunk - 1) / chunk;
  OBST_REQUIRE(ychunks <= 65535, "%s: last dim %d too large (max %d)", what, max_cols, 65535 * chunk);
  const dim3 grid((unsigned)nitems, (unsigned)ychunks);  // y: column chunks
#define OBST_OPT_NI(TS, V) \
  opt_update_kernel<TS, V, 4><<<grid, 256, 0, st>>>(table, item_tensor, item_row0, item_rows, norms, hp, wnext)
  if (dtype_state == OBST_BF16) {
    if (vec4) OBST_OPT_NI(bf16, 4);
    else OBST_OPT_NI(bf16, 1);
  } else {
    if (vec4) OBST_OPT_NI(float, 4);
    else OBST_OPT_NI(float, 1);
  }
#undef OBST_OPT_NI
  count_launch();
  return check_launch(what);
}

}  // namespace obst

using namespace obst;

extern "C" {

int obst_opt_norms(const obst_opt_tensor* table, int32_t ntensors, const int32_t* item_tensor,
                   const int32_t* item_row0, const int32_t* item_rows, int32_t nitems, int32_t dtype_act,
                   int32_t grad_bf16, double* norms, int32_t flags, double* wnext, int32_t first_tensor,
                   int32_t tensor_count, void* stream) {
  OBST_REQUIRE(table && norms && ntensors > 0, "obst_opt_norms: bad arguments");
  const int carry = (flags & 1) != 0;
  const int skip_g = (flags & 4) != 0;
  OBST_REQUIRE(!carry || wnext, "obst_opt_norms: carry mode needs wnext");
  OBST_REQUIRE(!(skip_g && !(flags & 2)), "obst_opt_norms: flags&4 (sum g^2 already present) needs flags&2 (no clear)");
  OBST_REQUIRE(first_tensor >= 0 && tensor_count >= 0 && first_tensor + tensor_count <= ntensors,
               "obst_opt_norms: tensor range [%d, %d) outside the table", first_tensor, first_tensor + tensor_count);
  cudaStream_t st = (cudaStream_t)stream;
  if (!(flags & 2)) {
    cudaError_t e = cudaMemsetAsync(norms + 2 * (size_t)first_tensor, 0, sizeof(double) * 2 * (size_t)tensor_count, st);
    if (e != cudaSuccess) {
      set_error("obst_opt_norms: memset failed: %s", cudaGetErrorString(e));
      return OBST_ECUDA;
    }
  }
  if (nitems <= 0) return OBST_OK;
  // carry + skip_g: only the carried weight sums have to be moved into place -- one block
  const unsigned blocks = (carry && skip_g) ? 1u : (unsigned)nitems;
  opt_norms_kernel<<<blocks, 256, 0, st>>>(table, item_tensor, item_row0, item_rows, dtype_act == OBST_BF16,
                                           grad_bf16, norms, carry, skip_g, wnext, first_tensor, tensor_count,
                                           (flags & 8) != 0);
  count_launch();
  return check_launch("obst_opt_norms");
}

int obst_grad_finalize(const obst_opt_tensor* table, int32_t ntensors, const int32_t* item_tensor,
                       const int32_t* item_row0, const int32_t* item_rows, int32_t nitems, double* norms,
                       int32_t flags, void* stream) {
  OBST_REQUIRE(table && ntensors > 0 && item_tensor && item_row0 && item_rows, "obst_grad_finalize: bad arguments");
  OBST_REQUIRE(!(flags & 1) || norms, "obst_grad_finalize: flags&1 needs norms");
  if (nitems <= 0) return OBST_OK;
  grad_finalize_kernel<<<(unsigned)nitems, 256, 0, (cudaStream_t)stream>>>(table, item_tensor, item_row0, item_rows,
                                                                            norms, flags & 1);
  count_launch();
  return check_launch("obst_grad_finalize");
}

int obst_opt_update(const obst_opt_tensor* table, int32_t ntensors, const int32_t* item_tensor,
                    const int32_t* item_row0, const int32_t* item_rows, int32_t nitems, int32_t max_cols,
                    int32_t vec4, int32_t dtype_act, int32_t dtype_state, int32_t grad_bf16, const double* norms,
                    double* wnext, float clip, float beta, int32_t nesterov, float lr, float weight_decay,
                    int32_t preview, void* stream) {
  OBST_REQUIRE(table && ntensors > 0 && max_cols > 0, "obst_opt_update: bad arguments");
  OBST_REQUIRE(clip <= 0.f || norms != nullptr, "obst_opt_update: clip > 0 needs norms");
  OBST_REQUIRE(dtype_state == OBST_BF16 || dtype_state == OBST_F32, "obst_opt_update: bad state dtype %d", dtype_state);
  if (nitems <= 0) return OBST_OK;
  OptHyper hp;
  hp.clip = clip; hp.beta = beta; hp.lr = lr; hp.weight_decay = weight_decay;
  hp.nesterov = nesterov; hp.act_bf16 = dtype_act == OBST_BF16; hp.grad_bf16 = grad_bf16;
  hp.sm3_only = 0;
  hp.preview = (preview & 1) != 0;
  hp.w_from_shadow = (preview & 2) != 0;
  return launch_update(table, item_tensor, item_row0, item_rows, nitems, max_cols, vec4, dtype_state, norms, hp,
                       hp.preview ? nullptr : wnext, (cudaStream_t)stream, "obst_opt_update");
}

/* SM3 as a stand-alone stage of a generic chain: g <- g * opt_rsqrt(nu), accumulators max-reduced into acc_new
 * (obst_opt_finalize commits them).  Same kernel as obst_opt_update with everything after SM3 switched off. */
int obst_opt_sm3_stage(const obst_opt_tensor* table, int32_t ntensors, const int32_t* item_tensor,
                       const int32_t* item_row0, const int32_t* item_rows, int32_t nitems, int32_t max_cols,
                       int32_t vec4, int32_t dtype_state, void* stream) {
  OBST_REQUIRE(table && item_tensor && item_row0 && item_rows, "obst_opt_sm3_stage: null argument");
  OBST_REQUIRE(max_cols > 0, "obst_opt_sm3_stage: last dim %d unsupported", max_cols);
  OBST_REQUIRE(dtype_state == OBST_BF16 || dtype_state == OBST_F32, "obst_opt_sm3_stage: bad state dtype %d",
               dtype_state);
  if (nitems <= 0) return OBST_OK;
  OptHyper hp;
  hp.clip = 0.f; hp.beta = 0.f; hp.lr = 1.f; hp.weight_decay = 0.f;
  hp.nesterov = 0; hp.act_bf16 = 0; hp.grad_bf16 = 0; hp.sm3_only = 1; hp.preview = 0; hp.w_from_shadow = 0;
  return launch_update(table, item_tensor, item_row0, item_rows, nitems, max_cols, vec4, dtype_state, nullptr, hp,
                       nullptr, (cudaStream_t)stream, "obst_opt_sm3_stage");
}

int obst_opt_finalize(const obst_opt_tensor* table, int32_t ntensors, int32_t dtype_state, void* stream) {
  OBST_REQUIRE(table && ntensors > 0, "obst_opt_finalize: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype_state == OBST_BF16) opt_finalize_kernel<bf16><<<(unsigned)ntensors, 256, 0, st>>>(table);
  else if (dtype_state == OBST_F32) opt_finalize_kernel<float><<<(unsigned)ntensors, 256, 0, st>>>(table);
  else OBST_REQUIRE(false, "obst_opt_finalize: bad state dtype %d", dtype_state);
  count_launch();
  return check_launch("obst_opt_finalize");
}

}  // extern "C"

// ================================================================================================
// Allocation-free single-tensor update
// ================================================================================================
// The entry points above expect the caller to have put a tensor table, a work list, the norm scratch and the
// accumulator staging into device memory.  Some hosts cannot do that: a plug-in living inside a process whose
// framework has already claimed the whole GPU for its own pool, or a runtime that forbids driver allocation calls
// from extension code.  For them the same three kernels are driven from a fixed block of device memory that is
// part of this library's image: nothing is allocated, by us or by the caller.
namespace obst {

static constexpr int NA_MAX_ITEMS = 4096;       // work items (row ranges) of one tensor
static constexpr int NA_MAX_DIM_SUM = 16384;    // sum of the tensor's dims (SM3 staging, one float per entry)
static constexpr size_t NA_OFF_TABLE = 0;                                  // 1 x obst_opt_tensor (padded to 256)
static constexpr size_t NA_OFF_NORMS = 256;                                // 2 doubles (padded to 256)
static constexpr size_t NA_OFF_ITEMS = 512;                                // 3 x int32[NA_MAX_ITEMS]
static constexpr size_t NA_OFF_ACC = NA_OFF_ITEMS + 3 * 4 * (size_t)NA_MAX_ITEMS;  // float[NA_MAX_DIM_SUM + 4*64]
static constexpr size_t NA_BYTES = NA_OFF_ACC + 4 * ((size_t)NA_MAX_DIM_SUM + 4 * 64);
static_assert(sizeof(obst_opt_tensor) <= 256, "table slot too small");
static_assert(NA_BYTES <= 128 * 1024, "keep the resident scratch modest");

// Lives in every GPU that loads the library, from load to exit (about 114 KB).  In HOST code this name is only a
// stand-in for the symbol: the device address must be asked for (cudaGetSymbolAddress) before it is handed to a
// kernel.
static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_BYTES];

struct NaFill {
  obst_opt_tensor t;  // acc_new[] is filled in on the device
  int32_t rows, rows_per_item, nitems, dim_sum_padded;
};

// Writes the one-entry table and the work list, and clears what the update expects to start at zero
// (norm sums, SM3 staging): the block keeps whatever the previous call left behind.
__global__ void __launch_bounds__(256) na_fill_kernel(unsigned char* scratch, NaFill f) {
  obst_opt_tensor* table = reinterpret_cast<obst_opt_tensor*>(scratch + NA_OFF_TABLE);
  double* norms = reinterpret_cast<double*>(scratch + NA_OFF_NORMS);
  int32_t* item_tensor = reinterpret_cast<int32_t*>(scratch + NA_OFF_ITEMS);
  int32_t* item_row0 = item_tensor + NA_MAX_ITEMS;
  int32_t* item_rows = item_row0 + NA_MAX_ITEMS;
  float* acc = reinterpret_cast<float*>(scratch + NA_OFF_ACC);
  if (threadIdx.x == 0) {
    obst_opt_tensor t = f.t;
    int off = 0;
    for (int i = 0; i < 4; ++i) {
      t.acc_new[i] = i < t.ndim ? acc + off : nullptr;
      if (i < t.ndim) off += t.dims[i] + ((64 - t.dims[i] % 64) % 64);
    }
    *table = t;
    norms[0] = 0.0;
    norms[1] = 0.0;
  }
  for (int i = threadIdx.x; i < f.nitems; i += blockDim.x) {
    const int r0 = i * f.rows_per_item;
    item_tensor[i] = 0;
    item_row0[i] = r0;
    item_rows[i] = min(f.rows_per_item, f.rows - r0);
  }
  for (int i = threadIdx.x; i < f.dim_sum_padded; i += blockDim.x) acc[i] = 0.f;
}

}  // namespace obst

extern "C" int obst_opt_step_noalloc(float* p, void* shadow, float* g, void* m, void* const* host_acc,
                                     const int32_t* host_dims, int32_t ndim, int32_t decay, int32_t tril_s,
                                     int32_t dtype_act, int32_t dtype_state, int32_t grad_bf16, float clip, float beta,
                                     int32_t nesterov, float lr, float weight_decay, void* stream) {
  OBST_REQUIRE(p && g && m && host_acc && host_dims, "obst_opt_step_noalloc: null argument");
  OBST_REQUIRE(ndim >= 1 && ndim <= 4, "obst_opt_step_noalloc: ndim %d outside 1..4", ndim);
  OBST_REQUIRE(dtype_state == OBST_BF16 || dtype_state == OBST_F32, "obst_opt_step_noalloc: bad state dtype %d",
               dtype_state);
  NaFill f;
  memset(&f, 0, sizeof(f));
  int64_t numel = 1;
  int64_t dim_sum = 0;
  for (int i = 0; i < ndim; ++i) {
    OBST_REQUIRE(host_dims[i] > 0 && host_acc[i], "obst_opt_step_noalloc: dim %d is empty or has no accumulator", i);
    numel *= host_dims[i];
    dim_sum += host_dims[i] + ((64 - host_dims[i] % 64) % 64);
    f.t.dims[i] = host_dims[i];
    f.t.acc[i] = host_acc[i];
  }
  const int cols = host_dims[ndim - 1];
  OBST_REQUIRE(dim_sum <= NA_MAX_DIM_SUM + 4 * 64,
               "obst_opt_step_noalloc: the dims (each rounded up to 64) add up to %lld, the built-in staging holds %d",
               (long long)dim_sum, NA_MAX_DIM_SUM + 4 * 64);
  const int vec4 = (cols % 4 == 0) ? 1 : 0;
  if (vec4)
    OBST_REQUIRE(((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
                   reinterpret_cast<uintptr_t>(shadow)) & 15) == 0,
                 "obst_opt_step_noalloc: p, g, m and shadow must be 16-byte aligned when the last dim is a multiple of 4");
  const int64_t rows64 = numel / cols;
  // rows per work item: ~32768 elements of one 512-column chunk (wider rows are split into chunks by the kernels)
  const int ccols = cols < 512 ? cols : 512;
  int rpi = 32768 / ccols;
  if (rpi < 1) rpi = 1;
  if (rpi > OPT_MAX_ITEM_ROWS) rpi = OPT_MAX_ITEM_ROWS;
  const int64_t nitems64 = (rows64 + rpi - 1) / rpi;
  OBST_REQUIRE(nitems64 <= NA_MAX_ITEMS,
               "obst_opt_step_noalloc: tensor of %lld elements needs %lld work items, the built-in list holds %d "
               "(about %lld elements at this last dim)",
               (long long)numel, (long long)nitems64, NA_MAX_ITEMS, (long long)NA_MAX_ITEMS * rpi * cols);
  f.t.p = p;
  f.t.shadow = shadow;
  f.t.g = g;
  f.t.m = m;
  f.t.numel = numel;
  f.t.ndim = ndim;
  f.t.decay = decay;
  f.t.tril_s = tril_s;
  f.rows = (int32_t)rows64;
  f.rows_per_item = rpi;
  f.nitems = (int32_t)nitems64;
  f.dim_sum_padded = (int32_t)dim_sum;

  // Set-up, once per device, with that device current: turn the symbol into a device address.
  static unsigned char* dev_ptr[64] = {nullptr};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= 64) {
    set_error("obst_opt_step_noalloc: cudaGetDevice failed or device index %d unsupported: %s", dev,
              cudaGetErrorString(e));
    return OBST_ECUDA;
  }
  if (dev_ptr[dev] == nullptr) {
    unsigned char* ptr = nullptr;
    e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
    if (e != cudaSuccess || ptr == nullptr) {
      set_error("obst_opt_step_noalloc: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
      return OBST_ECUDA;
    }
    dev_ptr[dev] = ptr;
  }
  unsigned char* scratch = dev_ptr[dev];
  cudaStream_t st = (cudaStream_t)stream;
  const obst_opt_tensor* table = reinterpret_cast<const obst_opt_tensor*>(scratch + NA_OFF_TABLE);
  double* norms = reinterpret_cast<double*>(scratch + NA_OFF_NORMS);
  const int32_t* item_tensor = reinterpret_cast<const int32_t*>(scratch + NA_OFF_ITEMS);
  const int32_t* item_row0 = item_tensor + NA_MAX_ITEMS;
  const int32_t* item_rows = item_row0 + NA_MAX_ITEMS;

  na_fill_kernel<<<1, 256, 0, st>>>(scratch, f);
  count_launch();
  if (clip > 0.f) {
    opt_norms_kernel<<<(unsigned)f.nitems, 256, 0, st>>>(table, item_tensor, item_row0, item_rows,
                                                         dtype_act == OBST_BF16, grad_bf16, norms, 0, 0, nullptr, 0, 0, 0);
    count_launch();
  }
  OptHyper hp;
  hp.clip = clip; hp.beta = beta; hp.lr = lr; hp.weight_decay = weight_decay;
  hp.nesterov = nesterov; hp.act_bf16 = dtype_act == OBST_BF16; hp.grad_bf16 = grad_bf16;
  hp.sm3_only = 0;
  hp.preview = 0;
  hp.w_from_shadow = 0;
  const int rc = launch_update(table, item_tensor, item_row0, item_rows, f.nitems, cols, vec4, dtype_state, norms, hp,
                               nullptr, st, "obst_opt_step_noalloc");
  if (rc != OBST_OK) return rc;
  if (dtype_state == OBST_BF16) opt_finalize_kernel<bf16><<<1, 256, 0, st>>>(table);
  else opt_finalize_kernel<float><<<1, 256, 0, st>>>(table);
  count_launch();
  return check_launch("obst_opt_step_noalloc");
}
