// Generic optimizer stages: every entry of the reference's OPTIMIZERS registry
// (src/optimizer/optimizers.py:32-151) as one multi-tensor launch per stage.
//
// The hot path's chain (adaptive_clip - sm3 - momentum - learning_rate) is ONE fused pass (optimizer.cu).  Any
// other optimizer string the reference accepts is executed stage by stage over the same tensor table and work
// list, on the flat fp32 gradient buffer (each stage rewrites it in place, as each stage function of the reference
// returns the new "grad"), with flat state buffers laid out like the gradient buffer.  Unfused = one HBM round trip
// per stage; that is the price of generality, not of the default configuration.
#include "obst_common.cuh"

namespace obst {

template <typename TS>
__device__ __forceinline__ float ld_st(const void* p, int64_t i) {
  return to_f32(reinterpret_cast<const TS*>(p)[i]);
}
template <typename TS>
__device__ __forceinline__ void st_st(void* p, int64_t i, float v) {
  reinterpret_cast<TS*>(p)[i] = from_f32<TS>(v);
}
// 1 / max(sqrt(x), 1e-5): opt_rsqrt, src/optimizer/optimizers.py:13-14
__device__ __forceinline__ float opt_rsqrt(float x) { return 1.0f / fmaxf(sqrtf(x), 1e-5f); }

__device__ __forceinline__ double block_sum_d(double v, double* s_part) {
  v = warp_sum_d(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) s_part[warp] = v;
  __syncthreads();
  double a = 0.0;
  for (int w = 0; w < (int)(blockDim.x >> 5); ++w) a += s_part[w];
  return a;
}

// One block per work item (a row range of one tensor = a contiguous element range).
template <typename TS>
__global__ void __launch_bounds__(256) opt_stage_kernel(const obst_opt_tensor* __restrict__ table, int ntensors,
                                                         const int32_t* __restrict__ item_tensor,
                                                         const int32_t* __restrict__ item_row0,
                                                         const int32_t* __restrict__ item_rows,
                                                         const obst_opt_stage_args a) {
  __shared__ double s_part[8];
  __shared__ float s_bcast;
  const int item = blockIdx.x;
  const int ti = item_tensor[item];
  const obst_opt_tensor T = table[ti];
  const int cols = T.dims[T.ndim - 1];
  const int64_t e0 = (int64_t)item_row0[item] * cols;
  const int64_t n = (int64_t)item_rows[item] * cols;
  const int64_t flat = (reinterpret_cast<const float*>(T.g) - a.grad_base) + e0;  // element offset inside every flat buffer
  float* __restrict__ g = a.work + flat;
  const float* __restrict__ p = T.p + e0;
  const bool act_bf16 = a.dtype_act == OBST_BF16;
  const double* sums = a.sums ? a.sums + 4 * (int64_t)ti : nullptr;
  const float numel = (float)T.numel;

  switch (a.stage) {
    case OBST_STAGE_REDUCE: {  // sums[t] += {sum g^2, sum g, sum w^2, sum w}, w = cast_AT(p)
      double sg2 = 0, sg = 0, sw2 = 0, sw = 0;
      for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
        const float gv = g[i];
        const float wv = act_bf16 ? round_to<bf16>(p[i]) : p[i];
        sg2 += (double)gv * gv;
        sg += gv;
        sw2 += (double)wv * wv;
        sw += wv;
      }
      sg2 = block_sum_d(sg2, s_part);
      sg = block_sum_d(sg, s_part);
      sw2 = block_sum_d(sw2, s_part);
      sw = block_sum_d(sw, s_part);
      if (threadIdx.x == 0) {
        double* o = a.sums + 4 * (int64_t)ti;
        atomicAdd(o + 0, sg2);
        atomicAdd(o + 1, sg);
        atomicAdd(o + 2, sw2);
        atomicAdd(o + 3, sw);
      }
      return;
    }
    case OBST_STAGE_ROUND_BF16:  // the graph hands the optimizer activation-dtype gradients (src/optimizer/__init__.py:180)
      for (int64_t i = threadIdx.x; i < n; i += blockDim.x) g[i] = round_to<bf16>(g[i]);
      return;
    case OBST_STAGE_COPY:  // work2 <- work (graft: the inner optimizer runs on a copy)
      for (int64_t i = threadIdx.x; i < n; i += blockDim.x) a.work2[flat + i] = g[i];
      return;
    case OBST_STAGE_ADAPTIVE_CLIP:   // optimizers.py:79-84
    case OBST_STAGE_L2NORM_CLIP:     // optimizers.py:87-90
    case OBST_STAGE_GLOBAL_L2NORM_CLIP:  // optimizers.py:93-98 (sum over the raw gradients of every variable)
    case OBST_STAGE_GRAD_CENTRALISE:     // optimizers.py:106-107
    case OBST_STAGE_WEIGHT_CENTRALISE:   // optimizers.py:110-111
    case OBST_STAGE_GRAFT_COMBINE: {     // optimizers.py:147-151
      if (threadIdx.x == 0) {
        const float clip = a.f[0];
        float mul = 1.f, add = 0.f;
        if (a.stage == OBST_STAGE_ADAPTIVE_CLIP) {
          const float gn = fminf(rsqrtf((float)sums[0]), 1e6f);
          const float wn = fmaxf(sqrtf((float)sums[2]), 1e-3f);
          mul = fminf(wn * gn * clip, 1.f);
        } else if (a.stage == OBST_STAGE_L2NORM_CLIP) {
          mul = clip * rsqrtf(fmaxf((float)sums[0], 1.0f / (clip * clip)));
        } else if (a.stage == OBST_STAGE_GLOBAL_L2NORM_CLIP) {
          double tot = 0.0;
          for (int t = 0; t < ntensors; ++t) tot += a.sums[4 * (int64_t)t];
          mul = clip * rsqrtf(fmaxf((float)tot, 1.0f / (clip * clip)));
        } else if (a.stage == OBST_STAGE_GRAD_CENTRALISE) {
          add = -(float)(sums[1] / (double)numel);
        } else if (a.stage == OBST_STAGE_WEIGHT_CENTRALISE) {
          add = (float)(sums[3] / (double)numel);
        } else {  // g * rsqrt(sum g^2) * sqrt(sum inner^2); the inner step's sums are in sums2
          mul = rsqrtf((float)sums[0]) * sqrtf((float)a.sums2[4 * (int64_t)ti]);
        }
        s_part[0] = mul;
        s_bcast = add;
      }
      __syncthreads();
      const float mul = (float)s_part[0], add = s_bcast;
      for (int64_t i = threadIdx.x; i < n; i += blockDim.x) g[i] = fmaf(g[i], mul, add);
      return;
    }
    case OBST_STAGE_VALUE_CLIP: {  // optimizers.py:101-103
      const float c = a.f[0];
      for (int64_t i = threadIdx.x; i < n; i += blockDim.x) g[i] = fmaxf(fminf(g[i], c), -c);
      return;
    }
    case OBST_STAGE_LEARNING_RATE: {  // optimizers.py:114-115
      const float lr = a.f[0];
      for (int64_t i = threadIdx.x; i < n; i += blockDim.x) g[i] *= lr;
      return;
    }
    case OBST_STAGE_MOMENTUM: {  // optimizers.py:118-128; f = {momentum_multiplier, gradient_multiplier}, flags&1 nesterov
      const float mm = a.f[0], gm = a.f[1];
      const bool nesterov = a.flags & 1;
      for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
        const float gv = g[i];
        const float ns = fmaf(mm, ld_st<TS>(a.state_a, flat + i), gv * gm);
        st_st<TS>(a.state_a, flat + i, ns);
        g[i] = nesterov ? fmaf(mm, ns, gv) : ns;
      }
      return;
    }
    case OBST_STAGE_ADAM: {  // optimizers.py:32-43; f = {beta1, beta2, 1/(1-beta1^t), 1/(1-beta2^t)}
      const float b1 = a.f[0], b2 = a.f[1], d1 = a.f[2], d2 = a.f[3];
      for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
        const float gv = g[i];
        const float p2 = ld_st<TS>(a.state_b, flat + i) * b2 + gv * gv * (1.f - b2);
        const float p1 = ld_st<TS>(a.state_a, flat + i) * b1 + gv * (1.f - b1);
        st_st<TS>(a.state_b, flat + i, p2);
        st_st<TS>(a.state_a, flat + i, p1);
        g[i] = opt_rsqrt(p2 * d2) * p1 * d1;
      }
      return;
    }
    case OBST_STAGE_NOVOGRAD: {  // optimizers.py:46-57; f = {beta1, beta2, -, 1/(1-beta2^t)}; scalar slot per tensor
      const float b1 = a.f[0], b2 = a.f[1], d2 = a.f[3];
      const float p2_old = to_f32(reinterpret_cast<const TS*>(a.scalars_in)[ti]);
      const float p2_new = p2_old * b2 + (float)sums[0] * (1.f - b2);
      if (item_row0[item] == 0 && threadIdx.x == 0) reinterpret_cast<TS*>(a.scalars_out)[ti] = from_f32<TS>(p2_new);
      const float r_old = opt_rsqrt(p2_old), r_new = opt_rsqrt(p2_new * d2);
      for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
        const float gv = g[i];
        const float p1 = fmaf(b1, ld_st<TS>(a.state_a, flat + i), gv * r_old);
        st_st<TS>(a.state_a, flat + i, p1);
        g[i] = fmaf(b1, p1, gv * r_new);
      }
      return;
    }
    case OBST_STAGE_APPLY: {  // weight decay + assign_sub, src/optimizer/__init__.py:49-66; f = {lr * weight_decay}
      const float wd = T.decay ? a.f[0] : 0.f;
      for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
        const float pv = p[i];
        float u = g[i];
        if (wd != 0.f) u = fmaf(act_bf16 ? round_to<bf16>(pv) : pv, wd, u);
        const float pn = pv - u;
        T.p[e0 + i] = pn;
        if (T.shadow) {
          float sh = pn;
          if (T.tril_s > 0) {
            const int64_t e = e0 + i;
            if ((int)(e % T.tril_s) > (int)((e / T.tril_s) % T.tril_s)) sh = 0.f;
          }
          if (act_bf16) reinterpret_cast<bf16*>(T.shadow)[e0 + i] = __float2bfloat16_rn(sh);
          else reinterpret_cast<float*>(T.shadow)[e0 + i] = sh;
        }
      }
      return;
    }
    default:
      return;
  }
}

}  // namespace obst

using namespace obst;

extern "C" int obst_opt_stage(const obst_opt_tensor* table, int32_t ntensors, const int32_t* item_tensor,
                              const int32_t* item_row0, const int32_t* item_rows, int32_t nitems,
                              const obst_opt_stage_args* args, void* stream) {
  OBST_REQUIRE(table && item_tensor && item_row0 && item_rows && args, "obst_opt_stage: null argument");
  OBST_REQUIRE(ntensors > 0, "obst_opt_stage: empty table");
  const obst_opt_stage_args a = *args;
  OBST_REQUIRE(a.stage >= OBST_STAGE_REDUCE && a.stage <= OBST_STAGE_APPLY, "obst_opt_stage: bad stage %d", a.stage);
  OBST_REQUIRE(a.grad_base && a.work, "obst_opt_stage: grad_base / work missing");
  const bool needs_sums = a.stage == OBST_STAGE_REDUCE || a.stage == OBST_STAGE_ADAPTIVE_CLIP ||
                          a.stage == OBST_STAGE_L2NORM_CLIP || a.stage == OBST_STAGE_GLOBAL_L2NORM_CLIP ||
                          a.stage == OBST_STAGE_GRAD_CENTRALISE || a.stage == OBST_STAGE_WEIGHT_CENTRALISE ||
                          a.stage == OBST_STAGE_GRAFT_COMBINE || a.stage == OBST_STAGE_NOVOGRAD;
  OBST_REQUIRE(!needs_sums || a.sums, "obst_opt_stage: stage %d needs sums", a.stage);
  OBST_REQUIRE(a.stage != OBST_STAGE_GRAFT_COMBINE || a.sums2, "obst_opt_stage: graft needs sums2");
  OBST_REQUIRE(a.stage != OBST_STAGE_COPY || a.work2, "obst_opt_stage: copy needs work2");
  OBST_REQUIRE(!(a.stage == OBST_STAGE_MOMENTUM || a.stage == OBST_STAGE_ADAM || a.stage == OBST_STAGE_NOVOGRAD) ||
                   a.state_a,
               "obst_opt_stage: stage %d needs state_a", a.stage);
  OBST_REQUIRE(a.stage != OBST_STAGE_ADAM || a.state_b, "obst_opt_stage: adam needs state_b");
  OBST_REQUIRE(a.stage != OBST_STAGE_NOVOGRAD || (a.scalars_in && a.scalars_out && a.scalars_in != a.scalars_out),
               "obst_opt_stage: novograd needs distinct scalars_in / scalars_out");
  OBST_REQUIRE(a.dtype_state == OBST_BF16 || a.dtype_state == OBST_F32, "obst_opt_stage: bad state dtype %d",
               a.dtype_state);
  if (nitems <= 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (a.stage == OBST_STAGE_REDUCE && (a.flags & 2)) {  // clear the sums first
    cudaError_t e = cudaMemsetAsync(a.sums, 0, sizeof(double) * 4 * (size_t)ntensors, st);
    if (e != cudaSuccess) {
      set_error("obst_opt_stage: memset failed: %s", cudaGetErrorString(e));
      return OBST_ECUDA;
    }
  }
  if (a.dtype_state == OBST_BF16)
    opt_stage_kernel<bf16><<<(unsigned)nitems, 256, 0, st>>>(table, ntensors, item_tensor, item_row0, item_rows, a);
  else
    opt_stage_kernel<float><<<(unsigned)nitems, 256, 0, st>>>(table, ntensors, item_tensor, item_row0, item_rows, a);
  count_launch();
  return check_launch("obst_opt_stage");
}
