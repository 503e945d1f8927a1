// Fused multi-tensor optimizer update for
//   adaptive_clip:<c> - sm3 - momentum:<beta>:1:<nesterov> - learning_rate  (+ weight decay, assign_sub)
// src/optimizer/__init__.py:31-66, src/optimizer/optimizers.py:13-14,60-84,114-128.
//
// HBM-bound: one pass computes the two L2 norms per tensor, one pass does everything else.  A tensor
// is viewed as [rows, cols] (cols = last dim).  A warp owns whole rows, lanes own fixed columns, so
//   - the SM3 "max over the other dims" of the last dim lives in registers across rows,
//   - the max for every other dim is one warp reduction per row + one atomicMax,
// and the new accumulators are staged in fp32 (acc_new) because an SM3 accumulator may DEcrease.
#include "obst_common.cuh"

namespace obst {

__device__ __forceinline__ void atomic_max_nonneg(float* addr, float v) {
  // v >= 0 and *addr >= 0: IEEE order == signed-int order
  atomicMax(reinterpret_cast<int*>(addr), __float_as_int(v));
}

template <typename TS>
__device__ __forceinline__ float ld_state(const void* p, int64_t i) {
  return to_f32(reinterpret_cast<const TS*>(p)[i]);
}

// ------------------------------------------------------------------------------------------------
// pass 1: per-tensor sum g^2 and sum cast_AT(p)^2
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) opt_norms_kernel(const obst_opt_tensor* __restrict__ table,
                                                         const int32_t* __restrict__ item_tensor,
                                                         const int32_t* __restrict__ item_row0,
                                                         const int32_t* __restrict__ item_rows, int act_bf16,
                                                         int grad_bf16, double* __restrict__ norms, int carry,
                                                         double* __restrict__ wnext, int carry_first,
                                                         int carry_count) {
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
  const int ti = item_tensor[item];
  const obst_opt_tensor T = table[ti];
  const int64_t cols = T.dims[T.ndim - 1];
  const int64_t e0 = (int64_t)item_row0[item] * cols;
  const int64_t n = (int64_t)item_rows[item] * cols;
  const float* __restrict__ g = T.g + e0;
  const float* __restrict__ p = T.p + e0;
  float sg = 0.f, sw = 0.f;
  const bool vec_ok = ((e0 & 3) == 0);
  const int64_t nv = vec_ok ? n / 4 : 0;
  for (int64_t i = threadIdx.x; i < nv; i += blockDim.x) {
    const float4 gv = reinterpret_cast<const float4*>(g)[i];
    const float4 pv = carry ? make_float4(0.f, 0.f, 0.f, 0.f) : reinterpret_cast<const float4*>(p)[i];
    float ga[4] = {gv.x, gv.y, gv.z, gv.w}, pa[4] = {pv.x, pv.y, pv.z, pv.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float gg = grad_bf16 ? round_to<bf16>(ga[j]) : ga[j];
      const float ww = act_bf16 ? round_to<bf16>(pa[j]) : pa[j];
      sg += gg * gg;
      sw += ww * ww;
    }
  }
  for (int64_t i = nv * 4 + threadIdx.x; i < n; i += blockDim.x) {
    const float gg = grad_bf16 ? round_to<bf16>(g[i]) : g[i];
    const float pw = carry ? 0.f : p[i];
    const float ww = act_bf16 ? round_to<bf16>(pw) : pw;
    sg += gg * gg;
    sw += ww * ww;
  }
  double dg = warp_sum_d((double)sg), dw = warp_sum_d((double)sw);
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
    atomicAdd(&norms[2 * ti], a);
    if (!carry) atomicAdd(&norms[2 * ti + 1], b);
  }
}

// ------------------------------------------------------------------------------------------------
// pass 2: the update
// ------------------------------------------------------------------------------------------------
static constexpr int OPT_MAX_ITEM_ROWS = 2048;  // rows one work item may cover (checked at launch)

struct OptHyper {
  float clip, beta, lr, weight_decay;
  int nesterov, act_bf16, grad_bf16;
  int sm3_only;  // generic optimizer chains (obst_opt_stage): write g * opt_rsqrt(nu) back to g, touch nothing else
};

// NI > 0: lanes keep their column maxima in registers (cols <= 32*VEC*NI).  NI == 0: any cols, slower.
template <typename TS, int VEC, int NI>
__global__ void __launch_bounds__(256) opt_update_kernel(const obst_opt_tensor* __restrict__ table,
                                                          const int32_t* __restrict__ item_tensor,
                                                          const int32_t* __restrict__ item_row0,
                                                          const int32_t* __restrict__ item_rows,
                                                          const double* __restrict__ norms, const OptHyper hp,
                                                          double* __restrict__ wnext) {
  extern __shared__ float s_col[];  // [cols] block-level column maxima
  __shared__ obst_opt_tensor sT;    // the table entry (dynamically indexed: keep it out of local memory)
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
  // Column chunking: blockIdx.y selects a chunk of 32 * VEC * NI columns, so a lane never owns more than NI vectors
  // of a row whatever the last dim is (a [.., 1024] weight used to need the NI = 8 instantiation: 248 registers,
  // one block per SM, 2.5 TB/s against 5.5 TB/s for the 512-wide tensors).
  const int cbase = NI > 0 ? (int)blockIdx.y * (32 * VEC * NI) : 0;
  if (cbase >= cols) return;  // whole block: this tensor is narrower than the widest one of the launch
  for (int i = threadIdx.x; i < cols; i += blockDim.x) s_col[i] = 0.f;
  // min over the non-last dims' accumulators, one ROW PER THREAD: the index arithmetic (a divide and a modulo per
  // dim) used to be done by every warp for its rows, i.e. 32 lanes computing the same thing.  The slot holds the
  // row's minimum until the row is processed and its maximum afterwards.
  const int row_begin = item_row0[item];
  for (int rl = threadIdx.x; rl < item_rows[item]; rl += blockDim.x) {
    int rem = row_begin + rl;
    float mn = INFINITY;
#pragma unroll
    for (int i = 2; i >= 0; --i) {
      if (i <= nd - 2) {
        const int d = T.dims[i];
        const int q = rem / d;
        mn = fminf(mn, ld_state<TS>(T.acc[i], rem - q * d));
        rem = q;
      }
    }
    s_rowmax[rl] = mn;
  }
  __syncthreads();
  // the table entry's hot fields in registers (sT is shared memory: every use would be a reload)
  float* const Tp = T.p;
  const float* const Tg = T.g;
  void* const Tm = T.m;
  void* const Tshadow = T.shadow;
  const int tril_s = T.tril_s;
  const float wd = (T.decay && hp.weight_decay > 0.f) ? hp.lr * hp.weight_decay : 0.f;
  const float nes_u = hp.nesterov ? 1.f : 0.f, nes_m = hp.nesterov ? hp.beta : 1.f;

  float scale = 1.f;
  if (hp.clip > 0.f) {
    const float sg = (float)norms[2 * ti], sw = (float)norms[2 * ti + 1];
    const float gn = fminf(rsqrtf(sg), 1e6f);  // rsqrt(0) = inf -> 1e6
    const float wn = fmaxf(sqrtf(sw), 1e-3f);
    scale = fminf(wn * gn * hp.clip, 1.f);
  }
  const void* acc_last = T.acc[nd - 1];
  constexpr int NIR = NI > 0 ? NI : 1;
  float al[NIR * VEC], cm[NIR * VEC];
  if (NI > 0) {
#pragma unroll
    for (int i = 0; i < NIR; ++i) {
      const int c = cbase + (lane + 32 * i) * VEC;
#pragma unroll
      for (int j = 0; j < VEC; ++j) {
        cm[i * VEC + j] = 0.f;
        al[i * VEC + j] = (c + j < cols) ? ld_state<TS>(acc_last, c + j) : 0.f;
      }
    }
  }
  const int row_end = row_begin + item_rows[item];
  const int iters = NI > 0 ? NI : (cols + 32 * VEC - 1) / (32 * VEC);
  const bool tril_fast = tril_s > 0 && tril_s == cols;  // [.., S, S] map: col = c + j, row = r % S
  float w2 = 0.f;  // sum cast_AT(p_new)^2 of this thread: next step's weight norm, for free
  // RG rows per warp per iteration: all their loads are issued before the first dependent instruction, which is
  // what gives the memory system enough to do (one row at a time left this kernel latency-bound at < 3 TB/s)
  constexpr int RG = 4;
  for (int r0 = row_begin + warp * RG; r0 < row_end; r0 += 8 * RG) {
    float rowmin[RG], rowmax[RG];
#pragma unroll
    for (int q = 0; q < RG; ++q) {
      rowmin[q] = r0 + q < row_end ? s_rowmax[r0 + q - row_begin] : INFINITY;
      rowmax[q] = 0.f;
    }
#pragma unroll(NI > 0 ? NI : 1)
    for (int i = 0; i < iters; ++i) {  // NI > 0: fully unrolled (iters == NI); NI == 0: a plain loop
      const int c = cbase + (lane + 32 * i) * VEC;
      if (c >= cols) continue;
      float pv[RG][VEC], gv[RG][VEC], mv[RG][VEC];
#pragma unroll
      for (int q = 0; q < RG; ++q) {
        if (r0 + q >= row_end) continue;
        const int64_t rbase = (int64_t)(r0 + q) * cols;
        if (VEC == 4) {
          const float4 p4 = *reinterpret_cast<const float4*>(Tp + rbase + c);
          const float4 g4 = *reinterpret_cast<const float4*>(Tg + rbase + c);
          pv[q][0] = p4.x; pv[q][1] = p4.y; pv[q][2] = p4.z; pv[q][3] = p4.w;
          gv[q][0] = g4.x; gv[q][1] = g4.y; gv[q][2] = g4.z; gv[q][3] = g4.w;
          if (sizeof(TS) == 2) {
            const uint2 m2 = *reinterpret_cast<const uint2*>(reinterpret_cast<const bf16*>(Tm) + rbase + c);
            const float2 a = unpack_bf16x2(m2.x), bb = unpack_bf16x2(m2.y);
            mv[q][0] = a.x; mv[q][1] = a.y; mv[q][2] = bb.x; mv[q][3] = bb.y;
          } else {
            const float4 m4 = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(Tm) + rbase + c);
            mv[q][0] = m4.x; mv[q][1] = m4.y; mv[q][2] = m4.z; mv[q][3] = m4.w;
          }
        } else {
#pragma unroll
          for (int j = 0; j < VEC; ++j) {
            pv[q][j] = Tp[rbase + c + j];
            gv[q][j] = Tg[rbase + c + j];
            mv[q][j] = ld_state<TS>(Tm, rbase + c + j);
          }
        }
      }
#pragma unroll
      for (int q = 0; q < RG; ++q) {
        if (r0 + q >= row_end) continue;
        const int r = r0 + q;
        const int64_t rbase = (int64_t)r * cols;
        float pn[VEC], mn[VEC], sh[VEC];
        const int row_in = tril_fast ? (r % cols) : 0;
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
          float g = hp.grad_bf16 ? round_to<bf16>(gv[q][j]) : gv[q][j];
          g = g * scale;
          const float a_last = NI > 0 ? al[i * VEC + j] : ld_state<TS>(acc_last, c + j);
          const float nu = fminf(rowmin[q], a_last) + g * g;
          if (NI > 0) cm[i * VEC + j] = fmaxf(cm[i * VEC + j], nu);
          else atomic_max_nonneg(&s_col[c + j], nu);
          rowmax[q] = fmaxf(rowmax[q], nu);
          // g / max(sqrt(nu), 1e-5)  (opt_rsqrt, src/optimizer/optimizers.py:13-14) without the IEEE sqrt + divide
          float u = g * (nu >= 1e-10f ? rsqrtf(nu) : 1e5f);
          if (hp.sm3_only) {
            pn[j] = u;
            continue;
          }
          const float m_new = fmaf(hp.beta, mv[q][j], u);
          u = fmaf(nes_m, m_new, nes_u * u) * hp.lr;  // nesterov: u + beta m_new; plain: m_new
          if (wd != 0.f) {
            const float w = hp.act_bf16 ? round_to<bf16>(pv[q][j]) : pv[q][j];
            u = fmaf(w, wd, u);
          }
          pn[j] = pv[q][j] - u;
          mn[j] = m_new;
          sh[j] = pn[j];
          if (tril_fast) {
            if (c + j > row_in) sh[j] = 0.f;
          } else if (tril_s > 0) {
            const int64_t e = rbase + c + j;
            const int col = (int)(e % tril_s);
            const int row = (int)((e / tril_s) % tril_s);
            if (col > row) sh[j] = 0.f;
          }
          const float wv = hp.act_bf16 ? round_to<bf16>(pn[j]) : pn[j];
          w2 += wv * wv;
        }
        if (hp.sm3_only) {
          float* gw = const_cast<float*>(Tg) + rbase + c;
          if (VEC == 4) {
            *reinterpret_cast<float4*>(gw) = make_float4(pn[0], pn[1], pn[2], pn[3]);
          } else {
#pragma unroll
            for (int j = 0; j < VEC; ++j) gw[j] = pn[j];
          }
          continue;
        }
        if (VEC == 4) {
          *reinterpret_cast<float4*>(Tp + rbase + c) = make_float4(pn[0], pn[1], pn[2], pn[3]);
          if (sizeof(TS) == 2) {
            uint2 m2;
            m2.x = pack_bf16x2(mn[0], mn[1]);
            m2.y = pack_bf16x2(mn[2], mn[3]);
            *reinterpret_cast<uint2*>(reinterpret_cast<bf16*>(Tm) + rbase + c) = m2;
          } else {
            *reinterpret_cast<float4*>(reinterpret_cast<float*>(Tm) + rbase + c) =
                make_float4(mn[0], mn[1], mn[2], mn[3]);
          }
          if (Tshadow) {
            if (hp.act_bf16) {
              uint2 s2;
              s2.x = pack_bf16x2(sh[0], sh[1]);
              s2.y = pack_bf16x2(sh[2], sh[3]);
              *reinterpret_cast<uint2*>(reinterpret_cast<bf16*>(Tshadow) + rbase + c) = s2;
            } else {
              *reinterpret_cast<float4*>(reinterpret_cast<float*>(Tshadow) + rbase + c) =
                  make_float4(sh[0], sh[1], sh[2], sh[3]);
            }
          }
        } else {
#pragma unroll
          for (int j = 0; j < VEC; ++j) {
            Tp[rbase + c + j] = pn[j];
            reinterpret_cast<TS*>(Tm)[rbase + c + j] = from_f32<TS>(mn[j]);
            if (Tshadow) {
              if (hp.act_bf16) reinterpret_cast<bf16*>(Tshadow)[rbase + c + j] = __float2bfloat16_rn(sh[j]);
              else reinterpret_cast<float*>(Tshadow)[rbase + c + j] = sh[j];
            }
          }
        }
      }
    }
#pragma unroll
    for (int q = 0; q < RG; ++q) {
      const float rm = warp_max(rowmax[q]);
      if (lane == 0 && r0 + q < row_end) s_rowmax[r0 + q - row_begin] = rm;
    }
  }
  __syncthreads();
  // Row maxima -> accumulators of the non-last dims.  Rows are consecutive, so for each dim the rows sharing an
  // index form runs; the first row of a run folds the run and issues ONE atomic (a [8, 512, 512] weight would
  // otherwise send 4096 same-address atomics to each of its 8 dim-0 entries, which serialise in L2).
  {
    const int nrows = item_rows[item];
    for (int rl = threadIdx.x; rl < nrows; rl += blockDim.x) {
      const int r = item_row0[item] + rl;
      int stride = 1;
      for (int i = nd - 2; i >= 0; --i) {
        const int v = (r / stride) % T.dims[i];
        const bool first = rl == 0 || ((r - 1) / stride) % T.dims[i] != v;
        if (first) {
          float mx = s_rowmax[rl];
          for (int k = rl + 1; k < nrows && ((item_row0[item] + k) / stride) % T.dims[i] == v; ++k)
            mx = fmaxf(mx, s_rowmax[k]);
          atomic_max_nonneg(T.acc_new[i] + v, mx);
        }
        stride *= T.dims[i];
      }
    }
  }
  if (NI > 0) {
#pragma unroll
    for (int i = 0; i < NIR; ++i) {
      const int c = cbase + (lane + 32 * i) * VEC;
#pragma unroll
      for (int j = 0; j < VEC; ++j)
        if (c + j < cols && cm[i * VEC + j] > 0.f) atomic_max_nonneg(&s_col[c + j], cm[i * VEC + j]);
    }
  }
  if (wnext) {
    const double dw = warp_sum_d((double)w2);
    if (lane == 0) s_w2[warp] = dw;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < cols; i += blockDim.x) {
    const float v = s_col[i];
    if (v > 0.f) atomic_max_nonneg(T.acc_new[nd - 1] + i, v);
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

}  // namespace obst

using namespace obst;

extern "C" {

int obst_opt_norms(const obst_opt_tensor* table, int32_t ntensors, const int32_t* item_tensor,
                   const int32_t* item_row0, const int32_t* item_rows, int32_t nitems, int32_t dtype_act,
                   int32_t grad_bf16, double* norms, int32_t flags, double* wnext, int32_t first_tensor,
                   int32_t tensor_count, void* stream) {
  OBST_REQUIRE(table && norms && ntensors > 0, "obst_opt_norms: bad arguments");
  const int carry = (flags & 1) != 0;
  OBST_REQUIRE(!carry || wnext, "obst_opt_norms: carry mode needs wnext");
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
  opt_norms_kernel<<<(unsigned)nitems, 256, 0, st>>>(table, item_tensor, item_row0, item_rows,
                                                     dtype_act == OBST_BF16, grad_bf16, norms, carry, wnext,
                                                     first_tensor, tensor_count);
  count_launch();
  return check_launch("obst_opt_norms");
}

// All tensors in one launch share the (VEC, NI) instantiation: max_cols = largest last dim in the
// table; vec4 != 0 promises every last dim is a multiple of 4 (so every row start is 16-byte aligned).
int obst_opt_update(const obst_opt_tensor* table, int32_t ntensors, const int32_t* item_tensor,
                    const int32_t* item_row0, const int32_t* item_rows, int32_t nitems, int32_t max_cols,
                    int32_t vec4, int32_t dtype_act, int32_t dtype_state, int32_t grad_bf16, const double* norms,
                    double* wnext, float clip, float beta, int32_t nesterov, float lr, float weight_decay,
                    void* stream) {
  OBST_REQUIRE(table && ntensors > 0 && max_cols > 0, "obst_opt_update: bad arguments");
  OBST_REQUIRE(clip <= 0.f || norms != nullptr, "obst_opt_update: clip > 0 needs norms");
  OBST_REQUIRE(max_cols <= 12288, "obst_opt_update: last dim %d too large (max 12288)", max_cols);
  if (nitems <= 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  OptHyper hp;
  hp.clip = clip; hp.beta = beta; hp.lr = lr; hp.weight_decay = weight_decay;
  hp.nesterov = nesterov; hp.act_bf16 = dtype_act == OBST_BF16; hp.grad_bf16 = grad_bf16;
  hp.sm3_only = 0;
  const size_t smem = (size_t)max_cols * sizeof(float);
  const int vec = vec4 ? 4 : 1;
  const dim3 grid((unsigned)nitems, (unsigned)((max_cols + 32 * vec * 4 - 1) / (32 * vec * 4)));  // y: column chunks
#define OBST_OPT_NI(TS, V) \
  opt_update_kernel<TS, V, 4><<<grid, 256, smem, st>>>(table, item_tensor, item_row0, item_rows, norms, hp, wnext)
  if (dtype_state == OBST_BF16) {
    if (vec4) OBST_OPT_NI(bf16, 4);
    else OBST_OPT_NI(bf16, 1);
  } else if (dtype_state == OBST_F32) {
    if (vec4) OBST_OPT_NI(float, 4);
    else OBST_OPT_NI(float, 1);
  } else {
    OBST_REQUIRE(false, "obst_opt_update: bad state dtype %d", dtype_state);
  }
#undef OBST_OPT_NI
  count_launch();
  return check_launch("obst_opt_update");
}

/* SM3 as a stand-alone stage of a generic chain: g <- g * opt_rsqrt(nu), accumulators max-reduced into acc_new
 * (obst_opt_finalize commits them).  Same kernel as obst_opt_update with everything after SM3 switched off. */
int obst_opt_sm3_stage(const obst_opt_tensor* table, int32_t ntensors, const int32_t* item_tensor,
                       const int32_t* item_row0, const int32_t* item_rows, int32_t nitems, int32_t max_cols,
                       int32_t vec4, int32_t dtype_state, void* stream) {
  OBST_REQUIRE(table && item_tensor && item_row0 && item_rows, "obst_opt_sm3_stage: null argument");
  OBST_REQUIRE(max_cols > 0 && max_cols <= 12288, "obst_opt_sm3_stage: last dim %d unsupported", max_cols);
  if (nitems <= 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  OptHyper hp;
  hp.clip = 0.f; hp.beta = 0.f; hp.lr = 1.f; hp.weight_decay = 0.f;
  hp.nesterov = 0; hp.act_bf16 = 0; hp.grad_bf16 = 0; hp.sm3_only = 1;
  const size_t smem = (size_t)max_cols * sizeof(float);
  const int vec = vec4 ? 4 : 1;
  const dim3 grid((unsigned)nitems, (unsigned)((max_cols + 32 * vec * 4 - 1) / (32 * vec * 4)));
#define OBST_SM3(TS, V) \
  opt_update_kernel<TS, V, 4><<<grid, 256, smem, st>>>(table, item_tensor, item_row0, item_rows, nullptr, hp, nullptr)
  if (dtype_state == OBST_BF16) {
    if (vec4) OBST_SM3(bf16, 4);
    else OBST_SM3(bf16, 1);
  } else if (dtype_state == OBST_F32) {
    if (vec4) OBST_SM3(float, 4);
    else OBST_SM3(float, 1);
  } else {
    OBST_REQUIRE(false, "obst_opt_sm3_stage: bad state dtype %d", dtype_state);
  }
#undef OBST_SM3
  count_launch();
  return check_launch("obst_opt_sm3_stage");
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

extern "C" int obst_opt_step_noalloc(float* p, void* shadow, float* g, void* m, void* const* acc,
                                     const int32_t* dims, int32_t ndim, int32_t decay, int32_t tril_s,
                                     int32_t dtype_act, int32_t dtype_state, int32_t grad_bf16, float clip, float beta,
                                     int32_t nesterov, float lr, float weight_decay, void* stream) {
  OBST_REQUIRE(p && g && m && acc && dims, "obst_opt_step_noalloc: null argument");
  OBST_REQUIRE(ndim >= 1 && ndim <= 4, "obst_opt_step_noalloc: ndim %d outside 1..4", ndim);
  OBST_REQUIRE(dtype_state == OBST_BF16 || dtype_state == OBST_F32, "obst_opt_step_noalloc: bad state dtype %d",
               dtype_state);
  NaFill f;
  memset(&f, 0, sizeof(f));
  int64_t numel = 1;
  int dim_sum = 0;
  for (int i = 0; i < ndim; ++i) {
    OBST_REQUIRE(dims[i] > 0 && acc[i], "obst_opt_step_noalloc: dim %d is empty or has no accumulator", i);
    numel *= dims[i];
    dim_sum += dims[i] + ((64 - dims[i] % 64) % 64);
    f.t.dims[i] = dims[i];
    f.t.acc[i] = acc[i];
  }
  const int cols = dims[ndim - 1];
  OBST_REQUIRE(dim_sum <= NA_MAX_DIM_SUM + 4 * 64,
               "obst_opt_step_noalloc: the dims (each rounded up to 64) add up to %d, the built-in staging holds %d",
               dim_sum, NA_MAX_DIM_SUM + 4 * 64);
  OBST_REQUIRE(cols <= 12288, "obst_opt_step_noalloc: last dim %d too large (max 12288)", cols);
  const int64_t rows64 = numel / cols;
  const int rpi = 32768 / cols > 0 ? (32768 / cols < OPT_MAX_ITEM_ROWS ? 32768 / cols : OPT_MAX_ITEM_ROWS) : 1;
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
  f.dim_sum_padded = dim_sum;

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
  if (clip > 0.f)
    opt_norms_kernel<<<(unsigned)f.nitems, 256, 0, st>>>(table, item_tensor, item_row0, item_rows,
                                                         dtype_act == OBST_BF16, grad_bf16, norms, 0, nullptr, 0, 0);
  OptHyper hp;
  hp.clip = clip; hp.beta = beta; hp.lr = lr; hp.weight_decay = weight_decay;
  hp.nesterov = nesterov; hp.act_bf16 = dtype_act == OBST_BF16; hp.grad_bf16 = grad_bf16;
  hp.sm3_only = 0;
  const size_t smem = (size_t)cols * sizeof(float);
  const int vec = (cols % 4 == 0) ? 4 : 1;
  const dim3 grid((unsigned)f.nitems, (unsigned)((cols + 32 * vec * 4 - 1) / (32 * vec * 4)));
#define OBST_NA_NI(TS, V) \
  opt_update_kernel<TS, V, 4><<<grid, 256, smem, st>>>(table, item_tensor, item_row0, item_rows, norms, hp, nullptr)
  if (dtype_state == OBST_BF16) {
    if (vec == 4) OBST_NA_NI(bf16, 4);
    else OBST_NA_NI(bf16, 1);
  } else {
    if (vec == 4) OBST_NA_NI(float, 4);
    else OBST_NA_NI(float, 1);
  }
#undef OBST_NA_NI
  if (dtype_state == OBST_BF16) opt_finalize_kernel<bf16><<<1, 256, 0, st>>>(table);
  else opt_finalize_kernel<float><<<1, 256, 0, st>>>(table);
  count_launch(clip > 0.f ? 4 : 3);
  return check_launch("obst_opt_step_noalloc");
}
