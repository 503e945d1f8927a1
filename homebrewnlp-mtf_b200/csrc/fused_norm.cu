// Helpers of the "group norm folded into the neighbouring GEMMs" path of bottleneck_group_linear
// (src/model/basic.py:47-57,122-126): mid = norm(relu(h1 @ W_mid)), out = mid @ W_out per head.
//
// With n = (h - mu) * r * gamma + beta (src/model/normalization.py:22-34; mu, r per token and head):
//   n @ W            = r * (h @ (gamma . W)) - r * mu * (gamma^T W) + beta^T W
//   d/dW sum(dy * (n @ W))[j, k]
//                    = gamma_j * ( sum_t h[t,j] * (r_t dy[t,k]) - sum_t r_t mu_t dy[t,k] ) + beta_j * sum_t dy[t,k]
// so neither the forward nor the weight gradient needs the normalised tensor n in memory: the GEMMs read the ReLU
// output h directly (obst_gemm epilogues 5 / 6 / 8), the statistics come out of the producing GEMM's epilogue
// (epilogue 7), and the kernels below supply the small pieces: the statistics' final form, gamma . W and the two
// vectors, r . dy with its two column sums, and the rank-one terms of the weight gradient.
#include "obst_common.cuh"

namespace obst {

static constexpr float FN_EPS = 1e-5f;  // src/model/normalization.py:28

// (sum, sum of squares) partials of obst_gemm epilogue 7 -> (mean, rstd) per (row, head), the format obst_gn_bwd reads
__global__ void __launch_bounds__(256) norm_stats_kernel(const float2* __restrict__ part, float2* __restrict__ stats,
                                                          int64_t rows, int heads, int per_head, float inv_w) {
  const int64_t total = rows * heads;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const float2* p = part + i * per_head;
    float s1 = 0.f, s2 = 0.f;
    for (int q = 0; q < per_head; ++q) {
      const float2 v = p[q];
      s1 += v.x;
      s2 += v.y;
    }
    const float mean = s1 * inv_w;
    const float var = fmaxf(s2 * inv_w - mean * mean, 0.f);
    stats[i] = make_float2(mean, rsqrtf(var + FN_EPS));
  }
}

// gamma . W (transposed for the forward GEMM's K-major B operand, and untransposed for the backward's), c = gamma^T W,
// d = beta^T W, and gamma as fp32.
// One block per (32 output features, head): it walks the whole contraction axis, so the two vectors need no atomics.
__global__ void __launch_bounds__(256) fold_norm_weight_kernel(const obst_fold_item* __restrict__ items, int heads,
                                                                int kin, int kout) {
  const obst_fold_item it = items[blockIdx.z];  // blockIdx.z: which weight of a batched refresh
  const bf16* __restrict__ w = reinterpret_cast<const bf16*>(it.w);
  const bf16* __restrict__ gamma = reinterpret_cast<const bf16*>(it.gamma);
  const bf16* __restrict__ beta = reinterpret_cast<const bf16*>(it.beta);
  bf16* __restrict__ wg_t = reinterpret_cast<bf16*>(it.wg_t);
  bf16* __restrict__ wg_n = reinterpret_cast<bf16*>(it.wg);  // gamma . W in W's own layout (may be NULL)
  float* __restrict__ cvec = it.cvec;
  float* __restrict__ dvec = it.dvec;
  float* __restrict__ gamma_f = it.gamma_f;
  __shared__ bf16 tile[32][33];
  __shared__ float s_c[8][32], s_d[8][32];
  const int h = blockIdx.y;
  const int k0 = blockIdx.x * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  float c_acc = 0.f, d_acc = 0.f;
  for (int j0 = 0; j0 < kin; j0 += 32) {
#pragma unroll
    for (int i = 0; i < 32; i += 8) {
      const int j = j0 + ty + i, k = k0 + tx;
      float wg = 0.f;
      if (j < kin && k < kout) {
        const float wv = to_f32(w[((int64_t)h * kin + j) * kout + k]);
        const float g = gamma ? to_f32(gamma[(int64_t)h * kin + j]) : 1.f;
        const float b = beta ? to_f32(beta[(int64_t)h * kin + j]) : 0.f;
        const bf16 wgb = __float2bfloat16_rn(g * wv);
        wg = to_f32(wgb);                 // c is the column sum of what the GEMM really multiplies with
        d_acc = fmaf(b, wv, d_acc);
        tile[ty + i][tx] = wgb;
        if (wg_n) wg_n[((int64_t)h * kin + j) * kout + k] = wgb;
      }
      c_acc += wg;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 32; i += 8) {
      const int k = k0 + ty + i, j = j0 + tx;
      if (j < kin && k < kout) wg_t[((int64_t)h * kout + k) * kin + j] = tile[tx][ty + i];
    }
    __syncthreads();
  }
  s_c[ty][tx] = c_acc;
  s_d[ty][tx] = d_acc;
  __syncthreads();
  if (ty == 0 && k0 + tx < kout) {
    float c = 0.f, d = 0.f;
    for (int q = 0; q < 8; ++q) {
      c += s_c[q][tx];
      d += s_d[q][tx];
    }
    cvec[(int64_t)h * kout + k0 + tx] = c;
    dvec[(int64_t)h * kout + k0 + tx] = d;
  }
  if (blockIdx.x == 0)
    for (int j = threadIdx.x; j < kin; j += blockDim.x)
      gamma_f[(int64_t)h * kin + j] = gamma ? to_f32(gamma[(int64_t)h * kin + j]) : 1.f;
}

// dyr[t,h,:] = rstd[t,h] * dy[t,h,:];  v1[h,:] += sum_t rstd*mean*dy[t,h,:];  v2[h,:] += sum_t dy[t,h,:]
// grid (token chunks, heads); a warp walks tokens of one head, lanes own fixed columns (16 bytes each per 256 columns)
template <int MAXC>
__global__ void __launch_bounds__(256) rowscale_colsum_kernel(const bf16* __restrict__ dy, const float2* __restrict__ stats,
                                                               bf16* __restrict__ dyr, float* __restrict__ v1,
                                                               float* __restrict__ v2, int64_t tokens, int heads,
                                                               int tokens_per_block) {
  constexpr int WIDTH = 256 * MAXC;
  __shared__ float s_red[2][8][WIDTH];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int h = blockIdx.y;
  const int64_t t0 = (int64_t)blockIdx.x * tokens_per_block;
  const int64_t t1 = t0 + tokens_per_block < tokens ? t0 + tokens_per_block : tokens;
  float a1[MAXC][8], a2[MAXC][8];
#pragma unroll
  for (int i = 0; i < MAXC; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) a1[i][j] = a2[i][j] = 0.f;
  for (int64_t t = t0 + warp; t < t1; t += 8) {
    const int64_t row = t * heads + h;
    const float2 st = stats[row];
    const float r = st.y, rm = st.y * st.x;
#pragma unroll
    for (int i = 0; i < MAXC; ++i) {
      const int c = (lane + 32 * i) * 8;
      Raw16<bf16> in, out;
      in.load(dy + row * WIDTH + c);
      float v[8], o[8];
      in.unpack(v);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        o[j] = v[j] * r;
        a1[i][j] = fmaf(v[j], rm, a1[i][j]);
        a2[i][j] += v[j];
      }
      out.pack(o);
      out.store(dyr + row * WIDTH + c);
    }
  }
#pragma unroll
  for (int i = 0; i < MAXC; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      s_red[0][warp][(lane + 32 * i) * 8 + j] = a1[i][j];
      s_red[1][warp][(lane + 32 * i) * 8 + j] = a2[i][j];
    }
  __syncthreads();
  for (int c = threadIdx.x; c < WIDTH; c += blockDim.x) {
    float x1 = 0.f, x2 = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
      x1 += s_red[0][w][c];
      x2 += s_red[1][w][c];
    }
    atomicAdd(&v1[(int64_t)h * WIDTH + c], x1);
    atomicAdd(&v2[(int64_t)h * WIDTH + c], x2);
  }
}

// dW[h,j,k] += -gamma[h,j] * v1[h,k] + beta[h,j] * v2[h,k]
__global__ void __launch_bounds__(256) norm_wgrad_rank1_kernel(float* __restrict__ dw, const bf16* __restrict__ gamma,
                                                                const bf16* __restrict__ beta, const float* __restrict__ v1,
                                                                const float* __restrict__ v2, int heads, int kin,
                                                                int kout) {
  const int64_t total = (int64_t)heads * kin * kout;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int k = (int)(i % kout);
    const int64_t hj = i / kout;
    const int h = (int)(hj / kin);
    const float g = gamma ? to_f32(gamma[hj]) : 1.f;
    const float b = beta ? to_f32(beta[hj]) : 0.f;
    dw[i] += fmaf(-g, v1[(int64_t)h * kout + k], b * v2[(int64_t)h * kout + k]);
  }
}

// The folded norm's backward, per (row, head): the two reductions over the head's features arrive as partial
// dot products from W_out's recompute epilogue (obst_gemm epilogues 9 / 10):
//   s1 = sum_j gamma_j dn_j        = sum_k dy_k (gamma^T W)_k
//   s2 = sum_j gamma_j h_j dn_j    = sum_k dy_k (h @ (gamma.W))_k
// and dh_j = r (gamma_j dn_j - s1/K - xhat_j * r (s2 - mu s1)/K), xhat_j = (h_j - mu) r, which the dgrad epilogue (11)
// evaluates as a * gamma_j * dn_j + b * h_j + c with  a = r,  b = -r^2 m2,  c = r^2 mu m2 - r m1,
// m1 = s1 / K,  m2 = r (s2 - mu s1) / K.
__global__ void __launch_bounds__(256) norm_bwd_coeffs_kernel(const float2* __restrict__ part,
                                                              const float2* __restrict__ stats,
                                                              float4* __restrict__ coeffs, int64_t total, int nparts,
                                                              float inv_width) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    float s1 = 0.f, s2 = 0.f;
    for (int k = 0; k < nparts; ++k) {
      const float2 v = part[i * nparts + k];
      s1 += v.x;
      s2 += v.y;
    }
    const float2 st = stats[i];
    const float mu = st.x, r = st.y;
    const float m1 = s1 * inv_width;
    const float m2 = r * (s2 - mu * s1) * inv_width;
    coeffs[i] = make_float4(r, -r * r * m2, r * r * mu * m2 - r * m1, 0.f);
  }
}

// dgamma[h,j] -= sum_k W[h,j,k] v1[h,k];  dbeta[h,j] += sum_k W[h,j,k] v2[h,k]   (one warp per (h, j))
// -- the parts of the norm's scale / shift gradients that do not need the weight gradient's accumulator:
//   dbeta_j  = sum_t dn_tj           = sum_k W_jk sum_t dy_tk
//   dgamma_j = sum_t dn_tj xhat_tj   = sum_k W_jk (G_jk - sum_t r_t mu_t dy_tk),  G = h^T (r . dy): epilogue 8's part
__global__ void __launch_bounds__(256) norm_affine_grad_kernel(const bf16* __restrict__ w, const float* __restrict__ v1,
                                                               const float* __restrict__ v2,
                                                               float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                               int heads, int kin, int kout) {
  const int lane = threadIdx.x & 31;
  const int64_t rows = (int64_t)heads * kin;
  for (int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < rows;
       row += (int64_t)gridDim.x * (blockDim.x >> 5)) {
    const int h = (int)(row / kin);
    const bf16* wr = w + row * kout;
    float a1 = 0.f, a2 = 0.f;
    for (int k = lane; k < kout; k += 32) {
      const float wv = to_f32(wr[k]);
      a1 = fmaf(wv, v1[(int64_t)h * kout + k], a1);
      a2 = fmaf(wv, v2[(int64_t)h * kout + k], a2);
    }
    a1 = warp_sum(a1);
    a2 = warp_sum(a2);
    if (lane == 0) {
      if (dgamma) dgamma[row] -= a1;
      if (dbeta) dbeta[row] += a2;
    }
  }
}

}  // namespace obst

using namespace obst;

extern "C" {

int obst_norm_stats_finalize(const float* partials, float* stats, int64_t rows, int32_t heads, int32_t per_head,
                             int32_t width, void* stream) {
  OBST_REQUIRE(partials && stats && heads > 0 && per_head > 0 && width > 0, "obst_norm_stats_finalize: bad arguments");
  if (rows <= 0) return OBST_OK;
  const int64_t total = rows * heads;
  int64_t blocks = (total + 255) / 256;
  if (blocks > (int64_t)num_sms() * 16) blocks = (int64_t)num_sms() * 16;
  norm_stats_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const float2*>(partials), reinterpret_cast<float2*>(stats), rows, heads, per_head,
      1.0f / (float)width);
  count_launch();
  return check_launch("obst_norm_stats_finalize");
}

int obst_fold_norm_weight(const obst_fold_item* items, int32_t nitems, int32_t heads, int32_t kin, int32_t kout,
                          void* stream) {
  OBST_REQUIRE(nitems >= 0 && nitems <= 65535 && heads > 0 && kin > 0 && kout > 0, "obst_fold_norm_weight: bad arguments");
  if (nitems == 0) return OBST_OK;
  OBST_REQUIRE(items != nullptr, "obst_fold_norm_weight: null item table");
  const dim3 grid((unsigned)((kout + 31) / 32), (unsigned)heads, (unsigned)nitems);
  fold_norm_weight_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(items, heads, kin, kout);
  count_launch();
  return check_launch("obst_fold_norm_weight");
}

int obst_rowscale_colsum(const void* dy, const float* stats, void* dyr, float* v1, float* v2, int64_t tokens,
                         int32_t heads, int32_t width, void* stream) {
  OBST_REQUIRE(dy && stats && dyr && v1 && v2 && heads > 0, "obst_rowscale_colsum: bad arguments");
  OBST_REQUIRE(width == 256 || width == 512, "obst_rowscale_colsum: width %d unsupported (256, 512)", width);
  if (tokens <= 0) return OBST_OK;
  int tpb = 256;
  while (tpb > 8 && ((tokens + tpb - 1) / tpb) * heads < (int64_t)num_sms() * 4) tpb >>= 1;
  const dim3 grid((unsigned)((tokens + tpb - 1) / tpb), (unsigned)heads);
  cudaStream_t st = (cudaStream_t)stream;
  if (width == 256)
    rowscale_colsum_kernel<1><<<grid, 256, 0, st>>>((const bf16*)dy, reinterpret_cast<const float2*>(stats), (bf16*)dyr,
                                                    v1, v2, tokens, heads, tpb);
  else
    rowscale_colsum_kernel<2><<<grid, 256, 0, st>>>((const bf16*)dy, reinterpret_cast<const float2*>(stats), (bf16*)dyr,
                                                    v1, v2, tokens, heads, tpb);
  count_launch();
  return check_launch("obst_rowscale_colsum");
}

int obst_norm_wgrad_rank1(float* dw, const void* gamma, const void* beta, const float* v1, const float* v2,
                          int32_t heads, int32_t kin, int32_t kout, void* stream) {
  OBST_REQUIRE(dw && v1 && v2 && heads > 0 && kin > 0 && kout > 0, "obst_norm_wgrad_rank1: bad arguments");
  const int64_t total = (int64_t)heads * kin * kout;
  int64_t blocks = (total + 255) / 256;
  if (blocks > (int64_t)num_sms() * 16) blocks = (int64_t)num_sms() * 16;
  norm_wgrad_rank1_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(dw, (const bf16*)gamma, (const bf16*)beta,
                                                                               v1, v2, heads, kin, kout);
  count_launch();
  return check_launch("obst_norm_wgrad_rank1");
}

int obst_norm_bwd_coeffs(const float* partials, const float* stats, float* coeffs, int64_t rows, int32_t heads,
                         int32_t nparts, int32_t width, void* stream) {
  OBST_REQUIRE(partials && stats && coeffs && heads > 0 && nparts > 0 && width > 0, "obst_norm_bwd_coeffs: bad arguments");
  OBST_REQUIRE((reinterpret_cast<uintptr_t>(coeffs) & 15) == 0, "obst_norm_bwd_coeffs: coeffs must be 16-byte aligned");
  if (rows <= 0) return OBST_OK;
  const int64_t total = rows * heads;
  int64_t blocks = (total + 255) / 256;
  if (blocks > (int64_t)num_sms() * 16) blocks = (int64_t)num_sms() * 16;
  norm_bwd_coeffs_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const float2*>(partials), reinterpret_cast<const float2*>(stats),
      reinterpret_cast<float4*>(coeffs), total, nparts, 1.0f / (float)width);
  count_launch();
  return check_launch("obst_norm_bwd_coeffs");
}

int obst_norm_affine_grad(const void* w, const float* v1, const float* v2, float* dgamma, float* dbeta, int32_t heads,
                          int32_t kin, int32_t kout, void* stream) {
  OBST_REQUIRE(w && v1 && v2 && heads > 0 && kin > 0 && kout > 0, "obst_norm_affine_grad: bad arguments");
  if (!dgamma && !dbeta) return OBST_OK;
  const int64_t rows = (int64_t)heads * kin;
  int64_t blocks = (rows + 7) / 8;
  if (blocks > (int64_t)num_sms() * 16) blocks = (int64_t)num_sms() * 16;
  norm_affine_grad_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const bf16*)w, v1, v2, dgamma, dbeta,
                                                                               heads, kin, kout);
  count_launch();
  return check_launch("obst_norm_affine_grad");
}

}  // extern "C"
