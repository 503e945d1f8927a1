// Group norm backward, bf16, rows of 256 / 512 / 1024 elements, with the (mean, rstd) the forward kernel saved.
//
// Plan (as the generic kernel in elementwise.cu): a warp walks the tokens of ONE head, so the dgamma / dbeta partials
// stay in registers and gamma / beta sit packed in registers; rows are streamed through a per-warp cp.async ring.
// Every option (fused gelu derivative, ReLU mask of a W_mid output, dx accumulation) is a template parameter.
//
// What bounds the kernel is the number of BYTES IN FLIGHT per SM, not instruction issue: ~6.5 TB/s x ~1 us of loaded
// HBM latency wants >= 44 KB outstanding per SM.  The generic kernel (128 registers -> two blocks per SM, one row of
// x and dy in flight per warp) holds 32 KB and runs at 4.7 TB/s; with dx accumulation it holds 48 KB and runs at 5.4.
// Here the ring is RING rows deep (RING - 1 in flight), sized so that two blocks per SM keep ~96 KB outstanding.
//
// A 1024-wide row is split over two warps (SPLIT = 2): per-lane state stays that of the 512-wide case and the two
// row sums are exchanged through shared memory under a 64-thread named barrier.
#include "obst_common.cuh"

namespace obst {

__device__ __forceinline__ void bcp16(void* smem_dst, const void* gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)),
               "l"(gmem_src)
               : "memory");
}
__device__ __forceinline__ void bcp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bcp_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

static constexpr int BWARPS = 8;  // warps per block

// MAXC: 16-byte chunks per lane (1: 256 columns per warp, 2: 512).  SPLIT: warps per row.  width = 256 MAXC SPLIT.
template <int MAXC, int SPLIT, int RING, int ACT, bool RELU, bool ACC>
__global__ void __launch_bounds__(BWARPS * 32, 2)
gn_bwd_ring_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ x, const bf16* __restrict__ gamma,
                   const bf16* __restrict__ beta, bf16* __restrict__ dx, float* __restrict__ dgamma,
                   float* __restrict__ dbeta, const float2* __restrict__ stats, int64_t tokens, int heads,
                   int tokens_per_block) {
  constexpr int NT = ACC ? 3 : 2;         // tensors staged per row: x, dy (, dx)
  constexpr int WCOLS = 256 * MAXC;       // columns per warp
  constexpr int WIDTH = WCOLS * SPLIT;    // row width
  constexpr int ROWB = WCOLS * 2;         // bytes of one warp's part of a row
  constexpr int SLOTS = BWARPS / SPLIT;   // rows a block works on at a time
  extern __shared__ __align__(16) uint8_t s_raw[];
  __shared__ float s_x[SLOTS][2][2][2];   // [row slot][row parity][which sum][half]: the SPLIT == 2 exchange
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int slot = warp / SPLIT;   // which row of the block's current group this warp works on
  const int half = warp % SPLIT;   // which part of that row
  const int col0 = half * WCOLS;
  const int h = blockIdx.y;
  uint8_t* ring = s_raw + (size_t)warp * RING * NT * ROWB;
  uint4 gp[MAXC], bp[MAXC];
#pragma unroll
  for (int i = 0; i < MAXC; ++i) {
    const int c = col0 + (lane + 32 * i) * 8;
    gp[i] = gamma ? *reinterpret_cast<const uint4*>(gamma + (int64_t)h * WIDTH + c)
                  : make_uint4(0x3F803F80u, 0x3F803F80u, 0x3F803F80u, 0x3F803F80u);  // bf16 1.0 pairs
    bp[i] = (ACT == 1 && beta) ? *reinterpret_cast<const uint4*>(beta + (int64_t)h * WIDTH + c)
                               : make_uint4(0u, 0u, 0u, 0u);
  }
  float pg[MAXC][8], pb[MAXC][8];
#pragma unroll
  for (int i = 0; i < MAXC; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) pg[i][j] = pb[i][j] = 0.f;

  const int64_t t0 = (int64_t)blockIdx.x * tokens_per_block;
  const int64_t t1 = t0 + tokens_per_block < tokens ? t0 + tokens_per_block : tokens;
  auto issue = [&](int64_t t, int stage) {
    const int64_t off = (t * heads + h) * WIDTH + col0;
    uint8_t* dst = ring + stage * NT * ROWB;
#pragma unroll
    for (int i = 0; i < MAXC; ++i) {
      const int c = lane + 32 * i;
      bcp16(dst + c * 16, x + off + c * 8);
      bcp16(dst + ROWB + c * 16, dy + off + c * 8);
      if (ACC) bcp16(dst + 2 * ROWB + c * 16, dx + off + c * 8);
    }
    bcp_commit();
  };
#pragma unroll
  for (int s = 0; s < RING - 1; ++s) {
    if (t0 + slot + (int64_t)s * SLOTS < t1) issue(t0 + slot + (int64_t)s * SLOTS, s);
    else bcp_commit();  // keep one group per ring slot so the wait_group counts stay aligned
  }
  float2 st_next = make_float2(0.f, 0.f);
  if (t0 + slot < t1) st_next = stats[(t0 + slot) * heads + h];
  int stage = 0, parity = 0;
  for (int64_t t = t0 + slot; t < t1; t += SLOTS, stage = (stage + 1 == RING ? 0 : stage + 1), parity ^= 1) {
    const int64_t tn = t + (int64_t)(RING - 1) * SLOTS;
    if (tn < t1) issue(tn, stage == 0 ? RING - 1 : stage - 1);
    else bcp_commit();
    const float2 st = st_next;
    if (t + SLOTS < t1) st_next = stats[(t + SLOTS) * heads + h];
    bcp_wait<RING - 1>();
    __syncwarp();
    const uint8_t* src = ring + stage * NT * ROWB;
    const float rstd = st.y, nmr = -st.x * st.y;
    float z[MAXC][8], g[MAXC][8];
    float sg = 0.f, sgz = 0.f;
    uint32_t maskbits = 0;
#pragma unroll
    for (int i = 0; i < MAXC; ++i) {
      const int c = lane + 32 * i;
      Raw16<bf16> rx, rd, rg;
      rx.r = *reinterpret_cast<const uint4*>(src + c * 16);
      rd.r = *reinterpret_cast<const uint4*>(src + ROWB + c * 16);
      rg.r = gp[i];
      float xv[8], dv[8], gam[8], bet[8];
      rx.unpack(xv);
      rd.unpack(dv);
      rg.unpack(gam);
      if (ACT == 1) {
        Raw16<bf16> rb;
        rb.r = bp[i];
        rb.unpack(bet);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (RELU && xv[j] > 0.f) maskbits |= 1u << (i * 8 + j);
        const float zz = fmaf(xv[j], rstd, nmr);
        float d = dv[j];
        if (ACT == 1) d *= gelu_obst_grad<true>(fmaf(zz, gam[j], bet[j]));
        pb[i][j] += d;
        pg[i][j] = fmaf(d, zz, pg[i][j]);
        const float gg = d * gam[j];
        z[i][j] = zz;
        g[i][j] = gg;
        sg += gg;
        sgz = fmaf(gg, zz, sgz);
      }
    }
    sg = warp_sum(sg);
    sgz = warp_sum(sgz);
    if (SPLIT == 2) {  // the other half of the row lives in the partner warp
      if (lane == 0) {
        s_x[slot][parity][0][half] = sg;
        s_x[slot][parity][1][half] = sgz;
      }
      asm volatile("bar.sync %0, 64;" ::"r"(1 + slot) : "memory");
      sg = s_x[slot][parity][0][0] + s_x[slot][parity][0][1];
      sgz = s_x[slot][parity][1][0] + s_x[slot][parity][1][1];
    }
    const float c0 = -sg * (1.0f / WIDTH) * rstd, c1 = -sgz * (1.0f / WIDTH) * rstd;
    bf16* dst = dx + (t * heads + h) * WIDTH + col0;
#pragma unroll
    for (int i = 0; i < MAXC; ++i) {
      const int c = lane + 32 * i;
      float o[8];
      if (ACC) {
        Raw16<bf16> ra;
        ra.r = *reinterpret_cast<const uint4*>(src + 2 * ROWB + c * 16);
        ra.unpack(o);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float r = fmaf(z[i][j], c1, fmaf(g[i][j], rstd, c0));  // rstd * (g - mean(g) - z * mean(g z))
        if (RELU && !((maskbits >> (i * 8 + j)) & 1u)) r = 0.f;
        o[j] = ACC ? o[j] + r : r;
      }
      Raw16<bf16> out;
      out.pack(o);
      out.store(dst + c * 8);
    }
    __syncwarp();  // everyone is done with this stage before the next iteration refills it
  }
  // Block reduce of the affine gradients without shared-memory atomics (fp32 atomicAdd on shared memory is a
  // compare-and-swap loop, 8-way contended here): every warp parks its partials in its own -- now idle -- ring,
  // then each thread sums one column over the warps that hold it and issues ONE global atomic.
  bcp_wait<0>();
  __syncwarp();
  static_assert(RING * NT * ROWB >= 2 * WCOLS * (int)sizeof(float), "ring too small to park the partials");
  float* park = reinterpret_cast<float*>(ring);
#pragma unroll
  for (int i = 0; i < MAXC; ++i) {
    const int c = (lane + 32 * i) * 8;
    *reinterpret_cast<float4*>(park + c) = make_float4(pg[i][0], pg[i][1], pg[i][2], pg[i][3]);
    *reinterpret_cast<float4*>(park + c + 4) = make_float4(pg[i][4], pg[i][5], pg[i][6], pg[i][7]);
    *reinterpret_cast<float4*>(park + WCOLS + c) = make_float4(pb[i][0], pb[i][1], pb[i][2], pb[i][3]);
    *reinterpret_cast<float4*>(park + WCOLS + c + 4) = make_float4(pb[i][4], pb[i][5], pb[i][6], pb[i][7]);
  }
  __syncthreads();
  const float* rings = reinterpret_cast<const float*>(s_raw);
  constexpr int RING_FLOATS = RING * NT * ROWB / (int)sizeof(float);
  for (int col = threadIdx.x; col < WIDTH; col += blockDim.x) {
    const int hf = col / WCOLS, cc = col - hf * WCOLS;
    float a = 0.f, b = 0.f;
#pragma unroll
    for (int sl = 0; sl < SLOTS; ++sl) {
      const float* pw = rings + (size_t)(sl * SPLIT + hf) * RING_FLOATS;
      a += pw[cc];
      b += pw[WCOLS + cc];
    }
    if (dgamma) atomicAdd(&dgamma[(int64_t)h * WIDTH + col], a);
    if (dbeta) atomicAdd(&dbeta[(int64_t)h * WIDTH + col], b);
  }
}

template <int MAXC, int SPLIT, int ACT, bool RELU, bool ACC>
static int gn_bwd_ring_typed(const void* dy, const void* x, const void* gamma, const void* beta, void* dx,
                             float* dgamma, float* dbeta, const float2* stats, int64_t tokens, int heads,
                             cudaStream_t st) {
  // ~8-12 KB of ring per warp whatever the row part: narrow parts get a deeper ring
  constexpr int RING = MAXC == 1 ? 8 : 4;
  constexpr int WIDTH = 256 * MAXC * SPLIT;
  constexpr int SMEM = BWARPS * RING * (ACC ? 3 : 2) * (256 * MAXC * 2);
  // 32 rows per warp (the ring's start-up is RING - 1 rows), fewer when the problem would not fill the machine
  int tpb = 32 * (BWARPS / SPLIT) * (MAXC == 1 ? 2 : 1);
  while (tpb > 32 && ((tokens + tpb - 1) / tpb) * heads < 4 * (int64_t)num_sms()) tpb >>= 1;
  const dim3 grid((unsigned)((tokens + tpb - 1) / tpb), (unsigned)heads);
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(gn_bwd_ring_kernel<MAXC, SPLIT, RING, ACT, RELU, ACC>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess) {
      set_error("obst_gn_bwd: cannot opt in to %d B of shared memory: %s", SMEM, cudaGetErrorString(e));
      return OBST_ECUDA;
    }
    attr = true;
  }
  gn_bwd_ring_kernel<MAXC, SPLIT, RING, ACT, RELU, ACC><<<grid, BWARPS * 32, SMEM, st>>>(
      (const bf16*)dy, (const bf16*)x, (const bf16*)gamma, (const bf16*)beta, (bf16*)dx, dgamma, dbeta, stats, tokens,
      heads, tpb);
  count_launch();
  count_path(OBST_PATH_GN_BWD_RING);
  return check_launch("obst_gn_bwd");
}

template <int MAXC, int SPLIT>
static int gn_bwd_ring_opts(const void* dy, const void* x, const void* gamma, const void* beta, void* dx, float* dgamma,
                            float* dbeta, const float2* stats, int64_t tokens, int heads, int act, int relu_mask,
                            int dx_acc, cudaStream_t st) {
#define OBST_BR(A, R, C) \
  return gn_bwd_ring_typed<MAXC, SPLIT, A, R, C>(dy, x, gamma, beta, dx, dgamma, dbeta, stats, tokens, heads, st)
  if (act == 1) {
    if (relu_mask) {
      if (dx_acc) OBST_BR(1, true, true);
      OBST_BR(1, true, false);
    }
    if (dx_acc) OBST_BR(1, false, true);
    OBST_BR(1, false, false);
  }
  if (relu_mask) {
    if (dx_acc) OBST_BR(0, true, true);
    OBST_BR(0, true, false);
  }
  if (dx_acc) OBST_BR(0, false, true);
  OBST_BR(0, false, false);
#undef OBST_BR
}

// Returns OBST_OK after launching, or 1 when this case is not covered (the caller falls back to the generic kernel).
int gn_bwd_ring_launch(const void* dy, const void* x, const void* gamma, const void* beta, void* dx, float* dgamma,
                       float* dbeta, const float2* stats, int64_t rows, int heads, int width, int act, int relu_mask,
                       int dx_acc, cudaStream_t st) {
  if (!stats || rows % heads != 0 || (width != 256 && width != 512 && width != 1024)) return 1;
  const int64_t tokens = rows / heads;
  if (tokens < 4096) return 1;  // small problems: the generic kernel's finer grid wins
  if (width == 256)
    return gn_bwd_ring_opts<1, 1>(dy, x, gamma, beta, dx, dgamma, dbeta, stats, tokens, heads, act, relu_mask, dx_acc, st);
  if (width == 512)
    return gn_bwd_ring_opts<2, 1>(dy, x, gamma, beta, dx, dgamma, dbeta, stats, tokens, heads, act, relu_mask, dx_acc, st);
  return gn_bwd_ring_opts<2, 2>(dy, x, gamma, beta, dx, dgamma, dbeta, stats, tokens, heads, act, relu_mask, dx_acc, st);
}

}  // namespace obst
