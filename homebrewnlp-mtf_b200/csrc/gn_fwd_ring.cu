// Group norm forward, bf16, rows of 256 * MAXC elements (256 / 512 / 1024 wide in the reference's configs).
//
// The generic forward kernel (elementwise.cu) gives every warp one row at a time: load, reduce, normalise, store,
// next row.  With 64 registers per thread an SM holds 32 warps = 32 KB of loads in flight, which at ~1.2 us of
// loaded HBM latency caps the kernel at ~5 TB/s.  Here a warp keeps a HEAD fixed (so gamma / beta sit in registers,
// packed) and streams its tokens through a private cp.async ring: the next two rows are in flight while the current
// one is reduced, without spending registers on them.
#include "obst_common.cuh"

namespace obst {

__device__ __forceinline__ void cp16(void* smem_dst, const void* gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)),
               "l"(gmem_src)
               : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

static constexpr float GN_EPS = 1e-5f;  // src/model/normalization.py:28
static constexpr int RING = 3;         // rows in flight per warp
static constexpr int RING_WARPS = 8;  // warps per block

template <int MAXC, int ACT, bool STATS>
__global__ void __launch_bounds__(RING_WARPS * 32) gn_fwd_ring_kernel(const bf16* __restrict__ x,
                                                                     const bf16* __restrict__ gamma,
                                                                     const bf16* __restrict__ beta,
                                                                     bf16* __restrict__ y, float2* __restrict__ stats,
                                                                     int64_t tokens, int heads, int tokens_per_block) {
  constexpr int WIDTH = 256 * MAXC;
  constexpr int ROWB = WIDTH * 2;
  extern __shared__ __align__(16) uint8_t s_ring_raw[];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int h = blockIdx.y;
  uint8_t* ring = s_ring_raw + (size_t)warp * RING * ROWB;
  const float inv_w = 1.0f / (float)WIDTH;

  // gamma / beta of this head: packed bf16, MAXC x 16 bytes each per lane
  uint4 gp[MAXC], bp[MAXC];
#pragma unroll
  for (int i = 0; i < MAXC; ++i) {
    const int c = (lane + 32 * i) * 8;
    gp[i] = gamma ? *reinterpret_cast<const uint4*>(gamma + (int64_t)h * WIDTH + c)
                  : make_uint4(0x3F803F80u, 0x3F803F80u, 0x3F803F80u, 0x3F803F80u);  // bf16 1.0 pairs
    bp[i] = beta ? *reinterpret_cast<const uint4*>(beta + (int64_t)h * WIDTH + c) : make_uint4(0u, 0u, 0u, 0u);
  }

  const int64_t t0 = (int64_t)blockIdx.x * tokens_per_block;
  const int64_t t1 = t0 + tokens_per_block < tokens ? t0 + tokens_per_block : tokens;
  auto issue = [&](int64_t t, int stage) {
    const bf16* src = x + (t * heads + h) * WIDTH;
    uint8_t* dst = ring + stage * ROWB;
#pragma unroll
    for (int i = 0; i < MAXC; ++i) cp16(dst + (lane + 32 * i) * 16, src + (lane + 32 * i) * 8);
    cp_commit();
  };
#pragma unroll
  for (int s = 0; s < RING - 1; ++s) {
    if (t0 + warp + (int64_t)s * RING_WARPS < t1) issue(t0 + warp + (int64_t)s * RING_WARPS, s);
    else cp_commit();
  }
  int stage = 0;
  for (int64_t t = t0 + warp; t < t1; t += RING_WARPS, stage = (stage + 1) % RING) {
    const int64_t tn = t + (int64_t)(RING - 1) * RING_WARPS;
    if (tn < t1) issue(tn, (stage + RING - 1) % RING);
    else cp_commit();
    cp_wait<RING - 1>();
    __syncwarp();
    const uint8_t* src = ring + stage * ROWB;
    float v[MAXC][8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < MAXC; ++i) {
      Raw16<bf16> r;
      r.r = *reinterpret_cast<const uint4*>(src + (lane + 32 * i) * 16);
      r.unpack(v[i]);
#pragma unroll
      for (int j = 0; j < 8; ++j) s += v[i][j];
    }
    const float mean = warp_sum(s) * inv_w;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < MAXC; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        v[i][j] -= mean;
        q = fmaf(v[i][j], v[i][j], q);
      }
    const float rstd = rsqrtf(warp_sum(q) * inv_w + GN_EPS);
    const int64_t row = t * heads + h;
    if (STATS && lane == 0) stats[row] = make_float2(mean, rstd);
    bf16* dst = y + row * WIDTH;
#pragma unroll
    for (int i = 0; i < MAXC; ++i) {
      float g[8], b[8];
      Raw16<bf16> rg, rb;
      rg.r = gp[i];
      rb.r = bp[i];
      rg.unpack(g);
      rb.unpack(b);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float o = fmaf(v[i][j] * rstd, g[j], b[j]);
        if (ACT == 1) o = gelu_obst<true>(o);
        v[i][j] = o;
      }
      Raw16<bf16> out;
      out.pack(v[i]);
      out.store(dst + (lane + 32 * i) * 8);
    }
    __syncwarp();  // the stage is refilled by the next iteration's issue
  }
}

// Returns OBST_OK after launching, or 1 when this shape is not covered (the caller falls back to the generic kernel).
int gn_fwd_ring_launch(const void* x, const void* gamma, const void* beta, void* y, float2* stats, int64_t rows,
                       int heads, int width, int act, cudaStream_t st) {
  if ((width != 256 && width != 512 && width != 1024) || rows % heads != 0) return 1;
  const int64_t tokens = rows / heads;
  if (tokens < 4096) return 1;  // small problems: the generic kernel's finer grid wins
  int tpb = 128;                // tokens per block: 16 rows per warp amortise the gamma / beta loads
  const dim3 grid((unsigned)((tokens + tpb - 1) / tpb), (unsigned)heads);
  const size_t smem = (size_t)RING_WARPS * RING * width * 2;
#define OBST_RING3(MC, A, S)                                                                                      \
  do {                                                                                                            \
    static bool attr = false;                                                                                     \
    if (!attr && smem > 48 * 1024) {                                                                              \
      cudaFuncSetAttribute(gn_fwd_ring_kernel<MC, A, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
      attr = true;                                                                                                \
    }                                                                                                             \
    gn_fwd_ring_kernel<MC, A, S><<<grid, RING_WARPS * 32, smem, st>>>((const bf16*)x, (const bf16*)gamma,         \
                                                                      (const bf16*)beta, (bf16*)y, stats, tokens, \
                                                                      heads, tpb);                                \
  } while (0)
#define OBST_RING2(MC, A)            \
  do {                               \
    if (stats) OBST_RING3(MC, A, true); \
    else OBST_RING3(MC, A, false);   \
  } while (0)
#define OBST_RING1(MC)            \
  do {                            \
    if (act == 1) OBST_RING2(MC, 1); \
    else OBST_RING2(MC, 0);       \
  } while (0)
  if (width == 256) OBST_RING1(1);
  else if (width == 512) OBST_RING1(2);
  else OBST_RING1(4);
#undef OBST_RING1
#undef OBST_RING2
#undef OBST_RING3
  count_launch();
  count_path(OBST_PATH_GN_FWD_RING);
  return check_launch("obst_gn_fwd");
}

}  // namespace obst
