// Bandwidth-bound kernels of the mixer block: group norm (+gelu) forward/backward, activations,
// residual arithmetic, embedding gather/scatter, softmax cross-entropy, parameter shadow casts.
// All are coalesced, 16-byte vectorised where the layout allows, with warp-shuffle reductions.
#include "obst_common.cuh"

namespace obst {

static constexpr float GN_EPS = 1e-5f;  // src/model/normalization.py:28

// ================================================================================================
// group norm forward: one warp per (token, head) row of `width` contiguous elements
// ================================================================================================
// ROWS rows are in flight per warp (all their loads are issued before the first reduction): with one 1 KB row per
// warp the kernel is latency-bound at ~60 % of HBM; 4 rows give the memory system 8 x 16 B per lane to chew on.
template <typename T, int MAXC, int ROWS, int ACT, bool STATS>
__global__ void __launch_bounds__(256) gn_fwd_kernel(const T* __restrict__ x, const T* __restrict__ gamma,
                                                      const T* __restrict__ beta, T* __restrict__ y,
                                                      float2* __restrict__ stats, int64_t rows, int heads, int width) {
  constexpr int N = Vec16<T>::N;
  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  const int chunks = width / N;
  const float inv_w = 1.0f / (float)width;
  const int64_t gstride = (int64_t)gridDim.x * warps_per_block * ROWS;
  for (int64_t base = ((int64_t)blockIdx.x * warps_per_block + (threadIdx.x >> 5)) * ROWS; base < rows;
       base += gstride) {
    Vec16<T> v[ROWS][MAXC];
    float s[ROWS];
#pragma unroll
    for (int r = 0; r < ROWS; ++r) {
      s[r] = 0.f;
      const int64_t row = base + r;
      if (row < rows) {
        const T* xr = x + row * width;
#pragma unroll
        for (int i = 0; i < MAXC; ++i) {
          const int c = lane + 32 * i;
          if (c < chunks) v[r][i].load(xr + c * N);
        }
      }
    }
#pragma unroll
    for (int r = 0; r < ROWS; ++r) {
#pragma unroll
      for (int i = 0; i < MAXC; ++i) {
        const int c = lane + 32 * i;
        if (c < chunks && base + r < rows) {
#pragma unroll
          for (int j = 0; j < N; ++j) s[r] += v[r][i].v[j];
        }
      }
      s[r] = warp_sum(s[r]) * inv_w;  // mean
    }
    float q[ROWS];
#pragma unroll
    for (int r = 0; r < ROWS; ++r) {
      q[r] = 0.f;
#pragma unroll
      for (int i = 0; i < MAXC; ++i) {
        const int c = lane + 32 * i;
        if (c < chunks && base + r < rows) {
#pragma unroll
          for (int j = 0; j < N; ++j) {
            v[r][i].v[j] -= s[r];
            q[r] += v[r][i].v[j] * v[r][i].v[j];
          }
        }
      }
      q[r] = rsqrtf(warp_sum(q[r]) * inv_w + GN_EPS);  // rstd
      if (STATS && lane == 0 && base + r < rows) stats[base + r] = make_float2(s[r], q[r]);
    }
#pragma unroll
    for (int r = 0; r < ROWS; ++r) {
      const int64_t row = base + r;
      if (row >= rows) continue;
      const int h = (int)(row % heads);
      T* yr = y + row * width;
#pragma unroll
      for (int i = 0; i < MAXC; ++i) {
        const int c = lane + 32 * i;
        if (c < chunks) {
          Vec16<T> g, b;
          if (gamma) g.load(gamma + (int64_t)h * width + c * N);
          if (beta) b.load(beta + (int64_t)h * width + c * N);
#pragma unroll
          for (int j = 0; j < N; ++j) {
            float o = v[r][i].v[j] * q[r];
            if (gamma) o *= g.v[j];
            if (beta) o += b.v[j];
            if (ACT == 1) o = gelu_obst<sizeof(T) == 2>(o);
            v[r][i].v[j] = o;
          }
          v[r][i].store(yr + c * N);
        }
      }
    }
  }
}

// ================================================================================================
// group norm backward.  grid = (token chunks, heads); a warp walks tokens of ONE head so the
// dgamma/dbeta partials stay in registers; block-level reduce in shared memory, then fp32 atomics.
// ================================================================================================
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)),
               "l"(gmem_src)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// Each warp streams its rows through a private BSTAGES-deep shared-memory ring with cp.async (no register staging): the
// next row's x / dy (/ dx) are in flight while the current row is reduced, and the only long-lived registers are
// the dgamma/dbeta partials -- occupancy AND memory-level parallelism stay high.
// SPLIT = warps sharing one row (1 or 2).  A 1024-wide row on one warp needs 32 elements per lane x (z, g, dgamma,
// dbeta) = 244 registers -> one block per SM; split over two warps it is the 512-wide case again (128 registers,
// two blocks), at the price of three 64-thread named-barrier exchanges per row for the row statistics.
template <typename T, int MAXC, int BSTAGES, int SPLIT, int ACT>
__global__ void __launch_bounds__(256) gn_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x,
                                                      const T* __restrict__ gamma, const T* __restrict__ beta,
                                                      T* __restrict__ dx, float* __restrict__ dgamma,
                                                      float* __restrict__ dbeta, const float2* __restrict__ stats,
                                                      int64_t tokens, int heads, int width, int tokens_per_block,
                                                      int relu_mask, int dx_acc) {
  constexpr int N = Raw16<T>::N;
  extern __shared__ __align__(16) uint8_t s_raw[];
  __shared__ float s_x[4][6][2];  // [row slot][exchange + 3 * row parity][half]
  __shared__ float s_y[4][2][2];  // second value of the third exchange
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int slots = (blockDim.x >> 5) / SPLIT;  // rows in flight per block iteration
  const int vw = warp / SPLIT;                  // row slot of this warp
  const int half = warp % SPLIT;                // which part of the row it owns
  const int wpart = width / SPLIT;              // columns per warp
  const int chunks = wpart / N;
  const int col0 = half * wpart;
  const int h = blockIdx.y;
  const float inv_w = 1.0f / (float)width;
  const int rowb = wpart * (int)sizeof(T);      // bytes of one warp's part of a row
  // layout: [2*width floats: block reduction][gamma row][beta row][per warp: BSTAGES x (x, dy, dx) row parts]
  float* s_red = reinterpret_cast<float*>(s_raw);
  T* s_gam = reinterpret_cast<T*>(s_raw + 2 * width * sizeof(float));
  T* s_bet = s_gam + width;
  uint8_t* s_ring = reinterpret_cast<uint8_t*>(s_bet + width) + (size_t)warp * BSTAGES * 3 * rowb;
  for (int i = threadIdx.x; i < 2 * width; i += blockDim.x) s_red[i] = 0.f;
  for (int i = threadIdx.x; i < width; i += blockDim.x) {
    s_gam[i] = gamma ? gamma[(int64_t)h * width + i] : from_f32<T>(1.f);
    s_bet[i] = beta ? beta[(int64_t)h * width + i] : from_f32<T>(0.f);
  }
  __syncthreads();

  // row-wide sum of a per-lane value: warp shuffle, then (SPLIT == 2) one exchange with the partner warp
  auto row_sum = [&](float v, int slot) -> float {
    v = warp_sum(v);
    if (SPLIT == 1) return v;
    if (lane == 0) s_x[vw][slot][half] = v;
    asm volatile("bar.sync %0, 64;" ::"r"(1 + vw) : "memory");
    return s_x[vw][slot][0] + s_x[vw][slot][1];
  };

  float pg[MAXC][N], pb[MAXC][N];
#pragma unroll
  for (int i = 0; i < MAXC; ++i)
#pragma unroll
    for (int j = 0; j < N; ++j) {
      pg[i][j] = 0.f;
      pb[i][j] = 0.f;
    }

  const int64_t t0 = (int64_t)blockIdx.x * tokens_per_block;
  const int64_t t1 = (t0 + tokens_per_block < tokens) ? t0 + tokens_per_block : tokens;
  auto issue = [&](int64_t t, int stage) {
    const int64_t off = (t * heads + h) * width + col0;
    uint8_t* base = s_ring + stage * 3 * rowb;
#pragma unroll
    for (int i = 0; i < MAXC; ++i) {
      const int c = lane + 32 * i;
      if (c < chunks) {
        cp_async16(base + c * 16, x + off + c * N);
        cp_async16(base + rowb + c * 16, dy + off + c * N);
        if (dx_acc) cp_async16(base + 2 * rowb + c * 16, dx + off + c * N);
      }
    }
    cp_async_commit();
  };
  int stage = 0;
  int parity = 0;
  // (mean, rstd) of the row saved by the forward kernel, fetched one row ahead: with them the two statistics passes
  // (and, for split rows, two of the three exchanges) disappear -- the kernel is instruction-issue bound
  float2 st_next = make_float2(0.f, 0.f);
  if (stats && t0 + vw < t1) st_next = stats[(t0 + vw) * heads + h];
#pragma unroll
  for (int s0 = 0; s0 < BSTAGES - 1; ++s0) {
    if (t0 + vw + (int64_t)s0 * slots < t1) issue(t0 + vw + (int64_t)s0 * slots, s0);
    else cp_async_commit();  // keep one group per slot so wait_group counts stay aligned
  }
  for (int64_t t = t0 + vw; t < t1; t += slots, stage = (stage + 1) % BSTAGES, parity ^= 1) {
    const int64_t tn = t + (int64_t)(BSTAGES - 1) * slots;
    if (tn < t1) issue(tn, (stage + BSTAGES - 1) % BSTAGES);
    else cp_async_commit();
    const float2 st_cur = st_next;
    if (stats && t + slots < t1) st_next = stats[(t + slots) * heads + h];
    cp_async_wait<BSTAGES - 1>();
    __syncwarp();
    const int64_t off = (t * heads + h) * width + col0;
    const T* sx = reinterpret_cast<const T*>(s_ring + stage * 3 * rowb);
    const T* sdy = reinterpret_cast<const T*>(s_ring + stage * 3 * rowb + rowb);
    const T* sdx = reinterpret_cast<const T*>(s_ring + stage * 3 * rowb + 2 * rowb);
    // The kernel is instruction-issue bound, not latency bound: unpack each operand ONCE into fp32 registers and
    // reuse (x -> z, dy -> g); gamma/beta come from shared memory when needed.
    float zv[MAXC][N], gv[MAXC][N];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < MAXC; ++i) {
      const int c = lane + 32 * i;
      if (c < chunks) {
        Vec16<T> v;
        v.load(sx + c * N);
#pragma unroll
        for (int j = 0; j < N; ++j) zv[i][j] = v.v[j];
      }
    }
    if (!stats) {
#pragma unroll
      for (int i = 0; i < MAXC; ++i) {
        if (lane + 32 * i < chunks) {
#pragma unroll
          for (int j = 0; j < N; ++j) s += zv[i][j];
        }
      }
    }
    float mean, rstd;
    if (stats) {
      mean = st_cur.x;
      rstd = st_cur.y;
    } else {
      mean = row_sum(s, 3 * parity + 0) * inv_w;
      float q = 0.f;
#pragma unroll
      for (int i = 0; i < MAXC; ++i) {
        if (lane + 32 * i < chunks) {
#pragma unroll
          for (int j = 0; j < N; ++j) {
            const float d = zv[i][j] - mean;
            q += d * d;
          }
        }
      }
      rstd = rsqrtf(row_sum(q, 3 * parity + 1) * inv_w + GN_EPS);
    }
    const float nmr = -mean * rstd;
    float sg = 0.f, sgz = 0.f;
    unsigned long long maskbits = 0;  // bit (i*N+j) <= 63: raw x > 0 (ReLU mask), taken before x becomes z
#pragma unroll
    for (int i = 0; i < MAXC; ++i) {
      const int c = lane + 32 * i;
      if (c < chunks) {
        Vec16<T> dv, gam;
        dv.load(sdy + c * N);
        gam.load(s_gam + col0 + c * N);
        Vec16<T> bet;
        if (ACT == 1) bet.load(s_bet + col0 + c * N);
#pragma unroll
        for (int j = 0; j < N; ++j) {
          if (relu_mask && zv[i][j] > 0.f) maskbits |= 1ull << (i * N + j);
          const float z = fmaf(zv[i][j], rstd, nmr);
          zv[i][j] = z;
          float d = dv.v[j];
          if (ACT == 1) d *= gelu_obst_grad<sizeof(T) == 2>(z * gam.v[j] + bet.v[j]);
          pb[i][j] += d;
          pg[i][j] += d * z;
          const float g = d * gam.v[j];
          gv[i][j] = g;
          sg += g;
          sgz += g * z;
        }
      }
    }
    // one exchange for both sums: pack them through two slots of the same barrier round
    float mg, mgz;
    if (SPLIT == 1) {
      mg = warp_sum(sg) * inv_w;
      mgz = warp_sum(sgz) * inv_w;
    } else {
      sg = warp_sum(sg);
      sgz = warp_sum(sgz);
      if (lane == 0) {
        s_x[vw][3 * parity + 2][half] = sg;
        s_y[vw][parity][half] = sgz;
      }
      asm volatile("bar.sync %0, 64;" ::"r"(1 + vw) : "memory");
      mg = (s_x[vw][3 * parity + 2][0] + s_x[vw][3 * parity + 2][1]) * inv_w;
      mgz = (s_y[vw][parity][0] + s_y[vw][parity][1]) * inv_w;
    }
    const float c0 = -mg * rstd, c1 = -mgz * rstd;
#pragma unroll
    for (int i = 0; i < MAXC; ++i) {
      const int c = lane + 32 * i;
      if (c < chunks) {
        Vec16<T> o;
        if (dx_acc) o.load(sdx + c * N);
#pragma unroll
        for (int j = 0; j < N; ++j) {
          float r = fmaf(zv[i][j], c1, fmaf(gv[i][j], rstd, c0));  // rstd * (g - mean(g) - z * mean(g z))
          if (relu_mask && !((maskbits >> (i * N + j)) & 1ull)) r = 0.f;
          o.v[j] = dx_acc ? (o.v[j] + r) : r;
        }
        o.store(dx + off + c * N);
      }
    }
    __syncwarp();  // everyone is done with this stage before the next iteration refills it
  }
  // block reduce of the affine gradients
#pragma unroll
  for (int i = 0; i < MAXC; ++i) {
    const int c = lane + 32 * i;
    if (c < chunks) {
#pragma unroll
      for (int j = 0; j < N; ++j) {
        if (dgamma) atomicAdd(&s_red[col0 + c * N + j], pg[i][j]);
        if (dbeta) atomicAdd(&s_red[width + col0 + c * N + j], pb[i][j]);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < width; i += blockDim.x) {
    if (dgamma) atomicAdd(&dgamma[(int64_t)h * width + i], s_red[i]);
    if (dbeta) atomicAdd(&dbeta[(int64_t)h * width + i], s_red[width + i]);
  }
}

template <typename T>
static int gn_fwd_launch(const void* x, const void* gamma, const void* beta, void* y, float2* stats, int64_t rows,
                         int heads, int width, int act, cudaStream_t st) {
  constexpr int N = Vec16<T>::N;
  OBST_REQUIRE(width % N == 0, "obst_gn_fwd: width %d must be a multiple of %d", width, N);
  const int per_lane = (width / N + 31) / 32;
  OBST_REQUIRE(per_lane <= 8, "obst_gn_fwd: width %d too large (max %d)", width, 256 * N);
  const int wpb = 8;
  const int rpw = 1;  // measured on B200: 1 row/warp at 64 warps/SM (4.0 TB/s) beats 4 rows/warp at 16 warps/SM (3.2)
  int64_t blocks = (rows + (int64_t)wpb * rpw - 1) / ((int64_t)wpb * rpw);
  const int64_t cap = (int64_t)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
#define OBST_GN_FWD2(MC, R, A, S)                                                                              \
  gn_fwd_kernel<T, MC, R, A, S><<<(unsigned)blocks, wpb * 32, 0, st>>>((const T*)x, (const T*)gamma, (const T*)beta, \
                                                                        (T*)y, stats, rows, heads, width)
#define OBST_GN_FWD(MC, R)                        \
  do {                                            \
    if (act == 1) {                               \
      if (stats) OBST_GN_FWD2(MC, R, 1, true);    \
      else OBST_GN_FWD2(MC, R, 1, false);         \
    } else {                                      \
      if (stats) OBST_GN_FWD2(MC, R, 0, true);    \
      else OBST_GN_FWD2(MC, R, 0, false);         \
    }                                             \
  } while (0)
  if (per_lane <= 1) OBST_GN_FWD(1, 1);
  else if (per_lane <= 2) OBST_GN_FWD(2, 1);
  else if (per_lane <= 4) OBST_GN_FWD(4, 1);
  else OBST_GN_FWD(8, 1);
#undef OBST_GN_FWD
#undef OBST_GN_FWD2
  count_launch();
  return check_launch("obst_gn_fwd");
}

template <typename T>
static int gn_bwd_launch(const void* dy, const void* x, const void* gamma, const void* beta, void* dx, float* dgamma,
                         float* dbeta, const float2* stats, int64_t rows, int heads, int width, int act, int relu_mask,
                         int dx_acc, cudaStream_t st) {
  constexpr int N = Vec16<T>::N;
  OBST_REQUIRE(width % N == 0, "obst_gn_bwd: width %d must be a multiple of %d", width, N);
  OBST_REQUIRE(rows % heads == 0, "obst_gn_bwd: rows %lld not a multiple of heads %d", (long long)rows, heads);
  const int per_lane = (width / N + 31) / 32;
  OBST_REQUIRE(per_lane <= 8, "obst_gn_bwd: width %d too large (max %d)", width, 256 * N);
  const int64_t tokens = rows / heads;
  int tpb = 256;
  // keep at least ~4 blocks per SM in flight when the problem is small
  while (tpb > 8 && ((tokens + tpb - 1) / tpb) * heads < (int64_t)num_sms() * 4) tpb >>= 1;
  dim3 grid((unsigned)((tokens + tpb - 1) / tpb), (unsigned)heads);
  const size_t rowb = (size_t)width * sizeof(T);
  // wide rows (more than 2 x 16 B chunks per lane) are split over two warps: per-lane state halves
  const int split = (per_lane > 2 && per_lane % 2 == 0 && (width / N) % 64 == 0) ? 2 : 1;
  const int pl = per_lane / split;
  // ring depth: 4 row parts per warp where two blocks still fit an SM, else 2
  const bool deep = (2 * (size_t)width * sizeof(float) + 2 * rowb + 8 * 4 * 3 * (rowb / split)) <= 100 * 1024;
  const int bst = deep ? 4 : 2;
  const size_t smem = 2 * (size_t)width * sizeof(float) + 2 * rowb + 8 * (size_t)bst * 3 * (rowb / split);
  OBST_REQUIRE(smem <= 220 * 1024, "obst_gn_bwd: width %d needs %zu B of shared memory", width, smem);
#define OBST_GN_BWD2(MC, ST, SP, AC)                                                                                  \
  do {                                                                                                                \
    if (smem > 48 * 1024)                                                                                             \
      cudaFuncSetAttribute(gn_bwd_kernel<T, MC, ST, SP, AC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    gn_bwd_kernel<T, MC, ST, SP, AC><<<grid, 256, smem, st>>>((const T*)dy, (const T*)x, (const T*)gamma,            \
                                                               (const T*)beta, (T*)dx, dgamma, dbeta, stats, tokens, \
                                                               heads, width, tpb, relu_mask, dx_acc);                \
  } while (0)
#define OBST_GN_BWD1(MC, ST, SP)               \
  do {                                         \
    if (act == 1) OBST_GN_BWD2(MC, ST, SP, 1); \
    else OBST_GN_BWD2(MC, ST, SP, 0);          \
  } while (0)
#define OBST_GN_BWD(MC, SP)            \
  do {                                 \
    if (deep) OBST_GN_BWD1(MC, 4, SP); \
    else OBST_GN_BWD1(MC, 2, SP);      \
  } while (0)
  if (split == 2) {
    if (pl <= 2) OBST_GN_BWD(2, 2);
    else OBST_GN_BWD(4, 2);
  } else {
    if (pl <= 1) OBST_GN_BWD(1, 1);
    else if (pl <= 2) OBST_GN_BWD(2, 1);
    else if (pl <= 4) OBST_GN_BWD(4, 1);
    else OBST_GN_BWD(8, 1);
  }
#undef OBST_GN_BWD2
#undef OBST_GN_BWD1
#undef OBST_GN_BWD
  count_launch();
  return check_launch("obst_gn_bwd");
}

// ================================================================================================
// activations and residual arithmetic: grid-stride, 16-byte vectors + scalar tail
// ================================================================================================
template <typename T, int KIND, bool BWD>
__global__ void __launch_bounds__(256) act_kernel(const T* __restrict__ x, const T* __restrict__ dy,
                                                   T* __restrict__ out, int64_t n) {
  constexpr int N = Vec16<T>::N;
  const int64_t nv = n / N;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += stride) {
    Vec16<T> a, g;
    a.load(x + i * N);
    if (BWD) g.load(dy + i * N);
#pragma unroll
    for (int j = 0; j < N; ++j) {
      const float v = a.v[j];
      float o;
      if (KIND == 0) o = BWD ? (v > 0.f ? g.v[j] : 0.f) : fmaxf(v, 0.f);
      else o = BWD ? g.v[j] * gelu_obst_grad<sizeof(T) == 2>(v) : gelu_obst<sizeof(T) == 2>(v);
      a.v[j] = o;
    }
    a.store(out + i * N);
  }
  for (int64_t i = nv * N + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float v = to_f32(x[i]);
    const float g = BWD ? to_f32(dy[i]) : 0.f;
    float o;
    if (KIND == 0) o = BWD ? (v > 0.f ? g : 0.f) : fmaxf(v, 0.f);
    else o = BWD ? g * gelu_obst_grad<sizeof(T) == 2>(v) : gelu_obst<sizeof(T) == 2>(v);
    out[i] = from_f32<T>(o);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) axpby_kernel(const T* a, const T* b, float sign, T* out, int64_t n) {
  constexpr int N = Vec16<T>::N;
  const int64_t nv = n / N;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += stride) {
    Vec16<T> va, vb;
    va.load(a + i * N);
    vb.load(b + i * N);
#pragma unroll
    for (int j = 0; j < N; ++j) va.v[j] = va.v[j] + sign * vb.v[j];
    va.store(out + i * N);
  }
  for (int64_t i = nv * N + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    out[i] = from_f32<T>(to_f32(a[i]) + sign * to_f32(b[i]));
}

static inline unsigned ew_blocks(int64_t work_items) {
  int64_t b = (work_items + 255) / 256;
  const int64_t cap = (int64_t)num_sms() * 16;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (unsigned)b;
}

// ================================================================================================
// embedding gather / scatter-add
// ================================================================================================
template <typename T>
__global__ void __launch_bounds__(256) gather_kernel(const int32_t* __restrict__ idx, const T* __restrict__ table,
                                                      T* __restrict__ out, int64_t tokens, int vocab, int width) {
  const int64_t total = tokens * width;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int64_t t = i / width;
    const int c = (int)(i - t * width);
    int v = idx[t];
    v = v < 0 ? 0 : (v >= vocab ? vocab - 1 : v);
    out[i] = table[(int64_t)v * width + c];
  }
}

// each block owns a token range, accumulates into a shared-memory copy of the table, then flushes
template <typename T>
__global__ void __launch_bounds__(256) scatter_kernel(const int32_t* __restrict__ idx, const T* __restrict__ dout,
                                                       float* __restrict__ dtable, int64_t tokens, int vocab,
                                                       int width, int tokens_per_block, int use_smem) {
  extern __shared__ float s_tab[];
  const int64_t t0 = (int64_t)blockIdx.x * tokens_per_block;
  const int64_t t1 = (t0 + tokens_per_block < tokens) ? t0 + tokens_per_block : tokens;
  const int tsz = vocab * width;
  if (use_smem) {
    for (int i = threadIdx.x; i < tsz; i += blockDim.x) s_tab[i] = 0.f;
    __syncthreads();
  }
  const int64_t total = (t1 - t0) * width;
  for (int64_t i = threadIdx.x; i < total; i += blockDim.x) {
    const int64_t t = t0 + i / width;
    const int c = (int)(i % width);
    int v = idx[t];
    v = v < 0 ? 0 : (v >= vocab ? vocab - 1 : v);
    const float g = to_f32(dout[t * width + c]);
    if (use_smem) atomicAdd(&s_tab[v * width + c], g);
    else atomicAdd(&dtable[(int64_t)v * width + c], g);
  }
  if (use_smem) {
    __syncthreads();
    for (int i = threadIdx.x; i < tsz; i += blockDim.x) {
      const float v = s_tab[i];
      if (v != 0.f) atomicAdd(&dtable[i], v);
    }
  }
}

// ================================================================================================
// softmax cross-entropy (+z-loss, accuracy), forward and backward: one warp per token
// ================================================================================================
template <typename T, int MAXI>
__global__ void __launch_bounds__(256) xent_kernel(const T* logits, const int32_t* __restrict__ targets, T* dlogits,
                                                    double* __restrict__ sums, int64_t tokens, int vocab, float inv_n,
                                                    float z_loss) {
  // logits / dlogits are NOT __restrict__: the model calls this in place (dlogits == logits), so everything that is
  // needed of a row -- including the target's logit -- is read before the row's gradient is stored.
  __shared__ double s_acc[3];
  if (threadIdx.x < 3) s_acc[threadIdx.x] = 0.0;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int wpb = blockDim.x >> 5;
  double a_nll = 0.0, a_z = 0.0, a_hit = 0.0;
  for (int64_t t = (int64_t)blockIdx.x * wpb + (threadIdx.x >> 5); t < tokens; t += (int64_t)gridDim.x * wpb) {
    const T* lr = logits + t * vocab;
    float v[MAXI];
    float mx = -INFINITY;
    int arg = 0x7fffffff;
#pragma unroll
    for (int i = 0; i < MAXI; ++i) {
      const int c = lane + 32 * i;
      v[i] = (c < vocab) ? to_f32(lr[c]) : -INFINITY;
      if (v[i] > mx) {  // strict: keeps the first index on ties (tf.argmax)
        mx = v[i];
        arg = c;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, mx, o);
      const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
      if (om > mx || (om == mx && oa < arg)) {
        mx = om;
        arg = oa;
      }
    }
    float se = 0.f;
#pragma unroll
    for (int i = 0; i < MAXI; ++i) {
      const int c = lane + 32 * i;
      if (c < vocab) {
        v[i] = expf(v[i] - mx);
        se += v[i];
      }
    }
    se = warp_sum(se);
    const float logz = logf(se) + mx;
    int tgt = targets[t];
    tgt = tgt < 0 ? 0 : (tgt >= vocab ? vocab - 1 : tgt);
    const float lt = to_f32(lr[tgt]);  // before the stores below: dlogits may alias logits
    __syncwarp();                      // ... by ANY lane of the warp that owns this row
    const float inv_se = 1.0f / se;
    const float zc = 2.0f * z_loss * inv_n * logz;
    T* dr = dlogits + t * vocab;
#pragma unroll
    for (int i = 0; i < MAXI; ++i) {
      const int c = lane + 32 * i;
      if (c < vocab) {
        const float p = v[i] * inv_se;
        const float d = (p - (c == tgt ? 1.f : 0.f)) * inv_n + zc * p;
        dr[c] = from_f32<T>(d);
      }
    }
    if (lane == 0) {
      a_nll += (double)(logz - lt);
      a_z += (double)logz * (double)logz;
      a_hit += (arg == tgt) ? 1.0 : 0.0;
    }
  }
  if (lane == 0) {
    atomicAdd(&s_acc[0], a_nll);
    atomicAdd(&s_acc[1], a_z);
    atomicAdd(&s_acc[2], a_hit);
  }
  __syncthreads();
  if (threadIdx.x < 3 && s_acc[threadIdx.x] != 0.0) atomicAdd(&sums[threadIdx.x], s_acc[threadIdx.x]);
}

// ================================================================================================
// parameter shadow cast (fp32 slice -> activation dtype, optional causal mask)
// ================================================================================================
template <typename T>
__global__ void __launch_bounds__(256) cast_param_kernel(const float* __restrict__ src, T* __restrict__ dst, int64_t n,
                                                          int tril_s) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    float v = src[i];
    if (tril_s > 0) {
      const int col = (int)(i % tril_s);
      const int row = (int)((i / tril_s) % tril_s);
      if (col > row) v = 0.f;
    }
    dst[i] = from_f32<T>(v);
  }
}

// ================================================================================================
// transposed parameter shadow: dst[b, c, r] = src[b, r, c].  The tcgen05 GEMM reads a K-major B operand
// 12-17 % faster than an MN-major one (measured), and the weights are tiny next to the activations, so
// every weight that is the B operand of a forward GEMM keeps a transposed copy, rewritten once per step.
// ================================================================================================
template <typename T>
__global__ void __launch_bounds__(256) transpose_kernel(const T* __restrict__ src, T* __restrict__ dst, int rows,
                                                         int cols) {
  __shared__ T tile[32][33];
  const int64_t mat = (int64_t)blockIdx.z * rows * cols;
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < 32; i += 8) {
    const int r = r0 + ty + i, c = c0 + tx;
    if (r < rows && c < cols) tile[ty + i][tx] = src[mat + (int64_t)r * cols + c];
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 32; i += 8) {
    const int c = c0 + ty + i, r = r0 + tx;
    if (r < rows && c < cols) dst[mat + (int64_t)c * rows + r] = tile[tx][ty + i];
  }
}

}  // namespace obst

using namespace obst;

extern "C" {

int obst_gn_fwd(int dtype, const void* x, const void* gamma, const void* beta, void* y, float* stats, int64_t rows,
                int32_t heads, int32_t width, int32_t act, void* stream) {
  OBST_REQUIRE(rows >= 0 && heads > 0 && width > 0, "obst_gn_fwd: bad shape");
  if (rows == 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  float2* s2 = reinterpret_cast<float2*>(stats);
  if (dtype == OBST_BF16) {
    static const bool ring_off = getenv("OBST_GN_FWD_RING") && atoi(getenv("OBST_GN_FWD_RING")) == 0;
    if (!ring_off) {
      const int rc = gn_fwd_ring_launch(x, gamma, beta, y, s2, rows, heads, width, act, st);
      if (rc != 1) return rc;
    }
    return gn_fwd_launch<bf16>(x, gamma, beta, y, s2, rows, heads, width, act, st);
  }
  if (dtype == OBST_F32) return gn_fwd_launch<float>(x, gamma, beta, y, s2, rows, heads, width, act, st);
  OBST_REQUIRE(false, "obst_gn_fwd: bad dtype %d", dtype);
}

int obst_gn_bwd(int dtype, const void* dy, const void* x, const void* gamma, const void* beta, void* dx,
                float* dgamma, float* dbeta, const float* stats, int64_t rows, int32_t heads, int32_t width,
                int32_t act, int32_t relu_mask, int32_t dx_accumulate, void* stream) {
  OBST_REQUIRE(rows >= 0 && heads > 0 && width > 0, "obst_gn_bwd: bad shape");
  if (rows == 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const float2* s2 = reinterpret_cast<const float2*>(stats);
  if (dtype == OBST_BF16) {
    static const bool ring_off = getenv("OBST_GN_BWD_RING") && atoi(getenv("OBST_GN_BWD_RING")) == 0;
    if (!ring_off) {
      const int rc = gn_bwd_ring_launch(dy, x, gamma, beta, dx, dgamma, dbeta, s2, rows, heads, width, act, relu_mask,
                                        dx_accumulate, st);
      if (rc != 1) return rc;
    }
    return gn_bwd_launch<bf16>(dy, x, gamma, beta, dx, dgamma, dbeta, s2, rows, heads, width, act, relu_mask,
                               dx_accumulate, st);
  }
  if (dtype == OBST_F32)
    return gn_bwd_launch<float>(dy, x, gamma, beta, dx, dgamma, dbeta, s2, rows, heads, width, act, relu_mask,
                                dx_accumulate, st);
  OBST_REQUIRE(false, "obst_gn_bwd: bad dtype %d", dtype);
}

int obst_act_fwd(int dtype, int32_t kind, const void* x, void* y, int64_t n, void* stream) {
  OBST_REQUIRE(kind >= 0 && kind <= OBST_ACT_EXP, "obst_act_fwd: bad kind %d", kind);
  if (n <= 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (kind >= 2) return act_any_launch(dtype, kind, false, x, nullptr, y, n, st);
  if (dtype == OBST_BF16) {
    const unsigned b = ew_blocks(n / 8 + 1);
    if (kind == 0) act_kernel<bf16, 0, false><<<b, 256, 0, st>>>((const bf16*)x, nullptr, (bf16*)y, n);
    else act_kernel<bf16, 1, false><<<b, 256, 0, st>>>((const bf16*)x, nullptr, (bf16*)y, n);
  } else if (dtype == OBST_F32) {
    const unsigned b = ew_blocks(n / 4 + 1);
    if (kind == 0) act_kernel<float, 0, false><<<b, 256, 0, st>>>((const float*)x, nullptr, (float*)y, n);
    else act_kernel<float, 1, false><<<b, 256, 0, st>>>((const float*)x, nullptr, (float*)y, n);
  } else {
    OBST_REQUIRE(false, "obst_act_fwd: bad dtype %d", dtype);
  }
  count_launch();
  return check_launch("obst_act_fwd");
}

int obst_act_bwd(int dtype, int32_t kind, const void* x, const void* dy, void* dx, int64_t n, void* stream) {
  OBST_REQUIRE(kind >= 0 && kind <= OBST_ACT_EXP, "obst_act_bwd: bad kind %d", kind);
  if (n <= 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (kind >= 2) return act_any_launch(dtype, kind, true, x, dy, dx, n, st);
  if (dtype == OBST_BF16) {
    const unsigned b = ew_blocks(n / 8 + 1);
    if (kind == 0) act_kernel<bf16, 0, true><<<b, 256, 0, st>>>((const bf16*)x, (const bf16*)dy, (bf16*)dx, n);
    else act_kernel<bf16, 1, true><<<b, 256, 0, st>>>((const bf16*)x, (const bf16*)dy, (bf16*)dx, n);
  } else if (dtype == OBST_F32) {
    const unsigned b = ew_blocks(n / 4 + 1);
    if (kind == 0) act_kernel<float, 0, true><<<b, 256, 0, st>>>((const float*)x, (const float*)dy, (float*)dx, n);
    else act_kernel<float, 1, true><<<b, 256, 0, st>>>((const float*)x, (const float*)dy, (float*)dx, n);
  } else {
    OBST_REQUIRE(false, "obst_act_bwd: bad dtype %d", dtype);
  }
  count_launch();
  return check_launch("obst_act_bwd");
}

int obst_axpby(int dtype, const void* a, const void* b, float sign, void* out, int64_t n, void* stream) {
  if (n <= 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == OBST_BF16)
    axpby_kernel<bf16><<<ew_blocks(n / 8 + 1), 256, 0, st>>>((const bf16*)a, (const bf16*)b, sign, (bf16*)out, n);
  else if (dtype == OBST_F32)
    axpby_kernel<float><<<ew_blocks(n / 4 + 1), 256, 0, st>>>((const float*)a, (const float*)b, sign, (float*)out, n);
  else
    OBST_REQUIRE(false, "obst_axpby: bad dtype %d", dtype);
  count_launch();
  return check_launch("obst_axpby");
}

int obst_embed_gather(int dtype, const int32_t* idx, const void* table, void* out, int64_t tokens, int32_t vocab,
                      int32_t width, void* stream) {
  OBST_REQUIRE(vocab > 0 && width > 0, "obst_embed_gather: bad shape");
  if (tokens <= 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned b = ew_blocks(tokens * width);
  if (dtype == OBST_BF16)
    gather_kernel<bf16><<<b, 256, 0, st>>>(idx, (const bf16*)table, (bf16*)out, tokens, vocab, width);
  else if (dtype == OBST_F32)
    gather_kernel<float><<<b, 256, 0, st>>>(idx, (const float*)table, (float*)out, tokens, vocab, width);
  else
    OBST_REQUIRE(false, "obst_embed_gather: bad dtype %d", dtype);
  count_launch();
  return check_launch("obst_embed_gather");
}

int obst_embed_scatter_add(int dtype, const int32_t* idx, const void* dout, float* dtable, int64_t tokens,
                           int32_t vocab, int32_t width, void* stream) {
  OBST_REQUIRE(vocab > 0 && width > 0, "obst_embed_scatter_add: bad shape");
  if (tokens <= 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t smem = (size_t)vocab * width * sizeof(float);
  const int use_smem = smem <= 160 * 1024 ? 1 : 0;
  int tpb = 1024;
  while (tpb > 32 && (tokens + tpb - 1) / tpb < num_sms()) tpb >>= 1;
  const unsigned blocks = (unsigned)((tokens + tpb - 1) / tpb);
  if (dtype == OBST_BF16) {
    if (use_smem && smem > 48 * 1024)
      cudaFuncSetAttribute(scatter_kernel<bf16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    scatter_kernel<bf16><<<blocks, 256, use_smem ? smem : 0, st>>>(idx, (const bf16*)dout, dtable, tokens, vocab,
                                                                    width, tpb, use_smem);
  } else if (dtype == OBST_F32) {
    if (use_smem && smem > 48 * 1024)
      cudaFuncSetAttribute(scatter_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    scatter_kernel<float><<<blocks, 256, use_smem ? smem : 0, st>>>(idx, (const float*)dout, dtable, tokens, vocab,
                                                                     width, tpb, use_smem);
  } else {
    OBST_REQUIRE(false, "obst_embed_scatter_add: bad dtype %d", dtype);
  }
  count_launch();
  return check_launch("obst_embed_scatter_add");
}

int obst_xent_fwd_bwd(int dtype, const void* logits, const int32_t* targets, void* dlogits, double* sums,
                      int64_t tokens, int32_t vocab, float inv_n, float z_loss, void* stream) {
  OBST_REQUIRE(vocab > 0 && vocab <= 1024, "obst_xent_fwd_bwd: vocab %d unsupported (1..1024)", vocab);
  if (tokens <= 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  int64_t blocks = (tokens + 7) / 8;
  const int64_t cap = (int64_t)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  const int per_lane = (vocab + 31) / 32;
#define OBST_XENT(T, MI)                                                                                        \
  xent_kernel<T, MI><<<(unsigned)blocks, 256, 0, st>>>((const T*)logits, targets, (T*)dlogits, sums, tokens, \
                                                        vocab, inv_n, z_loss)
  if (dtype == OBST_BF16) {
    if (per_lane <= 8) OBST_XENT(bf16, 8);
    else OBST_XENT(bf16, 32);
  } else if (dtype == OBST_F32) {
    if (per_lane <= 8) OBST_XENT(float, 8);
    else OBST_XENT(float, 32);
  } else {
    OBST_REQUIRE(false, "obst_xent_fwd_bwd: bad dtype %d", dtype);
  }
#undef OBST_XENT
  count_launch();
  return check_launch("obst_xent_fwd_bwd");
}

int obst_cast_param(int dtype_dst, const float* src, void* dst, int64_t n, int32_t tril_s, void* stream) {
  if (n <= 0) return OBST_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned b = ew_blocks(n);
  if (dtype_dst == OBST_BF16) cast_param_kernel<bf16><<<b, 256, 0, st>>>(src, (bf16*)dst, n, tril_s);
  else if (dtype_dst == OBST_F32) cast_param_kernel<float><<<b, 256, 0, st>>>(src, (float*)dst, n, tril_s);
  else OBST_REQUIRE(false, "obst_cast_param: bad dtype %d", dtype_dst);
  count_launch();
  return check_launch("obst_cast_param");
}

int obst_transpose_param(int dtype, const void* src, void* dst, int64_t batch, int32_t rows, int32_t cols,
                         void* stream) {
  OBST_REQUIRE(batch >= 0 && rows >= 0 && cols >= 0 && batch <= 65535, "obst_transpose_param: bad shape");
  if (batch == 0 || rows == 0 || cols == 0) return OBST_OK;
  OBST_REQUIRE(src != dst, "obst_transpose_param: in-place transpose is not supported");
  cudaStream_t st = (cudaStream_t)stream;
  const dim3 grid((unsigned)((cols + 31) / 32), (unsigned)((rows + 31) / 32), (unsigned)batch);
  OBST_REQUIRE(grid.y <= 65535, "obst_transpose_param: too many rows (%d)", rows);
  if (dtype == OBST_BF16) transpose_kernel<bf16><<<grid, 256, 0, st>>>((const bf16*)src, (bf16*)dst, rows, cols);
  else if (dtype == OBST_F32) transpose_kernel<float><<<grid, 256, 0, st>>>((const float*)src, (float*)dst, rows, cols);
  else OBST_REQUIRE(false, "obst_transpose_param: bad dtype %d", dtype);
  count_launch();
  return check_launch("obst_transpose_param");
}

}  // extern "C"
