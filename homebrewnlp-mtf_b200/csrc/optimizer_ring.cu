// Streaming form of the fused optimizer update (last dims that are multiples of 4), and the batched transposed-shadow
// refresh.
//
// The update is pure HBM traffic: per parameter it reads the master weight (4 B), the gradient (2 or 4 B) and the
// momentum (2 or 4 B) and writes the master, the momentum and the activation-dtype shadow.  The generic kernel
// (optimizer.cu) loads a row, computes, stores, and only then asks for the next row: at 128 registers per thread an
// SM holds 16 such warps and the memory system idles between rows (measured 3.7 TB/s).  Here every warp streams
// its rows through a private cp.async ring in shared memory -- two rows are always in flight while a third is being
// updated, without spending registers on them -- the same scheme that brought the group-norm forward kernel to the
// copy rate.
//
// A tensor is viewed as [rows, cols] (cols = last dim); a block takes one work item (a run of rows) and one chunk
// of 512 columns (block index = item * chunks + chunk); a warp owns whole rows of that chunk; lane l owns columns 128 g + 4 l + {0..3} of
// the chunk for g = 0..3, so every global access of a warp is one fully used, contiguous 512-byte (fp32) or
// 256-byte (bf16) run.  SM3 bookkeeping as in the generic kernel: the maxima over rows of the last dim's new
// accumulator live in registers, the maxima over columns go through one warp reduction per row.
#include "optimizer_math.cuh"

namespace obst {

static constexpr int RG = 4;                  // 128-column groups per chunk
static constexpr int RCHUNK = 128 * RG;       // columns per chunk
static constexpr int RWARPS = 8;              // warps per block

template <int BYTES>
__device__ __forceinline__ void cp_async(void* smem_dst, const void* gmem_src) {
  if (BYTES == 16)
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)),
                 "l"(gmem_src)
                 : "memory");
  else
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)),
                 "l"(gmem_src)
                 : "memory");
}
__device__ __forceinline__ void cp_commit_group() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_wait_group() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// four consecutive values of a staged row part (fp32: one 16-byte word; bf16: one 8-byte word)
__device__ __forceinline__ void ld4(const float* s, float (&v)[4]) {
  const float4 t = *reinterpret_cast<const float4*>(s);
  v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
__device__ __forceinline__ void ld4(const bf16* s, float (&v)[4]) {
  const uint2 t = *reinterpret_cast<const uint2*>(s);
  const float2 a = unpack_bf16x2(t.x), b = unpack_bf16x2(t.y);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}
__device__ __forceinline__ void st4(float* d, const float (&v)[4]) {
  *reinterpret_cast<float4*>(d) = make_float4(v[0], v[1], v[2], v[3]);
}
__device__ __forceinline__ void st4(bf16* d, const float (&v)[4]) {
  uint2 t;
  t.x = pack_bf16x2(v[0], v[1]);
  t.y = pack_bf16x2(v[2], v[3]);
  *reinterpret_cast<uint2*>(d) = t;
}

// TS = state dtype (momentum / accumulators), TG = gradient storage dtype, RING = rows in flight per warp + 1,
// ACT_BF16 = the graph reads the weights through a bf16 cast (else fp32)
template <typename TS, typename TG, int RING, bool ACT_BF16>
__global__ void __launch_bounds__(RWARPS * 32, 2) opt_update_ring_kernel(const obst_opt_tensor* __restrict__ table,
                                                                       const int32_t* __restrict__ item_tensor,
                                                                       const int32_t* __restrict__ item_row0,
                                                                       const int32_t* __restrict__ item_rows,
                                                                       const double* __restrict__ norms,
                                                                       const OptHyper hp, double* __restrict__ wnext,
                                                                       const unsigned ychunks) {
  constexpr int P_BYTES = RCHUNK * 4, G_BYTES = RCHUNK * (int)sizeof(TG), M_BYTES = RCHUNK * (int)sizeof(TS);
  constexpr int STAGE = P_BYTES + G_BYTES + M_BYTES;
  extern __shared__ __align__(16) uint8_t s_ring_raw[];
  __shared__ float s_col[RCHUNK];  // block-level column maxima of this chunk
  __shared__ double s_w2[RWARPS];
  __shared__ float s_rowmax[OPT_MAX_ITEM_ROWS];
  // item-major block order (the two 512-column chunks of a wide row run are neighbours): blocks are dispatched in
  // work-list order, so a list that ends in small items drains evenly
  const int item = (int)(blockIdx.x / ychunks);
  const int ti = item_tensor[item];
  const int nrows = item_rows[item];
  const int row_begin = item_row0[item];
  if (nrows > OPT_MAX_ITEM_ROWS) __trap();  // contract of the work list (include/obst_b200.h)
  // the table entry is read where it lies (read-only, L1/L2-resident after the first blocks): its fields arrive with one
  // memory latency for all threads, where a staged copy in shared memory would cost a latency plus a barrier
  const obst_opt_tensor& T = table[ti];
  if ((sizeof(TG) == 2) != (T.g_bf16 != 0)) __trap();  // the launch's gradient dtype must be the table's
  const int nd = T.ndim;
  const int cols = T.dims[nd - 1];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cbase = (int)(blockIdx.x % ychunks) * RCHUNK;
  if (cbase >= cols) return;  // whole block: this tensor is narrower than the widest one of the launch
  for (int i = threadIdx.x; i < RCHUNK; i += blockDim.x) s_col[i] = 0.f;
  // the table entry's hot fields in registers
  float* const Tp = T.p;
  const TG* const Tg = reinterpret_cast<const TG*>(T.g);
  TS* const Tm = reinterpret_cast<TS*>(T.m);
  void* const Tshadow = T.shadow;
  const int tril_s = T.tril_s;
  const bool wd_on = T.decay && hp.weight_decay > 0.f;
  const bool preview = hp.preview != 0;
  const float scale = hp.clip > 0.f ? clip_scale(norms[2 * ti], norms[2 * ti + 1], hp.clip) : 1.f;
  const bool tril_fast = tril_s > 0 && tril_s == cols;  // [.., S, S] map: col = c + j, row = r % S

  // which of my four column groups exist, and the last dim's accumulator for my columns
  bool live[RG];
  float al[RG][4], cm[RG][4];
#pragma unroll
  for (int k = 0; k < RG; ++k) {
    const int c = cbase + k * 128 + lane * 4;
    live[k] = c < cols;  // cols % 4 == 0: a group of four is in or out as a whole
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      cm[k][j] = 0.f;
      al[k][j] = live[k] ? to_f32(reinterpret_cast<const TS*>(T.acc[nd - 1])[c + j]) : 0.f;
    }
  }

  uint8_t* const ring = s_ring_raw + (size_t)warp * RING * STAGE;
  auto issue = [&](int r, int stage) {
    const int64_t rbase = (int64_t)r * cols + cbase + lane * 4;
    uint8_t* const dst = ring + stage * STAGE;
#pragma unroll
    for (int k = 0; k < RG; ++k) {
      if (!live[k]) continue;
      const int64_t e = rbase + k * 128;
      cp_async<16>(dst + k * 512 + lane * 16, Tp + e);
      if (sizeof(TG) == 4) cp_async<16>(dst + P_BYTES + k * 512 + lane * 16, Tg + e);
      else cp_async<8>(dst + P_BYTES + k * 256 + lane * 8, Tg + e);
      if (sizeof(TS) == 4) cp_async<16>(dst + P_BYTES + G_BYTES + k * 512 + lane * 16, Tm + e);
      else cp_async<8>(dst + P_BYTES + G_BYTES + k * 256 + lane * 8, Tm + e);
    }
    cp_commit_group();
  };

  const int row_end = row_begin + nrows;
  double w2 = 0.0;  // sum cast_AT(p_new)^2 of this thread: next step's weight norm, for free
#pragma unroll
  for (int s = 0; s < RING - 1; ++s) {
    if (row_begin + warp + s * RWARPS < row_end) issue(row_begin + warp + s * RWARPS, s);
    else cp_commit_group();  // keep one group per slot so the wait_group counts stay aligned
  }
  // the rows are on their way: only now the per-row minima of the other dims' accumulators (their loads overlap the
  // ring's first fill instead of preceding it)
  opt_row_minima<TS>(T, row_begin, nrows, s_rowmax);
  __syncthreads();
  int stage = 0;
  for (int r = row_begin + warp; r < row_end; r += RWARPS, stage = (stage + 1 == RING ? 0 : stage + 1)) {
    const int rn = r + (RING - 1) * RWARPS;
    if (rn < row_end) issue(rn, stage == 0 ? RING - 1 : stage - 1);
    else cp_commit_group();
    cp_wait_group<RING - 1>();
    __syncwarp();
    const uint8_t* const src = ring + stage * STAGE;
    const float rowmin = s_rowmax[r - row_begin];
    float rowmax = 0.f;
    const int64_t rbase = (int64_t)r * cols + cbase + lane * 4;
    const int row_in = tril_fast ? (r % cols) : 0;
#pragma unroll
    for (int k = 0; k < RG; ++k) {
      if (!live[k]) continue;
      float pv[4], g[4], mv[4], pn[4], mn[4], u[4];
      ld4(reinterpret_cast<const float*>(src + k * 512) + lane * 4, pv);
      ld4(reinterpret_cast<const TG*>(src + P_BYTES) + k * 128 + lane * 4, g);
      ld4(reinterpret_cast<const TS*>(src + P_BYTES + G_BYTES) + k * 128 + lane * 4, mv);
      const int64_t e = rbase + k * 128;
      // Every per-launch / per-tensor option below is a warp-uniform branch around a 4-element loop, never a test
      // inside one: the kernel is instruction-issue bound (about 40 instructions per element remain, half of them the
      // two correctly rounded special functions).
      if (sizeof(TG) == 4 && hp.grad_bf16) {
#pragma unroll
        for (int j = 0; j < 4; ++j) g[j] = round_to<bf16>(g[j]);
      }
      float nu[4], inv[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        g[j] = __fmul_rn(g[j], scale);
        nu[j] = __fadd_rn(fminf(rowmin, al[k][j]), __fmul_rn(g[j], g[j]));           // optimizers.py:70
        cm[k][j] = fmaxf(cm[k][j], nu[j]);
        rowmax = fmaxf(rowmax, nu[j]);
      }
      opt_rsqrt4(nu, inv);                                                           // opt_rsqrt, :13-14
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        u[j] = __fmul_rn(g[j], inv[j]);                                              // :76
        mn[j] = __fadd_rn(__fmul_rn(hp.beta, mv[j]), u[j]);                          // :124
      }
      if (hp.nesterov) {
#pragma unroll
        for (int j = 0; j < 4; ++j) u[j] = __fadd_rn(u[j], __fmul_rn(hp.beta, mn[j]));  // :126-128
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) u[j] = mn[j];
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) u[j] = __fmul_rn(u[j], hp.lr);                     // :114-115
      if (wd_on) {                                                                   // __init__.py:59-61
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float w = ACT_BF16 ? round_to<bf16>(pv[j]) : pv[j];
          u[j] = __fadd_rn(u[j], __fmul_rn(__fmul_rn(w, hp.lr), hp.weight_decay));
        }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) pn[j] = __fsub_rn(pv[j], u[j]);                    // assign_sub, __init__.py:64
      // the shadow: the new master in the activation dtype; its (unmasked) bf16 values also feed the weight norm
      if (ACT_BF16) {
        uint2 sb;
        sb.x = pack_bf16x2(pn[0], pn[1]);
        sb.y = pack_bf16x2(pn[2], pn[3]);
        if (!preview) {
          const float w0 = __uint_as_float(sb.x << 16), w1 = __uint_as_float(sb.x & 0xFFFF0000u);
          const float w2a = __uint_as_float(sb.y << 16), w3 = __uint_as_float(sb.y & 0xFFFF0000u);
          // squares of bf16 values are exact in fp32; they are summed in float64
          w2 += (double)__fmul_rn(w0, w0);
          w2 += (double)__fmul_rn(w1, w1);
          w2 += (double)__fmul_rn(w2a, w2a);
          w2 += (double)__fmul_rn(w3, w3);
        }
        if (tril_s > 0) {  // causal mask of a mixing map: entries above the diagonal read as zero
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            bool masked;
            if (tril_fast) {
              masked = cbase + k * 128 + lane * 4 + j > row_in;
            } else {
              const int64_t ee = e + j;
              masked = (int)(ee % tril_s) > (int)((ee / tril_s) % tril_s);
            }
            if (masked) {
              if (j == 0) sb.x &= 0xFFFF0000u;
              if (j == 1) sb.x &= 0x0000FFFFu;
              if (j == 2) sb.y &= 0xFFFF0000u;
              if (j == 3) sb.y &= 0x0000FFFFu;
            }
          }
        }
        if (Tshadow) *reinterpret_cast<uint2*>(reinterpret_cast<bf16*>(Tshadow) + e) = sb;
      } else {
        float sh[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          sh[j] = pn[j];
          if (!preview) w2 += (double)__fmul_rn(pn[j], pn[j]);
          if (tril_s > 0) {
            const int64_t ee = e + j;
            if ((int)(ee % tril_s) > (int)((ee / tril_s) % tril_s)) sh[j] = 0.f;
          }
        }
        if (Tshadow) st4(reinterpret_cast<float*>(Tshadow) + e, sh);
      }
      if (!preview) {
        st4(Tp + e, pn);
        st4(Tm + e, mn);
      }
    }
    rowmax = warp_max(rowmax);
    if (lane == 0) s_rowmax[r - row_begin] = rowmax;
    __syncwarp();  // everyone is done with this stage before the next iteration refills it
  }
  __syncthreads();
  if (preview) return;
  opt_row_maxima_commit(T, row_begin, nrows, s_rowmax);
#pragma unroll
  for (int k = 0; k < RG; ++k)
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (live[k] && cm[k][j] > 0.f) atomic_max_nonneg(&s_col[k * 128 + lane * 4 + j], cm[k][j]);
  if (wnext) {
    const double dw = warp_sum_d(w2);
    if (lane == 0) s_w2[warp] = dw;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < RCHUNK && cbase + i < cols; i += blockDim.x) {
    const float v = s_col[i];
    if (v > 0.f) atomic_max_nonneg(T.acc_new[nd - 1] + cbase + i, v);
  }
  if (wnext && threadIdx.x == 0) {
    double a = 0.0;
    for (int w = 0; w < RWARPS; ++w) a += s_w2[w];
    atomicAdd(&wnext[ti], a);
  }
}

template <typename TS, typename TG>
static int ring_launch_typed(const obst_opt_tensor* table, const int32_t* item_tensor, const int32_t* item_row0,
                             const int32_t* item_rows, int nitems, int max_cols, const double* norms,
                             const OptHyper& hp, double* wnext, cudaStream_t st) {
  // three rows per warp in the ring when two blocks still fit an SM (16-bit gradient and state: 96 KB), else two
  constexpr int STAGE = RCHUNK * (4 + (int)sizeof(TG) + (int)sizeof(TS));
  constexpr int RING = (STAGE * 3 * RWARPS <= 100 * 1024) ? 3 : 2;
  constexpr int SMEM = STAGE * RING * RWARPS;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(opt_update_ring_kernel<TS, TG, RING, true>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(opt_update_ring_kernel<TS, TG, RING, false>,
                               cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess) {
      set_error("obst_opt_update: cannot opt in to %d B of shared memory: %s", SMEM, cudaGetErrorString(e));
      return OBST_ECUDA;
    }
    attr = true;
  }
  const int64_t ychunks = ((int64_t)max_cols + RCHUNK - 1) / RCHUNK;
  OBST_REQUIRE(ychunks * (int64_t)nitems < (int64_t)1 << 31, "obst_opt_update: %d work items x %lld column chunks "
               "exceed the grid", nitems, (long long)ychunks);
  const dim3 grid((unsigned)(nitems * ychunks));
  if (hp.act_bf16)
    opt_update_ring_kernel<TS, TG, RING, true><<<grid, RWARPS * 32, SMEM, st>>>(table, item_tensor, item_row0,
                                                                                 item_rows, norms, hp, wnext,
                                                                                 (unsigned)ychunks);
  else
    opt_update_ring_kernel<TS, TG, RING, false><<<grid, RWARPS * 32, SMEM, st>>>(table, item_tensor, item_row0,
                                                                                  item_rows, norms, hp, wnext,
                                                                                  (unsigned)ychunks);
  count_launch();
  count_path(OBST_PATH_OPT_RING);
  return check_launch("obst_opt_update");
}

// All tensors of a launch share one gradient storage dtype: the ring is laid out for it.  The table says which per
// tensor; the host builds tables that are uniform, and a mixed table is refused by the kernel's caller below.
int opt_update_ring_launch(const obst_opt_tensor* table, const int32_t* item_tensor, const int32_t* item_row0,
                           const int32_t* item_rows, int nitems, int max_cols, int dtype_state, const double* norms,
                           const OptHyper& hp, double* wnext, cudaStream_t st) {
  // hp.grad_bf16 == 2 is the launch-level statement "g is stored as bf16" (set by obst_opt_update from the flag the
  // caller passes; the per-tensor g_bf16 field is what the generic kernels read)
  const bool g16 = hp.grad_bf16 == 2;
  if (dtype_state == OBST_BF16) {
    if (g16) return ring_launch_typed<bf16, bf16>(table, item_tensor, item_row0, item_rows, nitems, max_cols, norms, hp, wnext, st);
    return ring_launch_typed<bf16, float>(table, item_tensor, item_row0, item_rows, nitems, max_cols, norms, hp, wnext, st);
  }
  if (g16) return ring_launch_typed<float, bf16>(table, item_tensor, item_row0, item_rows, nitems, max_cols, norms, hp, wnext, st);
  return ring_launch_typed<float, float>(table, item_tensor, item_row0, item_rows, nitems, max_cols, norms, hp, wnext, st);
}

// ------------------------------------------------------------------------------------------------
// every transposed weight shadow in one launch: one block per 64 x 64 tile of any of the matrices
// ------------------------------------------------------------------------------------------------
// 16-bit elements move as 16-byte words on both sides (a 64-wide tile row is one full 128-byte line); other cases
// (fp32, odd sizes, unaligned bases) take the element-wise path of the same kernel.
static constexpr int TT = 64;

template <typename T>
__global__ void __launch_bounds__(256) transpose_batched_kernel(const obst_transpose_item* __restrict__ items,
                                                                 int nitems) {
  __shared__ T tile[TT][TT + 2];
  // which item owns this tile: binary search in the exclusive prefix sums
  int lo = 0, hi = nitems - 1;
  const int tile_id = blockIdx.x;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (items[mid].tile0 <= tile_id) lo = mid;
    else hi = mid - 1;
  }
  const obst_transpose_item it = items[lo];
  const int tiles_c = (it.cols + TT - 1) / TT, tiles_r = (it.rows + TT - 1) / TT;
  int local = tile_id - it.tile0;
  const int tc = local % tiles_c;
  local /= tiles_c;
  const int tr = local % tiles_r;
  const int b = local / tiles_r;
  if (b >= it.batch) return;
  const T* __restrict__ src = reinterpret_cast<const T*>(it.src) + (int64_t)b * it.rows * it.cols;
  T* __restrict__ dst = reinterpret_cast<T*>(it.dst) + (int64_t)b * it.rows * it.cols;
  const int c0 = tc * TT, r0 = tr * TT;
  const bool fast = sizeof(T) == 2 && (it.rows % 8 == 0) && (it.cols % 8 == 0) &&
                    ((reinterpret_cast<uintptr_t>(it.src) | reinterpret_cast<uintptr_t>(it.dst)) & 15) == 0;
  if (fast) {
    const int chunk = threadIdx.x & 7, rr = threadIdx.x >> 3;  // 8 chunks of 8 elements per tile row, 32 rows a pass
#pragma unroll
    for (int pass = 0; pass < 2; ++pass) {
      const int r = r0 + rr + 32 * pass, c = c0 + chunk * 8;
      if (r < it.rows && c < it.cols) {
        const uint4 v = *reinterpret_cast<const uint4*>(src + (int64_t)r * it.cols + c);
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
        uint16_t* trow = reinterpret_cast<uint16_t*>(&tile[rr + 32 * pass][chunk * 8]);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          trow[2 * j] = (uint16_t)(w[j] & 0xFFFFu);
          trow[2 * j + 1] = (uint16_t)(w[j] >> 16);
        }
      }
    }
    __syncthreads();
#pragma unroll
    for (int pass = 0; pass < 2; ++pass) {
      const int c = c0 + rr + 32 * pass, r = r0 + chunk * 8;  // output row c, 8 consecutive source rows
      if (c < it.cols && r < it.rows) {
        uint32_t w[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint16_t a = *reinterpret_cast<const uint16_t*>(&tile[chunk * 8 + 2 * j][rr + 32 * pass]);
          const uint16_t bb = *reinterpret_cast<const uint16_t*>(&tile[chunk * 8 + 2 * j + 1][rr + 32 * pass]);
          w[j] = (uint32_t)a | ((uint32_t)bb << 16);
        }
        *reinterpret_cast<uint4*>(dst + (int64_t)c * it.rows + r) = make_uint4(w[0], w[1], w[2], w[3]);
      }
    }
    return;
  }
  for (int i = threadIdx.x; i < TT * TT; i += blockDim.x) {
    const int r = r0 + i / TT, c = c0 + i % TT;
    if (r < it.rows && c < it.cols) tile[i / TT][i % TT] = src[(int64_t)r * it.cols + c];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < TT * TT; i += blockDim.x) {
    const int c = c0 + i / TT, r = r0 + i % TT;
    if (r < it.rows && c < it.cols) dst[(int64_t)c * it.rows + r] = tile[i % TT][i / TT];
  }
}

}  // namespace obst

using namespace obst;

extern "C" int obst_transpose_batched(int dtype, const obst_transpose_item* items, int32_t nitems, int32_t total_tiles,
                                      void* stream) {
  OBST_REQUIRE(nitems >= 0 && total_tiles >= 0, "obst_transpose_batched: bad sizes");
  if (nitems == 0 || total_tiles == 0) return OBST_OK;
  OBST_REQUIRE(items != nullptr, "obst_transpose_batched: null item table");
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == OBST_BF16) transpose_batched_kernel<bf16><<<(unsigned)total_tiles, 256, 0, st>>>(items, nitems);
  else if (dtype == OBST_F32) transpose_batched_kernel<float><<<(unsigned)total_tiles, 256, 0, st>>>(items, nitems);
  else OBST_REQUIRE(false, "obst_transpose_batched: bad dtype %d", dtype);
  count_launch();
  return check_launch("obst_transpose_batched");
}
