// tcgen05 / TMEM / TMA implementation of the obst_gemm descriptor for bf16 operands (sm_100a).
//
// One persistent, warp-specialised kernel (352 threads, one CTA per SM) serves every contraction of the mixer step.
// Default: CTA PAIRS (cta_group::2, cluster (2,1,1)): one 256-row x BN tile per pair, each CTA stages its 128 rows
// of A and half of B, the leader's MMA thread issues 256 x BN x 16 UMMAs over both CTAs' shared memory.
//   warp 0      : TMA producer for A (cp.async.bulk.tensor, 5-D maps, 128B swizzle, 5-stage mbarrier ring)
//   warp 10     : TMA producer for B (an MN-major operand needs one box per 64 elements: its own warp)
//   warp 1      : TMEM allocator + single-thread tcgen05.mma issuer (fp32 accumulate in TMEM, two 256-column
//                 accumulators: the epilogue of tile i overlaps the MMAs of tile i + 1)
//   warps 2..9  : epilogue, two warps per TMEM lane quarter: tcgen05.ld 32x32b -> registers (accumulator released at
//                 once) -> fused op -> 128B-swizzled shared-memory slab -> TMA store; the residual / ReLU-mask
//                 operand R arrives by TMA load into the same slab; fp32 (wgrad) outputs use red.global.add.v4.f32
// Producer and issuer loops are warp-uniform with one elected lane (elect.sync inside the asm statement).
// Operands may be K-major or MN-major in global memory (the smem/instruction descriptors carry the majorness), so
// forward, dgrad and wgrad all read the tensors where they lie: no activation is ever transposed.
// Triangular modes skip whole k-blocks / output tiles of the causal token-mixing matrices.
// Fused epilogues: ReLU, residual +/-, ReLU-backward mask, and the group norm around the bottleneck's W_mid / W_out
// pair (5-8: row statistics out, norm applied to the accumulator, row-scaled weight gradient; 9-11: the same norm's
// backward -- row dots of the accumulator with the upstream gradient, norm + ReLU backward on the dgrad accumulator --
// in two instantiations of their own, gemm_tc_kernel<2, 1> / <2, 2>), and the LM head's softmax cross-entropy (12,
// gemm_tc_kernel<2, 3>: a 256-wide vocabulary is one tile, the logits never leave the SM) -- include/obst_b200.h.
// Work items (tiles x split-K) are dealt round-robin to the CTA pairs: the split count is chosen for full rounds, and
// triangular outputs number only their live tiles.  The set-up precedes a griddepcontrol.wait (programmatic dependent
// launch): it overlaps the previous kernel's tail.
#include <cuda.h>
#include <stdlib.h>

#include <mutex>
#include <string>
#include <unordered_map>

#include "gemm_params.cuh"

namespace obst {

static constexpr int BM = 128;
static constexpr int BK = 64;
static constexpr int A_STAGE_BYTES = BM * BK * 2;        // 16 KB
static constexpr int TMEM_COLS = 512;
static constexpr int ACC_STRIDE = 256;                   // columns per accumulator buffer
static constexpr int TC_THREADS = 352;                   // TMA warp (A), MMA warp, 8 epilogue warps, TMA warp (B)
static constexpr int EPI_WARPS = 8;
static constexpr int EPI_BOX_ROWS = 32;                  // one epilogue warp's TMEM lane quarter
static constexpr int EPI_BOX_COLS = 64;                  // 64 bf16 = 128 B = one swizzle row
static constexpr int EPI_BUF_BYTES = EPI_BOX_ROWS * EPI_BOX_COLS * 2;  // 4 KB

struct TcParams {
  void* D;
  const void* R;
  int64_t d_ld, d_glo, d_ghi;
  int32_t M, N;
  int32_t nkb_inner, nkb;  // k-blocks per ko, total k-blocks
  int32_t GLO, G;
  int32_t BN, tiles_m, tiles_n, total_tiles;
  int32_t splits, tri, epi, accumulate, out_f32;
  int32_t tma_store;                   // bf16 output through smem staging + TMA store (and TMA load of R)
  int32_t tri_live;                    // tri == 3 with square tiles: tiles per group on or below the diagonal (else 0)
  int32_t m_first;                     // number the M tiles of one N tile first (the tiles that share B become neighbours)
  int32_t a_mn, b_mn;                  // 1 = MN-major operand
  int32_t a_uglo, a_ughi, a_uko;       // does the operand depend on g_lo / g_hi / ko
  int32_t b_uglo, b_ughi, b_uko;
  const void* aux0;                    // epilogues 5..8, see obst_gemm_desc
  const void* aux1;
  const void* aux2;
  const void* aux3;
  void* aux4;
  float scalar0, scalar1;              // epilogue 12: 1 / n, z_loss
};

// ------------------------------------------------------------------------------------------------
// PTX wrappers
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug must trap, never hang the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity, int which) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if ((++spins & 0x3ff) == 0 && clock64() - t0 > 4000000000ll) {
      printf("obst gemm_tc: mbarrier timeout (role %d, block %d, thread %d)\n", which, (int)blockIdx.x,
             (int)threadIdx.x);
      __trap();
    }
  }
}

// CG == 2: executed by both CTAs of a pair; the transaction bytes are credited to the LEADER CTA's barrier
// (same offset, CTA-rank bit cleared), which is the one the MMA thread waits on.
template <int CG>
__device__ __forceinline__ void tma_load_5d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2,
                                            int c3, int c4) {
  if (CG == 2) {
    asm volatile(
        "cp.async.bulk.tensor.5d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, "
        "%5, %6, %7}], [%2];" ::"r"(smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c0), "r"(c1), "r"(c2), "r"(c3),
        "r"(c4)
        : "memory");
  } else {
    asm volatile(
        "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], "
        "[%2];" ::"r"(smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
        : "memory");
  }
}

// Warp-uniform producer: the WHOLE warp runs the loop (so tile / k coordinates are uniform values the compiler
// keeps in uniform registers) and one elected lane issues.  With `if (lane == 0)` around the loop instead, every
// UTMALDG was preceded by an ELECT / R2UR.BROADCAST loop and the k-block coordinates by two integer divisions:
// the producer could not refill a stage in the 512 cycles the tensor core takes to drain one.
__device__ __forceinline__ void mbar_expect_tx_elect(uint64_t* bar, uint32_t bytes) {
  asm volatile(
      "{\n"
      ".reg .pred P;\n"
      "elect.sync _|P, 0xffffffff;\n"
      "@P mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(bytes)
      : "memory");
}
template <int CG>
__device__ __forceinline__ void tma_load_5d_elect(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1,
                                                  int c2, int c3, int c4) {
  if (CG == 2) {
    asm volatile(
        "{\n"
        ".reg .pred P;\n"
        "elect.sync _|P, 0xffffffff;\n"
        "@P cp.async.bulk.tensor.5d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, "
        "%4, %5, %6, %7}], [%2];\n"
        "}\n" ::"r"(dst),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(bar & 0xFEFFFFFFu), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
        : "memory");
  } else {
    asm volatile(
        "{\n"
        ".reg .pred P;\n"
        "elect.sync _|P, 0xffffffff;\n"
        "@P cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, "
        "%7}], [%2];\n"
        "}\n" ::"r"(dst),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
        : "memory");
  }
}

// Epilogue-side elected-lane helpers (called by a whole converged epilogue warp).  elect.sync names the same
// leader for the same mask every time, so the bulk-group bookkeeping (commit / wait_group are per thread) stays
// with one lane.
__device__ __forceinline__ void tma_store_commit_elect(const CUtensorMap* map, uint32_t src, int c0, int c1, int c2,
                                                       int c3, int c4) {
  asm volatile(
      "{\n"
      ".reg .pred P;\n"
      "elect.sync _|P, 0xffffffff;\n"
      "@P cp.async.bulk.tensor.5d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5, %6}], [%1];\n"
      "@P cp.async.bulk.commit_group;\n"
      "}\n" ::"l"(reinterpret_cast<uint64_t>(map)),
      "r"(src), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}
__device__ __forceinline__ void bulk_wait_read0_elect() {
  asm volatile(
      "{\n"
      ".reg .pred P;\n"
      "elect.sync _|P, 0xffffffff;\n"
      "@P cp.async.bulk.wait_group.read 0;\n"
      "}\n" ::
          : "memory");
}
__device__ __forceinline__ void bulk_wait_all_elect() {
  asm volatile(
      "{\n"
      ".reg .pred P;\n"
      "elect.sync _|P, 0xffffffff;\n"
      "@P cp.async.bulk.wait_group 0;\n"
      "}\n" ::
          : "memory");
}
// expect `bytes` on `bar` and request one R box into `dst` (CTA-local barrier, never the pair variant)
__device__ __forceinline__ void tma_load_r_elect(uint32_t dst, const CUtensorMap* map, uint32_t bar, uint32_t bytes,
                                                 int c0, int c1, int c2, int c3) {
  asm volatile(
      "{\n"
      ".reg .pred P;\n"
      "elect.sync _|P, 0xffffffff;\n"
      "@P mbarrier.arrive.expect_tx.shared::cta.b64 _, [%2], %3;\n"
      "@P cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%4, %5, %6, %7, "
      "%8}], [%2];\n"
      "}\n" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(bytes), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(0)
      : "memory");
}

__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// arrive on the barrier at the same offset in CTA `target` of the cluster (CG == 1: the local barrier)
template <int CG>
__device__ __forceinline__ void mbar_arrive_cta(uint64_t* bar, uint32_t target) {
  if (CG == 2) {
    asm volatile(
        "{\n"
        ".reg .b32 remote;\n"
        "mapa.shared::cluster.u32 remote, %0, %1;\n"
        "mbarrier.arrive.shared::cluster.b64 _, [remote];\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(target)
        : "memory");
  } else {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
  }
}

__device__ __forceinline__ void tma_store_5d(const CUtensorMap* map, const void* src, int c0, int c1, int c2, int c3,
                                             int c4) {
  asm volatile("cp.async.bulk.tensor.5d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5, %6}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(map)),
               "r"(smem_u32(src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

template <int CG>
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
  if (CG == 2) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)),
                 "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  } else {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)),
                 "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
}
template <int CG>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  if (CG == 2)
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
  else
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

// D[tmem] (+)= A[smem] * B[smem], bf16 inputs, fp32 accumulate.  Called by the whole (converged) MMA warp; one
// elected lane issues (see the producer note: a `lane == 0` region costs an ELECT/R2UR loop per instruction).
template <int CG>
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
  if (CG == 2) {
    asm volatile(
        "{\n"
        ".reg .pred p, e;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "elect.sync _|e, 0xffffffff;\n"
        "@e tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  } else {
    asm volatile(
        "{\n"
        ".reg .pred p, e;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "elect.sync _|e, 0xffffffff;\n"
        "@e tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  }
}
// mbarrier arrive once all previously issued tcgen05 ops of the issuing thread have completed
// (CG == 2: the arrive is multicast to the barrier at the same offset in BOTH CTAs of the pair).
// elect.sync picks the same lane every time for a full mask, so the commit tracks the MMAs above.
template <int CG>
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  if (CG == 2) {
    asm volatile(
        "{\n"
        ".reg .pred e;\n"
        "elect.sync _|e, 0xffffffff;\n"
        "@e tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n"
        "}\n" ::"r"(smem_u32(bar)),
        "h"((uint16_t)3)
        : "memory");
  } else {
    asm volatile(
        "{\n"
        ".reg .pred e;\n"
        "elect.sync _|e, 0xffffffff;\n"
        "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n"
        "}\n" ::"r"(smem_u32(bar))
        : "memory");
  }
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}

// shared-memory matrix descriptor (SWIZZLE_128B, descriptor version 1)
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3fff);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3fff) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32;
  d |= (uint64_t)1 << 46;  // version
  d |= (uint64_t)2 << 61;  // SWIZZLE_128B
  return d;
}

struct Tile {
  int m0, n0, glo, ghi, kb0, kb1;
};

// CG = CTAs per tile (1, or 2 = a CTA pair sharing one 256-row tile through tcgen05 cta_group::2)
template <int CG>
__device__ __forceinline__ bool get_tile(const TcParams& p, int tile, Tile& t) {
  constexpr int TM = BM * CG;
  int nt, mt, r;
  if (p.tri_live) {
    // only the tiles on or below the diagonal are numbered (row by row: (0,0) (1,0) (1,1) (2,0) ...): with all tiles
    // numbered and the dead ones skipped, the CTA pairs that drew them idled while the others set the kernel's time
    int idx = tile % p.tri_live;
    r = tile / p.tri_live;
    mt = 0;
    while (idx > mt) {
      idx -= mt + 1;
      ++mt;
    }
    nt = idx;
  } else if (p.m_first) {
    mt = tile % p.tiles_m;
    r = tile / p.tiles_m;
    nt = r % p.tiles_n;
    r /= p.tiles_n;
  } else {
    nt = tile % p.tiles_n;
    r = tile / p.tiles_n;
    mt = r % p.tiles_m;
    r /= p.tiles_m;
  }
  const int g = r % p.G;
  const int sp = r / p.G;
  t.m0 = mt * TM;
  t.n0 = nt * p.BN;
  t.glo = g % p.GLO;
  t.ghi = g / p.GLO;
  if (p.tri == 3 && t.n0 > t.m0 + TM - 1) return false;
  int kb0 = 0, kb1 = p.nkb;
  if (p.tri == 1) kb1 = min(p.nkb, (t.m0 + TM + BK - 1) / BK);
  if (p.tri == 2) kb0 = t.m0 / BK;
  if (p.splits > 1) {
    const int len = kb1 - kb0;
    const int per = (len + p.splits - 1) / p.splits;
    kb0 += sp * per;
    kb1 = min(kb1, kb0 + per);
  }
  t.kb0 = kb0;
  t.kb1 = kb1;
  return kb1 > kb0;
}

// ------------------------------------------------------------------------------------------------
// epilogue helpers: 8 consecutive output columns of one row
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void epi_apply(float (&v)[8], const float (&r)[8], int epi) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    switch (epi) {
      case 1: v[j] = fmaxf(v[j], 0.f); break;
      case 2: v[j] = r[j] + v[j]; break;
      case 3: v[j] = r[j] - v[j]; break;
      case 4: v[j] = r[j] > 0.f ? v[j] : 0.f; break;
      default: break;
    }
  }
}

// One epilogue chunk of one thread: 64 accumulator columns of its row -> fused op -> bf16 -> the row's 128-byte
// line of the swizzled slab (16-byte slot j lives at j ^ (row & 7): conflict-free for a quarter-warp).
template <int EPI>
__device__ __forceinline__ void epi_chunk(uint8_t* buf, int lane, const uint32_t (&r)[64]) {
  uint8_t* rowp = buf + lane * 128;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    uint4* slot = reinterpret_cast<uint4*>(rowp + ((j ^ (lane & 7)) << 4));
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(r[j * 8 + e]);
    if (EPI == 1) {
#pragma unroll
      for (int e = 0; e < 8; ++e) v[e] = fmaxf(v[e], 0.f);
    } else if (EPI >= 2) {
      const uint4 ru = *slot;
      const float2 f0 = unpack_bf16x2(ru.x), f1 = unpack_bf16x2(ru.y), f2 = unpack_bf16x2(ru.z),
                   f3 = unpack_bf16x2(ru.w);
      const float rr[8] = {f0.x, f0.y, f1.x, f1.y, f2.x, f2.y, f3.x, f3.y};
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        if (EPI == 2) v[e] = rr[e] + v[e];
        else if (EPI == 3) v[e] = rr[e] - v[e];
        else v[e] = rr[e] > 0.f ? v[e] : 0.f;
      }
    }
    uint4 o;
    o.x = pack_bf16x2(v[0], v[1]);
    o.y = pack_bf16x2(v[2], v[3]);
    o.z = pack_bf16x2(v[4], v[5]);
    o.w = pack_bf16x2(v[6], v[7]);
    *slot = o;
  }
}

// Epilogues 5 / 6: the group norm of the A operand folded into this GEMM (include/obst_b200.h):
//   D = R +/- ( rstd[m] * acc - rstd[m] * mean[m] * c[n] + d[n] )
// rs = rstd, rm = rstd * mean of this thread's row; cvec / dvec point at the chunk's 64 columns (every lane reads
// the same addresses: one transaction per load).
template <int EPI>
__device__ __forceinline__ void epi_chunk_affine(uint8_t* buf, int lane, const uint32_t (&r)[64], float rs, float rm,
                                                 const float* __restrict__ cvec, const float* __restrict__ dvec) {
  uint8_t* rowp = buf + lane * 128;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    uint4* slot = reinterpret_cast<uint4*>(rowp + ((j ^ (lane & 7)) << 4));
    const float4 c0 = __ldg(reinterpret_cast<const float4*>(cvec + j * 8));
    const float4 c1 = __ldg(reinterpret_cast<const float4*>(cvec + j * 8 + 4));
    const float4 d0 = __ldg(reinterpret_cast<const float4*>(dvec + j * 8));
    const float4 d1 = __ldg(reinterpret_cast<const float4*>(dvec + j * 8 + 4));
    const float cc[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
    const float dd[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
    const uint4 ru = *slot;
    const float2 f0 = unpack_bf16x2(ru.x), f1 = unpack_bf16x2(ru.y), f2 = unpack_bf16x2(ru.z), f3 = unpack_bf16x2(ru.w);
    const float rr[8] = {f0.x, f0.y, f1.x, f1.y, f2.x, f2.y, f3.x, f3.y};
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float f = fmaf(rs, __uint_as_float(r[j * 8 + e]), fmaf(-rm, cc[e], dd[e]));
      v[e] = EPI == 5 ? rr[e] + f : rr[e] - f;
    }
    uint4 o;
    o.x = pack_bf16x2(v[0], v[1]);
    o.y = pack_bf16x2(v[2], v[3]);
    o.z = pack_bf16x2(v[4], v[5]);
    o.w = pack_bf16x2(v[6], v[7]);
    *slot = o;
  }
}

// Epilogues 9 / 10 = 5 / 6, and on the way the two reductions the folded norm's BACKWARD needs, taken over this
// chunk's 64 columns of the thread's row: p1 += sum dy * c, p2 += sum dy * acc.  gy = the row's 64 dy values (the
// gradient that will reach this output) as 32 packed bf16 pairs, fetched into registers BEFORE the accumulator was
// waited for: a load issued here, between the uses, would expose an L2 round trip per 8 columns.
template <int EPI>
__device__ __forceinline__ void epi_chunk_affine_dots(uint8_t* buf, int lane, const uint32_t (&r)[64], float rs,
                                                      float rm, float2 cv, float2 dv, const uint32_t (&gy)[32],
                                                      float& p1, float& p2) {
  // cv / dv: c and d of the chunk's 64 columns, two columns per lane (as gamma in epilogue 11), read through shuffles:
  // with the 32 registers of dy next to the accumulator there is no room for vectors of them
  uint8_t* rowp = buf + lane * 128;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    uint4* slot = reinterpret_cast<uint4*>(rowp + ((j ^ (lane & 7)) << 4));
    const uint4 ru = *slot;
    const uint32_t rw[4] = {ru.x, ru.y, ru.z, ru.w};
    uint32_t ow[4];
#pragma unroll
    for (int h = 0; h < 4; ++h) {
      const float2 rr = unpack_bf16x2(rw[h]);
      const float2 gg = unpack_bf16x2(gy[j * 4 + h]);
      const int src = j * 4 + h;  // the lane that holds columns 2 src and 2 src + 1
      const float c0 = __shfl_sync(0xffffffffu, cv.x, src), c1 = __shfl_sync(0xffffffffu, cv.y, src);
      const float d0 = __shfl_sync(0xffffffffu, dv.x, src), d1 = __shfl_sync(0xffffffffu, dv.y, src);
      const float a0 = __uint_as_float(r[j * 8 + 2 * h]), a1 = __uint_as_float(r[j * 8 + 2 * h + 1]);
      p1 = fmaf(gg.x, c0, fmaf(gg.y, c1, p1));
      p2 = fmaf(gg.x, a0, fmaf(gg.y, a1, p2));
      const float f0 = fmaf(rs, a0, fmaf(-rm, c0, d0)), f1 = fmaf(rs, a1, fmaf(-rm, c1, d1));
      ow[h] = EPI == 9 ? pack_bf16x2(rr.x + f0, rr.y + f1) : pack_bf16x2(rr.x - f0, rr.y - f1);
    }
    *slot = make_uint4(ow[0], ow[1], ow[2], ow[3]);
  }
}

// Epilogue 11: the folded norm's backward: D = [h > 0] * (a * acc + b * h + c), h from the slab, acc = gamma * dn
// (the GEMM's B operand is gamma . W: a scale per column applied here cost a shuffle per element, and this epilogue
// belongs to a K = 512 GEMM whose main loop is shorter than its epilogue).
__device__ __forceinline__ void epi_chunk_norm_bwd(uint8_t* buf, int lane, const uint32_t (&r)[64], float ca, float cb,
                                                   float cc) {
  uint8_t* rowp = buf + lane * 128;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    uint4* slot = reinterpret_cast<uint4*>(rowp + ((j ^ (lane & 7)) << 4));
    const uint4 ru = *slot;
    const float2 f0 = unpack_bf16x2(ru.x), f1 = unpack_bf16x2(ru.y), f2 = unpack_bf16x2(ru.z), f3 = unpack_bf16x2(ru.w);
    const float hh[8] = {f0.x, f0.y, f1.x, f1.y, f2.x, f2.y, f3.x, f3.y};
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float t = fmaf(ca, __uint_as_float(r[j * 8 + e]), fmaf(cb, hh[e], cc));
      v[e] = hh[e] > 0.f ? t : 0.f;
    }
    uint4 o;
    o.x = pack_bf16x2(v[0], v[1]);
    o.y = pack_bf16x2(v[2], v[3]);
    o.z = pack_bf16x2(v[4], v[5]);
    o.w = pack_bf16x2(v[6], v[7]);
    *slot = o;
  }
}

// Epilogue 7: D = relu(acc), and the row's sum and sum of squares of the values AS STORED (bf16-rounded) are added to
// (s1, s2): the statistics of the group norm that follows (src/model/basic.py:55-56), taken where the data is made.
__device__ __forceinline__ void epi_chunk_relu_stats(uint8_t* buf, int lane, const uint32_t (&r)[64], float& s1,
                                                     float& s2) {
  uint8_t* rowp = buf + lane * 128;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    uint4* slot = reinterpret_cast<uint4*>(rowp + ((j ^ (lane & 7)) << 4));
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = fmaxf(__uint_as_float(r[j * 8 + e]), 0.f);
    uint4 o;
    o.x = pack_bf16x2(v[0], v[1]);
    o.y = pack_bf16x2(v[2], v[3]);
    o.z = pack_bf16x2(v[4], v[5]);
    o.w = pack_bf16x2(v[6], v[7]);
    *slot = o;
    const uint32_t w[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float lo = __uint_as_float(w[q] << 16), hi = __uint_as_float(w[q] & 0xFFFF0000u);
      s1 += lo + hi;
      s2 = fmaf(lo, lo, fmaf(hi, hi, s2));
    }
  }
}

// Epilogue 12: the softmax cross-entropy's gradient for 64 columns of one row (columns col .. col + 63 of the
// vocabulary), from the logits AS THEY WOULD HAVE BEEN STORED (bf16-rounded accumulator):
//   d = (p - [n == tgt]) / n_tokens + (2 z logZ / n_tokens) p,   p = exp(l - max) / sum
__device__ __forceinline__ float bf16_round(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

__device__ __forceinline__ void epi_chunk_xent(uint8_t* buf, int lane, const uint32_t (&r)[64], int col, int tgt,
                                               float mx, float inv_se, float inv_n, float zc) {
  uint8_t* rowp = buf + lane * 128;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    uint4* slot = reinterpret_cast<uint4*>(rowp + ((j ^ (lane & 7)) << 4));
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float pr = expf(bf16_round(__uint_as_float(r[j * 8 + e])) - mx) * inv_se;
      v[e] = (pr - (col + j * 8 + e == tgt ? 1.f : 0.f)) * inv_n + zc * pr;
    }
    uint4 o;
    o.x = pack_bf16x2(v[0], v[1]);
    o.y = pack_bf16x2(v[2], v[3]);
    o.z = pack_bf16x2(v[4], v[5]);
    o.w = pack_bf16x2(v[6], v[7]);
    *slot = o;
  }
}

template <int CG>
struct Cfg {
  static constexpr int STAGES = CG == 2 ? 5 : 3;
  static constexpr int B_STAGE = 256 / CG * BK * 2;  // bytes of this CTA's share of the B tile
  static constexpr int EPI_OFF = STAGES * (A_STAGE_BYTES + B_STAGE);
  static constexpr int BAR_OFF = EPI_OFF + EPI_WARPS * 2 * EPI_BUF_BYTES;
  static constexpr int SMEM = BAR_OFF + 512 + 1024;
};

// SEQ: the instantiation for epilogues 9 / 10 (the row dots with a second tile): its epilogue warps hold that tile's
// rows in registers and take the two column chunks one after the other.  A kernel of its own so that the register
// allocation of every other epilogue stays what it was.
// MODE 2: the instantiation for epilogue 11 (the folded norm's backward), split off for the same reason.
// MODE 3: the instantiation for epilogue 12 (LM head with the softmax cross-entropy): three passes over the row's
// accumulator in TMEM (max / arg-max, sum of exponentials, gradient) -- the logits never leave the SM.
template <int CG, int MODE = 0>
__global__ void __launch_bounds__(TC_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmD, const __grid_constant__ CUtensorMap tmR, const TcParams p) {
  constexpr int STAGES = Cfg<CG>::STAGES;
  constexpr int B_STAGE = Cfg<CG>::B_STAGE;
  constexpr bool SEQ = MODE == 1;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + STAGES * A_STAGE_BYTES;
  uint8_t* smem_epi = smem + Cfg<CG>::EPI_OFF;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg<CG>::BAR_OFF);
  uint64_t* full_bar = bars;                     // [STAGES]  (CG==2: only the leader's are waited on)
  uint64_t* empty_bar = bars + STAGES;           // [STAGES]
  uint64_t* tfull_bar = bars + 2 * STAGES;       // [2]
  uint64_t* tempty_bar = bars + 2 * STAGES + 2;  // [2]       (CG==2: only the leader's are used)
  uint64_t* rbar = bars + 2 * STAGES + 4;        // [8 warps][2 buffers] R-tile arrival
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4 + 2 * EPI_WARPS);

  // warp index through a shuffle and CTA rank from blockIdx (clusters are (CG,1,1) on a 1-D grid): both are then
  // warp-uniform values for the compiler, and everything derived from them can live in uniform registers
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const uint32_t rank = CG == 2 ? (blockIdx.x & 1u) : 0u;
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x / CG;
  const int num_clusters = gridDim.x / CG;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmA)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmB)) : "memory");
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);
      mbar_init(&tempty_bar[s], EPI_WARPS * 32 * CG);
    }
    for (int s = 0; s < 2 * EPI_WARPS; ++s) mbar_init(&rbar[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc<CG>(tmem_ptr, TMEM_COLS);
  tcgen05_fence_before();
  if (CG == 2) cluster_sync();
  else __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  // Programmatic dependent launch: everything above (barrier init, TMEM allocation, descriptor prefetch, the cluster
  // barrier) touches no global data and may run while the previous kernel of the stream drains its last tiles; from
  // here on its results are needed.  A no-op when the launch did not ask for the overlap.  The trigger right after
  // lets the NEXT tensor-core GEMM do the same as this grid's CTAs retire (one wave: nobody is kept off an SM).
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

  const int bn_cta = p.BN / CG;  // B rows (N) this CTA stages
  const uint32_t stage_bytes = (uint32_t)(A_STAGE_BYTES + bn_cta * BK * 2) * CG;  // what the (leader's) barrier sees

  if (warp == 0 || warp == 2 + EPI_WARPS) {
    // ===================== TMA producers (whole warp, one elected lane issues) =====================
    // warp 0 requests the A boxes of every stage (and posts the stage's byte count), warp 10 the B boxes: an
    // MN-major operand needs a box per 64 elements, and one warp issuing up to four boxes per stage was the limit.
    {
      const bool is_a = warp == 0;
      int stage = 0;
      uint32_t phase = 0;
      Tile t;
      const uint32_t s0 = is_a ? smem_u32(smem_a) : smem_u32(smem_b), fb0 = smem_u32(full_bar);
      const uint32_t s_stride = is_a ? (uint32_t)A_STAGE_BYTES : (uint32_t)B_STAGE;
      const CUtensorMap* tm = is_a ? &tmA : &tmB;
      const int mn = is_a ? p.a_mn : p.b_mn;
      const int boxes = mn ? (is_a ? BM : bn_cta) / 64 : 1;
      const int uglo = is_a ? p.a_uglo : p.b_uglo, ughi = is_a ? p.a_ughi : p.b_ughi, uko = is_a ? p.a_uko : p.b_uko;
      for (int tile = cluster_id; tile < p.total_tiles; tile += num_clusters) {
        if (!get_tile<CG>(p, tile, t)) continue;
        const int c2 = uglo ? t.glo : 0, c3 = ughi ? t.ghi : 0;
        const int mn0 = is_a ? t.m0 + (int)rank * BM : t.n0 + (int)rank * bn_cta;
        int ki = t.kb0 % p.nkb_inner, ko = t.kb0 / p.nkb_inner;  // once per tile; advanced incrementally below
        for (int kb = t.kb0; kb < t.kb1; ++kb) {
          const int k0 = ki * BK;
          const int c4 = uko ? ko : 0;
          const uint32_t fb = fb0 + (uint32_t)stage * 8u;
          const uint32_t dst = s0 + (uint32_t)stage * s_stride;
          mbar_wait(&empty_bar[stage], phase ^ 1, 0);
          if (is_a && leader) mbar_expect_tx_elect(&full_bar[stage], stage_bytes);
          // K-major: one box (64 k x rows).  MN-major: one box per 64 MN elements (64 mn x 64 k each).
          for (int j = 0; j < boxes; ++j)
            tma_load_5d_elect<CG>(dst + j * (64 * BK * 2), tm, fb, mn ? mn0 + 64 * j : k0, mn ? k0 : mn0, c2, c3, c4);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
          if (++ki == p.nkb_inner) {
            ki = 0;
            ++ko;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (warp 1 of the leader CTA; one elected lane issues) =====================
    if (leader) {
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)p.a_mn << 15) | ((uint32_t)p.b_mn << 16) |
                             ((uint32_t)(p.BN >> 3) << 17) | ((uint32_t)((BM * CG) >> 4) << 24);
      // K-major: 8-row groups 1024 B apart (SBO), 32 B per UMMA_K step.
      // MN-major: 64-element MN groups 8192 B apart (LBO), 8-row k groups 1024 B apart (SBO), 2048 B per UMMA_K.
      const uint32_t a_lbo = p.a_mn ? 64 * BK * 2 : 16, b_lbo = p.b_mn ? 64 * BK * 2 : 16;
      const uint32_t a_kadv = p.a_mn ? 2048 : 32, b_kadv = p.b_mn ? 2048 : 32;
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      Tile t;
      for (int tile = cluster_id; tile < p.total_tiles; tile += num_clusters) {
        if (!get_tile<CG>(p, tile, t)) continue;
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1, 1);
        tcgen05_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)(acc * ACC_STRIDE);
        for (int kb = t.kb0; kb < t.kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase, 2);
          tcgen05_fence_after();
          const uint32_t sa = smem_u32(smem_a + stage * A_STAGE_BYTES);
          const uint32_t sb = smem_u32(smem_b + stage * B_STAGE);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t adesc = make_smem_desc(sa + k * a_kadv, a_lbo, 1024);
            const uint64_t bdesc = make_smem_desc(sb + k * b_kadv, b_lbo, 1024);
            umma_bf16<CG>(tmem_d, adesc, bdesc, idesc, (kb > t.kb0 || k > 0) ? 1u : 0u);
          }
          umma_commit<CG>(&empty_bar[stage]);  // frees the smem slot (in both CTAs) when these MMAs retire
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit<CG>(&tfull_bar[acc]);  // accumulator ready for the epilogue warps (of both CTAs)
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else {
    // ===================== epilogue (warps 2..9) =====================
    // Two warps per TMEM lane quarter; they take alternate column chunks of the tile.
    const int quarter = warp & 3;  // TMEM lanes [32*quarter, 32*quarter+32) are the ones this warp may read
    const int ew = warp - 2;       // 0..7
    const int half = ew >> 2;      // which of the two warps of this quarter
    int acc = 0;
    uint32_t acc_phase = 0;
    Tile t;
    const int cta_m = (int)rank * BM;
    if (p.tma_store) {
      // bf16 output: TMEM -> registers -> fused op -> swizzled smem slab (32 rows x 128 B) -> TMA store.
      // R (residual / ReLU mask) arrives the same way (TMA load into the slab), so global traffic is full lines.
      // A warp owns up to two 64-column chunks of its 32 rows (c = half and half + 2); chunk slot s uses slab s.
      // Per tile: (1) R tiles are requested first, so they fly in while the MMAs finish; (2) BOTH chunks are read
      // from TMEM into registers and the accumulator is handed back at once -- the conversion / store work then
      // runs while the tensor core is already on the tile after next.
      uint8_t* const buf0 = smem_epi + ew * 2 * EPI_BUF_BYTES;
      uint64_t* const rb0 = &rbar[ew * 2];
      uint32_t rphase = 0;  // bit s = phase of slab s's R barrier
      [[maybe_unused]] double x_nll = 0.0, x_z = 0.0, x_hit = 0.0;  // epilogue 12: this thread's rows, all its tiles
      const int nchunks = (p.BN + EPI_BOX_COLS - 1) / EPI_BOX_COLS;
      for (int tile = cluster_id; tile < p.total_tiles; tile += num_clusters) {
        if (!get_tile<CG>(p, tile, t)) continue;
        const int row0 = t.m0 + cta_m + quarter * 32;
        const int c0 = half, c1 = half + 2;
        const int col0 = t.n0 + c0 * EPI_BOX_COLS, col1 = t.n0 + c1 * EPI_BOX_COLS;
        const bool live0 = c0 < nchunks && row0 < p.M && col0 < p.N;  // warp-uniform
        const bool live1 = c1 < nchunks && row0 < p.M && col1 < p.N;
        bulk_wait_read0_elect();  // last tile's stores have read both slabs (they were issued a whole tile ago)
        __syncwarp();
        const bool has_r = (p.epi >= 2 && p.epi <= 6) || (p.epi >= 9 && p.epi <= 11);
        if (has_r) {
          if (live0) tma_load_r_elect(smem_u32(buf0), &tmR, smem_u32(rb0), EPI_BUF_BYTES, col0, row0, t.glo, t.ghi);
          if (live1)
            tma_load_r_elect(smem_u32(buf0 + EPI_BUF_BYTES), &tmR, smem_u32(rb0 + 1), EPI_BUF_BYTES, col1, row0, t.glo,
                             t.ghi);
        }
        // epilogues 5 / 6: this thread's row statistics (mean, rstd) of the folded group norm
        float rs = 0.f, rm = 0.f;
        [[maybe_unused]] float rk = 0.f;
        const bool row_live = row0 + lane < p.M;
        if constexpr (MODE == 2) {  // (a, b, c) of the folded norm's backward for this thread's row
          if (row_live) {
            const float4 st = __ldg(reinterpret_cast<const float4*>(p.aux0) + (int64_t)(row0 + lane) * p.GLO + t.glo);
            rs = st.x;
            rm = st.y;
            rk = st.z;
          }
        } else if (SEQ || p.epi == 5 || p.epi == 6) {
          if (row_live) {
            const float2 st = __ldg(reinterpret_cast<const float2*>(p.aux0) + (int64_t)(row0 + lane) * p.GLO + t.glo);
            rs = st.y;
            rm = st.y * st.x;
          }
        }
        // epilogues 9 / 10: this thread's row of the second tile (the upstream gradient, D's layout): the first chunk's
        // 64 values into registers now -- the loads fly while the MMAs finish -- and the second chunk's line asked into
        // L2; it is fetched when the first chunk is done.  The two chunks are taken one after the other (`seq`): both
        // chunks' accumulators next to both rows of dy would not fit the register file.
        constexpr bool seq = SEQ || MODE == 3;  // (epilogue 12 keeps the row state of its loss next to the accumulator)
        uint32_t gy[SEQ ? 32 : 1];
        const bf16* dy_row = nullptr;
        // epilogues 9 / 10: c (gv) and d (dv) of the two chunks, two columns per lane
        float2 cdv[SEQ ? 4 : 1];  // c of chunk 0 / 1, d of chunk 0 / 1
        if constexpr (SEQ) {
          dy_row = reinterpret_cast<const bf16*>(p.aux3) + (int64_t)t.glo * p.d_glo + (int64_t)t.ghi * p.d_ghi +
                   (int64_t)(row0 + lane) * p.d_ld;
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            uint4 a = make_uint4(0u, 0u, 0u, 0u);
            if (row_live && live0) a = __ldg(reinterpret_cast<const uint4*>(dy_row + col0) + q);
            gy[q * 4] = a.x; gy[q * 4 + 1] = a.y; gy[q * 4 + 2] = a.z; gy[q * 4 + 3] = a.w;
          }
          if (row_live && live1) asm volatile("prefetch.global.L2 [%0];" ::"l"(dy_row + col1));
          // c (in gv0 / gv1) and d (dv0 / dv1) of the two chunks, two columns per lane
          const float* const cvec = reinterpret_cast<const float*>(p.aux1) + (int64_t)t.glo * p.N;
          const float* const dvec = reinterpret_cast<const float*>(p.aux2) + (int64_t)t.glo * p.N;
          cdv[0] = cdv[1] = cdv[2] = cdv[3] = make_float2(0.f, 0.f);
          if (live0) {
            cdv[0] = __ldg(reinterpret_cast<const float2*>(cvec + col0) + lane);
            cdv[2] = __ldg(reinterpret_cast<const float2*>(dvec + col0) + lane);
          }
          if (live1) {
            cdv[1] = __ldg(reinterpret_cast<const float2*>(cvec + col1) + lane);
            cdv[3] = __ldg(reinterpret_cast<const float2*>(dvec + col1) + lane);
          }
        }
        float s1 = 0.f, s2 = 0.f;  // epilogue 7 (sum, sum of squares); 9 / 10 (the two dot products)
        [[maybe_unused]] int tgt = 0;
        if constexpr (MODE == 3) {
          if (row_live) {
            tgt = __ldg(reinterpret_cast<const int32_t*>(p.aux0) + row0 + lane);
            tgt = tgt < 0 ? 0 : (tgt >= p.N ? p.N - 1 : tgt);
          }
        }
        mbar_wait(&tfull_bar[acc], acc_phase, 3);
        tcgen05_fence_after();
        const uint32_t taddr0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * ACC_STRIDE);
        [[maybe_unused]] float x_mx = 0.f, x_inv = 0.f, x_zc = 0.f;
        if constexpr (MODE == 3) {
          // the whole row (N == BN == 256) is read by BOTH warps of the quarter -- TMEM reads are cheap, an exchange
          // between the warps is not -- each then writes its own two chunks; the `half == 0` warp counts the loss
          float mx = -INFINITY;
          int arg = 0;
          for (int c = 0; c < 8; ++c) {
            uint32_t r[32];
            tmem_ld32(taddr0 + (uint32_t)(c * 32), r);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const float v = bf16_round(__uint_as_float(r[j]));
              if (v > mx) {  // strict: the first index on ties (tf.argmax)
                mx = v;
                arg = c * 32 + j;
              }
            }
          }
          float se = 0.f, lt = 0.f;
          for (int c = 0; c < 8; ++c) {
            uint32_t r[32];
            tmem_ld32(taddr0 + (uint32_t)(c * 32), r);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const float v = bf16_round(__uint_as_float(r[j]));
              se += expf(v - mx);
              if (c * 32 + j == tgt) lt = v;
            }
          }
          const float logz = logf(se) + mx;
          x_mx = mx;
          x_inv = 1.0f / se;
          x_zc = 2.0f * p.scalar1 * p.scalar0 * logz;
          if (half == 0 && row_live) {
            x_nll += (double)(logz - lt);
            x_z += (double)logz * (double)logz;
            x_hit += (arg == tgt) ? 1.0 : 0.0;
          }
        }
        uint32_t ra[64], rc[64];
        if (c0 < nchunks) {
          tmem_ld32(taddr0 + (uint32_t)(c0 * EPI_BOX_COLS), *reinterpret_cast<uint32_t (*)[32]>(&ra[0]));
          tmem_ld32(taddr0 + (uint32_t)(c0 * EPI_BOX_COLS + 32), *reinterpret_cast<uint32_t (*)[32]>(&ra[32]));
        }
        if (!seq && c1 < nchunks) {
          tmem_ld32(taddr0 + (uint32_t)(c1 * EPI_BOX_COLS), *reinterpret_cast<uint32_t (*)[32]>(&rc[0]));
          tmem_ld32(taddr0 + (uint32_t)(c1 * EPI_BOX_COLS + 32), *reinterpret_cast<uint32_t (*)[32]>(&rc[32]));
        }
        tmem_ld_wait();
        if (!seq) {
          tcgen05_fence_before();
          mbar_arrive_cta<CG>(&tempty_bar[acc], 0);  // accumulator is in registers: release it
        }
#pragma unroll
        for (int sl = 0; sl < 2; ++sl) {
          if (seq && sl == 1) {  // the second chunk's dy row and accumulator only now (see above), then release
            if constexpr (SEQ) {
#pragma unroll
              for (int q = 0; q < 8; ++q) {
                uint4 a = make_uint4(0u, 0u, 0u, 0u);
                if (row_live && live1) a = __ldg(reinterpret_cast<const uint4*>(dy_row + col1) + q);
                gy[q * 4] = a.x; gy[q * 4 + 1] = a.y; gy[q * 4 + 2] = a.z; gy[q * 4 + 3] = a.w;
              }
            }
            if (c1 < nchunks) {  // into the first chunk's registers: they are free now
              tmem_ld32(taddr0 + (uint32_t)(c1 * EPI_BOX_COLS), *reinterpret_cast<uint32_t (*)[32]>(&ra[0]));
              tmem_ld32(taddr0 + (uint32_t)(c1 * EPI_BOX_COLS + 32), *reinterpret_cast<uint32_t (*)[32]>(&ra[32]));
            }
            tmem_ld_wait();
            tcgen05_fence_before();
            mbar_arrive_cta<CG>(&tempty_bar[acc], 0);
          }
          const bool live = sl == 0 ? live0 : live1;
          if (!live) continue;
          uint8_t* const buf = buf0 + sl * EPI_BUF_BYTES;
          const uint32_t (&r)[64] = (seq || sl == 0) ? ra : rc;
          if (has_r) {
            mbar_wait(rb0 + sl, (rphase >> sl) & 1u, 4);
            rphase ^= 1u << sl;
          }
          const int colx = sl == 0 ? col0 : col1;
          if constexpr (SEQ) {
            if (p.epi == 9) epi_chunk_affine_dots<9>(buf, lane, r, rs, rm, cdv[sl], cdv[2 + sl], gy, s1, s2);
            else epi_chunk_affine_dots<10>(buf, lane, r, rs, rm, cdv[sl], cdv[2 + sl], gy, s1, s2);
          } else if constexpr (MODE == 2) {
            epi_chunk_norm_bwd(buf, lane, r, rs, rm, rk);
          } else if constexpr (MODE == 3) {
            epi_chunk_xent(buf, lane, r, colx, tgt, x_mx, x_inv, p.scalar0, x_zc);
          } else
          switch (p.epi) {
            case 0: epi_chunk<0>(buf, lane, r); break;
            case 1: epi_chunk<1>(buf, lane, r); break;
            case 2: epi_chunk<2>(buf, lane, r); break;
            case 3: epi_chunk<3>(buf, lane, r); break;
            case 4: epi_chunk<4>(buf, lane, r); break;
            case 5:
              epi_chunk_affine<5>(buf, lane, r, rs, rm,
                                  reinterpret_cast<const float*>(p.aux1) + (int64_t)t.glo * p.N + colx,
                                  reinterpret_cast<const float*>(p.aux2) + (int64_t)t.glo * p.N + colx);
              break;
            case 6:
              epi_chunk_affine<6>(buf, lane, r, rs, rm,
                                  reinterpret_cast<const float*>(p.aux1) + (int64_t)t.glo * p.N + colx,
                                  reinterpret_cast<const float*>(p.aux2) + (int64_t)t.glo * p.N + colx);
              break;
            default: epi_chunk_relu_stats(buf, lane, r, s1, s2); break;
          }
          fence_async_smem();
          __syncwarp();
          tma_store_commit_elect(&tmD, smem_u32(buf), sl == 0 ? col0 : col1, row0, t.glo, t.ghi, 0);
        }
        if (SEQ && row_live) {
          // one pair per row, group, 256-column tile and epilogue warp: [M][GLO][2 * tiles_n] float2
          float2* part = reinterpret_cast<float2*>(p.aux4);
          part[((int64_t)(row0 + lane) * p.GLO + t.glo) * (2 * p.tiles_n) + 2 * (t.n0 / p.BN) + half] =
              make_float2(s1, s2);
        }
        if (MODE == 0 && p.epi == 7 && row0 + lane < p.M) {
          // one (sum, sum of squares) per row, tile and epilogue warp: [M][2 * tiles_n] float2, summed by the consumer
          float2* part = reinterpret_cast<float2*>(const_cast<void*>(p.aux0));
          part[(int64_t)(row0 + lane) * (2 * p.tiles_n) + 2 * (t.n0 / p.BN) + half] = make_float2(s1, s2);
        }
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
      if constexpr (MODE == 3) {
        if (half == 0) {  // warp-uniform
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            x_nll += __shfl_xor_sync(0xffffffffu, x_nll, o);
            x_z += __shfl_xor_sync(0xffffffffu, x_z, o);
            x_hit += __shfl_xor_sync(0xffffffffu, x_hit, o);
          }
          if (lane == 0) {
            double* sums = reinterpret_cast<double*>(p.aux4);
            atomicAdd(sums, x_nll);
            atomicAdd(sums + 1, x_z);
            atomicAdd(sums + 2, x_hit);
          }
        }
      }
      bulk_wait_all_elect();
      __syncwarp();
    } else {
      // fp32 output (plain store or split-K / cross-depth accumulation with red.global.add)
      for (int tile = cluster_id; tile < p.total_tiles; tile += num_clusters) {
        if (!get_tile<CG>(p, tile, t)) continue;
        mbar_wait(&tfull_bar[acc], acc_phase, 3);
        tcgen05_fence_after();
        const int m = t.m0 + cta_m + quarter * 32 + lane;
        const bool row_ok = m < p.M;
        const int64_t row_off = (int64_t)t.glo * p.d_glo + (int64_t)t.ghi * p.d_ghi + (int64_t)m * p.d_ld;
        // epilogue 8: every output row is scaled by aux0[g_lo, m] before it is accumulated
        const float row_scale =
            (p.epi == 8 && row_ok) ? __ldg(reinterpret_cast<const float*>(p.aux0) + (int64_t)t.glo * p.M + m) : 1.f;
        const uint32_t taddr0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * ACC_STRIDE);
        // epilogue 8 with aux1 / aux2: this row's dot product of the raw accumulator with a bf16 tile in D's layout
        const bool row_dot = p.epi == 8 && p.aux1 != nullptr && row_ok;
        float wdot = 0.f;
        for (int c = half; c < p.BN / 32; c += 2) {
          uint32_t r[32];
          tmem_ld32(taddr0 + (uint32_t)(c * 32), r);
          tmem_ld_wait();
          if (row_ok) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const int n = t.n0 + c * 32 + q * 8;
              if (n >= p.N) continue;
              float v[8], rr[8];
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                v[j] = __uint_as_float(r[q * 8 + j]);
                rr[j] = 0.f;
              }
              if (p.out_f32) {
                float* dp = reinterpret_cast<float*>(p.D) + row_off + n;
                if (p.epi >= 2 && p.epi <= 4) {
                  const float4 r0 =
                      *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.R) + row_off + n);
                  const float4 r1 =
                      *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.R) + row_off + n + 4);
                  rr[0] = r0.x; rr[1] = r0.y; rr[2] = r0.z; rr[3] = r0.w;
                  rr[4] = r1.x; rr[5] = r1.y; rr[6] = r1.z; rr[7] = r1.w;
                }
                epi_apply(v, rr, p.epi);
                if (p.epi == 8) {
                  if (row_dot) {
                    const uint4 wu = __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const bf16*>(p.aux1) + row_off + n));
                    const float2 w0 = unpack_bf16x2(wu.x), w1 = unpack_bf16x2(wu.y), w2 = unpack_bf16x2(wu.z),
                                 w3 = unpack_bf16x2(wu.w);
                    wdot = fmaf(w0.x, v[0], fmaf(w0.y, v[1], fmaf(w1.x, v[2], fmaf(w1.y, v[3], wdot))));
                    wdot = fmaf(w2.x, v[4], fmaf(w2.y, v[5], fmaf(w3.x, v[6], fmaf(w3.y, v[7], wdot))));
                  }
#pragma unroll
                  for (int j = 0; j < 8; ++j) v[j] *= row_scale;
                }
                if (p.accumulate) {
                  if (p.tri == 3 && n + 7 > m) {
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                      if (n + j <= m) atomicAdd(dp + j, v[j]);
                  } else {
                    red_add_v4(dp, v[0], v[1], v[2], v[3]);
                    red_add_v4(dp + 4, v[4], v[5], v[6], v[7]);
                  }
                } else {
                  *reinterpret_cast<float4*>(dp) = make_float4(v[0], v[1], v[2], v[3]);
                  *reinterpret_cast<float4*>(dp + 4) = make_float4(v[4], v[5], v[6], v[7]);
                }
              } else {
                bf16* dp = reinterpret_cast<bf16*>(p.D) + row_off + n;
                if (p.epi >= 2 && p.epi <= 4) {
                  const uint4 ru = *reinterpret_cast<const uint4*>(reinterpret_cast<const bf16*>(p.R) + row_off + n);
                  const float2 f0 = unpack_bf16x2(ru.x), f1 = unpack_bf16x2(ru.y), f2 = unpack_bf16x2(ru.z),
                               f3 = unpack_bf16x2(ru.w);
                  rr[0] = f0.x; rr[1] = f0.y; rr[2] = f1.x; rr[3] = f1.y;
                  rr[4] = f2.x; rr[5] = f2.y; rr[6] = f3.x; rr[7] = f3.y;
                }
                epi_apply(v, rr, p.epi);
                uint4 o;
                o.x = pack_bf16x2(v[0], v[1]);
                o.y = pack_bf16x2(v[2], v[3]);
                o.z = pack_bf16x2(v[4], v[5]);
                o.w = pack_bf16x2(v[6], v[7]);
                *reinterpret_cast<uint4*>(dp) = o;
              }
            }
          }
        }
        if (row_dot) atomicAdd(reinterpret_cast<float*>(const_cast<void*>(p.aux2)) + (int64_t)t.glo * p.M + m, wdot);
        tcgen05_fence_before();
        mbar_arrive_cta<CG>(&tempty_bar[acc], 0);  // 256*CG arrivals release the accumulator to the MMA warp
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  }

  tcgen05_fence_before();
  if (CG == 2) cluster_sync();  // the peer's smem / barriers / TMEM must outlive every remote access
  else __syncthreads();
  if (warp == 1) {
    __syncwarp();
    tcgen05_fence_after();
    tmem_dealloc<CG>(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------
// host side: tensor-map construction (cached) and launch
// ------------------------------------------------------------------------------------------------
typedef CUresult (*PFN_tmapEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                        const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                        CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                        CUtensorMapFloatOOBfill);

static PFN_tmapEncodeTiled get_encode_fn() {
  static PFN_tmapEncodeTiled fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_tmapEncodeTiled>(ptr);
  }
  return fn;
}

struct MapKey {
  const void* ptr;
  uint64_t dims[5];
  uint64_t strides[4];
  uint32_t box[2];
};

static std::mutex g_map_mu;
static std::unordered_map<std::string, CUtensorMap> g_map_cache;

// operand view: 5-D (inner, outer, g_lo, g_hi, ko); rows = M (A) or N (B); box_rows = rows per TMA box
static int make_operand_map(const void* ptr, int64_t s_mn, int64_t s_k, int64_t s_ko, int64_t s_glo, int64_t s_ghi,
                            int rows, int KI, int KO, int GLO, int GHI, bool mn_major, int box_rows, CUtensorMap* out,
                            bool strict) {
  MapKey key;
  memset(&key, 0, sizeof(key));
  key.ptr = ptr;
  const uint64_t es = 2;
  if (mn_major) {
    key.dims[0] = (uint64_t)rows;
    key.dims[1] = (uint64_t)KI;
    key.strides[0] = (uint64_t)s_k * es;
    key.box[0] = 64;
    key.box[1] = 64;
  } else {
    key.dims[0] = (uint64_t)KI;
    key.dims[1] = (uint64_t)rows;
    key.strides[0] = (uint64_t)s_mn * es;
    key.box[0] = 64;
    key.box[1] = (uint32_t)box_rows;
  }
  key.dims[2] = s_glo ? (uint64_t)GLO : 1;
  key.dims[3] = s_ghi ? (uint64_t)GHI : 1;
  key.dims[4] = (s_ko && KO > 1) ? (uint64_t)KO : 1;
  // strides of size-1 dims are never used for addressing; give them a legal (multiple of 16 B) value
  key.strides[1] = key.dims[2] > 1 ? (uint64_t)s_glo * es : 16;
  key.strides[2] = key.dims[3] > 1 ? (uint64_t)s_ghi * es : 16;
  key.strides[3] = key.dims[4] > 1 ? (uint64_t)s_ko * es : 16;
  if (key.dims[1] == 1 && (key.strides[0] % 16) != 0) key.strides[0] = 16;
  for (int i = 0; i < 4; ++i) {
    if (key.strides[i] % 16 != 0 || key.strides[i] == 0) {
      if (strict) set_error("obst_gemm(tc): operand stride %llu B not a multiple of 16", (unsigned long long)key.strides[i]);
      return OBST_ENOTSUP;
    }
  }
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0) {
    if (strict) set_error("obst_gemm(tc): operand pointer not 16-byte aligned");
    return OBST_ENOTSUP;
  }
  std::string skey(reinterpret_cast<const char*>(&key), sizeof(key));
  {
    std::lock_guard<std::mutex> lk(g_map_mu);
    auto it = g_map_cache.find(skey);
    if (it != g_map_cache.end()) {
      *out = it->second;
      return OBST_OK;
    }
  }
  PFN_tmapEncodeTiled enc = get_encode_fn();
  if (!enc) {
    set_error("obst_gemm(tc): cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
    return OBST_ECUDA;
  }
  cuuint64_t gdim[5], gstr[4];
  cuuint32_t box[5] = {key.box[0], key.box[1], 1, 1, 1}, estr[5] = {1, 1, 1, 1, 1};
  for (int i = 0; i < 5; ++i) gdim[i] = key.dims[i];
  for (int i = 0; i < 4; ++i) gstr[i] = key.strides[i];
  CUtensorMap m;
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(ptr), gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("obst_gemm(tc): cuTensorMapEncodeTiled failed (%d) dims=[%llu,%llu,%llu,%llu,%llu] strides=[%llu,%llu,%llu,%llu]",
              (int)r, (unsigned long long)gdim[0], (unsigned long long)gdim[1], (unsigned long long)gdim[2],
              (unsigned long long)gdim[3], (unsigned long long)gdim[4], (unsigned long long)gstr[0],
              (unsigned long long)gstr[1], (unsigned long long)gstr[2], (unsigned long long)gstr[3]);
    return OBST_ECUDA;
  }
  {
    std::lock_guard<std::mutex> lk(g_map_mu);
    if (g_map_cache.size() > 65536) g_map_cache.clear();
    g_map_cache.emplace(skey, m);
  }
  *out = m;
  return OBST_OK;
}

// output / residual view: 5-D (n, m, g_lo, g_hi, 1), box = 64 columns x 32 rows (one epilogue warp's slab)
static int make_out_map(const void* ptr, int N, int M, int GLO, int GHI, int64_t d_ld, int64_t d_glo, int64_t d_ghi,
                        CUtensorMap* out, bool strict) {
  MapKey key;
  memset(&key, 0, sizeof(key));
  key.ptr = ptr;
  const uint64_t es = 2;
  key.dims[0] = (uint64_t)N;
  key.dims[1] = (uint64_t)M;
  key.dims[2] = (GLO > 1) ? (uint64_t)GLO : 1;
  key.dims[3] = (GHI > 1) ? (uint64_t)GHI : 1;
  key.dims[4] = 1;
  key.strides[0] = (uint64_t)d_ld * es;
  key.strides[1] = key.dims[2] > 1 ? (uint64_t)d_glo * es : 16;
  key.strides[2] = key.dims[3] > 1 ? (uint64_t)d_ghi * es : 16;
  key.strides[3] = 16;
  key.box[0] = EPI_BOX_COLS;
  key.box[1] = EPI_BOX_ROWS;
  if (key.dims[1] == 1 && (key.strides[0] % 16) != 0) key.strides[0] = 16;
  for (int i = 0; i < 4; ++i)
    if (key.strides[i] % 16 != 0 || key.strides[i] == 0) return OBST_ENOTSUP;
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0) return OBST_ENOTSUP;
  std::string skey(reinterpret_cast<const char*>(&key), sizeof(key));
  skey.push_back('D');
  {
    std::lock_guard<std::mutex> lk(g_map_mu);
    auto it = g_map_cache.find(skey);
    if (it != g_map_cache.end()) {
      *out = it->second;
      return OBST_OK;
    }
  }
  PFN_tmapEncodeTiled enc = get_encode_fn();
  if (!enc) {
    set_error("obst_gemm(tc): cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
    return OBST_ECUDA;
  }
  cuuint64_t gdim[5], gstr[4];
  cuuint32_t box[5] = {key.box[0], key.box[1], 1, 1, 1}, estr[5] = {1, 1, 1, 1, 1};
  for (int i = 0; i < 5; ++i) gdim[i] = key.dims[i];
  for (int i = 0; i < 4; ++i) gstr[i] = key.strides[i];
  CUtensorMap m;
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(ptr), gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    if (strict) set_error("obst_gemm(tc): cuTensorMapEncodeTiled(output) failed (%d)", (int)r);
    return OBST_ENOTSUP;
  }
  {
    std::lock_guard<std::mutex> lk(g_map_mu);
    g_map_cache.emplace(skey, m);
  }
  *out = m;
  return OBST_OK;
}

int gemm_tc_launch(const GemmParams& g, cudaStream_t st, bool strict) {
#define TC_NOTSUP(...)                  \
  do {                                  \
    if (strict) set_error(__VA_ARGS__); \
    return OBST_ENOTSUP;                \
  } while (0)
  if (g.dtype_in != OBST_BF16) TC_NOTSUP("obst_gemm(tc): inputs must be bf16");
  const bool a_mn = g.a_mn == 1 && g.a_k != 1;
  const bool b_mn = g.b_mn == 1 && g.b_k != 1;
  if (!a_mn && g.a_k != 1) TC_NOTSUP("obst_gemm(tc): operand A has no unit stride");
  if (!b_mn && g.b_k != 1) TC_NOTSUP("obst_gemm(tc): operand B has no unit stride");
  if ((g.tri == 1 || g.tri == 2) && g.KO != 1) TC_NOTSUP("obst_gemm(tc): triangular modes need KO == 1");
  const bool out_f32 = g.dtype_out == OBST_F32;
  const int oal = out_f32 ? 4 : 8;
  if (g.N % 8 != 0 || g.d_ld % oal != 0 || g.d_glo % oal != 0 || g.d_ghi % oal != 0 ||
      (reinterpret_cast<uintptr_t>(g.D) & 15) != 0 || (g.R && (reinterpret_cast<uintptr_t>(g.R) & 15) != 0))
    TC_NOTSUP("obst_gemm(tc): output not 16-byte addressable (N=%d ld=%lld)", g.N, (long long)g.d_ld);
  if (g.accumulate && !out_f32) TC_NOTSUP("obst_gemm(tc): accumulate needs fp32 output");
  const bool affine = g.epi == 5 || g.epi == 6 || g.epi == 9 || g.epi == 10;
  if (((g.epi >= 2 && g.epi <= 6) || (g.epi >= 9 && g.epi <= 11)) && !g.R)
    TC_NOTSUP("obst_gemm(tc): epilogue %d needs R", g.epi);
  if (affine && (!g.aux0 || !g.aux1 || !g.aux2 || g.GHI != 1 || out_f32 || g.N % 64 != 0))
    TC_NOTSUP("obst_gemm(tc): epilogue %d needs aux0..2, GHI == 1, bf16 output and N %% 64 == 0", g.epi);
  if ((g.epi == 9 || g.epi == 10) && (!g.aux3 || !g.aux4 || (reinterpret_cast<uintptr_t>(g.aux3) & 15) != 0))
    TC_NOTSUP("obst_gemm(tc): epilogue %d needs aux3 (16-byte aligned, D's layout) and aux4", g.epi);
  if (g.epi == 11 && (!g.aux0 || g.GHI != 1 || out_f32 || g.N % 64 != 0 ||
                      (reinterpret_cast<uintptr_t>(g.aux0) & 15) != 0))
    TC_NOTSUP("obst_gemm(tc): epilogue 11 needs aux0 (float4 rows), GHI == 1, bf16 output and N %% 64 == 0");
  if (g.epi == 8 && g.aux1 && (!g.aux2 || (reinterpret_cast<uintptr_t>(g.aux1) & 15) != 0))
    TC_NOTSUP("obst_gemm(tc): epilogue 8 with aux1 needs aux2 and a 16-byte aligned aux1");
  if (g.epi > 12 || g.epi < 0) TC_NOTSUP("obst_gemm(tc): unknown epilogue %d", g.epi);
  if (g.epi == 12 && (!g.aux0 || !g.aux4 || out_f32 || g.N != 256 || g.GLO * g.GHI != 1 || g.accumulate ||
                      (reinterpret_cast<uintptr_t>(g.aux4) & 7) != 0))
    TC_NOTSUP("obst_gemm(tc): epilogue 12 needs aux0 (targets), aux4 (double[3]), bf16 output, one group and N == 256");
  if (g.epi == 7 && (!g.aux0 || out_f32 || g.N % 256 != 0 || g.GLO * g.GHI != 1))
    TC_NOTSUP("obst_gemm(tc): epilogue 7 needs aux0, bf16 output, one group and N %% 256 == 0");
  if (g.epi == 8 && (!g.aux0 || !g.accumulate || g.GHI != 1))
    TC_NOTSUP("obst_gemm(tc): epilogue 8 needs aux0, accumulate and GHI == 1");
#undef TC_NOTSUP

  TcParams p;
  memset(&p, 0, sizeof(p));
  p.D = g.D;
  p.R = g.R;
  p.aux0 = g.aux0;
  p.aux1 = g.aux1;
  p.aux2 = g.aux2;
  p.aux3 = g.aux3;
  p.aux4 = g.aux4;
  p.scalar0 = g.scalar0;
  p.scalar1 = g.scalar1;
  p.d_ld = g.d_ld;
  p.d_glo = g.d_glo;
  p.d_ghi = g.d_ghi;
  p.M = g.M;
  p.N = g.N;
  p.nkb_inner = (g.KI + BK - 1) / BK;
  p.nkb = p.nkb_inner * g.KO;
  p.GLO = g.GLO;
  p.G = g.GLO * g.GHI;
  p.BN = g.N > 128 ? 256 : (g.N > 64 ? 128 : 64);
  // CTA pairs (tcgen05 cta_group::2): one 256-row tile per pair, each CTA stages half of B, so every byte of B
  // crosses L2->SM once per 256 rows instead of once per 128 (the 1-CTA kernel is L2-bandwidth-bound at
  // 128x256 tiles).  Needs >= 64 B rows per CTA and more than one 128-row tile of M.
  static const int force_cg = [] {
    const char* e = getenv("OBST_GEMM_CTA_GROUP");
    return e ? atoi(e) : 0;
  }();
  int cg = (p.BN >= 128 && g.M > BM) ? 2 : 1;
  if (force_cg == 1) cg = 1;
  // causal (triangular) products: a 256-row pair tile straddling the diagonal executes MACs a 128-row tile skips;
  // OBST_GEMM_TRI_CTA_GROUP=1 runs those GEMMs with 128-row tiles (A/B switch)
  static const int tri_cg = [] {
    const char* e = getenv("OBST_GEMM_TRI_CTA_GROUP");
    return e ? atoi(e) : 0;
  }();
  if (tri_cg == 1 && (g.tri == 1 || g.tri == 2)) cg = 1;
  if (g.epi >= 9 && g.epi <= 12 && cg != 2) {
    if (strict) set_error("obst_gemm(tc): epilogue %d exists in the CTA-pair kernel only (M > 128, N > 64)", g.epi);
    return OBST_ENOTSUP;
  }
  p.tiles_m = (g.M + BM * cg - 1) / (BM * cg);
  p.tiles_n = (g.N + p.BN - 1) / p.BN;
  p.tri = g.tri;
  p.epi = g.epi;
  p.accumulate = g.accumulate;
  p.out_f32 = out_f32 ? 1 : 0;
  p.a_mn = a_mn;
  p.b_mn = b_mn;
  p.a_uglo = g.a_glo != 0;
  p.a_ughi = g.a_ghi != 0;
  p.a_uko = (g.a_ko != 0 && g.KO > 1);
  p.b_uglo = g.b_glo != 0;
  p.b_ughi = g.b_ghi != 0;
  p.b_uko = (g.b_ko != 0 && g.KO > 1);
  // lower-triangular output (gradient of a causal mixing map) with square tiles: number the live tiles only
  p.tri_live = (g.tri == 3 && p.tiles_m == p.tiles_n && p.BN == BM * cg) ? p.tiles_m * (p.tiles_m + 1) / 2 : 0;
  const int64_t base_tiles = (p.tri_live ? (int64_t)p.tri_live : (int64_t)p.tiles_m * p.tiles_n) * p.G;
  int splits = 1;
  if (g.accumulate) {
    splits = g.splits;
    if (splits <= 0) {
      // The work items (tiles x splits) are dealt round-robin to the resident CTAs (CTA pairs), all of the same length:
      // the kernel takes ceil(items / slots) rounds, and a last round that is mostly empty costs a whole round
      // (the weight gradients of the step have 32 or 64 tiles: 10 / 5 splits = 320 items on 74 CTA pairs ran 5 rounds
      // for 4.3 rounds of work).  Pick the split count whose last round is fullest, preferring fewer splits (each one
      // adds a tile of fp32 reduction traffic) and keeping >= 8 k-blocks per split.
      const int slots = num_sms() / cg;
      const int max_by_k = p.nkb / 8 > 0 ? p.nkb / 8 : 1;
      int64_t hi = (8 * (int64_t)slots + base_tiles - 1) / base_tiles;
      if (hi > max_by_k) hi = max_by_k;
      if (hi < 1 || base_tiles >= 4 * (int64_t)slots) hi = 1;  // plenty of tiles already: no split
      static const bool legacy = [] {  // OBST_GEMM_SPLIT_LEGACY=1: round 1's rule (2 x SMs / tiles), for A/B runs
        const char* e = getenv("OBST_GEMM_SPLIT_LEGACY");
        return e && e[0] == '1';
      }();
      if (legacy) {
        splits = (int)((2 * num_sms() + base_tiles - 1) / base_tiles);
        if (splits > max_by_k) splits = max_by_k;
        hi = 0;
      }
      double best = -1.0;
      for (int s = 1; s <= (int)hi; ++s) {
        const int64_t items = base_tiles * s;
        const int64_t rounds = (items + slots - 1) / slots;
        const double score = (double)items / (double)(rounds * slots) - 0.002 * s;
        if (score > best) {
          best = score;
          splits = s;
        }
      }
    }
    if (splits > p.nkb) splits = p.nkb;
  }
  p.splits = splits;
  const int64_t total = base_tiles * splits;
  OBST_REQUIRE(total < (1ll << 31), "obst_gemm(tc): too many tiles");
  p.total_tiles = (int)total;
  // A weight gradient whose big operand is B (W_mid: 2 M tiles x 32 N tiles, each slab of dh serves the 2 M tiles):
  // with the M tiles numbered first the two tiles that read the same slab run on neighbouring CTA pairs; numbered
  // N-first they ran 32 pairs apart and half of the slab was fetched from DRAM twice (ncu).  OBST_GEMM_M_FIRST=0: off.
  static const bool m_first_off = [] {
    const char* e = getenv("OBST_GEMM_M_FIRST");
    return e && e[0] == '0';
  }();
  p.m_first = (!m_first_off && g.accumulate && g.tri == 0 && p.tiles_m > 1 && p.tiles_m < p.tiles_n) ? 1 : 0;

  CUtensorMap tmA, tmB;
  int rc = make_operand_map(g.A, g.a_mn, g.a_k, g.a_ko, g.a_glo, g.a_ghi, g.M, g.KI, g.KO, g.GLO, g.GHI, a_mn, BM, &tmA,
                            strict);
  if (rc != OBST_OK) return rc;
  rc = make_operand_map(g.B, g.b_mn, g.b_k, g.b_ko, g.b_glo, g.b_ghi, g.N, g.KI, g.KO, g.GLO, g.GHI, b_mn, p.BN / cg,
                        &tmB, strict);
  if (rc != OBST_OK) return rc;

  CUtensorMap tmD = tmA, tmR = tmA;  // placeholders when the direct-store epilogue is used
  p.tma_store = 0;
  if (!out_f32 && !g.accumulate) {
    int rd = make_out_map(g.D, g.N, g.M, g.GLO, g.GHI, g.d_ld, g.d_glo, g.d_ghi, &tmD, strict);
    int rr = OBST_OK;
    if (rd == OBST_OK && ((g.epi >= 2 && g.epi <= 6) || (g.epi >= 9 && g.epi <= 11)))
      rr = make_out_map(g.R, g.N, g.M, g.GLO, g.GHI, g.d_ld, g.d_glo, g.d_ghi, &tmR, strict);
    if (rd == OBST_ECUDA || rr == OBST_ECUDA) return OBST_ECUDA;
    p.tma_store = (rd == OBST_OK && rr == OBST_OK) ? 1 : 0;
  }
  if (((g.epi >= 5 && g.epi <= 7) || (g.epi >= 9 && g.epi <= 12)) && !p.tma_store) {
    if (strict) set_error("obst_gemm(tc): epilogue %d needs a TMA-addressable output", g.epi);
    return OBST_ENOTSUP;
  }

  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         Cfg<1>::SMEM);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(gemm_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<2>::SMEM);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(gemm_tc_kernel<2, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<2>::SMEM);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(gemm_tc_kernel<2, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<2>::SMEM);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(gemm_tc_kernel<2, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<2>::SMEM);
    if (e != cudaSuccess) {
      set_error("obst_gemm(tc): cannot opt in to %d B of shared memory: %s", Cfg<2>::SMEM, cudaGetErrorString(e));
      return OBST_ECUDA;
    }
    attr_set = true;
  }
  if (cg == 2) {
    int clusters = num_sms() / 2;
    if (clusters > p.total_tiles) clusters = p.total_tiles;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)(2 * clusters));
    cfg.blockDim = dim3(TC_THREADS);
    cfg.dynamicSmemBytes = Cfg<2>::SMEM;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    // the kernel waits (griddepcontrol.wait) before its first global access: its prologue may overlap the tail of
    // the kernel before it.  OBST_GEMM_PDL=0 launches fully serialised (A/B switch).
    static const bool pdl = [] {
      const char* e = getenv("OBST_GEMM_PDL");
      return !(e && e[0] == '0');
    }();
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 2 : 1;
    cudaError_t e = (g.epi == 9 || g.epi == 10) ? cudaLaunchKernelEx(&cfg, gemm_tc_kernel<2, 1>, tmA, tmB, tmD, tmR, p)
                    : g.epi == 11               ? cudaLaunchKernelEx(&cfg, gemm_tc_kernel<2, 2>, tmA, tmB, tmD, tmR, p)
                    : g.epi == 12               ? cudaLaunchKernelEx(&cfg, gemm_tc_kernel<2, 3>, tmA, tmB, tmD, tmR, p)
                                                : cudaLaunchKernelEx(&cfg, gemm_tc_kernel<2>, tmA, tmB, tmD, tmR, p);
    if (e != cudaSuccess) {
      set_error("obst_gemm(tc): cluster launch failed: %s", cudaGetErrorString(e));
      return OBST_ECUDA;
    }
  } else {
    int grid = num_sms();
    if (grid > p.total_tiles) grid = p.total_tiles;
    gemm_tc_kernel<1><<<grid, TC_THREADS, Cfg<1>::SMEM, st>>>(tmA, tmB, tmD, tmR, p);
  }
  count_launch();
  count_path(OBST_PATH_GEMM_TC);
  return check_launch("obst_gemm(tc)");
}

}  // namespace obst

using namespace obst;

extern "C" int obst_gemm(const obst_gemm_desc* d, int backend, void* stream) {
  OBST_REQUIRE(d != nullptr, "obst_gemm: null descriptor");
  OBST_REQUIRE(d->M > 0 && d->N > 0 && d->KI > 0 && d->KO > 0 && d->GLO > 0 && d->GHI > 0,
               "obst_gemm: bad sizes M=%d N=%d KI=%d KO=%d GLO=%d GHI=%d", d->M, d->N, d->KI, d->KO, d->GLO, d->GHI);
  OBST_REQUIRE(d->A.ptr && d->B.ptr && d->D, "obst_gemm: null operand");
  OBST_REQUIRE(d->tri >= 0 && d->tri <= 3 && d->epi >= 0 && d->epi <= 12, "obst_gemm: bad tri/epi %d/%d", d->tri, d->epi);
  OBST_REQUIRE(!(((d->epi >= 2 && d->epi <= 6) || (d->epi >= 9 && d->epi <= 11)) && d->R == nullptr),
               "obst_gemm: epilogue %d needs R",
               d->epi);
  OBST_REQUIRE(!(d->accumulate && d->dtype_out != OBST_F32), "obst_gemm: accumulate needs fp32 output");
  OBST_REQUIRE(!(d->accumulate && d->epi >= 2 && d->epi != 8),
               "obst_gemm: accumulate cannot be combined with an R epilogue");
  OBST_REQUIRE(!(d->epi >= 5 && d->aux0 == nullptr), "obst_gemm: epilogue %d needs aux0", d->epi);
  OBST_REQUIRE(!((d->tri == 1 || d->tri == 2) && d->KO != 1), "obst_gemm: tri 1/2 need KO == 1");
  OBST_REQUIRE(!(d->tri == 3 && !d->accumulate), "obst_gemm: tri 3 needs accumulate");
  GemmParams p;
  p.A = d->A.ptr; p.B = d->B.ptr; p.D = d->D; p.R = d->R;
  p.a_mn = d->A.s_mn; p.a_k = d->A.s_k; p.a_ko = d->A.s_ko; p.a_glo = d->A.s_glo; p.a_ghi = d->A.s_ghi;
  p.b_mn = d->B.s_mn; p.b_k = d->B.s_k; p.b_ko = d->B.s_ko; p.b_glo = d->B.s_glo; p.b_ghi = d->B.s_ghi;
  p.d_ld = d->d_ld; p.d_glo = d->d_glo; p.d_ghi = d->d_ghi;
  p.M = d->M; p.N = d->N; p.KI = d->KI; p.KO = d->KO; p.GLO = d->GLO; p.GHI = d->GHI;
  p.dtype_in = d->dtype_in; p.dtype_out = d->dtype_out;
  p.tri = d->tri; p.epi = d->epi; p.accumulate = d->accumulate; p.splits = d->splits;
  p.aux0 = d->aux0; p.aux1 = d->aux1; p.aux2 = d->aux2; p.aux3 = d->aux3; p.aux4 = d->aux4;
  p.scalar0 = d->scalar0; p.scalar1 = d->scalar1;
  cudaStream_t st = (cudaStream_t)stream;
  if (d->epi >= 5) {  // the fused-norm epilogues exist in the tcgen05 kernel only
    OBST_REQUIRE(backend != 1, "obst_gemm: epilogue %d is not available in the SIMT kernel", d->epi);
    return gemm_tc_launch(p, st, true);
  }
  if (backend == 1) return gemm_simt_launch(p, st);
  if (backend == 2) return gemm_tc_launch(p, st, true);
  OBST_REQUIRE(backend == 0, "obst_gemm: bad backend %d", backend);
  if (p.dtype_in == OBST_BF16) {
    const int rc = gemm_tc_launch(p, st, false);
    if (rc != OBST_ENOTSUP) return rc;
  }
  return gemm_simt_launch(p, st);
}
