// Flattened, validated form of obst_gemm_desc shared by the SIMT and tcgen05 back ends.
#pragma once
#include "obst_common.cuh"

namespace obst {

struct GemmParams {
  const void* A;
  const void* B;
  void* D;
  const void* R;
  const void* aux0;
  const void* aux1;
  const void* aux2;
  const void* aux3;
  void* aux4;
  int64_t a_mn, a_k, a_ko, a_glo, a_ghi;
  int64_t b_mn, b_k, b_ko, b_glo, b_ghi;
  int64_t d_ld, d_glo, d_ghi;
  int32_t M, N, KI, KO, GLO, GHI;
  int32_t dtype_in, dtype_out;
  int32_t tri, epi, accumulate, splits;
  float scalar0, scalar1;
};

int gemm_simt_launch(const GemmParams& p, cudaStream_t st);
// returns OBST_ENOTSUP (without touching the error text unless `strict`) when the geometry is not TMA-legal
int gemm_tc_launch(const GemmParams& p, cudaStream_t st, bool strict);

}  // namespace obst
