// Shared device/host helpers for the obst_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/obst_b200.h"

namespace obst {

// ---- host-side error + launch bookkeeping (defined in api.cu) ------------------------------
void set_error(const char* fmt, ...);
void count_launch(int n = 1);
void count_path(int which);  // OBST_PATH_*: which specialised kernel a call was routed to
int check_launch(const char* what);  // cudaGetLastError -> OBST_ECUDA

#define OBST_REQUIRE(cond, ...)            \
  do {                                     \
    if (!(cond)) {                         \
      ::obst::set_error(__VA_ARGS__);      \
      return OBST_EINVAL;                  \
    }                                      \
  } while (0)

typedef __nv_bfloat16 bf16;

// ---- scalar conversions ----------------------------------------------------------------------
__device__ __forceinline__ float to_f32(float v) { return v; }
__device__ __forceinline__ float to_f32(bf16 v) { return __bfloat162float(v); }
template <typename T>
__device__ __forceinline__ T from_f32(float v);
template <>
__device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ bf16 from_f32<bf16>(float v) { return __float2bfloat16_rn(v); }

// round-trip through T: the value the graph "sees" when a fp32 slice is cast to the activation dtype
template <typename T>
__device__ __forceinline__ float round_to(float v) { return to_f32(from_f32<T>(v)); }

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t u) {
  __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
  return __bfloat1622float2(v);
}

// ---- vector of 8 (bf16) / 4 (fp32) elements = one 16-byte access ------------------------------
template <typename T>
struct Vec16;
template <>
struct Vec16<float> {
  static constexpr int N = 4;
  float v[4];
  __device__ __forceinline__ void load(const float* p) {
    float4 t = *reinterpret_cast<const float4*>(p);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
  __device__ __forceinline__ void store(float* p) const {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
};
template <>
struct Vec16<bf16> {
  static constexpr int N = 8;
  float v[8];
  __device__ __forceinline__ void load(const bf16* p) {
    uint4 t = *reinterpret_cast<const uint4*>(p);
    float2 a = unpack_bf16x2(t.x), b = unpack_bf16x2(t.y), c = unpack_bf16x2(t.z), d = unpack_bf16x2(t.w);
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
  }
  __device__ __forceinline__ void store(bf16* p) const {
    uint4 t;
    t.x = pack_bf16x2(v[0], v[1]); t.y = pack_bf16x2(v[2], v[3]);
    t.z = pack_bf16x2(v[4], v[5]); t.w = pack_bf16x2(v[6], v[7]);
    *reinterpret_cast<uint4*>(p) = t;
  }
};

// ---- the same 16 bytes kept PACKED in registers (4 regs) and unpacked on use: for kernels that must hold
// several rows / operands at once and would otherwise lose occupancy to fp32 copies
template <typename T>
struct Raw16;
template <>
struct Raw16<float> {
  static constexpr int N = 4;
  float4 r;
  __device__ __forceinline__ void load(const float* p) { r = *reinterpret_cast<const float4*>(p); }
  __device__ __forceinline__ void store(float* p) const { *reinterpret_cast<float4*>(p) = r; }
  __device__ __forceinline__ void unpack(float (&v)[4]) const { v[0] = r.x; v[1] = r.y; v[2] = r.z; v[3] = r.w; }
  __device__ __forceinline__ void pack(const float (&v)[4]) { r = make_float4(v[0], v[1], v[2], v[3]); }
};
template <>
struct Raw16<bf16> {
  static constexpr int N = 8;
  uint4 r;
  __device__ __forceinline__ void load(const bf16* p) { r = *reinterpret_cast<const uint4*>(p); }
  __device__ __forceinline__ void store(bf16* p) const { *reinterpret_cast<uint4*>(p) = r; }
  __device__ __forceinline__ void unpack(float (&v)[8]) const {
    const float2 a = unpack_bf16x2(r.x), b = unpack_bf16x2(r.y), c = unpack_bf16x2(r.z), d = unpack_bf16x2(r.w);
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
  }
  __device__ __forceinline__ void pack(const float (&v)[8]) {
    r.x = pack_bf16x2(v[0], v[1]); r.y = pack_bf16x2(v[2], v[3]);
    r.z = pack_bf16x2(v[4], v[5]); r.w = pack_bf16x2(v[6], v[7]);
  }
};

// ---- warp reductions ---------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---- the reference's activation (src/model/activation.py:158-161): NOT the textbook tanh-GELU --
// 0.5 x (1 + tanh(u)), u = 0.044715 x^3 + sqrt(2/pi) x, is x * sigmoid(2u) exactly; written that way it costs one
// MUFU.EX2 and one MUFU.RCP instead of libdevice's tanhf (the GN kernels with the fused activation ran at half the
// bandwidth of the plain ones with tanhf).  ex2.approx / rcp.approx are ~1e-7 relative: inside the fp32 parity bar.
__device__ __forceinline__ float gelu_obst_sigmoid(float x) {
  // exp(-2u) = 2^(x * (a x^2 + b)),  a = -2 * 0.044715 * log2(e),  b = -2 * sqrt(2/pi) * log2(e)
  const float t = x * fmaf(x * x, -0.12902022f, -2.3022082f);
  float e;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(t));
  return __fdividef(1.0f, 1.0f + e);
}
// FAST (16-bit activations): tanh.approx.f32 is one MUFU op with ~2^-11 relative error, a quarter of a bf16 ulp --
// the GN kernels with the fused activation are MUFU/issue bound, not HBM bound, so the second MUFU op matters.
__device__ __forceinline__ float tanh_approx(float u) {
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
  return t;
}
template <bool FAST>
__device__ __forceinline__ float gelu_obst(float x) {
  if (FAST) {
    const float hx = 0.5f * x;
    return fmaf(hx, tanh_approx(x * fmaf(x * x, 0.044715f, 0.7978845608028654f)), hx);
  }
  return x * gelu_obst_sigmoid(x);
}
template <bool FAST>
__device__ __forceinline__ float gelu_obst_grad(float x) {
  if (FAST) {
    const float x2 = x * x;
    const float th = tanh_approx(x * fmaf(x2, 0.044715f, 0.7978845608028654f));
    // 0.5 (1 + th) + 0.5 x (1 - th^2) u',  u' = 3 * 0.044715 x^2 + sqrt(2/pi)
    const float a = fmaf(0.5f, th, 0.5f);
    return fmaf(x * fmaf(-th, th, 1.0f), fmaf(x2, 0.0670725f, 0.3989422804014327f), a);
  }
  const float s = gelu_obst_sigmoid(x);
  // d/dx [x s(2u)] = s + x s (1 - s) 2u',  2u' = 6 * 0.044715 x^2 + 2 sqrt(2/pi)
  return s * fmaf(x * (1.0f - s), fmaf(x * x, 0.26829f, 1.5957691216057308f), 1.0f);
}

// pointwise.cu: activations other than relu / gelu (obst_act_fwd / obst_act_bwd route kinds >= 2 here)
int act_any_launch(int dtype, int kind, bool bwd, const void* x, const void* dy, void* out, int64_t n, cudaStream_t st);

// gn_fwd_ring.cu: bf16 rows of 256 / 512 / 1024 elements; returns 1 when the shape is not covered
int gn_fwd_ring_launch(const void* x, const void* gamma, const void* beta, void* y, float2* stats, int64_t rows,
                       int heads, int width, int act, cudaStream_t st);

// gn_bwd_ring.cu: bf16 rows of 256 / 512 / 1024 elements with saved statistics; returns 1 when not covered
int gn_bwd_ring_launch(const void* dy, const void* x, const void* gamma, const void* beta, void* dx, float* dgamma,
                       float* dbeta, const float2* stats, int64_t rows, int heads, int width, int act, int relu_mask,
                       int dx_acc, cudaStream_t st);

static inline int num_sms() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
  }
  return sms;
}

}  // namespace obst
