// The activations the reference registers beyond relu / gelu (src/model/activation.py:13-155,168-200) and the
// sigmoid gate of the GLU variants of activated_linear (src/model/basic.py:51-54).  None of them is on the hot path
// of the three benchmark configs (those use relu and the fused norm+gelu), so this is one plain grid-stride kernel
// per direction with the kind as a runtime switch.
//
// silu and mish carry the reference's OWN backward formulas (its custom MTF operations, activation.py:35-47,71-81),
// which are not the analytic derivatives; mtf_mish is differentiated by MTF itself, i.e. analytically.
#include "obst_common.cuh"

namespace obst {

__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }
__device__ __forceinline__ float softplusf_(float x) { return x > 20.f ? x : log1pf(expf(x)); }

__device__ __forceinline__ float act_fwd_any(int kind, float x) {
  switch (kind) {
    case OBST_ACT_SIGMOID: return sigmoidf_(x);
    case OBST_ACT_TANH: return tanhf(x);
    case OBST_ACT_LECUN_TANH: return tanhf(x) + x * 0.1f;                  // activation.py:94-95
    case OBST_ACT_SILU: return x * sigmoidf_(x);                           // activation.py:59-60
    case OBST_ACT_MISH:
    case OBST_ACT_MTF_MISH: return x * tanhf(softplusf_(x));               // activation.py:24-25,180-181
    case OBST_ACT_SOFTSIGN: return x / (1.0f + fabsf(x));                  // activation.py:129-130
    case OBST_ACT_EXP: return expf(x);
    default: return x;
  }
}

__device__ __forceinline__ float act_bwd_any(int kind, float x, float dy) {
  switch (kind) {
    case OBST_ACT_SIGMOID: {
      const float s = sigmoidf_(x);
      return dy * s * (1.0f - s);
    }
    case OBST_ACT_TANH: {
      const float t = tanhf(x);
      return dy * (1.0f - t * t);
    }
    case OBST_ACT_LECUN_TANH: {  // activation.py:112-113
      const float t = tanhf(x);
      return dy * (1.1f - t * t);
    }
    case OBST_ACT_SILU:  // activation.py:77-79: dy * ((x - 1) * sigmoid(x) + 1)
      return dy * ((x - 1.0f) * sigmoidf_(x) + 1.0f);
    case OBST_ACT_MISH: {  // activation.py:41-44: gte = tanh(softplus(x)); gte += 1 - gte^2 * x * sigmoid(x)
      float gte = tanhf(softplusf_(x));
      gte += 1.0f - gte * gte * x * sigmoidf_(x);
      return dy * gte;
    }
    case OBST_ACT_MTF_MISH: {  // analytic: tanh(sp) + x * (1 - tanh(sp)^2) * sigmoid(x)
      const float t = tanhf(softplusf_(x));
      return dy * (t + x * (1.0f - t * t) * sigmoidf_(x));
    }
    case OBST_ACT_SOFTSIGN: {  // activation.py:147-148
      const float d = 1.0f + fabsf(x);
      return dy / (d * d);
    }
    case OBST_ACT_EXP: return dy * expf(x);
    default: return dy;
  }
}

template <typename T, bool BWD>
__global__ void __launch_bounds__(256) act_any_kernel(int kind, const T* __restrict__ x, const T* __restrict__ dy,
                                                       T* __restrict__ out, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float v = to_f32(x[i]);
    out[i] = from_f32<T>(BWD ? act_bwd_any(kind, v, to_f32(dy[i])) : act_fwd_any(kind, v));
  }
}

// out = a * sigmoid(b)   (GLU gate);   da = dy * sigmoid(b),  db = dy * a * sigmoid(b) * (1 - sigmoid(b))
template <typename T, bool BWD>
__global__ void __launch_bounds__(256) gate_kernel(const T* __restrict__ a, const T* __restrict__ b,
                                                    const T* __restrict__ dy, T* __restrict__ out0,
                                                    T* __restrict__ out1, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float av = to_f32(a[i]);
    const float s = sigmoidf_(to_f32(b[i]));
    if (!BWD) {
      out0[i] = from_f32<T>(av * s);
    } else {
      const float g = to_f32(dy[i]);
      out0[i] = from_f32<T>(g * s);
      out1[i] = from_f32<T>(g * av * s * (1.0f - s));
    }
  }
}

static unsigned pw_blocks(int64_t n) {
  int64_t b = (n + 255) / 256;
  const int64_t cap = (int64_t)num_sms() * 16;
  return (unsigned)(b < 1 ? 1 : (b > cap ? cap : b));
}

int act_any_launch(int dtype, int kind, bool bwd, const void* x, const void* dy, void* out, int64_t n,
                   cudaStream_t st) {
  const unsigned blocks = pw_blocks(n);
  if (dtype == OBST_BF16) {
    if (bwd) act_any_kernel<bf16, true><<<blocks, 256, 0, st>>>(kind, (const bf16*)x, (const bf16*)dy, (bf16*)out, n);
    else act_any_kernel<bf16, false><<<blocks, 256, 0, st>>>(kind, (const bf16*)x, nullptr, (bf16*)out, n);
  } else if (dtype == OBST_F32) {
    if (bwd) act_any_kernel<float, true><<<blocks, 256, 0, st>>>(kind, (const float*)x, (const float*)dy, (float*)out, n);
    else act_any_kernel<float, false><<<blocks, 256, 0, st>>>(kind, (const float*)x, nullptr, (float*)out, n);
  } else {
    set_error("obst_act: bad dtype %d", dtype);
    return OBST_EINVAL;
  }
  count_launch();
  return check_launch("obst_act");
}

}  // namespace obst

using namespace obst;

extern "C" {

int obst_gate_fwd(int dtype, const void* a, const void* b, void* out, int64_t n, void* stream) {
  OBST_REQUIRE(n >= 0, "obst_gate_fwd: bad size");
  if (n == 0) return OBST_OK;
  OBST_REQUIRE(a && b && out, "obst_gate_fwd: null argument");
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned blocks = pw_blocks(n);
  if (dtype == OBST_BF16)
    gate_kernel<bf16, false><<<blocks, 256, 0, st>>>((const bf16*)a, (const bf16*)b, nullptr, (bf16*)out, nullptr, n);
  else if (dtype == OBST_F32)
    gate_kernel<float, false><<<blocks, 256, 0, st>>>((const float*)a, (const float*)b, nullptr, (float*)out, nullptr,
                                                       n);
  else
    OBST_REQUIRE(false, "obst_gate_fwd: bad dtype %d", dtype);
  count_launch();
  return check_launch("obst_gate_fwd");
}

int obst_gate_bwd(int dtype, const void* a, const void* b, const void* dy, void* da, void* db, int64_t n,
                  void* stream) {
  OBST_REQUIRE(n >= 0, "obst_gate_bwd: bad size");
  if (n == 0) return OBST_OK;
  OBST_REQUIRE(a && b && dy && da && db, "obst_gate_bwd: null argument");
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned blocks = pw_blocks(n);
  if (dtype == OBST_BF16)
    gate_kernel<bf16, true><<<blocks, 256, 0, st>>>((const bf16*)a, (const bf16*)b, (const bf16*)dy, (bf16*)da,
                                                     (bf16*)db, n);
  else if (dtype == OBST_F32)
    gate_kernel<float, true><<<blocks, 256, 0, st>>>((const float*)a, (const float*)b, (const float*)dy, (float*)da,
                                                      (float*)db, n);
  else
    OBST_REQUIRE(false, "obst_gate_bwd: bad dtype %d", dtype);
  count_launch();
  return check_launch("obst_gate_bwd");
}

}  // extern "C"
