// Fused SGD weight update (K9): w = fmaf(-lr, g, w), the 16-lane ffma array of
// hw/weight_update.sv:440-452 with a runtime learning rate instead of the hard-wired -0.1
// (hw/weight_update.sv:439-444).  Optionally emits the bf16 shadow the tcgen05 GEMMs
// read, so no separate cast kernel runs.  HBM-bound: 12 B/value (+2 B with the shadow).
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "../../include/ainic.h"

namespace ainic {

template <bool SHADOW>
__global__ void __launch_bounds__(256) sgd_kernel(float* __restrict__ w, const float* __restrict__ g,
                                                  __nv_bfloat16* __restrict__ wb, size_t count, float lr) {
  const size_t n4 = count >> 2;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 wv = reinterpret_cast<float4*>(w)[i];
    const float4 gv = __ldcs(reinterpret_cast<const float4*>(g) + i);
    wv.x = fmaf(-lr, gv.x, wv.x);
    wv.y = fmaf(-lr, gv.y, wv.y);
    wv.z = fmaf(-lr, gv.z, wv.z);
    wv.w = fmaf(-lr, gv.w, wv.w);
    reinterpret_cast<float4*>(w)[i] = wv;
    if (SHADOW) {
      __nv_bfloat162 a = __floats2bfloat162_rn(wv.x, wv.y);
      __nv_bfloat162 b = __floats2bfloat162_rn(wv.z, wv.w);
      reinterpret_cast<uint2*>(wb)[i] = make_uint2(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b));
    }
  }
  // tail (count % 4)
  const size_t t = (n4 << 2) + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < count) {
    const float nw = fmaf(-lr, g[t], w[t]);
    w[t] = nw;
    if (SHADOW) wb[t] = __float2bfloat16_rn(nw);
  }
}

__global__ void __launch_bounds__(256) cast_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y, size_t count) {
  const size_t n4 = count >> 2;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const float4 v = reinterpret_cast<const float4*>(x)[i];
    __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y);
    __nv_bfloat162 b = __floats2bfloat162_rn(v.z, v.w);
    reinterpret_cast<uint2*>(y)[i] = make_uint2(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b));
  }
  const size_t t = (n4 << 2) + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < count) y[t] = __float2bfloat16_rn(x[t]);
}

static int ew_grid(size_t count) {
  size_t b = (count / 4 + 255) / 256;
  const size_t cap = 148 * 8;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace ainic

using namespace ainic;

extern "C" int ainic_sgd(float* w, const float* g, void* w_bf16, size_t count, float lr, void* stream) {
  if (!w || !g || count == 0 || ((uintptr_t)w & 15) || ((uintptr_t)g & 15) || ((uintptr_t)w_bf16 & 7)) return AINIC_ERR_ARG;
  cudaStream_t s = (cudaStream_t)stream;
  if (w_bf16) sgd_kernel<true><<<ew_grid(count), 256, 0, s>>>(w, g, (__nv_bfloat16*)w_bf16, count, lr);
  else sgd_kernel<false><<<ew_grid(count), 256, 0, s>>>(w, g, nullptr, count, lr);
  return cudaGetLastError() == cudaSuccess ? AINIC_OK : AINIC_ERR_CUDA;
}

extern "C" int ainic_cast_f32_bf16(const float* x, void* y, size_t count, void* stream) {
  if (!x || !y || count == 0 || ((uintptr_t)x & 15) || ((uintptr_t)y & 7)) return AINIC_ERR_ARG;
  cast_kernel<<<ew_grid(count), 256, 0, (cudaStream_t)stream>>>(x, (__nv_bfloat16*)y, count);
  return cudaGetLastError() == cudaSuccess ? AINIC_OK : AINIC_ERR_CUDA;
}
