// Softmax + cross-entropy forward/backward (K4, K5) and bias gradient.
// Replaces libxsmm_dnn_smax_fwd_exec_f32 / libxsmm_dnn_smax_bwd_exec_f32
// (sw/mlp_mpi_example_f32.cpp:718-728, loss_weight = 0.1 at :250) with one fused kernel:
// one warp per row, the row is read three times out of L1/L2 (max, sum, write), fp32 math.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math_constants.h>

#include "../../include/ainic.h"

namespace ainic {

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(w[i] << 16);
    f[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
  }
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint32_t w[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    __nv_bfloat162 p = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
    w[i] = *reinterpret_cast<uint32_t*>(&p);
  }
  return make_uint4(w[0], w[1], w[2], w[3]);
}

__global__ void __launch_bounds__(128) softmax_xent_kernel(const __nv_bfloat16* __restrict__ logits,
                                                           const int32_t* __restrict__ labels,
                                                           __nv_bfloat16* __restrict__ probs,
                                                           __nv_bfloat16* __restrict__ dlogits,
                                                           float* __restrict__ loss_sum, int mb, int n, int ld,
                                                           float loss_weight) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= mb) return;
  const uint4* src = reinterpret_cast<const uint4*>(logits + (size_t)row * ld);
  const int n8 = n >> 3;
  float m = -CUDART_INF_F;
  for (int i = lane; i < n8; i += 32) {
    float f[8];
    unpack8(src[i], f);
#pragma unroll
    for (int j = 0; j < 8; ++j) m = fmaxf(m, f[j]);
  }
  m = warp_max(m);
  float s = 0.0f;
  for (int i = lane; i < n8; i += 32) {
    float f[8];
    unpack8(src[i], f);
#pragma unroll
    for (int j = 0; j < 8; ++j) s += __expf(f[j] - m);
  }
  s = warp_sum(s);
  const float inv = 1.0f / s;
  const int lab = labels[row];
  const float scale = loss_weight / (float)mb;
  for (int i = lane; i < n8; i += 32) {
    float f[8], p[8], d[8];
    unpack8(src[i], f);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      p[j] = __expf(f[j] - m) * inv;
      d[j] = scale * (p[j] - ((i * 8 + j) == lab ? 1.0f : 0.0f));
    }
    if (probs) reinterpret_cast<uint4*>(probs + (size_t)row * ld)[i] = pack8(p);
    reinterpret_cast<uint4*>(dlogits + (size_t)row * ld)[i] = pack8(d);
  }
  if (lane == 0 && loss_sum) {
    const float l = __bfloat162float(logits[(size_t)row * ld + lab]);
    atomicAdd(loss_sum, -(l - m - __logf(s)));
  }
}

// Same kernel for the all-fp32 ("f32" / tf32) configuration: fp32 logits in, fp32 dlogits out.
__global__ void __launch_bounds__(128) softmax_xent_f32_kernel(const float* __restrict__ logits, const int32_t* __restrict__ labels,
                                                               float* __restrict__ dlogits, float* __restrict__ loss_sum, int mb,
                                                               int n, int ld, float loss_weight) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= mb) return;
  const float4* src = reinterpret_cast<const float4*>(logits + (size_t)row * ld);
  const int n4 = n >> 2;
  float m = -CUDART_INF_F;
  for (int i = lane; i < n4; i += 32) {
    const float4 f = src[i];
    m = fmaxf(fmaxf(m, fmaxf(f.x, f.y)), fmaxf(f.z, f.w));
  }
  m = warp_max(m);
  float s = 0.0f;
  for (int i = lane; i < n4; i += 32) {
    const float4 f = src[i];
    s += __expf(f.x - m) + __expf(f.y - m) + __expf(f.z - m) + __expf(f.w - m);
  }
  s = warp_sum(s);
  const float inv = 1.0f / s;
  const int lab = labels[row];
  const float scale = loss_weight / (float)mb;
  for (int i = lane; i < n4; i += 32) {
    const float4 f = src[i];
    float4 d;
    d.x = scale * (__expf(f.x - m) * inv - ((i * 4 + 0) == lab ? 1.0f : 0.0f));
    d.y = scale * (__expf(f.y - m) * inv - ((i * 4 + 1) == lab ? 1.0f : 0.0f));
    d.z = scale * (__expf(f.z - m) * inv - ((i * 4 + 2) == lab ? 1.0f : 0.0f));
    d.w = scale * (__expf(f.w - m) * inv - ((i * 4 + 3) == lab ? 1.0f : 0.0f));
    reinterpret_cast<float4*>(dlogits + (size_t)row * ld)[i] = d;
  }
  if (lane == 0 && loss_sum) atomicAdd(loss_sum, -(logits[(size_t)row * ld + lab] - m - __logf(s)));
}

// dbias[c] = sum_r dy[r][c]; block = 32 column-pairs x 8 row groups, atomics across row chunks.
__global__ void __launch_bounds__(256) bias_grad_kernel(const __nv_bfloat16* __restrict__ dy, float* __restrict__ dbias,
                                                        int mb, int n, int ld, int rows_per_block) {
  __shared__ float2 red[8][32];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int col = (blockIdx.x * 32 + cx) * 2;
  const int r0 = blockIdx.y * rows_per_block;
  const int r1 = min(mb, r0 + rows_per_block);
  float2 acc = make_float2(0.f, 0.f);
  if (col < n) {
    for (int r = r0 + ry; r < r1; r += 8) {
      const uint32_t u = *reinterpret_cast<const uint32_t*>(dy + (size_t)r * ld + col);
      acc.x += __uint_as_float(u << 16);
      acc.y += __uint_as_float(u & 0xFFFF0000u);
    }
  }
  red[ry][cx] = acc;
  __syncthreads();
  if (ry == 0 && col < n) {
#pragma unroll
    for (int k = 1; k < 8; ++k) {
      acc.x += red[k][cx].x;
      acc.y += red[k][cx].y;
    }
    atomicAdd(dbias + col, acc.x);
    atomicAdd(dbias + col + 1, acc.y);
  }
}

// x[i] = mask[i] ? x[i] : 0  (ReLU backward on the loss gradient when the last layer has a ReLU)
__global__ void __launch_bounds__(256) apply_mask_kernel(__nv_bfloat16* __restrict__ x, const uint8_t* __restrict__ mask, size_t count) {
  const size_t n8 = count >> 3;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
    uint4 v = reinterpret_cast<uint4*>(x)[i];
    const uint2 m = reinterpret_cast<const uint2*>(mask)[i];
    uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const uint32_t mb = ((j < 4 ? m.x : m.y) >> (8 * (j & 3))) & 0xFFu;
      if (!mb) w[j >> 1] &= (j & 1) ? 0x0000FFFFu : 0xFFFF0000u;
    }
    reinterpret_cast<uint4*>(x)[i] = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

}  // namespace ainic

using namespace ainic;

extern "C" int ainic_apply_mask_bf16(void* x, const uint8_t* mask, size_t count, void* stream) {
  if (!x || !mask || count == 0 || (count % 8) || ((uintptr_t)x & 15) || ((uintptr_t)mask & 7)) return AINIC_ERR_ARG;
  size_t blocks = (count / 8 + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  apply_mask_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((__nv_bfloat16*)x, mask, count);
  return cudaGetLastError() == cudaSuccess ? AINIC_OK : AINIC_ERR_CUDA;
}

extern "C" int ainic_softmax_xent(const void* logits, const int32_t* labels, void* probs, void* dlogits,
                                  float* loss_sum, int mb, int n, int ld, float loss_weight, void* stream) {
  if (!logits || !labels || !dlogits || mb <= 0 || n <= 0 || (n % 8) || (ld % 8) || ((uintptr_t)logits & 15) ||
      ((uintptr_t)dlogits & 15) || ((uintptr_t)probs & 15))
    return AINIC_ERR_ARG;
  const int rows_per_block = 4;
  softmax_xent_kernel<<<(mb + rows_per_block - 1) / rows_per_block, 128, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)logits, labels, (__nv_bfloat16*)probs, (__nv_bfloat16*)dlogits, loss_sum, mb, n, ld,
      loss_weight);
  return cudaGetLastError() == cudaSuccess ? AINIC_OK : AINIC_ERR_CUDA;
}

extern "C" int ainic_softmax_xent_f32(const float* logits, const int32_t* labels, float* dlogits, float* loss_sum, int mb, int n,
                                      int ld, float loss_weight, void* stream) {
  if (!logits || !labels || !dlogits || mb <= 0 || n <= 0 || (n % 4) || (ld % 4) || ((uintptr_t)logits & 15) || ((uintptr_t)dlogits & 15))
    return AINIC_ERR_ARG;
  softmax_xent_f32_kernel<<<(mb + 3) / 4, 128, 0, (cudaStream_t)stream>>>(logits, labels, dlogits, loss_sum, mb, n, ld, loss_weight);
  return cudaGetLastError() == cudaSuccess ? AINIC_OK : AINIC_ERR_CUDA;
}

extern "C" int ainic_bias_grad(const void* dy, float* dbias, int mb, int n, int ld, void* stream) {
  if (!dy || !dbias || mb <= 0 || n <= 0 || (n % 2) || (ld % 2)) return AINIC_ERR_ARG;
  cudaStream_t s = (cudaStream_t)stream;
  if (cudaMemsetAsync(dbias, 0, (size_t)n * sizeof(float), s) != cudaSuccess) return AINIC_ERR_CUDA;
  const int rows_per_block = 256;
  dim3 grid((n / 2 + 31) / 32, (mb + rows_per_block - 1) / rows_per_block);
  bias_grad_kernel<<<grid, 256, 0, s>>>((const __nv_bfloat16*)dy, dbias, mb, n, ld, rows_per_block);
  return cudaGetLastError() == cudaSuccess ? AINIC_OK : AINIC_ERR_CUDA;
}
