// Standalone compress / decompress kernels (K8, K8', K14): the GPU form of the reference's BFP
// adapter pipe hw/bfp_adapter.sv (fp32 flits -> 16 mantissa flits + 1 exponent flit per frame and
// back).  HBM-bound: 4 B read + 1.0625 B written per value.  One warp per frame, grid-stride.
#include <cuda_runtime.h>

#include "../../include/ainic.h"
#include "../common/frame.cuh"

namespace ainic {

template <int CODEC, int MODE>
__global__ void __launch_bounds__(256) compress_kernel(const float* __restrict__ x, size_t count, uint8_t* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const size_t frames = bfp_frames(count);
  const size_t warps = (size_t)gridDim.x * (blockDim.x >> 5);
  for (size_t f = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); f < frames; f += warps) {
    float v[16];
    const size_t base = f * kFrameVals;
    const int valid = (int)(count - base < (size_t)kFrameVals ? count - base : (size_t)kFrameVals);
    frame_load_f32(x + base, valid, v, lane);
    FrameCodec<CODEC, MODE>::encode(v, out + f * FrameCodec<CODEC, MODE>::kBytes, lane);
  }
}

template <int CODEC, int MODE>
__global__ void __launch_bounds__(256) decompress_kernel(const uint8_t* __restrict__ in, size_t count, float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const size_t frames = bfp_frames(count);
  const size_t warps = (size_t)gridDim.x * (blockDim.x >> 5);
  for (size_t f = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); f < frames; f += warps) {
    float v[16];
    frame_decode<CODEC, MODE>(in + f * FrameCodec<CODEC, MODE>::kBytes, v, lane);
    const size_t base = f * kFrameVals;
    const int valid = (int)(count - base < (size_t)kFrameVals ? count - base : (size_t)kFrameVals);
    frame_store_f32(out + base, valid, v, lane);
  }
}

static int grid_for(size_t count) {
  const size_t frames = bfp_frames(count);
  size_t blocks = (frames + 7) / 8;
  const size_t cap = 148 * 8;
  return (int)(blocks < 1 ? 1 : (blocks > cap ? cap : blocks));
}

}  // namespace ainic

using namespace ainic;

extern "C" size_t ainic_bfp_bytes(size_t count) { return bfp_bytes(count); }
extern "C" size_t ainic_mx_bytes(size_t count) { return mx_bytes(count); }

#define CHECK_LAUNCH() (cudaGetLastError() == cudaSuccess ? AINIC_OK : AINIC_ERR_CUDA)

extern "C" int ainic_bfp_compress(const float* x, size_t count, uint8_t* out, int mode, void* stream) {
  if (!x || !out || count == 0 || ((uintptr_t)x & 15) || ((uintptr_t)out & 15)) return AINIC_ERR_ARG;
  cudaStream_t s = (cudaStream_t)stream;
  if (mode == AINIC_BFP_HW_EXACT) compress_kernel<1, 0><<<grid_for(count), 256, 0, s>>>(x, count, out);
  else compress_kernel<1, 1><<<grid_for(count), 256, 0, s>>>(x, count, out);
  return CHECK_LAUNCH();
}
extern "C" int ainic_bfp_decompress(const uint8_t* in, size_t count, float* out, int mode, void* stream) {
  if (!in || !out || count == 0 || ((uintptr_t)in & 15) || ((uintptr_t)out & 15)) return AINIC_ERR_ARG;
  cudaStream_t s = (cudaStream_t)stream;
  if (mode == AINIC_BFP_HW_EXACT) decompress_kernel<1, 0><<<grid_for(count), 256, 0, s>>>(in, count, out);
  else decompress_kernel<1, 1><<<grid_for(count), 256, 0, s>>>(in, count, out);
  return CHECK_LAUNCH();
}
extern "C" int ainic_mx_compress(const float* x, size_t count, uint8_t* out, void* stream) {
  if (!x || !out || count == 0 || ((uintptr_t)x & 15) || ((uintptr_t)out & 15)) return AINIC_ERR_ARG;
  compress_kernel<2, 0><<<grid_for(count), 256, 0, (cudaStream_t)stream>>>(x, count, out);
  return CHECK_LAUNCH();
}
extern "C" int ainic_mx_decompress(const uint8_t* in, size_t count, float* out, void* stream) {
  if (!in || !out || count == 0 || ((uintptr_t)in & 15) || ((uintptr_t)out & 15)) return AINIC_ERR_ARG;
  decompress_kernel<2, 0><<<grid_for(count), 256, 0, (cudaStream_t)stream>>>(in, count, out);
  return CHECK_LAUNCH();
}
