// Warp-level frame codecs.  A frame is 512 consecutive gradient values; one warp moves one frame
// at a time with fully coalesced accesses on the fp32 side:
//   row i (0..3), lane l  <->  elements i*128 + 4*l .. +3      (v[4*i + j])
// so a 16-value BFP block is 4 adjacent lanes and a 32-value MX block is 8 adjacent lanes; the
// shared exponent is a 2- or 3-step xor-shuffle max - the GPU form of the reference's pipelined
// max tree (hw/max_u.sv:13-103) - and the barrel shifter / leading-zero counter
// (hw/barrel_shifter.sv, hw/a10_leading_zero_counter.sv) become integer shifts and __clz.
//
// Every codec exposes   Raw / load(frame, raw, lane) / unpack(raw, v, lane) / encode(v, frame, lane)
// so callers can issue the loads of several peers back to back (NVLink latency is ~2 us) before
// any of them is decoded.
#pragma once
#include <cuda_bf16.h>

#include "bfp_codec.h"
#include "ptx.cuh"

namespace ainic {

constexpr unsigned kFullMask = 0xffffffffu;

template <int CODEC, int MODE>
struct FrameCodec;

// ---------------------------------------------------------------- fp32 (no compression)
template <int MODE>
struct FrameCodec<0, MODE> {
  static constexpr int kBytes = 2048;
  struct Raw { uint4 q[4]; };
  __device__ static __forceinline__ void load(const uint8_t* frame, Raw& r, int lane) {
#pragma unroll
    for (int i = 0; i < 4; ++i) r.q[i] = ptx::ld_stream_v4(frame + i * 512 + lane * 16);
  }
  __device__ static __forceinline__ void unpack(const Raw& r, float (&v)[16], int lane) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[4 * i] = __uint_as_float(r.q[i].x);
      v[4 * i + 1] = __uint_as_float(r.q[i].y);
      v[4 * i + 2] = __uint_as_float(r.q[i].z);
      v[4 * i + 3] = __uint_as_float(r.q[i].w);
    }
  }
  __device__ static __forceinline__ void encode(const float (&v)[16], uint8_t* frame, int lane) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
      ptx::st_stream_v4(frame + i * 512 + lane * 16,
                        make_uint4(__float_as_uint(v[4 * i]), __float_as_uint(v[4 * i + 1]),
                                   __float_as_uint(v[4 * i + 2]), __float_as_uint(v[4 * i + 3])));
  }
};

// ---------------------------------------------------------------- BFP16x8 (reference layout)
template <int MODE>
struct FrameCodec<1, MODE> {
  static constexpr int kBytes = kBfpFrameBytes;
  struct Raw { uint32_t w[4]; uint32_t e; };
  __device__ static __forceinline__ void load(const uint8_t* frame, Raw& r, int lane) {
#pragma unroll
    for (int i = 0; i < 4; ++i) r.w[i] = ptx::ld_stream_u32(frame + i * 128 + lane * 4);
    r.e = ptx::ld_stream_u8(frame + kBfpExpOff + lane);
  }
  __device__ static __forceinline__ void unpack(const Raw& r, float (&v)[16], int lane) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint32_t E = __shfl_sync(kFullMask, r.e, i * 8 + (lane >> 2));
#pragma unroll
      for (int j = 0; j < 4; ++j) v[4 * i + j] = bfp_dequant((int32_t)(int8_t)(r.w[i] >> (8 * j)), E, MODE);
    }
  }
  __device__ static __forceinline__ void encode(const float (&v)[16], uint8_t* frame, int lane) {
    uint32_t e_lane = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint32_t E = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) E = max(E, bfp_exp_field(__float_as_uint(v[4 * i + j])));
      E = max(E, __shfl_xor_sync(kFullMask, E, 1));
      E = max(E, __shfl_xor_sync(kFullMask, E, 2));
      uint32_t word = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) word |= ((uint32_t)bfp_quant(v[4 * i + j], E, MODE) & 0xFFu) << (8 * j);
      *reinterpret_cast<uint32_t*>(frame + i * 128 + lane * 4) = word;
      // block b = i*8 + lane/4 -> exponent byte b is stored by lane b
      const uint32_t t = __shfl_sync(kFullMask, E, (lane & 7) * 4);
      if ((lane >> 3) == i) e_lane = t;
    }
    frame[kBfpExpOff + lane] = (uint8_t)e_lane;
  }
  // encode into registers (same bytes encode() writes) / store such a frame; used to publish one
  // reduced frame to several ranks without quantising it more than once
  __device__ static __forceinline__ void pack(const float (&v)[16], Raw& r, int lane) {
    uint32_t e_lane = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint32_t E = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) E = max(E, bfp_exp_field(__float_as_uint(v[4 * i + j])));
      E = max(E, __shfl_xor_sync(kFullMask, E, 1));
      E = max(E, __shfl_xor_sync(kFullMask, E, 2));
      uint32_t word = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) word |= ((uint32_t)bfp_quant(v[4 * i + j], E, MODE) & 0xFFu) << (8 * j);
      r.w[i] = word;
      const uint32_t t = __shfl_sync(kFullMask, E, (lane & 7) * 4);
      if ((lane >> 3) == i) e_lane = t;
    }
    r.e = e_lane;
  }
  __device__ static __forceinline__ void store(const Raw& r, uint8_t* frame, int lane) {
#pragma unroll
    for (int i = 0; i < 4; ++i) *reinterpret_cast<uint32_t*>(frame + i * 128 + lane * 4) = r.w[i];
    frame[kBfpExpOff + lane] = (uint8_t)r.e;
  }
  // encode + replace v by what a reader of the frame will decode (no trip through memory)
  __device__ static __forceinline__ void encode_rt(float (&v)[16], uint8_t* frame, int lane) {
    uint32_t e_lane = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint32_t E = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) E = max(E, bfp_exp_field(__float_as_uint(v[4 * i + j])));
      E = max(E, __shfl_xor_sync(kFullMask, E, 1));
      E = max(E, __shfl_xor_sync(kFullMask, E, 2));
      uint32_t word = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int32_t q = bfp_quant(v[4 * i + j], E, MODE);
        word |= ((uint32_t)q & 0xFFu) << (8 * j);
        v[4 * i + j] = bfp_dequant((int32_t)(int8_t)q, E, MODE);
      }
      *reinterpret_cast<uint32_t*>(frame + i * 128 + lane * 4) = word;
      const uint32_t t = __shfl_sync(kFullMask, E, (lane & 7) * 4);
      if ((lane >> 3) == i) e_lane = t;
    }
    frame[kBfpExpOff + lane] = (uint8_t)e_lane;
  }
};

// ---------------------------------------------------------------- MXFP8 (e4m3 + ue8m0 / 32)
template <int MODE>
struct FrameCodec<2, MODE> {
  static constexpr int kBytes = kMxFrameBytes;
  struct Raw { uint32_t w[4]; uint32_t s; };
  __device__ static __forceinline__ void load(const uint8_t* frame, Raw& r, int lane) {
#pragma unroll
    for (int i = 0; i < 4; ++i) r.w[i] = ptx::ld_stream_u32(frame + i * 128 + lane * 4);
    r.s = ptx::ld_stream_u8(frame + kMxScaleOff + (lane & 15));
  }
  __device__ static __forceinline__ void unpack(const Raw& r, float (&v)[16], int lane) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float s = mx_scale(__shfl_sync(kFullMask, r.s, i * 4 + (lane >> 3)));
#pragma unroll
      for (int j = 0; j < 4; ++j) v[4 * i + j] = e4m3_to_f32((r.w[i] >> (8 * j)) & 0xFFu) * s;
    }
  }
  __device__ static __forceinline__ void encode(const float (&v)[16], uint8_t* frame, int lane) {
    uint32_t s_lane = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint32_t E = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) E = max(E, bfp_exp_field(__float_as_uint(v[4 * i + j])));
      E = max(E, __shfl_xor_sync(kFullMask, E, 1));
      E = max(E, __shfl_xor_sync(kFullMask, E, 2));
      E = max(E, __shfl_xor_sync(kFullMask, E, 4));
      const uint32_t sb = mx_scale_byte(E);
      const float inv = mx_inv_scale(sb);
      uint32_t word = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) word |= f32_to_e4m3(v[4 * i + j] * inv) << (8 * j);
      *reinterpret_cast<uint32_t*>(frame + i * 128 + lane * 4) = word;
      // block b = i*4 + lane/8 -> scale byte b is stored by lane b (b < 16)
      const uint32_t t = __shfl_sync(kFullMask, sb, (lane & 3) * 8);
      if ((lane >> 2) == i) s_lane = t;
    }
    if (lane < 16) frame[kMxScaleOff + lane] = (uint8_t)s_lane;
  }
  __device__ static __forceinline__ void pack(const float (&v)[16], Raw& r, int lane) {
    uint32_t s_lane = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint32_t E = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) E = max(E, bfp_exp_field(__float_as_uint(v[4 * i + j])));
      E = max(E, __shfl_xor_sync(kFullMask, E, 1));
      E = max(E, __shfl_xor_sync(kFullMask, E, 2));
      E = max(E, __shfl_xor_sync(kFullMask, E, 4));
      const uint32_t sb = mx_scale_byte(E);
      const float inv = mx_inv_scale(sb);
      uint32_t word = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) word |= f32_to_e4m3(v[4 * i + j] * inv) << (8 * j);
      r.w[i] = word;
      const uint32_t t = __shfl_sync(kFullMask, sb, (lane & 3) * 8);
      if ((lane >> 2) == i) s_lane = t;
    }
    r.s = __shfl_sync(kFullMask, s_lane, lane & 15);   // same lane -> byte mapping as load()
  }
  __device__ static __forceinline__ void store(const Raw& r, uint8_t* frame, int lane) {
#pragma unroll
    for (int i = 0; i < 4; ++i) *reinterpret_cast<uint32_t*>(frame + i * 128 + lane * 4) = r.w[i];
    if (lane < 16) frame[kMxScaleOff + lane] = (uint8_t)r.s;
  }
};

// ---------------------------------------------------------------- bf16
template <int MODE>
struct FrameCodec<3, MODE> {
  static constexpr int kBytes = 1024;
  struct Raw { uint2 q[4]; };
  __device__ static __forceinline__ void load(const uint8_t* frame, Raw& r, int lane) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      r.q[i].x = ptx::ld_stream_u32(frame + i * 256 + lane * 8);
      r.q[i].y = ptx::ld_stream_u32(frame + i * 256 + lane * 8 + 4);
    }
  }
  __device__ static __forceinline__ void unpack(const Raw& r, float (&v)[16], int lane) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[4 * i] = __uint_as_float(r.q[i].x << 16);
      v[4 * i + 1] = __uint_as_float(r.q[i].x & 0xFFFF0000u);
      v[4 * i + 2] = __uint_as_float(r.q[i].y << 16);
      v[4 * i + 3] = __uint_as_float(r.q[i].y & 0xFFFF0000u);
    }
  }
  __device__ static __forceinline__ void pack(const float (&v)[16], Raw& r, int lane) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __nv_bfloat162 a = __floats2bfloat162_rn(v[4 * i], v[4 * i + 1]);
      __nv_bfloat162 b = __floats2bfloat162_rn(v[4 * i + 2], v[4 * i + 3]);
      r.q[i] = make_uint2(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b));
    }
  }
  __device__ static __forceinline__ void store(const Raw& r, uint8_t* frame, int lane) {
#pragma unroll
    for (int i = 0; i < 4; ++i) *reinterpret_cast<uint2*>(frame + i * 256 + lane * 8) = r.q[i];
  }
  __device__ static __forceinline__ void encode(const float (&v)[16], uint8_t* frame, int lane) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __nv_bfloat162 a = __floats2bfloat162_rn(v[4 * i], v[4 * i + 1]);
      __nv_bfloat162 b = __floats2bfloat162_rn(v[4 * i + 2], v[4 * i + 3]);
      *reinterpret_cast<uint2*>(frame + i * 256 + lane * 8) =
          make_uint2(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b));
    }
  }
};

template <int CODEC, int MODE>
__device__ __forceinline__ void frame_decode(const uint8_t* frame, float (&v)[16], int lane) {
  typename FrameCodec<CODEC, MODE>::Raw r;
  FrameCodec<CODEC, MODE>::load(frame, r, lane);
  FrameCodec<CODEC, MODE>::unpack(r, v, lane);
}

// Encode v into `frame` and leave in v exactly what a reader of the frame will decode.  BFP does it
// in registers; the other codecs read their own bytes back (L2 hit).
template <int CODEC, int MODE>
__device__ __forceinline__ void frame_encode_rt(float (&v)[16], uint8_t* frame, int lane) {
  if constexpr (CODEC == 1) {
    FrameCodec<1, MODE>::encode_rt(v, frame, lane);
  } else {
    FrameCodec<CODEC, MODE>::encode(v, frame, lane);
    __syncwarp();
    frame_decode<CODEC, MODE>(frame, v, lane);
  }
}

// ---------------------------------------------------------------- fp32 frame load / store with tail predication
// The reference pads with dummy reads and suppresses the writes (hw/all_reduce.sv:403-409,
// :1306-1310); here the tail is predicated.  `valid` = number of valid values in this frame.
__device__ __forceinline__ void frame_load_f32(const float* base, int valid, float (&v)[16], int lane) {
  if (valid >= kFrameVals) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint4 u = ptx::ld_stream_v4(base + i * 128 + lane * 4);
      v[4 * i] = __uint_as_float(u.x);
      v[4 * i + 1] = __uint_as_float(u.y);
      v[4 * i + 2] = __uint_as_float(u.z);
      v[4 * i + 3] = __uint_as_float(u.w);
    }
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int idx = i * 128 + lane * 4 + j;
        v[4 * i + j] = idx < valid ? base[idx] : 0.0f;
      }
  }
}
__device__ __forceinline__ void frame_store_f32(float* base, int valid, const float (&v)[16], int lane) {
  if (valid >= kFrameVals) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
      *reinterpret_cast<float4*>(base + i * 128 + lane * 4) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int idx = i * 128 + lane * 4 + j;
        if (idx < valid) base[idx] = v[4 * i + j];
      }
  }
}
__device__ __forceinline__ void frame_store_bf16(__nv_bfloat16* base, int valid, const float (&v)[16], int lane) {
  if (valid >= kFrameVals) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __nv_bfloat162 a = __floats2bfloat162_rn(v[4 * i], v[4 * i + 1]);
      __nv_bfloat162 b = __floats2bfloat162_rn(v[4 * i + 2], v[4 * i + 3]);
      *reinterpret_cast<uint2*>(base + i * 128 + lane * 4) =
          make_uint2(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b));
    }
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int idx = i * 128 + lane * 4 + j;
        if (idx < valid) base[idx] = __float2bfloat16_rn(v[4 * i + j]);
      }
  }
}

}  // namespace ainic
