// Pieces shared by the 1-CTA and 2-CTA (cta_group::2) tcgen05 GEMM kernels: tile constants,
// shared-memory / instruction descriptors, the fused epilogue, TMA tensor-map construction.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../include/ainic.h"
#include "../common/bfp_codec.h"
#include "../common/ptx.cuh"

namespace ainic {
namespace gemm {

using namespace ptx;

constexpr int BLOCK_M = 128;   // rows per CTA (TMEM has 128 lanes)
constexpr int BLOCK_K = 64;    // 64 bf16 = 128 B = one swizzle span
constexpr int UMMA_K = 16;
// warp 0 TMA, warp 1 MMA + TMEM alloc, warps 2-9 epilogue.  Most of the MLP's GEMMs have one tile
// per CTA, so the epilogue is not hidden behind a next main loop: eight warps (two per TMEM lane
// quarter, each draining half of the columns) halve that exposed tail.
constexpr int kThreads = 320;
constexpr int kEpilogueWarps = 8;
constexpr int kSmemBudget = 220 * 1024;

struct EpiParams {
  void* c;                 // output, row-major [M][ldc]; bf16 or fp32
  const float* bias;       // [N] or null (added before activation)
  uint8_t* mask_out;       // [M][ldc] or null: 1 where the (biased) result is > 0 (ReLU mask)
  const uint8_t* mask_in;  // [M][ldc] or null: output zeroed where mask_in == 0 (ReLU backward)
  float alpha;             // result scale (fp32 / compressed output)
  int relu;                // apply max(0, x)
  int accumulate;          // fp32 output only: c += result
  int M, N, K, ldc;
  uint8_t* cgrad;          // compressed output (OUT_COMPRESSED): frames of the flat [M][ldc] bucket
  int codec, bfp_mode;     // AINIC_CODEC_BFP / AINIC_CODEC_MXFP8, AINIC_BFP_*
  int epi_sleep;           // ns the epilogue warps sleep between polls of the accumulator barrier (0 = spin)
  // reduce-scatter in the epilogue (ainic_gemm_scatter_t): frame f -> inbox of rank f / shard
  uint8_t* push[AINIC_MAX_RANKS];
  int push_n, push_rank;   // push_n <= 1: frames go to cgrad
  uint32_t push_shard, push_frame0;
  int stage_frames;        // compressed output leaves through shared memory (epilogue_store_frames)
  // "ready" doorbell for a consumer kernel that is already running (the flow all-reduce): the last CTA to
  // finish bumps *done_flag once every tile of this launch has been stored (release, gpu scope)
  uint32_t* done_flag;
  uint32_t* done_ticket;
  // plain bf16 / fp32 output of a CTA's LAST tile leaves through shared memory and TMA stores
  // (epilogue_store_tma); set by the host when no mask / accumulate / doorbell needs per-element access
  int tma_out;
  // diagnostics (ainic_gemm_trace): 8 time stamps per CTA - entry, after griddepcontrol.wait, first operand
  // stage landed, last MMA issued, first accumulator complete, epilogue done, exit; null = off
  unsigned long long* trace;
};

// Output selector of the epilogue.
// OUT_COMPRESSED_STAGED: same bytes as OUT_COMPRESSED, written through shared memory (epilogue_store_frames);
// a separate instantiation so that the direct variant keeps its register count and shared-memory size.
constexpr int OUT_BF16 = 0, OUT_FP32 = 1, OUT_COMPRESSED = 2, OUT_COMPRESSED_STAGED = 3;

// 32 fp32 values of one row -> two BFP16x8 blocks: w[0..7] = int8 mantissas, return = the two shared
// exponents (low byte first).  Same bits as bfp_exp_field / bfp_quant value by value; the rounding mode
// is a template parameter so the per-value code has no branch, and the shared exponent comes from one
// integer max over the magnitudes (the exponent field is their top byte).
template <int MODE>
__device__ __forceinline__ uint32_t bfp_pack32(const float (&v)[32], uint32_t (&w)[8]) {
  uint32_t exps = 0;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    uint32_t mag = 0;
#pragma unroll
    for (int j = 0; j < 16; ++j) mag = max(mag, __float_as_uint(v[16 * h + j]) & 0x7FFFFFFFu);
    const uint32_t E = mag >> 23;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int32_t q = MODE == AINIC_BFP_HW_EXACT ? bfp_quant_hw(__float_as_uint(v[16 * h + j]), E) : bfp_quant_clean(v[16 * h + j], E);
      w[4 * h + (j >> 2)] |= ((uint32_t)q & 0xFFu) << (8 * (j & 3));
    }
    exps |= E << (8 * h);
  }
  return exps;
}

// Where frame `frame` of this GEMM's output goes: the local buffer, or (reduce-scatter epilogue) the
// inbox of the rank that owns it.
__device__ __forceinline__ uint8_t* frame_dst(const EpiParams& ep, size_t frame, size_t fbytes) {
  if (ep.push_n > 1) {
    const uint32_t F = ep.push_frame0 + (uint32_t)frame;
    const uint32_t owner = F / ep.push_shard;
    return ep.push[owner] + ((size_t)ep.push_rank * ep.push_shard + (F - owner * ep.push_shard)) * fbytes;
  }
  return ep.cgrad + frame * fbytes;
}

// 32 consecutive values of one row -> two BFP16x8 blocks or one MXFP8 block, written straight into
// the frame layout the all-reduce kernels read (512 mantissa bytes + 32 exponent / 16 scale bytes
// per 512 values, csrc/common/bfp_codec.h).  This is the reference's BFP adapter sitting directly
// behind the weight-gradient GEMM: the fp32 gradient never goes to memory and the all-reduce
// kernel skips its compression pass.  idx = flat index of v[0] (a multiple of 32).
__device__ __forceinline__ void store_compressed32(const EpiParams& ep, size_t idx, const float (&v)[32]) {
  const size_t frame = idx >> 9;
  const uint32_t off = (uint32_t)(idx & 511);
  const size_t fbytes = ep.codec == AINIC_CODEC_BFP ? (size_t)kBfpFrameBytes : (size_t)kMxFrameBytes;
  uint8_t* fr = frame_dst(ep, frame, fbytes);
  if (ep.codec == AINIC_CODEC_BFP) {
    uint32_t w[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const uint32_t exps = ep.bfp_mode == AINIC_BFP_HW_EXACT ? bfp_pack32<AINIC_BFP_HW_EXACT>(v, w) : bfp_pack32<AINIC_BFP_CLEAN>(v, w);
    *reinterpret_cast<uint4*>(fr + off) = make_uint4(w[0], w[1], w[2], w[3]);
    *reinterpret_cast<uint4*>(fr + off + 16) = make_uint4(w[4], w[5], w[6], w[7]);
    *reinterpret_cast<uint16_t*>(fr + kBfpExpOff + (off >> 4)) = (uint16_t)exps;
  } else {  // MXFP8
    uint32_t E = 0;
#pragma unroll
    for (int j = 0; j < 32; ++j) E = max(E, bfp_exp_field(__float_as_uint(v[j])));
    const uint32_t sb = mx_scale_byte(E);
    const float inv = mx_inv_scale(sb);
    uint32_t w[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int j = 0; j < 32; ++j) w[j >> 2] |= f32_to_e4m3(v[j] * inv) << (8 * (j & 3));
    *reinterpret_cast<uint4*>(fr + off) = make_uint4(w[0], w[1], w[2], w[3]);
    *reinterpret_cast<uint4*>(fr + off + 16) = make_uint4(w[4], w[5], w[6], w[7]);
    fr[kMxScaleOff + (off >> 5)] = (uint8_t)sb;
  }
}

// Compressed epilogue through shared memory.  A tcgen05.ld lane owns ONE ROW of the accumulator, so
// writing frames straight from registers makes every store instruction of a warp touch 32 different
// frames, 16 bytes each - tolerable in local HBM, hopeless as NVLink write packets.  Here each
// epilogue warp first parks the quantised bytes of its 32 rows x BLOCK_N columns in shared memory
// (16-byte units XOR-swizzled by row: conflict-free both ways), then writes them out with
// consecutive lanes on consecutive 16-byte units of the same row: BLOCK_N contiguous bytes per row
// (half a frame for BLOCK_N = 256).  The exponent / scale bytes of a row's chunks are collected in
// registers and leave as one vector store per row when rows are whole frames (ldc % 512 == 0).
// stage = this warp's 32 * BLOCK_N bytes; row0 = first row of this warp; 512 % BLOCK_N == 0.
template <int BLOCK_N>
__device__ __forceinline__ void epilogue_store_frames(const EpiParams& ep, uint32_t taddr, int row0, int n0, uint8_t* stage, int lane) {
  static_assert(512 % BLOCK_N == 0 && BLOCK_N >= 128, "row spans must not straddle frames");
  constexpr int kChunks = BLOCK_N / 32;
  constexpr int kUnits = BLOCK_N / 16;          // 16-byte units per row
  const bool bfp = ep.codec == AINIC_CODEC_BFP;
  const size_t fbytes = bfp ? (size_t)kBfpFrameBytes : (size_t)kMxFrameBytes;
  uint32_t side[kChunks / 2];   // BFP: 2 exponent bytes per chunk; MX: 1 scale byte per chunk (half of it used)
#pragma unroll
  for (int i = 0; i < kChunks / 2; ++i) side[i] = 0;
  // one chunk (32 columns of this lane's row): quantise, park the 32 bytes in shared memory
  auto process = [&](const uint32_t (&r)[32], int c) {
    float v[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * ep.alpha;
    uint32_t w[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    uint32_t sb;
    if (bfp) {
      sb = ep.bfp_mode == AINIC_BFP_HW_EXACT ? bfp_pack32<AINIC_BFP_HW_EXACT>(v, w) : bfp_pack32<AINIC_BFP_CLEAN>(v, w);
      // kChunks * 2 bytes per row, packed little-endian in chunk order
      if (c & 1) side[c >> 1] |= sb << 16; else side[c >> 1] |= sb;
    } else {
      uint32_t mag = 0;
#pragma unroll
      for (int j = 0; j < 32; ++j) mag = max(mag, __float_as_uint(v[j]) & 0x7FFFFFFFu);
      sb = mx_scale_byte(mag >> 23);
      const float inv = mx_inv_scale(sb);
#pragma unroll
      for (int j = 0; j < 32; ++j) w[j >> 2] |= f32_to_e4m3(v[j] * inv) << (8 * (j & 3));
      side[c >> 2] |= sb << (8 * (c & 3));
    }
    uint4* srow = reinterpret_cast<uint4*>(stage) + lane * kUnits;
    srow[(2 * c) ^ (lane & (kUnits - 1))] = make_uint4(w[0], w[1], w[2], w[3]);
    srow[(2 * c + 1) ^ (lane & (kUnits - 1))] = make_uint4(w[4], w[5], w[6], w[7]);
  };
  // the TMEM read of chunk c + 1 is in flight while chunk c is quantised (two register buffers)
  uint32_t ra[32], rb[32];
  tmem_ld_32x32b_x32(taddr, ra);
#pragma unroll
  for (int c = 0; c < kChunks; c += 2) {
    tmem_ld_wait();
    tmem_ld_32x32b_x32(taddr + (c + 1) * 32, rb);
    process(ra, c);
    tmem_ld_wait();
    if (c + 2 < kChunks) tmem_ld_32x32b_x32(taddr + (c + 2) * 32, ra);
    process(rb, c + 1);
  }
  __syncwarp();
  // ---- mantissas: 32 lanes x 16 B per instruction = 32 / kUnits whole row segments
  if (n0 < ep.N) {
#pragma unroll 1
    for (int i = 0; i < kUnits; ++i) {
      const int idx = i * 32 + lane;
      const int rr = idx / kUnits, u = idx % kUnits;
      const int grow = row0 + rr;
      if (grow < ep.M && n0 + u * 16 < ep.N) {
        const uint4 q = reinterpret_cast<const uint4*>(stage)[rr * kUnits + (u ^ (rr & (kUnits - 1)))];
        const size_t flat = (size_t)grow * ep.ldc + n0 + u * 16;
        *reinterpret_cast<uint4*>(frame_dst(ep, flat >> 9, fbytes) + (flat & 511)) = q;
      }
    }
    // ---- exponents / scales of this lane's own row
    const int grow = row0 + lane;
    if (grow < ep.M) {
      const size_t flat = (size_t)grow * ep.ldc + n0;
      if ((ep.ldc & 511) == 0 && n0 + BLOCK_N <= ep.N) {
        uint8_t* fr = frame_dst(ep, flat >> 9, fbytes);
        const uint32_t off = (uint32_t)(flat & 511);
        if (bfp) {  // kChunks * 2 bytes at exponent offset off / 16
          uint8_t* d = fr + kBfpExpOff + (off >> 4);
          if constexpr (kChunks == 8) *reinterpret_cast<uint4*>(d) = make_uint4(side[0], side[1], side[2], side[3]);
          else if constexpr (kChunks == 4) *reinterpret_cast<uint2*>(d) = make_uint2(side[0], side[1]);
          else *reinterpret_cast<uint32_t*>(d) = side[0];
        } else {    // kChunks bytes at scale offset off / 32
          uint8_t* d = fr + kMxScaleOff + (off >> 5);
          if constexpr (kChunks == 8) *reinterpret_cast<uint2*>(d) = make_uint2(side[0], side[1]);
          else if constexpr (kChunks == 4) *reinterpret_cast<uint32_t*>(d) = side[0];
          else *reinterpret_cast<uint16_t*>(d) = (uint16_t)side[0];
        }
      } else {
#pragma unroll
        for (int c = 0; c < kChunks; ++c) {
          if (n0 + c * 32 < ep.N) {
            const size_t f2 = flat + c * 32;
            uint8_t* fr = frame_dst(ep, f2 >> 9, fbytes);
            const uint32_t off = (uint32_t)(f2 & 511);
            if (bfp) *reinterpret_cast<uint16_t*>(fr + kBfpExpOff + (off >> 4)) = (uint16_t)(side[c >> 1] >> (16 * (c & 1)));
            else fr[kMxScaleOff + (off >> 5)] = (uint8_t)(side[c >> 2] >> (8 * (c & 3)));
          }
        }
      }
    }
  }
  __syncwarp();   // the staging area is reused by this warp's next tile
}

// Shared-memory matrix descriptor (sm_100 format, version 1, SWIZZLE_128B).
//   K-major operand : rows of 128 B, 8-row core-matrix groups SBO = 1024 B apart (LBO unused).
//   MN-major operand: 64-element (128 B) spans along MN, one span per k row; k groups of 8 rows are
//                     SBO = 1024 B apart, the next 64 MN elements start LBO = 8192 B further.
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;  // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;  // SWIZZLE_128B
  return d;
}
template <bool MN_MAJOR>
__device__ __forceinline__ uint64_t operand_desc(uint32_t tile_addr, int k) {
  // advance by one UMMA_K = 16 elements: 32 B inside the swizzle span (K-major) or 16 k-rows (MN-major)
  return MN_MAJOR ? make_desc(tile_addr + k * (UMMA_K * 128), 8192, 1024) : make_desc(tile_addr + k * (UMMA_K * 2), 16, 1024);
}

// kind::f16 instruction descriptor: bf16 x bf16 -> fp32, M x N, operand majors.
template <int M, int N, bool A_MN, bool B_MN>
__host__ __device__ constexpr uint32_t make_idesc() {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((A_MN ? 1u : 0u) << 15) | ((B_MN ? 1u : 0u) << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

// Drain one 128 x BLOCK_N accumulator: thread `lane` of the warp owning TMEM lane quarter q reads
// row (32 q + lane) 32 columns at a time and writes it with the fused epilogue.
template <int BLOCK_N, int OUT>
__device__ __forceinline__ void epilogue_store(const EpiParams& ep, uint32_t taddr, int row, int n0, int half, int nsplit,
                                               int c_first = 0) {
  constexpr bool OUT_F32 = OUT == OUT_FP32;
  static_assert(OUT != OUT_COMPRESSED || BLOCK_N % 32 == 0, "compressed output works on whole 32-column chunks");
  const bool row_ok = row < ep.M;
  // with 8 epilogue warps two of them share each TMEM lane quarter: `half` 0 drains the first half of the
  // column chunks, `half` 1 the rest; with 4 warps (nsplit == 1) each drains all of them
  constexpr int kChunks = BLOCK_N / 32 + (BLOCK_N % 32 ? 1 : 0);
  int c_begin = (nsplit == 2 && half) ? kChunks / 2 : 0;
  const int c_end = (nsplit == 2 && !half) ? kChunks / 2 : kChunks;
  if (c_begin < c_first) c_begin = c_first;   // the chunks before c_first left through epilogue_store_tma
#pragma unroll 1
  for (int c = c_begin; c < c_end; ++c) {
    uint32_t r[32];
    if (BLOCK_N % 32 == 0 || c < BLOCK_N / 32) {
      tmem_ld_32x32b_x32(taddr + c * 32, r);
    } else {
      // 16-column tail (BLOCK_N % 32 == 16): read the last 32 columns, keep the upper half.
      tmem_ld_32x32b_x32(taddr + c * 32 - 16, r);
#pragma unroll
      for (int j = 0; j < 16; ++j) r[j] = r[j + 16];
    }
    tmem_ld_wait();
    const int ncols = (BLOCK_N % 32 == 0 || c < BLOCK_N / 32) ? 32 : 16;
    const int col0 = n0 + c * 32;
    if constexpr (OUT == OUT_COMPRESSED) {
      if (row_ok && col0 < ep.N) {   // host guarantees N % 32 == 0 and ldc % 32 == 0
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * ep.alpha;
        store_compressed32(ep, (size_t)row * ep.ldc + col0, v);
      }
      continue;
    }
    if (row_ok && col0 < ep.N) {
      const size_t off = (size_t)row * ep.ldc + col0;
#pragma unroll
      for (int g = 0; g < 4; ++g) {  // groups of 8 columns
        if (g * 8 < ncols && col0 + g * 8 < ep.N) {
          float v[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r[g * 8 + j]);
          if constexpr (OUT_F32) {
            float4* dst = reinterpret_cast<float4*>(reinterpret_cast<float*>(ep.c) + off + g * 8);
            if (ep.alpha != 1.0f) {
#pragma unroll
              for (int j = 0; j < 8; ++j) v[j] *= ep.alpha;
            }
            if (ep.accumulate) {
              float4 o0 = dst[0], o1 = dst[1];
              v[0] += o0.x; v[1] += o0.y; v[2] += o0.z; v[3] += o0.w;
              v[4] += o1.x; v[5] += o1.y; v[6] += o1.z; v[7] += o1.w;
            }
            dst[0] = make_float4(v[0], v[1], v[2], v[3]);
            dst[1] = make_float4(v[4], v[5], v[6], v[7]);
          } else {
            if (ep.bias) {
              const float4 b0 = *reinterpret_cast<const float4*>(ep.bias + col0 + g * 8);
              const float4 b1 = *reinterpret_cast<const float4*>(ep.bias + col0 + g * 8 + 4);
              v[0] += b0.x; v[1] += b0.y; v[2] += b0.z; v[3] += b0.w;
              v[4] += b1.x; v[5] += b1.y; v[6] += b1.z; v[7] += b1.w;
            }
            if (ep.mask_out) {
              uint32_t lo = 0, hi = 0;
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                lo |= (v[j] > 0.0f ? 1u : 0u) << (8 * j);
                hi |= (v[4 + j] > 0.0f ? 1u : 0u) << (8 * j);
              }
              *reinterpret_cast<uint2*>(ep.mask_out + off + g * 8) = make_uint2(lo, hi);
            }
            if (ep.relu) {
#pragma unroll
              for (int j = 0; j < 8; ++j) v[j] = fmaxf(v[j], 0.0f);
            }
            if (ep.mask_in) {
              const uint2 mk = *reinterpret_cast<const uint2*>(ep.mask_in + off + g * 8);
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                if (((mk.x >> (8 * j)) & 0xFFu) == 0u) v[j] = 0.0f;
                if (((mk.y >> (8 * j)) & 0xFFu) == 0u) v[4 + j] = 0.0f;
              }
            }
            uint4 o;
            o.x = pack_bf16(v[0], v[1]);
            o.y = pack_bf16(v[2], v[3]);
            o.z = pack_bf16(v[4], v[5]);
            o.w = pack_bf16(v[6], v[7]);
            *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(ep.c) + off + g * 8) = o;
          }
        }
      }
    }
  }
}

// Epilogue of a CTA's LAST tile through shared memory and TMA stores.  epilogue_store above has every
// lane write 16 bytes of its own row: 32 rows, 32 half-filled sectors per store instruction - fine
// while the next tile's main loop hides it, but most GEMMs of the MLP have one tile per CTA, so this
// drain is exposed.  After the last tile's accumulator is complete the operand ring is idle (every
// MMA that read it has retired, the producer has nothing left to load), so it becomes the staging
// area: each epilogue warp parks its 32 rows in 128-byte-wide, 32-row boxes (32 fp32 or 64 bf16
// columns; 16-byte units XOR-swizzled by row = the SWIZZLE_128B pattern of the output tensor map, and
// conflict-free for the row-per-lane writes) and one lane hands every finished box to the TMA unit,
// which writes whole 128-byte lines and clips at the matrix edge.  No box is reused, so nothing
// waits until the final wait_group.read before the CTA lets go of its shared memory.
// The compressed (BFP / MXFP8) epilogues do not use this: their time is the quantiser's ALU work, which hides
// the scattered 16-byte stores; a 3-D TMA-store version of them measured 0.7 us SLOWER (3.8 vs 3.1 us).
// slab = this TMEM lane quarter's BLOCK_N / box-columns boxes (1024-byte aligned); row0 = first row.
template <int BLOCK_N, int OUT>
__device__ __forceinline__ void epilogue_store_tma(const EpiParams& ep, const CUtensorMap* tmap_c, uint32_t taddr, int row0,
                                                   int n0, uint8_t* slab, int lane, int half, int nsplit) {
  static_assert(OUT == OUT_BF16 || OUT == OUT_FP32, "plain outputs only");
  constexpr bool F32 = OUT == OUT_FP32;
  constexpr int kCPB = F32 ? 2 : 4;              // 16-column accumulator chunks per box
  constexpr int kBoxCols = 16 * kCPB;
  constexpr int kBoxes = BLOCK_N / kBoxCols;     // a 16-column remainder (BLOCK_N = 208) goes out directly
  const int c_begin = ((nsplit == 2 && half) ? kBoxes / 2 : 0) * kCPB;
  const int c_end = ((nsplit == 2 && !half) ? kBoxes / 2 : kBoxes) * kCPB;
  const uint32_t sw = (uint32_t)(lane & 7);
  // one 16-column chunk of this lane's row: convert, park its two (bf16) / four (fp32) 16-byte units; when
  // it completes a box: generic-proxy writes -> async proxy, then one lane hands the box to the TMA unit
  auto park = [&](const uint32_t (&r)[16], int c) {
    const int b = c / kCPB;
    uint8_t* box = slab + b * 4096;
    uint4* srow = reinterpret_cast<uint4*>(box + lane * 128);
    if constexpr (F32) {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        uint4 o;
        o.x = __float_as_uint(__uint_as_float(r[4 * u]) * ep.alpha);
        o.y = __float_as_uint(__uint_as_float(r[4 * u + 1]) * ep.alpha);
        o.z = __float_as_uint(__uint_as_float(r[4 * u + 2]) * ep.alpha);
        o.w = __float_as_uint(__uint_as_float(r[4 * u + 3]) * ep.alpha);
        srow[(uint32_t)((c % kCPB) * 4 + u) ^ sw] = o;
      }
    } else {
      const int col0 = n0 + c * 16;
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r[g * 8 + j]);
        if (ep.bias && col0 + g * 8 < ep.N) {
          const float4 b0 = *reinterpret_cast<const float4*>(ep.bias + col0 + g * 8);
          const float4 b1 = *reinterpret_cast<const float4*>(ep.bias + col0 + g * 8 + 4);
          v[0] += b0.x; v[1] += b0.y; v[2] += b0.z; v[3] += b0.w;
          v[4] += b1.x; v[5] += b1.y; v[6] += b1.z; v[7] += b1.w;
        }
        if (ep.relu) {
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = fmaxf(v[j], 0.0f);
        }
        uint4 o;
        o.x = pack_bf16(v[0], v[1]);
        o.y = pack_bf16(v[2], v[3]);
        o.z = pack_bf16(v[4], v[5]);
        o.w = pack_bf16(v[6], v[7]);
        srow[(uint32_t)((c % kCPB) * 2 + g) ^ sw] = o;
      }
    }
    if (c % kCPB == kCPB - 1) {
      fence_proxy_async();
      __syncwarp();
      if (lane == 0 && row0 < ep.M && n0 + b * kBoxCols < ep.N) tma_store_2d(tmap_c, box, n0 + b * kBoxCols, row0);
    }
  };
  // Two 16-register buffers: the TMEM read of the next chunk is in flight while this one is converted and
  // parked.  (With 32-column chunks the second buffer costs enough registers to stop an all-reduce CTA from
  // sharing the SM with this one; without a second buffer the bf16 epilogue is ~1 us slower.)
  uint32_t ra[16], rb[16];
  if (c_begin < c_end) tmem_ld_32x32b_x16(taddr + c_begin * 16, ra);
#pragma unroll 1
  for (int c = c_begin; c < c_end; c += 2) {   // chunk counts per warp are even
    tmem_ld_wait();
    tmem_ld_32x32b_x16(taddr + (c + 1) * 16, rb);
    park(ra, c);
    tmem_ld_wait();
    if (c + 2 < c_end) tmem_ld_32x32b_x16(taddr + (c + 2) * 16, ra);
    park(rb, c + 1);
  }
  if (lane == 0) tma_store_commit();
  if constexpr (BLOCK_N % kBoxCols != 0) epilogue_store<BLOCK_N, OUT>(ep, taddr, row0 + lane, n0, half, nsplit, kBoxes * kBoxCols / 32);
  if (lane == 0) tma_store_wait_read<0>();
  __syncwarp();
}

// ------------------------------------------------------------------------------------------ host
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      return nullptr;
    fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// bf16 row-major [rows][cols] (cols contiguous, leading dimension ld elements); box = box_c x box_r.
inline int make_tmap(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_c,
                     uint32_t box_r, bool f32 = false) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return AINIC_ERR_CUDA;
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {ld * (f32 ? 4 : 2)};
  cuuint32_t box[2] = {box_c, box_r};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? AINIC_OK : AINIC_ERR_CUDA;
}

// Encoding a tensor map costs a driver call (~1-2 us); a training step re-uses the same ~60 maps
// every iteration, so keep them (direct-mapped by a hash of the key; collisions just re-encode).
struct TmapKey {
  const void* base;
  uint64_t rows, cols, ld;
  uint32_t box_c, box_r;
  bool f32;
  bool operator==(const TmapKey& o) const {
    return base == o.base && rows == o.rows && cols == o.cols && ld == o.ld && box_c == o.box_c && box_r == o.box_r && f32 == o.f32;
  }
};
inline int cached_tmap(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_c,
                       uint32_t box_r, bool f32 = false) {
  constexpr int kSlots = 256;
  struct Slot { TmapKey key; CUtensorMap map; bool valid; };
  static thread_local Slot slots[kSlots];
  const TmapKey key{base, rows, cols, ld, box_c, box_r, f32};
  uint64_t h = reinterpret_cast<uint64_t>(base) >> 8;
  h ^= (uint64_t)f32 << 40;
  h ^= rows * 0x9E3779B97F4A7C15ull ^ cols * 0xC2B2AE3D27D4EB4Full ^ (uint64_t)box_r * 0x165667B19E3779F9ull ^ ld << 17 ^ box_c;
  Slot& s = slots[(h ^ (h >> 29)) % kSlots];
  if (!(s.valid && s.key == key)) {
    const int rc = make_tmap(&s.map, base, rows, cols, ld, box_c, box_r, f32);
    if (rc) { s.valid = false; return rc; }
    s.key = key;
    s.valid = true;
  }
  *out = s.map;
  return AINIC_OK;
}

inline int num_sms() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
  }
  return n;
}

// Threads per CTA: 2 non-epilogue warps + 4 (default) or 8 (AINIC_GEMM_EPI_WARPS=8) epilogue warps (two per TMEM
// lane quarter, each draining half of the columns).  8 warps do not pay: with the plain-output kernels (up to 88
// registers) a third warp per SM sub-partition pushes an all-reduce CTA off the SM (3 x 32 x 88 + 2 x 32 x 144 >
// 16 K registers per sub-partition: 0.815 vs 0.768 ms/step at 2 GPUs), and in the compressed-output kernel
// (67 registers, fits) the epilogue takes 3.3 us with 8 warps against 3.1 us with 4 (profiles/README.md R2.7) -
// it is not bound by per-warp work.
inline int gemm_threads() {
  const char* e = getenv("AINIC_GEMM_EPI_WARPS");
  return (e && atoi(e) == 8) ? kThreads : 192;
}
inline int gemm_epi_sleep_ns() {
  static const int v = getenv("AINIC_GEMM_EPI_SLEEP") ? atoi(getenv("AINIC_GEMM_EPI_SLEEP")) : 0;
  return v;
}

// Last-tile epilogue through shared memory + TMA stores (epilogue_store_tma); AINIC_GEMM_TMA_STORE=0 keeps
// the direct row-per-lane stores everywhere.
inline bool tma_store_enabled() {
  const char* e = getenv("AINIC_GEMM_TMA_STORE");   // read per call (cheap): A/B runs switch it inside one process
  return !(e && atoi(e) == 0);
}

// ainic_gemm_trace state: the next `left` launches write their per-CTA time stamps to consecutive [148][8]
// blocks of `buf` (diagnostics, scripts/gemm_timeline.py).  One instance for both kernel files.
struct TraceState { unsigned long long* buf = nullptr; int left = 0; };
inline TraceState& trace_state() {
  static TraceState t;
  return t;
}
inline void trace_attach(EpiParams& ep) {
  TraceState& t = trace_state();
  if (t.left > 0) {
    ep.trace = t.buf;
    t.buf += 148 * 8;
    --t.left;
  }
}

inline bool pdl_enabled() {
  static const bool on = !(getenv("AINIC_GEMM_PDL") && atoi(getenv("AINIC_GEMM_PDL")) == 0);
  return on;
}

// Launch entry of the 2-CTA kernels (gemm_tcgen05_2cta.cu); returns AINIC_ERR_UNSUPPORTED for
// combinations that have no 2-CTA instantiation.
int launch_2cta(int layout, const void* a, int lda, const void* b, int ldb, const EpiParams& ep, int block_n,
                int max_ctas, cudaStream_t stream);

}  // namespace gemm
}  // namespace ainic
