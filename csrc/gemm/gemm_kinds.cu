// tcgen05 GEMMs for the two operand kinds the bf16 kernel (gemm_tcgen05.cu) does not cover:
//
//   kind::tf32               fp32 operands read as tf32 (10-bit mantissa), fp32 accumulate.  The reference
//                            computes its MLP in fp32 end to end (libxsmm_dnn_fc_*_exec_f32,
//                            sw/mlp_mpi_example_f32.cpp:708-709, :741-742, :770-771); this is the tensor-core
//                            form of that "f32" configuration: no bf16 rounding of activations or weights.
//   kind::mxf8f6f4.block_scale   e4m3 operands with one ue8m0 scale per 32 values along K (the OCP MX format,
//                            the same block geometry as the MXFP8 wire codec of bfp_codec.h and the
//                            generalisation of the reference's 16-value shared-exponent blocks,
//                            hw/bfp_adapter.sv:59-76): the tensor core applies the scales itself, reading
//                            them from tensor memory where tcgen05.cp put them.
//
// Same structure as the bf16 kernel: persistent, one CTA per SM, warp 0 TMA producer, warp 1 TMEM
// allocator + single-thread MMA issuer, warps 2-5 epilogue, 128B-swizzled shared-memory ring, two
// TMEM accumulator stages.  A pipeline stage is always 128 bytes of K per row: 32 fp32 or 128 e4m3
// values; one MMA consumes 32 bytes of K (8 tf32 / 32 e4m3).  The three layouts NT / NN / TN are
// expressed through the descriptors' major bits like in the bf16 kernel.
//
// MX scale factors.  For a 128-row operand tile and the 4 K-blocks (4 x 32 values) of one stage the
// 512 scale bytes are stored as one "atom": byte (r % 32) * 16 + (r / 32) * 4 + kb for row r, block kb.
// That is exactly the 32 x 128-bit matrix tcgen05.cp.32x128b.warpx4 copies into 4 TMEM columns (lane
// r % 32, column r / 32, byte kb; replicated over the four lane quarters), which is where the MMA of
// K-block kb looks for the scale of row r when its instruction descriptor says sf_id = kb.  Atoms of
// an operand with R rows and K values are laid out [R / 128][K / 128]; ainic_mx_quantize_* produce them.
#include <stdio.h>

#include "gemm_common.cuh"

namespace ainic {
namespace gemm {

constexpr int KIND_TF32 = 1, KIND_MXF8 = 2;
constexpr int kStageKBytes = 128;   // bytes of K per row and stage = one swizzle span

template <int KIND> struct Kind;
template <> struct Kind<KIND_TF32> { static constexpr int kElem = 4, kBlockK = 32, kUmmaK = 8; };
template <> struct Kind<KIND_MXF8> { static constexpr int kElem = 1, kBlockK = 128, kUmmaK = 32; };

// instruction descriptors (cute/arch/mma_sm100_desc.hpp bit layout)
template <int M, int N, bool A_MN, bool B_MN>
__host__ __device__ constexpr uint32_t make_idesc_tf32() {
  return (1u << 4) /*D = f32*/ | (2u << 7) /*A = tf32*/ | (2u << 10) /*B = tf32*/ | ((A_MN ? 1u : 0u) << 15) |
         ((B_MN ? 1u : 0u) << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// block-scaled: A = B = e4m3 (format 0), scale format ue8m0 (bit 23), scale-factor ids in [4,6) (B) and [29,31) (A)
template <int M, int N, bool A_MN, bool B_MN>
__host__ __device__ constexpr uint32_t make_idesc_mxf8(uint32_t sf_id) {
  return (sf_id << 4) | ((A_MN ? 1u : 0u) << 15) | ((B_MN ? 1u : 0u) << 16) | ((uint32_t)(N >> 3) << 17) | (1u << 23) |
         ((uint32_t)(M >> 4) << 24) | (sf_id << 29);
}

__device__ __forceinline__ void umma_mxf8(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t tmem_sfa,
                                          uint32_t tmem_sfb, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf8f6f4.block_scale [%0], %1, %2, %3, [%5], [%6], p;\n\t}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(tmem_sfa), "r"(tmem_sfb)
      : "memory");
}
// shared memory (32 rows x 16 bytes, no swizzle) -> 4 TMEM columns, replicated over the four lane quarters
__device__ __forceinline__ void tmem_cp_32x128b_warpx4(uint32_t tmem_dst, uint64_t smem_desc) {
  asm volatile("tcgen05.cp.cta_group::1.32x128b.warpx4 [%0], %1;\n" ::"r"(tmem_dst), "l"(smem_desc) : "memory");
}
// descriptor of one scale-factor atom in shared memory: no swizzle, 8-row core matrices 128 bytes apart
__device__ __forceinline__ uint64_t sf_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((128u >> 4) & 0x3FFFu) << 32;   // SBO
  d |= (uint64_t)1 << 46;                          // descriptor version (Blackwell)
  return d;
}
// 1-D bulk copy global -> shared, completion on an mbarrier (the scale-factor atoms: 512 contiguous bytes)
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// MN-major operands of 32-bit elements cannot use the ordinary 128-byte swizzle (its 16-byte units would
// keep 4 neighbouring MN elements together where the tensor core needs to transpose single 32-bit
// values): they use the "128B swizzle, 32-byte atom" mode instead - TMA CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B,
// descriptor layout type 1 - whose swizzle pattern repeats every 4 rows of 128 bytes (SBO = 512).
__device__ __forceinline__ uint64_t make_desc_base32(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;  // descriptor version (Blackwell)
  d |= (uint64_t)1 << 61;  // SWIZZLE_128B_BASE32B
  return d;
}
template <int KIND, bool MN_MAJOR>
__device__ __forceinline__ uint64_t kind_desc(uint32_t tile_addr, int k) {
  // K-major: 32 bytes further inside the 128-byte row.  MN-major: kUmmaK rows of 128 bytes further; the
  // next 128-byte span along MN starts one TMA box (kBlockK rows x 128 B) later (LBO).
  using T = Kind<KIND>;
  if constexpr (MN_MAJOR && KIND == KIND_TF32) return make_desc_base32(tile_addr + k * (T::kUmmaK * 128), T::kBlockK * 128, 512);
  return MN_MAJOR ? make_desc(tile_addr + k * (T::kUmmaK * 128), T::kBlockK * 128, 1024) : make_desc(tile_addr + k * 32, 16, 1024);
}

struct MxParams {
  const uint8_t* sfa;   // scale atoms of A: [ceil(M / 128)][K / 128][512]
  const uint8_t* sfb;   // scale atoms of B: [ceil(N / 128)][K / 128][512]
};

__host__ __device__ constexpr int kinds_sf_atoms_b(int block_n) { return (block_n + 127) / 128; }
__host__ __device__ constexpr int kinds_stage_bytes(int block_n, int kind) {
  return BLOCK_M * kStageKBytes + block_n * kStageKBytes + (kind == KIND_MXF8 ? 512 * (1 + kinds_sf_atoms_b(block_n)) : 0);
}
__host__ __device__ constexpr int kinds_num_stages(int block_n, int kind) {
  return (kSmemBudget / kinds_stage_bytes(block_n, kind)) > 8 ? 8 : (kSmemBudget / kinds_stage_bytes(block_n, kind));
}

template <int KIND, int BLOCK_N, bool A_MN, bool B_MN, int OUT>
__global__ void __launch_bounds__(kThreads, 1)
gemm_kind_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, EpiParams ep, MxParams mx) {
  using T = Kind<KIND>;
  constexpr bool MX = KIND == KIND_MXF8;
  constexpr int kStages = kinds_num_stages(BLOCK_N, KIND);
  constexpr int kABytes = BLOCK_M * kStageKBytes;
  constexpr int kBBytes = BLOCK_N * kStageKBytes;
  constexpr int kSfBAtoms = kinds_sf_atoms_b(BLOCK_N);
  constexpr int kSfBytes = MX ? 512 * (1 + kSfBAtoms) : 0;
  constexpr int kSpan = kStageKBytes / T::kElem;          // elements of one 128-byte span along MN
  constexpr int kBoxBytes = T::kBlockK * 128;             // one MN-major TMA box
  // two accumulator stages unless the scale factors would not fit beside them (MX, 256 columns wide)
  constexpr int kAccStages = (MX && 2 * BLOCK_N + 4 + 4 * kSfBAtoms > 512) ? 1 : 2;
  constexpr uint32_t kColsNeeded = kAccStages * BLOCK_N + (MX ? 4 + 4 * kSfBAtoms : 0);
  constexpr uint32_t kTmemCols = kColsNeeded <= 32 ? 32u : kColsNeeded <= 64 ? 64u : kColsNeeded <= 128 ? 128u : kColsNeeded <= 256 ? 256u : 512u;
  static_assert(BLOCK_N % 16 == 0 && BLOCK_N >= 16 && BLOCK_N <= 256, "invalid UMMA N");
  static_assert(!B_MN || BLOCK_N % kSpan == 0, "MN-major B tiles are built from whole 128-byte spans");
  static_assert(!MX || BLOCK_N % 128 == 0, "scale-factor atoms cover 128 rows");
  static_assert(kStages >= 2, "pipeline too shallow");

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kStages * kABytes;
  uint8_t* smem_sf = smem_b + kStages * kBBytes;          // per stage: SFA atom, then the SFB atoms
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_sf + kStages * kSfBytes);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full = empty_bar + kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int epi_warps = (int)(blockDim.x >> 5) - 2;

  const int m_tiles = (ep.M + BLOCK_M - 1) / BLOCK_M;
  const int n_tiles = (ep.N + BLOCK_N - 1) / BLOCK_N;
  const int num_tiles = m_tiles * n_tiles;
  const int k_blocks = (ep.K + T::kBlockK - 1) / T::kBlockK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], epi_warps);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<kTmemCols>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  griddep_wait();
  griddep_launch_dependents();

  if (warp == 0) {
    // ===================================================== TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int mt = tile % m_tiles, nt = tile / m_tiles;
        const int m0 = mt * BLOCK_M, n0 = nt * BLOCK_N;
        for (int kb = 0; kb < k_blocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          mbar_expect_tx(&full_bar[stage], kABytes + kBBytes + kSfBytes);
          uint8_t* sa = smem_a + stage * kABytes;
          uint8_t* sb = smem_b + stage * kBBytes;
          const int k0 = kb * T::kBlockK;
          if constexpr (A_MN) {
#pragma unroll
            for (int j = 0; j < BLOCK_M / kSpan; ++j) tma_load_2d(sa + j * kBoxBytes, &tmap_a, &full_bar[stage], m0 + kSpan * j, k0);
          } else {
            tma_load_2d(sa, &tmap_a, &full_bar[stage], k0, m0);
          }
          if constexpr (B_MN) {
#pragma unroll
            for (int j = 0; j < BLOCK_N / kSpan; ++j) tma_load_2d(sb + j * kBoxBytes, &tmap_b, &full_bar[stage], n0 + kSpan * j, k0);
          } else {
            tma_load_2d(sb, &tmap_b, &full_bar[stage], k0, n0);
          }
          if constexpr (MX) {
            uint8_t* ssf = smem_sf + stage * kSfBytes;
            bulk_load(ssf, mx.sfa + ((size_t)mt * k_blocks + kb) * 512, 512, &full_bar[stage]);
#pragma unroll
            for (int a = 0; a < kSfBAtoms; ++a)
              bulk_load(ssf + 512 * (1 + a), mx.sfb + ((size_t)(nt * kSfBAtoms + a) * k_blocks + kb) * 512, 512, &full_bar[stage]);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      const uint32_t tmem_sfa = tmem_base + (uint32_t)(kAccStages * BLOCK_N);
      const uint32_t tmem_sfb = tmem_sfa + 4;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)(acc * BLOCK_N);
        for (int kb = 0; kb < k_blocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem_a + stage * kABytes);
          const uint32_t b_addr = smem_u32(smem_b + stage * kBBytes);
          if constexpr (MX) {
            // scale factors of this stage: shared memory -> tensor memory.  tcgen05.cp and tcgen05.mma execute in
            // issue order, so the copy waits for the previous stage's MMAs and these MMAs wait for the copy.
            const uint32_t sf_addr = smem_u32(smem_sf + stage * kSfBytes);
            tmem_cp_32x128b_warpx4(tmem_sfa, sf_desc(sf_addr));
#pragma unroll
            for (int a = 0; a < kSfBAtoms; ++a) tmem_cp_32x128b_warpx4(tmem_sfb + 4 * a, sf_desc(sf_addr + 512 * (1 + a)));
          }
#pragma unroll
          for (int k = 0; k < T::kBlockK / T::kUmmaK; ++k) {
            const uint64_t da = kind_desc<KIND, A_MN>(a_addr, k), db = kind_desc<KIND, B_MN>(b_addr, k);
            const uint32_t accumulate = (uint32_t)((kb | k) != 0);
            if constexpr (MX) umma_mxf8(tmem_d, da, db, make_idesc_mxf8<BLOCK_M, BLOCK_N, A_MN, B_MN>((uint32_t)k), tmem_sfa, tmem_sfb, accumulate);
            else umma_tf32(tmem_d, da, db, make_idesc_tf32<BLOCK_M, BLOCK_N, A_MN, B_MN>(), accumulate);
          }
          umma_commit(&empty_bar[stage]);
          if (kb == k_blocks - 1) umma_commit(&tmem_full[acc]);
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
        if (++acc == kAccStages) { acc = 0; acc_phase ^= 1u; }
      }
    }
  } else {
    // ===================================================== epilogue (warps 2..5)
    const int quarter = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int m0 = (tile % m_tiles) * BLOCK_M;
      const int n0 = (tile / m_tiles) * BLOCK_N;
      mbar_wait_backoff(&tmem_full[acc], acc_phase, (uint32_t)ep.epi_sleep);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BLOCK_N);
      epilogue_store<BLOCK_N, OUT>(ep, taddr, m0 + quarter * 32 + lane, n0, warp >= 6, epi_warps / 4);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
      if (++acc == kAccStages) { acc = 0; acc_phase ^= 1u; }
    }
  }

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) {
    __syncwarp();
    tmem_dealloc<kTmemCols>(tmem_base);
  }
}

// ------------------------------------------------------------------------------------------ MX quantisers
// byte offset of the scale of (row r, K-block kblock) inside the atom-tiled scale array of an operand with K values per row
__host__ __device__ inline size_t sf_offset(int r, int kblock, int K) {
  const size_t atom = (size_t)(r >> 7) * (size_t)(K >> 7) + (size_t)(kblock >> 2);
  return atom * 512 + (size_t)(r & 31) * 16 + (size_t)((r & 127) >> 5) * 4 + (size_t)(kblock & 3);
}
__device__ __forceinline__ float load_as_f32(const void* p, size_t i, bool bf16) {
  return bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i]) : reinterpret_cast<const float*>(p)[i];
}
// Blocks of 32 consecutive values along the contiguous dimension (K-major operands).  Four adjacent lanes share
// a block, eight values each: a warp reads 256 consecutive values per load instruction (16 bytes per lane for
// bf16, 2 x 16 for fp32) and writes 256 consecutive e4m3 bytes; the block maximum takes two xor-shuffles.
// (One thread per block - 32 scalar loads 64 bytes apart between lanes - ran at 0.7-1.1 TB/s.)
__global__ void mx_quantize_rows_kernel(const void* x, bool bf16, int rows, int K, int ld, uint8_t* q, uint8_t* sf) {
  const int kblocks = K >> 5;
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t blk = t >> 2;
  const int part = (int)(t & 3);
  const bool live = blk < (size_t)rows * kblocks;      // whole groups of four are live or dead together
  const int r = live ? (int)(blk / kblocks) : 0, kb = live ? (int)(blk % kblocks) : 0;
  const size_t src = (size_t)r * ld + kb * 32 + part * 8;
  float v[8];
  if (bf16) {
    const uint4 p = live ? *reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(x) + src) : make_uint4(0, 0, 0, 0);
    const uint32_t w[4] = {p.x, p.y, p.z, p.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      v[2 * j] = __uint_as_float(w[j] << 16);
      v[2 * j + 1] = __uint_as_float(w[j] & 0xFFFF0000u);
    }
  } else {
    const float4 a = live ? *reinterpret_cast<const float4*>(static_cast<const float*>(x) + src) : make_float4(0, 0, 0, 0);
    const float4 c = live ? *reinterpret_cast<const float4*>(static_cast<const float*>(x) + src + 4) : make_float4(0, 0, 0, 0);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = c.x; v[5] = c.y; v[6] = c.z; v[7] = c.w;
  }
  uint32_t E = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) E = max(E, bfp_exp_field(__float_as_uint(v[j])));
  E = max(E, __shfl_xor_sync(0xFFFFFFFFu, E, 1));
  E = max(E, __shfl_xor_sync(0xFFFFFFFFu, E, 2));
  if (!live) return;
  const uint32_t sb = mx_scale_byte(E);
  const float inv = mx_inv_scale(sb);
  uint32_t w0 = 0, w1 = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    w0 |= f32_to_e4m3(v[j] * inv) << (8 * j);
    w1 |= f32_to_e4m3(v[4 + j] * inv) << (8 * j);
  }
  *reinterpret_cast<uint2*>(q + (size_t)r * K + kb * 32 + part * 8) = make_uint2(w0, w1);
  if (part == 0) sf[sf_offset(r, kb, K)] = (uint8_t)sb;
}
// Blocks of 32 consecutive ROWS of one column (MN-major operands: x is [K][cols], K is the slow dimension).
// One thread per (8 adjacent columns, row block): every row access of a warp is 512 (bf16) / 1024 (fp32)
// contiguous bytes, the 8 e4m3 results of a row leave as one 8-byte store.  Two passes over the 32 rows (maximum,
// then conversion; the second one hits L1 / L2) keep the thread at 8 live values instead of 256.
// (One thread per single column - 2-byte loads, 1-byte stores - ran at 0.7 TB/s.)
__device__ __forceinline__ void load8_as_f32(const void* p, size_t i, bool bf16, float (&v)[8]) {
  if (bf16) {
    const uint4 q = *reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(p) + i);
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      v[2 * j] = __uint_as_float(w[j] << 16);
      v[2 * j + 1] = __uint_as_float(w[j] & 0xFFFF0000u);
    }
  } else {
    const float4 a = *reinterpret_cast<const float4*>(static_cast<const float*>(p) + i);
    const float4 c = *reinterpret_cast<const float4*>(static_cast<const float*>(p) + i + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = c.x; v[5] = c.y; v[6] = c.z; v[7] = c.w;
  }
}
__global__ void mx_quantize_cols_kernel(const void* x, bool bf16, int K, int cols, int ld, uint8_t* q, uint8_t* sf) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int kblocks = K >> 5, cgroups = cols >> 3;
  if (t >= (size_t)cgroups * kblocks) return;
  const int c0 = (int)(t % cgroups) * 8, kb = (int)(t / cgroups);
  uint32_t E[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll 4
  for (int j = 0; j < 32; ++j) {
    float v[8];
    load8_as_f32(x, (size_t)(kb * 32 + j) * ld + c0, bf16, v);
#pragma unroll
    for (int i = 0; i < 8; ++i) E[i] = max(E[i], bfp_exp_field(__float_as_uint(v[i])));
  }
  float inv[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const uint32_t sb = mx_scale_byte(E[i]);
    inv[i] = mx_inv_scale(sb);
    sf[sf_offset(c0 + i, kb, K)] = (uint8_t)sb;
  }
#pragma unroll 4
  for (int j = 0; j < 32; ++j) {
    float v[8];
    load8_as_f32(x, (size_t)(kb * 32 + j) * ld + c0, bf16, v);
    uint32_t w0 = 0, w1 = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      w0 |= f32_to_e4m3(v[i] * inv[i]) << (8 * i);
      w1 |= f32_to_e4m3(v[4 + i] * inv[4 + i]) << (8 * i);
    }
    *reinterpret_cast<uint2*>(q + (size_t)(kb * 32 + j) * cols + c0) = make_uint2(w0, w1);
  }
}

// ------------------------------------------------------------------------------------------ host
template <int KIND>
static int kind_tmap(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_c, uint32_t box_r,
                     bool mn_major) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return AINIC_ERR_CUDA;
  constexpr int elem = Kind<KIND>::kElem;
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {ld * elem};
  cuuint32_t box[2] = {box_c, box_r};
  cuuint32_t estr[2] = {1, 1};
  const CUtensorMapDataType dt = KIND == KIND_TF32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_UINT8;
  const CUtensorMapSwizzle sw = (KIND == KIND_TF32 && mn_major) ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B;
  CUresult r = fn(map, dt, 2, const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? AINIC_OK : AINIC_ERR_CUDA;
}

template <int KIND, int BLOCK_N, bool A_MN, bool B_MN, int OUT>
static int launch_kind(const void* a, int lda, const void* b, int ldb, const EpiParams& ep, const MxParams& mx, int max_ctas,
                       cudaStream_t stream) {
  using T = Kind<KIND>;
  constexpr int kSpan = kStageKBytes / T::kElem;
  CUtensorMap ta, tb;
  int rc;
  // K-major: global [rows][K], box = kBlockK x rows.  MN-major: global [K][rows], box = one 128-byte span x kBlockK.
  if (A_MN) rc = kind_tmap<KIND>(&ta, a, ep.K, ep.M, lda, kSpan, T::kBlockK, true);
  else rc = kind_tmap<KIND>(&ta, a, ep.M, ep.K, lda, T::kBlockK, BLOCK_M, false);
  if (rc) return rc;
  if (B_MN) rc = kind_tmap<KIND>(&tb, b, ep.K, ep.N, ldb, kSpan, T::kBlockK, true);
  else rc = kind_tmap<KIND>(&tb, b, ep.N, ep.K, ldb, T::kBlockK, BLOCK_N, false);
  if (rc) return rc;
  auto kern = gemm_kind_kernel<KIND, BLOCK_N, A_MN, B_MN, OUT>;
  constexpr int smem = kinds_num_stages(BLOCK_N, KIND) * kinds_stage_bytes(BLOCK_N, KIND) + 1024 + 256;
  static_assert(smem <= 227 * 1024, "shared memory budget");
  static bool configured = false;
  if (!configured) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return AINIC_ERR_CUDA;
    configured = true;
  }
  const int tiles = ((ep.M + BLOCK_M - 1) / BLOCK_M) * ((ep.N + BLOCK_N - 1) / BLOCK_N);
  int grid = num_sms();
  if (max_ctas > 0 && max_ctas < grid) grid = max_ctas;
  if (tiles < grid) grid = tiles;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(192);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  if (cudaLaunchKernelEx(&cfg, kern, ta, tb, ep, mx) != cudaSuccess) return AINIC_ERR_CUDA;
  return cudaGetLastError() == cudaSuccess ? AINIC_OK : AINIC_ERR_CUDA;
}

static int pick_block_n(int M, int N, const int* cand, int ncand, int sms) {
  double best = -1.0;
  int bn = cand[0];
  for (int i = 0; i < ncand; ++i) {
    const long tiles = (long)((M + 127) / 128) * ((N + cand[i] - 1) / cand[i]);
    const long waves = (tiles + sms - 1) / sms;
    const double e = ((double)M * N) / ((double)waves * sms * 128 * (cand[i] + 24.0));
    if (e > best * 1.02) { best = e; bn = cand[i]; }
  }
  return bn;
}

}  // namespace gemm
}  // namespace ainic

using namespace ainic::gemm;

static int fill_epi(EpiParams& ep, void* c, int ldc, int M, int N, int K, const ainic_gemm_epilogue_t* epi) {
  ep.c = c;
  ep.M = M; ep.N = N; ep.K = K; ep.ldc = ldc;
  ep.alpha = 1.0f;
  ep.epi_sleep = gemm_epi_sleep_ns();
  if (epi) {
    if (epi->cgrad || epi->scatter) return AINIC_ERR_UNSUPPORTED;   // compressed epilogues: bf16 kernel only
    ep.bias = epi->bias;
    ep.mask_out = epi->mask_out;
    ep.mask_in = epi->mask_in;
    ep.relu = epi->relu;
    ep.accumulate = epi->accumulate;
    ep.alpha = epi->alpha;
    ep.done_flag = epi->done_flag;
    ep.done_ticket = epi->done_ticket;
    if (ep.done_flag) return AINIC_ERR_UNSUPPORTED;                 // doorbells: bf16 kernel only
  }
  return AINIC_OK;
}

#define KIND_CASE(KIND, BN, AMN, BMN, OUT) \
  case BN: return launch_kind<KIND, BN, AMN, BMN, OUT>(a, lda, b, ldb, ep, mx, max_ctas, (cudaStream_t)stream);

// fp32 operands on the tensor cores as tf32, fp32 result (alpha / accumulate epilogue).  Layouts as ainic_gemm_bf16.
extern "C" int ainic_gemm_tf32(int layout, const float* a, int lda, const float* b, int ldb, float* c, int ldc, int M, int N,
                               int K, const ainic_gemm_epilogue_t* epi, int block_n, int max_ctas, void* stream) {
  if (M <= 0 || N <= 0 || K <= 0 || !a || !b || !c) return AINIC_ERR_ARG;
  if ((N % 8) || (K % 4) || (lda % 4) || (ldb % 4) || (ldc % 4)) return AINIC_ERR_ARG;
  if (((uintptr_t)a & 15) || ((uintptr_t)b & 15) || ((uintptr_t)c & 15)) return AINIC_ERR_ARG;
  EpiParams ep{};
  int rc = fill_epi(ep, c, ldc, M, N, K, epi);
  if (rc) return rc;
  if (ep.bias || ep.relu || ep.mask_in || ep.mask_out) return AINIC_ERR_UNSUPPORTED;   // fp32 outputs: scale / accumulate only
  MxParams mx{};
  const int sms = (max_ctas > 0 && max_ctas < num_sms()) ? max_ctas : num_sms();
  if (block_n == 0) {
    const int cand[] = {256, 192, 128};
    block_n = pick_block_n(M, N, cand, 3, sms);
  }
  switch (layout) {
    case AINIC_GEMM_NT:
      switch (block_n) { KIND_CASE(KIND_TF32, 256, false, false, OUT_FP32) KIND_CASE(KIND_TF32, 192, false, false, OUT_FP32)
                         KIND_CASE(KIND_TF32, 128, false, false, OUT_FP32) default: return AINIC_ERR_ARG; }
    case AINIC_GEMM_NN:
      switch (block_n) { KIND_CASE(KIND_TF32, 256, false, true, OUT_FP32) KIND_CASE(KIND_TF32, 192, false, true, OUT_FP32)
                         KIND_CASE(KIND_TF32, 128, false, true, OUT_FP32) default: return AINIC_ERR_ARG; }
    case AINIC_GEMM_TN:
      switch (block_n) { KIND_CASE(KIND_TF32, 256, true, true, OUT_FP32) KIND_CASE(KIND_TF32, 192, true, true, OUT_FP32)
                         KIND_CASE(KIND_TF32, 128, true, true, OUT_FP32) default: return AINIC_ERR_ARG; }
  }
  return AINIC_ERR_ARG;
}

extern "C" size_t ainic_mx_sf_bytes(size_t rows, size_t K) { return ((rows + 127) / 128) * (K / 128) * 512; }

// x: [rows][ld] (K contiguous) fp32 or bf16 -> q: [rows][K] e4m3 + scale atoms (one ue8m0 per 32 values along K)
extern "C" int ainic_mx_quantize_rows(const void* x, int x_is_bf16, int rows, int K, int ld, void* q, void* sf, void* stream) {
  if (!x || !q || !sf || rows <= 0 || K <= 0 || (K % 128) || ld < K || ((uintptr_t)q & 15)) return AINIC_ERR_ARG;
  if (((uintptr_t)x & 15) || (ld % (x_is_bf16 ? 8 : 4))) return AINIC_ERR_ARG;   // 16-byte vector loads
  cudaStream_t s = (cudaStream_t)stream;
  if (cudaMemsetAsync(sf, 0, ainic_mx_sf_bytes(rows, K), s) != cudaSuccess) return AINIC_ERR_CUDA;
  const size_t n = (size_t)rows * (K / 32) * 4;   // four lanes per 32-value block
  mx_quantize_rows_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(x, x_is_bf16 != 0, rows, K, ld, static_cast<uint8_t*>(q),
                                                                       static_cast<uint8_t*>(sf));
  return cudaGetLastError() == cudaSuccess ? AINIC_OK : AINIC_ERR_CUDA;
}
// x: [K][ld] (cols contiguous) -> q: [K][cols] e4m3 + scale atoms indexed (column, K-block): blocks of 32 consecutive rows
extern "C" int ainic_mx_quantize_cols(const void* x, int x_is_bf16, int K, int cols, int ld, void* q, void* sf, void* stream) {
  if (!x || !q || !sf || cols <= 0 || K <= 0 || (K % 128) || ld < cols) return AINIC_ERR_ARG;
  // 8 columns per thread: 16-byte loads and 8-byte stores
  if ((cols % 8) || ((uintptr_t)x & 15) || ((uintptr_t)q & 7) || (ld % (x_is_bf16 ? 8 : 4))) return AINIC_ERR_ARG;
  cudaStream_t s = (cudaStream_t)stream;
  if (cudaMemsetAsync(sf, 0, ainic_mx_sf_bytes(cols, K), s) != cudaSuccess) return AINIC_ERR_CUDA;
  const size_t n = (size_t)(cols / 8) * (K / 32);
  mx_quantize_cols_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(x, x_is_bf16 != 0, K, cols, ld, static_cast<uint8_t*>(q),
                                                                       static_cast<uint8_t*>(sf));
  return cudaGetLastError() == cudaSuccess ? AINIC_OK : AINIC_ERR_CUDA;
}

// e4m3 operands with ue8m0 block scales (32 values along K), scales applied by the tensor core.
// NT: a [M][lda], b [N][ldb] (K contiguous).  NN: a [M][lda], b [K][ldb].  TN: a [K][lda] (M contiguous), b [K][ldb].
// sfa / sfb: scale atoms from ainic_mx_quantize_rows (K-contiguous operands) / _cols (K-slow operands).
extern "C" int ainic_gemm_mxfp8(int layout, const void* a, int lda, const void* sfa, const void* b, int ldb, const void* sfb,
                                void* c, int ldc, int out_fp32, int M, int N, int K, const ainic_gemm_epilogue_t* epi,
                                int block_n, int max_ctas, void* stream) {
  if (M <= 0 || N <= 0 || K <= 0 || !a || !b || !c || !sfa || !sfb) return AINIC_ERR_ARG;
  if ((N % 8) || (K % 128) || (lda % 16) || (ldb % 16) || (ldc % 8)) return AINIC_ERR_ARG;
  if (((uintptr_t)a & 15) || ((uintptr_t)b & 15) || ((uintptr_t)c & 15) || ((uintptr_t)sfa & 15) || ((uintptr_t)sfb & 15)) return AINIC_ERR_ARG;
  EpiParams ep{};
  int rc = fill_epi(ep, c, ldc, M, N, K, epi);
  if (rc) return rc;
  if (out_fp32 && (ep.bias || ep.relu || ep.mask_in || ep.mask_out)) return AINIC_ERR_UNSUPPORTED;
  MxParams mx{static_cast<const uint8_t*>(sfa), static_cast<const uint8_t*>(sfb)};
  // 256-wide tiles halve the shared-memory bytes read per MMA (the 1-CTA fp8 limiter: a 128 x 128 x 32 MMA reads
  // 8 KB in the 64 cycles it takes) at the price of a single accumulator stage; measured on the reference layer
  // 12.9 vs 17.5 us forward, 14.6 vs 16.9 us backward-weight, and 1990 vs 1391 TFLOPS at 1792 x 8192 x 8192
  if (block_n == 0) block_n = 256;
  if (out_fp32) {
    switch (layout) {
      case AINIC_GEMM_NT: switch (block_n) { KIND_CASE(KIND_MXF8, 128, false, false, OUT_FP32) KIND_CASE(KIND_MXF8, 256, false, false, OUT_FP32) default: return AINIC_ERR_ARG; }
      case AINIC_GEMM_NN: switch (block_n) { KIND_CASE(KIND_MXF8, 128, false, true, OUT_FP32) KIND_CASE(KIND_MXF8, 256, false, true, OUT_FP32) default: return AINIC_ERR_ARG; }
      case AINIC_GEMM_TN: switch (block_n) { KIND_CASE(KIND_MXF8, 128, true, true, OUT_FP32) KIND_CASE(KIND_MXF8, 256, true, true, OUT_FP32) default: return AINIC_ERR_ARG; }
    }
  } else {
    switch (layout) {
      case AINIC_GEMM_NT: switch (block_n) { KIND_CASE(KIND_MXF8, 128, false, false, OUT_BF16) KIND_CASE(KIND_MXF8, 256, false, false, OUT_BF16) default: return AINIC_ERR_ARG; }
      case AINIC_GEMM_NN: switch (block_n) { KIND_CASE(KIND_MXF8, 128, false, true, OUT_BF16) KIND_CASE(KIND_MXF8, 256, false, true, OUT_BF16) default: return AINIC_ERR_ARG; }
      case AINIC_GEMM_TN: switch (block_n) { KIND_CASE(KIND_MXF8, 128, true, true, OUT_BF16) KIND_CASE(KIND_MXF8, 256, true, true, OUT_BF16) default: return AINIC_ERR_ARG; }
    }
  }
  return AINIC_ERR_ARG;
}
