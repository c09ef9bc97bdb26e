// Hand-written tcgen05 / TMEM / TMA GEMM for the MLP's three products (bf16 in, fp32 accumulate):
//   forward        Y[mb,n]  = X[mb,k] . W[n,k]^T          (A K-major,  B K-major)   K1
//   backward data  dX[mb,k] = dY[mb,n] . W[n,k]           (A K-major,  B MN-major)  K2
//   backward wgt   dW[n,k]  = dY[mb,n]^T . X[mb,k]        (A MN-major, B MN-major)  K3
// These replace the libxsmm_dnn_fc_{fwd,bwd}_exec_f32 calls of the reference
// (sw/mlp_mpi_example_f32.cpp:708-709, :741-742, :770-771).  The bwd-weight epilogue writes fp32
// straight into the flat gradient bucket the all-reduce engine consumes (the buffer the reference
// hands to the NIC, sw/mlp_mpi_example_f32.cpp:754), optionally accumulating or pre-scaling.
//
// Structure: persistent, warp-specialised, one CTA per SM.
//   warp 0   : TMA producer (one elected lane)         global -> 128B-swizzled smem ring
//   warp 1   : TMEM allocator + MMA issuer (one lane)  tcgen05.mma.cta_group::1.kind::f16, 128xBLOCK_Nx16
//   warps 2-5: epilogue                                tcgen05.ld 32x32b -> registers -> global
// Two TMEM accumulator stages let the epilogue of tile i overlap the main loop of tile i+1.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../common/ptx.cuh"
#include "../../include/ainic.h"

namespace ainic {
namespace gemm {

using namespace ptx;

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;  // 64 bf16 = 128 B = one swizzle span
constexpr int UMMA_K = 16;
constexpr int kThreads = 192;
constexpr int kSmemBudget = 220 * 1024;

struct EpiParams {
  void* c;              // output, row-major [M][ldc]; bf16 or fp32
  const float* bias;    // [N] or null (added before activation)
  uint8_t* mask_out;    // [M][ldc] or null: 1 where the (biased) result is > 0 (ReLU mask)
  const uint8_t* mask_in;  // [M][ldc] or null: output zeroed where mask_in == 0 (ReLU backward)
  float alpha;          // result scale (fp32 output only)
  int relu;             // apply max(0, x)
  int accumulate;       // fp32 output only: c += result
  int M, N, K, ldc;
};

__host__ __device__ constexpr int stage_bytes(int block_n) { return BLOCK_M * BLOCK_K * 2 + block_n * BLOCK_K * 2; }
__host__ __device__ constexpr int num_stages(int block_n) {
  return (kSmemBudget / stage_bytes(block_n)) > 8 ? 8 : (kSmemBudget / stage_bytes(block_n));
}
__host__ __device__ constexpr uint32_t tmem_cols(int block_n) {
  return 2 * block_n <= 32 ? 32u : 2 * block_n <= 64 ? 64u : 2 * block_n <= 128 ? 128u : 2 * block_n <= 256 ? 256u : 512u;
}

// Shared-memory matrix descriptor (sm_100 format, version 1, SWIZZLE_128B).
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;  // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;  // SWIZZLE_128B
  return d;
}

template <int BLOCK_N, bool A_MN, bool B_MN>
__host__ __device__ constexpr uint32_t make_idesc() {
  return (1u << 4)                          // D format: F32
         | (1u << 7)                        // A format: BF16
         | (1u << 10)                       // B format: BF16
         | ((A_MN ? 1u : 0u) << 15)         // A major
         | ((B_MN ? 1u : 0u) << 16)         // B major
         | ((uint32_t)(BLOCK_N >> 3) << 17)  // N
         | ((uint32_t)(BLOCK_M >> 4) << 24); // M
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

template <int BLOCK_N, bool A_MN, bool B_MN, bool OUT_F32>
__global__ void __launch_bounds__(kThreads, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, EpiParams ep) {
  constexpr int kStages = num_stages(BLOCK_N);
  constexpr int kABytes = BLOCK_M * BLOCK_K * 2;
  constexpr int kBBytes = BLOCK_N * BLOCK_K * 2;
  constexpr uint32_t kTmemCols = tmem_cols(BLOCK_N);
  static_assert(BLOCK_N % 16 == 0 && BLOCK_N >= 16 && BLOCK_N <= 256, "invalid UMMA N");
  static_assert(!B_MN || BLOCK_N % 64 == 0, "MN-major B tiles are built from 64-wide TMA boxes");
  static_assert(kStages >= 2, "pipeline too shallow");

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kStages * kABytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kStages * (kABytes + kBBytes));
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full = empty_bar + kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const int m_tiles = (ep.M + BLOCK_M - 1) / BLOCK_M;
  const int n_tiles = (ep.N + BLOCK_N - 1) / BLOCK_N;
  const int num_tiles = m_tiles * n_tiles;
  const int k_blocks = (ep.K + BLOCK_K - 1) / BLOCK_K;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], 4);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<kTmemCols>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ===================================================== TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m0 = (tile % m_tiles) * BLOCK_M;
        const int n0 = (tile / m_tiles) * BLOCK_N;
        for (int kb = 0; kb < k_blocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          mbar_expect_tx(&full_bar[stage], kABytes + kBBytes);
          uint8_t* sa = smem_a + stage * kABytes;
          uint8_t* sb = smem_b + stage * kBBytes;
          const int k0 = kb * BLOCK_K;
          if constexpr (A_MN) {
#pragma unroll
            for (int j = 0; j < BLOCK_M / 64; ++j) tma_load_2d(sa + j * 8192, &tmap_a, &full_bar[stage], m0 + 64 * j, k0);
          } else {
            tma_load_2d(sa, &tmap_a, &full_bar[stage], k0, m0);
          }
          if constexpr (B_MN) {
#pragma unroll
            for (int j = 0; j < BLOCK_N / 64; ++j) tma_load_2d(sb + j * 8192, &tmap_b, &full_bar[stage], n0 + 64 * j, k0);
          } else {
            tma_load_2d(sb, &tmap_b, &full_bar[stage], k0, n0);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc<BLOCK_N, A_MN, B_MN>();
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)(acc * BLOCK_N);
        for (int kb = 0; kb < k_blocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem_a + stage * kABytes);
          const uint32_t b_addr = smem_u32(smem_b + stage * kBBytes);
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
            // K-major: 16 elements = 32 B inside the 128 B swizzle span.
            // MN-major: 16 k-rows of 128 B = two 1024 B core-matrix groups.
            const uint64_t da = A_MN ? make_desc(a_addr + k * (UMMA_K * 128), 8192, 1024)
                                     : make_desc(a_addr + k * (UMMA_K * 2), 16, 1024);
            const uint64_t db = B_MN ? make_desc(b_addr + k * (UMMA_K * 128), 8192, 1024)
                                     : make_desc(b_addr + k * (UMMA_K * 2), 16, 1024);
            umma_f16(tmem_d, da, db, idesc, (uint32_t)((kb | k) != 0));
          }
          umma_commit(&empty_bar[stage]);
          if (kb == k_blocks - 1) umma_commit(&tmem_full[acc]);
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
      }
    }
  } else {
    // ===================================================== epilogue (warps 2..5)
    const int quarter = warp & 3;  // TMEM lane quarter this warp may touch
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int m0 = (tile % m_tiles) * BLOCK_M;
      const int n0 = (tile / m_tiles) * BLOCK_N;
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const int row = m0 + quarter * 32 + lane;
      const bool row_ok = row < ep.M;
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BLOCK_N);
#pragma unroll 1
      for (int c = 0; c < BLOCK_N / 32 + (BLOCK_N % 32 ? 1 : 0); ++c) {
        uint32_t r[32];
        if (BLOCK_N % 32 == 0 || c < BLOCK_N / 32) {
          tmem_ld_32x32b_x32(taddr + c * 32, r);
        } else {
          // 16-column tail (BLOCK_N % 32 == 16): read 32 columns, the upper half is ignored below.
          tmem_ld_32x32b_x32(taddr + c * 32 - 16, r);
#pragma unroll
          for (int j = 0; j < 16; ++j) r[j] = r[j + 16];
        }
        tmem_ld_wait();
        const int ncols = (BLOCK_N % 32 == 0 || c < BLOCK_N / 32) ? 32 : 16;
        const int col0 = n0 + c * 32;
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
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
  }

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) tmem_dealloc<kTmemCols>(tmem_base);
}

// ------------------------------------------------------------------------------------------ host
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
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
static int make_tmap(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_c,
                     uint32_t box_r) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return AINIC_ERR_CUDA;
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {ld * 2};
  cuuint32_t box[2] = {box_c, box_r};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? AINIC_OK : AINIC_ERR_CUDA;
}

static int g_num_sms = 0;
static int num_sms() {
  if (!g_num_sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
  }
  return g_num_sms;
}

template <int BLOCK_N, bool A_MN, bool B_MN, bool OUT_F32>
static int launch(const void* a, int lda, const void* b, int ldb, const EpiParams& ep, int max_ctas,
                  cudaStream_t stream) {
  CUtensorMap ta, tb;
  int rc;
  // A: K-major -> global [M][K]; MN-major -> global [K][M].
  if (A_MN) rc = make_tmap(&ta, a, ep.K, ep.M, lda, 64, BLOCK_K);
  else rc = make_tmap(&ta, a, ep.M, ep.K, lda, BLOCK_K, BLOCK_M);
  if (rc) return rc;
  if (B_MN) rc = make_tmap(&tb, b, ep.K, ep.N, ldb, 64, BLOCK_K);
  else rc = make_tmap(&tb, b, ep.N, ep.K, ldb, BLOCK_K, BLOCK_N);
  if (rc) return rc;
  auto kern = gemm_kernel<BLOCK_N, A_MN, B_MN, OUT_F32>;
  constexpr int smem = num_stages(BLOCK_N) * stage_bytes(BLOCK_N) + 1024 + 256;
  static bool configured = false;
  if (!configured) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return AINIC_ERR_CUDA;
    configured = true;
  }
  const int tiles = ((ep.M + BLOCK_M - 1) / BLOCK_M) * ((ep.N + BLOCK_N - 1) / BLOCK_N);
  int grid = num_sms();
  if (max_ctas > 0 && max_ctas < grid) grid = max_ctas;
  if (tiles < grid) grid = tiles;
  kern<<<grid, kThreads, smem, stream>>>(ta, tb, ep);
  return cudaGetLastError() == cudaSuccess ? AINIC_OK : AINIC_ERR_CUDA;
}

// Wave efficiency of a BLOCK_N choice: useful work / (waves * grid * tile work).
static double wave_eff(int M, int N, int bn, int sms) {
  const long tiles = (long)((M + BLOCK_M - 1) / BLOCK_M) * ((N + bn - 1) / bn);
  const long waves = (tiles + sms - 1) / sms;
  // fixed per-tile cost (prologue/epilogue) favours fewer, larger tiles a little
  return ((double)M * N) / ((double)waves * sms * BLOCK_M * (bn + 24.0));
}

}  // namespace gemm
}  // namespace ainic

using namespace ainic::gemm;

#define DISPATCH_BN(BN, AMN, BMN, F32)                                                     \
  case BN:                                                                                 \
    return launch<BN, AMN, BMN, F32>(a, lda, b, ldb, ep, max_ctas, (cudaStream_t)stream);

extern "C" int ainic_gemm_bf16(int layout, const void* a, int lda, const void* b, int ldb, void* c, int ldc, int M,
                               int N, int K, const ainic_gemm_epilogue_t* epi, int block_n, int max_ctas,
                               void* stream) {
  if (M <= 0 || N <= 0 || K <= 0) return AINIC_ERR_ARG;
  if ((N % 8) || (K % 8) || (lda % 8) || (ldb % 8) || (ldc % 8)) return AINIC_ERR_ARG;
  EpiParams ep{};
  ep.c = c;
  ep.M = M; ep.N = N; ep.K = K; ep.ldc = ldc;
  ep.alpha = 1.0f;
  if (epi) {
    ep.bias = epi->bias;
    ep.mask_out = epi->mask_out;
    ep.mask_in = epi->mask_in;
    ep.relu = epi->relu;
    ep.accumulate = epi->accumulate;
    ep.alpha = epi->alpha;
  }
  const int sms = (max_ctas > 0 && max_ctas < num_sms()) ? max_ctas : num_sms();
  if (block_n <= 0) {
    const int cand_kk[] = {256, 208, 128};
    const int cand_mn[] = {256, 192, 128};
    const int* cand = (layout == AINIC_GEMM_NT) ? cand_kk : cand_mn;
    double best = -1.0;
    for (int i = 0; i < 3; ++i) {
      double e = wave_eff(M, N, cand[i], sms);
      if (e > best * 1.02) { best = e; block_n = cand[i]; }
    }
  }
  switch (layout) {
    case AINIC_GEMM_NT:  // C = A[M,K] . B[N,K]^T, bf16 out
      switch (block_n) {
        DISPATCH_BN(256, false, false, false)
        DISPATCH_BN(208, false, false, false)
        DISPATCH_BN(128, false, false, false)
        default: return AINIC_ERR_ARG;
      }
    case AINIC_GEMM_NN:  // C = A[M,K] . B[K,N], bf16 out
      switch (block_n) {
        DISPATCH_BN(256, false, true, false)
        DISPATCH_BN(192, false, true, false)
        DISPATCH_BN(128, false, true, false)
        default: return AINIC_ERR_ARG;
      }
    case AINIC_GEMM_TN:  // C = A[K,M]^T . B[K,N], fp32 out
      switch (block_n) {
        DISPATCH_BN(256, true, true, true)
        DISPATCH_BN(192, true, true, true)
        DISPATCH_BN(128, true, true, true)
        default: return AINIC_ERR_ARG;
      }
    default:
      return AINIC_ERR_ARG;
  }
}
