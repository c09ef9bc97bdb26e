// Hand-written tcgen05 / TMEM / TMA GEMM for the MLP's three products (bf16 in, fp32 accumulate):
//   forward        Y[mb,n]  = X[mb,k] . W[n,k]^T          (A K-major,  B K-major)   K1
//   backward data  dX[mb,k] = dY[mb,n] . W[n,k]           (A K-major,  B MN-major)  K2
//   backward wgt   dW[n,k]  = dY[mb,n]^T . X[mb,k]        (A MN-major, B MN-major)  K3
// These replace the libxsmm_dnn_fc_{fwd,bwd}_exec_f32 calls of the reference
// (sw/mlp_mpi_example_f32.cpp:708-709, :741-742, :770-771).  The bwd-weight epilogue writes fp32
// straight into the flat gradient bucket the all-reduce engine consumes (the buffer the reference
// hands to the NIC, sw/mlp_mpi_example_f32.cpp:754), optionally accumulating or pre-scaling.
//
// This file: the 1-CTA kernel (tcgen05.mma.cta_group::1, 128 x BLOCK_N tiles) and the public entry
// point, which prefers the 2-CTA kernel (gemm_tcgen05_2cta.cu, 256 x BLOCK_N tiles over an SM pair)
// when the shape allows it.  Structure of both: persistent, warp-specialised, one CTA per SM.
//   warp 0   : TMA producer (one elected lane)         global -> 128B-swizzled smem ring
//   warp 1   : TMEM allocator + MMA issuer (one lane)  tcgen05.mma kind::f16
//   warps 2-5: epilogue                                tcgen05.ld 32x32b -> registers -> global (a CTA's last
//              tile: -> the idle operand ring -> TMA stores, epilogue_store_tma)
// Two TMEM accumulator stages let the epilogue of tile i overlap the main loop of tile i+1; the
// prologue overlaps the previous kernel's tail through programmatic dependent launch.
#include <stdio.h>

#include "gemm_common.cuh"

namespace ainic {
namespace gemm {

__host__ __device__ constexpr int stage_bytes(int block_n) { return BLOCK_M * BLOCK_K * 2 + block_n * BLOCK_K * 2; }
__host__ __device__ constexpr int num_stages(int block_n) {
  return (kSmemBudget / stage_bytes(block_n)) > 8 ? 8 : (kSmemBudget / stage_bytes(block_n));
}
// Compressed output of tiles whose rows never straddle a frame is staged through shared memory
// (epilogue_store_frames): 32 * block_n bytes for each of the four epilogue warps.
__host__ __device__ constexpr bool staged_frames(int block_n, int out) {
  return out == OUT_COMPRESSED_STAGED && 512 % block_n == 0 && block_n >= 128;
}
__host__ __device__ constexpr int staging_bytes(int block_n, int out) { return staged_frames(block_n, out) ? 4 * 32 * block_n : 0; }
__host__ __device__ constexpr uint32_t tmem_cols(int block_n) {
  return 2 * block_n <= 32 ? 32u : 2 * block_n <= 64 ? 64u : 2 * block_n <= 128 ? 128u : 2 * block_n <= 256 ? 256u : 512u;
}

// Register cap: one CTA per SM runs (shared memory), but an all-reduce CTA must fit beside it, and registers
// are allocated per SM sub-partition (16 K each): the all-reduce CTA's 8 warps put two on each sub-partition
// (2 x 32 x 144 registers for the overlap kernels, x 168 for the flow kernel), this kernel's 6 warps up to
// two more, which leaves 112 / 88 registers per thread.  With 96 the flow kernel's doorbells were never
// rung: the GEMM CTAs could not become resident beside it.  The compressed-output kernels are held at 72 so that
// they still fit with eight epilogue warps (AINIC_GEMM_EPI_WARPS=8: three warps per sub-partition).
template <int BLOCK_N, bool A_MN, bool B_MN, int OUT_F32>
__global__ void __maxnreg__(OUT_F32 == OUT_COMPRESSED ? 72 : 88)
gemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
            const __grid_constant__ CUtensorMap tmap_c, EpiParams ep) {
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
  const int epi_warps = (int)(blockDim.x >> 5) - 2;   // 4 or 8 (two per TMEM lane quarter)
  unsigned long long* const trace = ep.trace ? ep.trace + (size_t)blockIdx.x * 8 : nullptr;
  if (trace && threadIdx.x == 0) trace[0] = globaltimer_ns();

  const int m_tiles = (ep.M + BLOCK_M - 1) / BLOCK_M;
  const int n_tiles = (ep.N + BLOCK_N - 1) / BLOCK_N;
  const int num_tiles = m_tiles * n_tiles;
  const int k_blocks = (ep.K + BLOCK_K - 1) / BLOCK_K;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    if (ep.tma_out) tma_prefetch_desc(&tmap_c);
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
  // Programmatic dependent launch: everything above (barrier init, TMEM allocation, tensor-map
  // prefetch) overlaps the tail of the previous kernel in the stream; from here on we touch memory
  // it may have written.
  griddep_wait();
  griddep_launch_dependents();
  if (trace && threadIdx.x == 0) trace[1] = globaltimer_ns();

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
      constexpr uint32_t idesc = make_idesc<BLOCK_M, BLOCK_N, A_MN, B_MN>();
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
          if (trace && kb == 0 && tile == (int)blockIdx.x) trace[2] = globaltimer_ns();
          const uint32_t a_addr = smem_u32(smem_a + stage * kABytes);
          const uint32_t b_addr = smem_u32(smem_b + stage * kBBytes);
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
            umma_f16(tmem_d, operand_desc<A_MN>(a_addr, k), operand_desc<B_MN>(b_addr, k), idesc, (uint32_t)((kb | k) != 0));
          umma_commit(&empty_bar[stage]);
          if (kb == k_blocks - 1) umma_commit(&tmem_full[acc]);
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
      }
      if (trace) trace[3] = globaltimer_ns();
    }
  } else {
    // ===================================================== epilogue (warps 2..5)
    const int quarter = warp & 3;  // TMEM lane quarter this warp may touch
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int m0 = (tile % m_tiles) * BLOCK_M;
      const int n0 = (tile / m_tiles) * BLOCK_N;
      mbar_wait_backoff(&tmem_full[acc], acc_phase, (uint32_t)ep.epi_sleep);
      tc_fence_after();
      if (trace && warp == 2 && lane == 0 && tile == (int)blockIdx.x) trace[4] = globaltimer_ns();
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BLOCK_N);
      if constexpr (staged_frames(BLOCK_N, OUT_F32)) {
        if (epi_warps == 4) {   // compressed frames leave through shared memory, whole row segments at a time
          uint8_t* stage = smem + kStages * (kABytes + kBBytes) + 256 + (warp - 2) * (32 * BLOCK_N);
          epilogue_store_frames<BLOCK_N>(ep, taddr, m0 + quarter * 32, n0, stage, lane);
        } else {
          epilogue_store<BLOCK_N, OUT_COMPRESSED>(ep, taddr, m0 + quarter * 32 + lane, n0, warp >= 6, epi_warps / 4);
        }
      } else if constexpr (OUT_F32 == OUT_BF16 || OUT_F32 == OUT_FP32) {
        if (ep.tma_out && tile + (int)gridDim.x >= num_tiles) {
          // last tile of this CTA: the operand ring is idle, stage through it and let the TMA unit write
          constexpr int kSlab = (BLOCK_N / (OUT_F32 == OUT_FP32 ? 32 : 64)) * 4096;
          static_assert(4 * kSlab <= kStages * (kABytes + kBBytes), "staging boxes must fit in the operand ring");
          epilogue_store_tma<BLOCK_N, OUT_F32>(ep, &tmap_c, taddr, m0 + quarter * 32, n0, smem + quarter * kSlab, lane, warp >= 6,
                                               epi_warps / 4);
        } else {
          epilogue_store<BLOCK_N, OUT_F32>(ep, taddr, m0 + quarter * 32 + lane, n0, warp >= 6, epi_warps / 4);
        }
      } else {
        epilogue_store<BLOCK_N, OUT_F32>(ep, taddr, m0 + quarter * 32 + lane, n0, warp >= 6, epi_warps / 4);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
    if (trace && warp == 2 && lane == 0) trace[5] = globaltimer_ns();
  }

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) {
    __syncwarp();
    tmem_dealloc<kTmemCols>(tmem_base);
  }
  if (trace && threadIdx.x == 0) trace[6] = globaltimer_ns();
  if (ep.done_flag && threadIdx.x == 0) {
    // last-CTA ticket (the grid-barrier pattern: CTA barrier, then ONE thread fences and takes a ticket - the
    // barrier orders the epilogue warps' stores before this thread, the fence makes them visible device-wide
    // before the increment), so whoever observes the flag - ld.acquire.gpu in flow.cuh, then the peers it
    // notifies - sees the complete output
    __threadfence();
    if (atomicAdd(ep.done_ticket, 1u) == gridDim.x - 1) {
      *ep.done_ticket = 0u;
      __threadfence();
      atomicAdd(ep.done_flag, 1u);
    }
  }
}

template <int BLOCK_N, bool A_MN, bool B_MN, int OUT_F32>
static int launch(const void* a, int lda, const void* b, int ldb, const EpiParams& ep, int max_ctas,
                  cudaStream_t stream) {
  CUtensorMap ta, tb, tc;
  int rc;
  // A: K-major -> global [M][K]; MN-major -> global [K][M].
  if (A_MN) rc = cached_tmap(&ta, a, ep.K, ep.M, lda, 64, BLOCK_K);
  else rc = cached_tmap(&ta, a, ep.M, ep.K, lda, BLOCK_K, BLOCK_M);
  if (rc) return rc;
  if (B_MN) rc = cached_tmap(&tb, b, ep.K, ep.N, ldb, 64, BLOCK_K);
  else rc = cached_tmap(&tb, b, ep.N, ep.K, ldb, BLOCK_K, BLOCK_N);
  if (rc) return rc;
  // output map of the TMA-store epilogue: 128-byte-wide, 32-row boxes (one per epilogue warp and column group)
  if (ep.tma_out) rc = cached_tmap(&tc, ep.c, ep.M, ep.N, ep.ldc, OUT_F32 == OUT_FP32 ? 32 : 64, 32, OUT_F32 == OUT_FP32);
  else tc = ta;
  if (rc) return rc;
  auto kern = gemm_kernel<BLOCK_N, A_MN, B_MN, OUT_F32>;
  constexpr int smem = num_stages(BLOCK_N) * stage_bytes(BLOCK_N) + 1024 + 256 + staging_bytes(BLOCK_N, OUT_F32);
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
  cfg.blockDim = dim3(gemm_threads());
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  EpiParams e2 = ep;
  trace_attach(e2);
  if (cudaLaunchKernelEx(&cfg, kern, ta, tb, tc, e2) != cudaSuccess) return AINIC_ERR_CUDA;
  return cudaGetLastError() == cudaSuccess ? AINIC_OK : AINIC_ERR_CUDA;
}

// Wave efficiency of a tile choice: useful work / (waves * slots * tile work).  `rows` is the tile
// height (128 for 1-CTA, 256 for a CTA pair), `slots` the number of tiles that run concurrently.
static double wave_eff(int M, int N, int rows, int bn, int slots) {
  const long tiles = (long)((M + rows - 1) / rows) * ((N + bn - 1) / bn);
  const long waves = (tiles + slots - 1) / slots;
  // fixed per-tile cost (prologue/epilogue) favours fewer, larger tiles a little
  return ((double)M * N) / ((double)waves * slots * rows * (bn + 24.0));
}

}  // namespace gemm
}  // namespace ainic

using namespace ainic::gemm;

#define DISPATCH_BN(BN, AMN, BMN, F32)                                                     \
  case BN:                                                                                 \
    return launch<BN, AMN, BMN, F32>(a, lda, b, ldb, ep, max_ctas, (cudaStream_t)stream);

extern "C" int ainic_gemm_trace(void* buf, int launches) {
  trace_state().buf = static_cast<unsigned long long*>(buf);
  trace_state().left = buf ? launches : 0;
  return AINIC_OK;
}

// block_n: 0 = choose; > 0 = 1-CTA kernel with that tile width; < 0 = 2-CTA kernel with width -block_n.
extern "C" int ainic_gemm_bf16(int layout, const void* a, int lda, const void* b, int ldb, void* c, int ldc, int M,
                               int N, int K, const ainic_gemm_epilogue_t* epi, int block_n, int max_ctas,
                               void* stream) {
  if (M <= 0 || N <= 0 || K <= 0) return AINIC_ERR_ARG;
  if ((N % 8) || (K % 8) || (lda % 8) || (ldb % 8) || (ldc % 8)) return AINIC_ERR_ARG;
  EpiParams ep{};
  ep.c = c;
  ep.M = M; ep.N = N; ep.K = K; ep.ldc = ldc;
  ep.alpha = 1.0f;
  ep.epi_sleep = gemm_epi_sleep_ns();
  if (epi) {
    ep.bias = epi->bias;
    ep.mask_out = epi->mask_out;
    ep.mask_in = epi->mask_in;
    ep.relu = epi->relu;
    ep.accumulate = epi->accumulate;
    ep.alpha = epi->alpha;
    ep.cgrad = static_cast<uint8_t*>(epi->cgrad);
    ep.codec = epi->codec;
    ep.bfp_mode = epi->bfp_mode;
    ep.done_flag = epi->done_flag;
    ep.done_ticket = epi->done_ticket;
    if (ep.done_flag && (!ep.done_ticket || block_n < 0)) return AINIC_ERR_ARG;   // 1-CTA kernels only
    if (const ainic_gemm_scatter_t* sc = epi->scatter) {
      if (sc->nranks < 1 || sc->nranks > AINIC_MAX_RANKS || sc->rank < 0 || sc->rank >= sc->nranks || sc->shard_frames == 0)
        return AINIC_ERR_ARG;
      const size_t frames = ((size_t)M * ldc + 511) / 512;
      if (((size_t)sc->frame0 + frames + sc->shard_frames - 1) / sc->shard_frames > (size_t)sc->nranks) return AINIC_ERR_ARG;
      for (int r = 0; r < sc->nranks; ++r) {
        if (!sc->inbox[r] || ((uintptr_t)sc->inbox[r] & 15)) return AINIC_ERR_ARG;
        ep.push[r] = static_cast<uint8_t*>(sc->inbox[r]);
      }
      ep.push_n = sc->nranks;
      ep.push_rank = sc->rank;
      ep.push_shard = sc->shard_frames;
      ep.push_frame0 = sc->frame0;
      ep.cgrad = ep.push[sc->rank];   // selects the compressed epilogue
      if (sc->nranks == 1) { ep.cgrad += (size_t)sc->frame0 * (ep.codec == AINIC_CODEC_BFP ? 544 : 528); }
    }
  }
  // Staging through shared memory pays for itself when the stores cross NVLink (scatter).  For local
  // frames it is opt-in (AINIC_GEMM_STAGE_FRAMES=1): alone on the GPU it is as fast as the direct path
  // (24.6 vs 25.6 us on the reference layer), inside the training step it measured slower (2 ranks,
  // same box, A/B/A/B: 0.680 vs 0.647 ms/step; cause not isolated).  The staged variant is its own
  // kernel instantiation (143 registers, 32 KB more shared memory), so the direct one stays as small
  // as it was when the headline numbers were taken.
  {
    static const int stage_env = getenv("AINIC_GEMM_STAGE_FRAMES") ? atoi(getenv("AINIC_GEMM_STAGE_FRAMES")) : -1;
    ep.stage_frames = stage_env >= 0 ? stage_env : (ep.push_n > 1 ? 1 : 0);
  }
  ep.tma_out = tma_store_enabled() && c && !ep.cgrad && !ep.accumulate && !ep.mask_out && !ep.mask_in && !ep.done_flag &&
               ((uintptr_t)c & 15) == 0 && (layout != AINIC_GEMM_TN || (ldc % 4) == 0);
  if (ep.cgrad) {  // compressed gradient output: backward-weight layout only, whole 32-value groups
    if (layout != AINIC_GEMM_TN || (N % 32) || (ldc % 32) || ((uintptr_t)ep.cgrad & 15) ||
        (ep.codec != AINIC_CODEC_BFP && ep.codec != AINIC_CODEC_MXFP8))
      return AINIC_ERR_ARG;
    if (block_n < 0) return AINIC_ERR_UNSUPPORTED;
  }
  const int sms = (max_ctas > 0 && max_ctas < num_sms()) ? max_ctas : num_sms();
  if (block_n == 0) {
    // The 2-CTA kernels are correct but currently slower than the 1-CTA ones on the reference shape
    // (profiles/gemm_2cta_ncu.md), so the heuristic only considers them on request.
    static const bool allow_2cta = getenv("AINIC_GEMM_2CTA") && atoi(getenv("AINIC_GEMM_2CTA")) == 1;
    const int cand1_kk[] = {256, 208, 128}, cand1_mn[] = {256, 192, 128};
    const int cand2_kk[] = {256, 208}, cand2_mn[] = {256, 128};
    const bool kk = layout == AINIC_GEMM_NT;
    double best = -1.0;
    for (int i = 0; i < 3; ++i) {
      const int bn = kk ? cand1_kk[i] : cand1_mn[i];
      const double e = wave_eff(M, N, 128, bn, sms) * 0.80;   // 1-CTA main loop is smem-bandwidth bound
      if (e > best * 1.02) { best = e; block_n = bn; }
    }
    if (allow_2cta && M >= 256 && sms >= 2 && !ep.done_flag && !ep.cgrad) {   // no doorbell / compressed epilogue there
      for (int i = 0; i < 2; ++i) {
        const int bn = kk ? cand2_kk[i] : cand2_mn[i];
        const double e = wave_eff(M, N, 256, bn, sms / 2);
        if (e > best * 1.02) { best = e; block_n = -bn; }
      }
    }
  }
  if (block_n < 0) return launch_2cta(layout, a, lda, b, ldb, ep, -block_n, max_ctas, (cudaStream_t)stream);
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
    case AINIC_GEMM_TN:  // C = A[K,M]^T . B[K,N], fp32 out (or compressed frames)
      if (ep.cgrad && ep.stage_frames) {
        switch (block_n) {   // 192-wide tiles straddle frames: direct stores below
          DISPATCH_BN(256, true, true, OUT_COMPRESSED_STAGED)
          DISPATCH_BN(128, true, true, OUT_COMPRESSED_STAGED)
          default: break;
        }
      }
      if (ep.cgrad) {
        switch (block_n) {
          DISPATCH_BN(256, true, true, OUT_COMPRESSED)
          DISPATCH_BN(192, true, true, OUT_COMPRESSED)
          DISPATCH_BN(128, true, true, OUT_COMPRESSED)
          default: return AINIC_ERR_ARG;
        }
      }
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
