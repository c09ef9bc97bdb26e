// 2-CTA tcgen05 GEMM: a thread-block cluster of two CTAs (an SM pair) computes one 256 x BLOCK_N
// output tile with tcgen05.mma.cta_group::2.  Each CTA stages its own 128 rows of A and HALF of the
// B tile (BLOCK_N / 2 columns); the tensor cores of the pair read both halves, so per SM the
// shared-memory traffic per FLOP (TMA writes + operand reads share one 128 B/clk port) drops by a
// third against the 1-CTA 128 x 256 kernel - the measured limiter of that kernel
// (profiles/gemm_1cta_ncu.md: tensor pipe ~50-60 % active while L2 sits at 25 %).
//
// Roles per CTA (192 threads): warp 0 TMA producer, warp 1 TMEM allocation (+ MMA issue in the
// leader CTA only), warps 2-5 epilogue over the CTA's own 128 accumulator rows.
// Barriers: the leader's full[] collect the bytes of both CTAs' TMA loads; tcgen05.commit
// multicasts to empty[] / tmem_full[] in both CTAs; the peer's epilogue releases the accumulator
// with a remote arrive on the leader's tmem_empty[].
// Same math, same epilogue and same layouts (NT / NN / TN) as gemm_tcgen05.cu; reference
// counterpart: libxsmm_dnn_fc_{fwd,bwd}_exec_f32 (sw/mlp_mpi_example_f32.cpp:708-709, :741-742).
#include "gemm_common.cuh"

namespace ainic {
namespace gemm {

__host__ __device__ constexpr int stage_bytes2(int block_n) { return BLOCK_M * BLOCK_K * 2 + (block_n / 2) * BLOCK_K * 2; }
__host__ __device__ constexpr int num_stages2(int block_n) {
  return (kSmemBudget / stage_bytes2(block_n)) > 8 ? 8 : (kSmemBudget / stage_bytes2(block_n));
}
__host__ __device__ constexpr uint32_t tmem_cols2(int block_n) { return 2 * block_n <= 128 ? 128u : 2 * block_n <= 256 ? 256u : 512u; }

template <int BLOCK_N, bool A_MN, bool B_MN, int OUT_F32>
__global__ void __launch_bounds__(kThreads, 1)
gemm2_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, EpiParams ep) {
  constexpr int HALF_N = BLOCK_N / 2;
  constexpr int kStages = num_stages2(BLOCK_N);
  constexpr int kABytes = BLOCK_M * BLOCK_K * 2;
  constexpr int kBBytes = HALF_N * BLOCK_K * 2;
  constexpr uint32_t kTmemCols = tmem_cols2(BLOCK_N);
  static_assert(BLOCK_N % 16 == 0 && BLOCK_N >= 32 && BLOCK_N <= 256, "invalid UMMA N for M = 256");
  static_assert(HALF_N % 8 == 0, "B half tile must be whole 8-row core matrices");
  static_assert(!B_MN || HALF_N % 64 == 0, "MN-major B halves are built from 64-wide TMA boxes");

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
  const uint32_t cta_rank = cluster_ctarank();   // 0 = leader (issues the MMAs), 1 = peer
  const bool leader = cta_rank == 0;
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;
  unsigned long long* const trace = ep.trace ? ep.trace + (size_t)blockIdx.x * 8 : nullptr;
  if (trace && threadIdx.x == 0) trace[0] = globaltimer_ns();

  const int m_tiles = (ep.M + 2 * BLOCK_M - 1) / (2 * BLOCK_M);
  const int n_tiles = (ep.N + BLOCK_N - 1) / BLOCK_N;
  const int num_tiles = m_tiles * n_tiles;
  const int k_blocks = (ep.K + BLOCK_K - 1) / BLOCK_K;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);    // leader's arrive.expect_tx; the byte count covers both CTAs' loads
      mbar_init(&empty_bar[s], 1);   // multicast tcgen05.commit
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);   // multicast tcgen05.commit
      mbar_init(&tmem_empty[s], 2 * epi_warps);  // leader: the epilogue warps of both CTAs
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc_2cta<kTmemCols>(tmem_ptr);
  tc_fence_before();
  cluster_sync();                    // both CTAs' barriers are initialised before any remote arrive
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  griddep_wait();
  griddep_launch_dependents();
  if (trace && threadIdx.x == 0) trace[1] = globaltimer_ns();

  if (warp == 0) {
    // ===================================================== TMA producer (both CTAs)
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        const int m0 = (tile % m_tiles) * (2 * BLOCK_M) + (int)cta_rank * BLOCK_M;
        const int n0 = (tile / m_tiles) * BLOCK_N + (int)cta_rank * HALF_N;
        for (int kb = 0; kb < k_blocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          // Only the leader arms its barrier, with the bytes of BOTH CTAs.  The peer needs no arrive of
          // its own: its loads can only be issued after the MMAs of the previous round on this stage
          // retired, i.e. inside the same barrier phase, and a complete_tx that lands before the
          // expect_tx just drives the tx-count negative for a while.  (A per-stage remote
          // mbarrier.arrive.release.cluster made the peer's producer wait on a membar for its own
          // in-flight TMA loads - 30 % tensor-pipe utilisation, profiles/gemm_2cta_ncu.md.)
          if (leader) mbar_expect_tx(&full_bar[stage], 2 * (kABytes + kBBytes));
          uint8_t* sa = smem_a + stage * kABytes;
          uint8_t* sb = smem_b + stage * kBBytes;
          const int k0 = kb * BLOCK_K;
          if constexpr (A_MN) {
#pragma unroll
            for (int j = 0; j < BLOCK_M / 64; ++j) tma_load_2d_2cta(sa + j * 8192, &tmap_a, &full_bar[stage], m0 + 64 * j, k0);
          } else {
            tma_load_2d_2cta(sa, &tmap_a, &full_bar[stage], k0, m0);
          }
          if constexpr (B_MN) {
#pragma unroll
            for (int j = 0; j < HALF_N / 64; ++j) tma_load_2d_2cta(sb + j * 8192, &tmap_b, &full_bar[stage], n0 + 64 * j, k0);
          } else {
            tma_load_2d_2cta(sb, &tmap_b, &full_bar[stage], k0, n0);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer (leader CTA only)
    if (leader && lane == 0) {
      constexpr uint32_t idesc = make_idesc<2 * BLOCK_M, BLOCK_N, A_MN, B_MN>();
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)(acc * BLOCK_N);
        for (int kb = 0; kb < k_blocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (trace && kb == 0 && tile == cluster_id) trace[2] = globaltimer_ns();
          const uint32_t a_addr = smem_u32(smem_a + stage * kABytes);
          const uint32_t b_addr = smem_u32(smem_b + stage * kBBytes);
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
            umma_f16_2cta(tmem_d, operand_desc<A_MN>(a_addr, k), operand_desc<B_MN>(b_addr, k), idesc, (uint32_t)((kb | k) != 0));
          umma_commit_2cta(&empty_bar[stage], 3);                       // frees the slot in both CTAs
          if (kb == k_blocks - 1) umma_commit_2cta(&tmem_full[acc], 3);  // accumulator ready in both CTAs
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
      }
      if (trace) trace[3] = globaltimer_ns();
    }
  } else {
    // ===================================================== epilogue (warps 2..5, both CTAs)
    const int quarter = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
      const int m0 = (tile % m_tiles) * (2 * BLOCK_M) + (int)cta_rank * BLOCK_M;
      const int n0 = (tile / m_tiles) * BLOCK_N;
      mbar_wait_backoff(&tmem_full[acc], acc_phase, (uint32_t)ep.epi_sleep);
      tc_fence_after();
      if (trace && warp == 2 && lane == 0 && tile == cluster_id) trace[4] = globaltimer_ns();
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BLOCK_N);
      epilogue_store<BLOCK_N, OUT_F32>(ep, taddr, m0 + quarter * 32 + lane, n0, warp >= 6, epi_warps / 4);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (leader) mbar_arrive(&tmem_empty[acc]);
        else mbar_arrive_remote(&tmem_empty[acc], 0);
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
    if (trace && warp == 2 && lane == 0) trace[5] = globaltimer_ns();
  }

  tc_fence_before();
  cluster_sync();                    // the peer's smem / TMEM stay valid until the leader's last MMA retired
  tc_fence_after();
  if (warp == 1) {
    __syncwarp();
    tmem_dealloc_2cta<kTmemCols>(tmem_base);
  }
  if (trace && threadIdx.x == 0) trace[6] = globaltimer_ns();
}

template <int BLOCK_N, bool A_MN, bool B_MN, int OUT_F32>
static int launch2(const void* a, int lda, const void* b, int ldb, const EpiParams& ep, int max_ctas, cudaStream_t stream) {
  constexpr int HALF_N = BLOCK_N / 2;
  CUtensorMap ta, tb;
  int rc;
  if (A_MN) rc = cached_tmap(&ta, a, ep.K, ep.M, lda, 64, BLOCK_K);
  else rc = cached_tmap(&ta, a, ep.M, ep.K, lda, BLOCK_K, BLOCK_M);
  if (rc) return rc;
  if (B_MN) rc = cached_tmap(&tb, b, ep.K, ep.N, ldb, 64, BLOCK_K);
  else rc = cached_tmap(&tb, b, ep.N, ep.K, ldb, BLOCK_K, HALF_N);
  if (rc) return rc;
  auto kern = gemm2_kernel<BLOCK_N, A_MN, B_MN, OUT_F32>;
  constexpr int smem = num_stages2(BLOCK_N) * stage_bytes2(BLOCK_N) + 1024 + 256;
  static bool configured = false;
  if (!configured) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return AINIC_ERR_CUDA;
    configured = true;
  }
  const int tiles = ((ep.M + 2 * BLOCK_M - 1) / (2 * BLOCK_M)) * ((ep.N + BLOCK_N - 1) / BLOCK_N);
  int sms = num_sms();
  if (max_ctas > 0 && max_ctas < sms) sms = max_ctas;
  int clusters = sms / 2;
  if (clusters < 1) clusters = 1;
  if (tiles < clusters) clusters = tiles;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(2 * clusters);
  cfg.blockDim = dim3(gemm_threads());
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  EpiParams e2 = ep;
  trace_attach(e2);
  if (cudaLaunchKernelEx(&cfg, kern, ta, tb, e2) != cudaSuccess) return AINIC_ERR_CUDA;
  return cudaGetLastError() == cudaSuccess ? AINIC_OK : AINIC_ERR_CUDA;
}

int launch_2cta(int layout, const void* a, int lda, const void* b, int ldb, const EpiParams& ep, int block_n,
                int max_ctas, cudaStream_t stream) {
  switch (layout) {
    case AINIC_GEMM_NT:
      if (block_n == 256) return launch2<256, false, false, false>(a, lda, b, ldb, ep, max_ctas, stream);
      if (block_n == 208) return launch2<208, false, false, false>(a, lda, b, ldb, ep, max_ctas, stream);
      if (block_n == 128) return launch2<128, false, false, false>(a, lda, b, ldb, ep, max_ctas, stream);
      break;
    case AINIC_GEMM_NN:
      if (block_n == 256) return launch2<256, false, true, false>(a, lda, b, ldb, ep, max_ctas, stream);
      if (block_n == 128) return launch2<128, false, true, false>(a, lda, b, ldb, ep, max_ctas, stream);
      break;
    case AINIC_GEMM_TN:
      if (block_n == 256) return launch2<256, true, true, true>(a, lda, b, ldb, ep, max_ctas, stream);
      if (block_n == 128) return launch2<128, true, true, true>(a, lda, b, ldb, ep, max_ctas, stream);
      break;
  }
  return AINIC_ERR_UNSUPPORTED;
}

}  // namespace gemm
}  // namespace ainic
