// Communicator: symmetric heap over CUDA IPC + host-side dispatch of the fused all-reduce kernels.
//
// Replaces the reference's platform glue: DMA-able buffers from fpga.allocBuffer
// (sw/mlp_mpi_example_f32.cpp:345-347, :396-401), MPF virtual addressing
// (hw/cci_mpf_app_conf.vh:36-39) and the IKL ring routing of sw/setup_route.sh.  NVSwitch is
// all-to-all, so "routing" is just mapping every peer's heap into this process.
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>

#include <mutex>
#include <string>
#include <vector>

#include "allreduce.cuh"
#include "heap_alloc.h"
#include "flow.cuh"

using namespace ainic;
using namespace ainic::comm;

static thread_local std::string g_last_error;
static void set_error(const std::string& s) { g_last_error = s; }
extern "C" const char* ainic_last_error(void) { return g_last_error.c_str(); }
extern "C" const char* ainic_version(void) { return "b200-ainic 0.1 (sm_100a)"; }

#define CUDA_TRY(expr)                                                                         \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess) {                                                                   \
      set_error(std::string(#expr) + ": " + cudaGetErrorString(_e));                           \
      return AINIC_ERR_CUDA;                                                                   \
    }                                                                                          \
  } while (0)

struct ainic_comm {
  int rank = 0, nranks = 1, device = 0;
  size_t heap_bytes = 0;
  ainic::HeapAlloc alloc;                 // offsets inside heap: same sequence of calls = same offsets on every rank
  std::vector<uint8_t*> retired;          // outgrown staging areas of a multi-rank communicator (see ensure_staging_frames)
  uint8_t* heap = nullptr;
  uint8_t* peer_base[kMaxRanks] = {};
  bool connected = false;
  uint32_t* flags = nullptr;              // in heap
  unsigned long long* counters = nullptr; // in heap
  uint8_t* stage_c = nullptr;             // in heap, 2 halves
  uint8_t* stage_d = nullptr;             // in heap
  size_t stage_frames = 0;
  uint32_t* epoch_dev = nullptr;   // device word: number of completed fused calls
  unsigned long long timeout_ns = 10ull * 1000 * 1000 * 1000;
  volatile uint32_t* next_done_slot = nullptr;  // set by the NIC layer for the next launch
};

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
// How many frames ahead of its loads a warp asks the L2 for local weights / gradients (AINIC_PF_DEPTH, 0 = off).
static uint32_t pf_depth() {
  const char* e = getenv("AINIC_PF_DEPTH");   // read per call: cheap, and lets one process compare settings
  const uint32_t d = e ? (uint32_t)atoi(e) : 8u;
  return d > 64 ? 64 : d;
}
// The ring inbox holds ceil(frames/(16N)) blocks x (2N-2) slices; rounding the block count up costs
// at most one block = 14 raw slices of 32 KiB.
static constexpr size_t kRingSlack = 512 * 1024;

extern "C" int ainic_comm_create(ainic_comm_t** out, int rank, int nranks, size_t heap_bytes) {
  if (!out || nranks < 1 || nranks > kMaxRanks || rank < 0 || rank >= nranks) return AINIC_ERR_ARG;
  ainic_comm* c = new ainic_comm();
  c->rank = rank;
  c->nranks = nranks;
  if (cudaGetDevice(&c->device) != cudaSuccess) { delete c; set_error("cudaGetDevice failed"); return AINIC_ERR_CUDA; }
  const size_t reserved = align_up(kFlagWords * sizeof(uint32_t), 4096) + 4096;
  c->heap_bytes = align_up(heap_bytes + reserved, 2u << 20);
  cudaError_t e = cudaMalloc(&c->heap, c->heap_bytes);
  if (e != cudaSuccess) { set_error(std::string("cudaMalloc heap: ") + cudaGetErrorString(e)); delete c; return AINIC_ERR_CUDA; }
  cudaMemset(c->heap, 0, reserved);
  cudaDeviceSynchronize();
  c->flags = reinterpret_cast<uint32_t*>(c->heap);
  c->counters = reinterpret_cast<unsigned long long*>(c->heap + align_up(kFlagWords * sizeof(uint32_t), 4096));
  c->epoch_dev = reinterpret_cast<uint32_t*>(c->counters + 32);   // same zeroed page, past the counters
  c->alloc.reset(c->heap_bytes);
  c->alloc.alloc(reserved, 4096);          // flags / counters / call number: offset 0, never freed
  c->peer_base[rank] = c->heap;
  c->connected = (nranks == 1);
  if (const char* t = getenv("AINIC_TIMEOUT_MS")) c->timeout_ns = strtoull(t, nullptr, 10) * 1000000ull;
  *out = c;
  return AINIC_OK;
}

extern "C" int ainic_comm_export(ainic_comm_t* c, void* handle_out) {
  if (!c || !handle_out) return AINIC_ERR_ARG;
  static_assert(sizeof(cudaIpcMemHandle_t) == AINIC_IPC_HANDLE_BYTES, "IPC handle size");
  cudaIpcMemHandle_t h;
  CUDA_TRY(cudaIpcGetMemHandle(&h, c->heap));
  memcpy(handle_out, &h, sizeof(h));
  return AINIC_OK;
}

extern "C" int ainic_comm_connect(ainic_comm_t* c, const void* all_handles) {
  if (!c || (!all_handles && c->nranks > 1)) return AINIC_ERR_ARG;
  for (int p = 0; p < c->nranks; ++p) {
    if (p == c->rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, static_cast<const uint8_t*>(all_handles) + (size_t)p * sizeof(h), sizeof(h));
    void* ptr = nullptr;
    CUDA_TRY(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
    c->peer_base[p] = static_cast<uint8_t*>(ptr);
  }
  c->connected = true;
  return AINIC_OK;
}

extern "C" int ainic_comm_destroy(ainic_comm_t* c) {
  if (!c) return AINIC_ERR_ARG;
  cudaDeviceSynchronize();
  for (int p = 0; p < c->nranks; ++p)
    if (p != c->rank && c->peer_base[p]) cudaIpcCloseMemHandle(c->peer_base[p]);
  if (c->heap) cudaFree(c->heap);
  delete c;
  return AINIC_OK;
}

extern "C" void* ainic_comm_alloc(ainic_comm_t* c, size_t bytes, size_t align) {
  if (!c || bytes == 0) return nullptr;
  if (align < 256) align = 256;
  if (align & (align - 1)) { set_error("alignment must be a power of two"); return nullptr; }
  const size_t off = c->alloc.alloc(bytes, align);
  if (off == ainic::HeapAlloc::kNone) { set_error("symmetric heap exhausted"); return nullptr; }
  return c->heap + off;
}

// Returns a block to the heap (same call, in the same order, on every rank).  The caller guarantees that no
// kernel that uses the block - on this rank or, through peer pointers, on another - is still running.
extern "C" int ainic_comm_free(ainic_comm_t* c, void* ptr) {
  if (!c || !ptr) return AINIC_ERR_ARG;
  const uint8_t* p = static_cast<const uint8_t*>(ptr);
  if (p <= c->heap || p >= c->heap + c->heap_bytes || !c->alloc.free((size_t)(p - c->heap))) {
    set_error("ainic_comm_free: not a block of this heap");
    return AINIC_ERR_ARG;
  }
  return AINIC_OK;
}
// Collective by contract: call on every rank after a barrier behind which no fused call is in flight anywhere.
// Returns the outgrown staging areas to the heap; the number of blocks released.
extern "C" int ainic_comm_release_retired(ainic_comm_t* c) {
  if (!c) return AINIC_ERR_ARG;
  cudaDeviceSynchronize();
  const int n = (int)c->retired.size();
  for (uint8_t* p : c->retired) ainic_comm_free(c, p);
  c->retired.clear();
  return n;
}
// out[0] bytes in use, out[1] high-water offset, out[2] free blocks below it, out[3] capacity
extern "C" int ainic_comm_heap_stats(ainic_comm_t* c, size_t out[4]) {
  if (!c || !out) return AINIC_ERR_ARG;
  out[0] = c->alloc.bytes_in_use();
  out[1] = c->alloc.high_water();
  out[2] = c->alloc.free_blocks();
  out[3] = c->heap_bytes;
  return AINIC_OK;
}

extern "C" void* ainic_comm_peer_ptr(ainic_comm_t* c, const void* local_ptr, int peer) {
  if (!c || peer < 0 || peer >= c->nranks || !c->peer_base[peer]) return nullptr;
  const uint8_t* p = static_cast<const uint8_t*>(local_ptr);
  if (p < c->heap || p >= c->heap + c->heap_bytes) return nullptr;
  return c->peer_base[peer] + (p - c->heap);
}
extern "C" int ainic_comm_rank(const ainic_comm_t* c) { return c ? c->rank : -1; }
extern "C" int ainic_comm_size(const ainic_comm_t* c) { return c ? c->nranks : -1; }

extern "C" int ainic_comm_counters(ainic_comm_t* c, uint64_t out[8]) {
  if (!c || !out) return AINIC_ERR_ARG;
  CUDA_TRY(cudaMemcpy(out, c->counters, 8 * sizeof(uint64_t), cudaMemcpyDeviceToHost));
  return AINIC_OK;
}

// Time stamps of the last flow call made with AINIC_FLOW_TRACE set: out[5][32] = %globaltimer (ns) when, for chunk k,
// the doorbell was seen / the reduce warps were released / finished / the gather warps were released / finished.
extern "C" int ainic_comm_flow_trace(ainic_comm_t* c, uint64_t out[160]) {
  if (!c || !out) return AINIC_ERR_ARG;
  CUDA_TRY(cudaMemcpy(out, reinterpret_cast<uint8_t*>(c->counters) + 1024, 160 * sizeof(uint64_t), cudaMemcpyDeviceToHost));
  return AINIC_OK;
}

// Internal hooks for the NIC layer (api/nic.cu).
extern "C" void ainic_comm_set_done_slot_(ainic_comm_t* c, volatile uint32_t* dev_visible_slot) { c->next_done_slot = dev_visible_slot; }
extern "C" void ainic_comm_set_timeout_ns_(ainic_comm_t* c, unsigned long long ns) { c->timeout_ns = ns; }

// The all-reduce kernels use (next to) no shared memory, but they share SMs with the persistent GEMM CTAs,
// which need ~200 KB of it.  The L1 / shared-memory split of an SM is fixed while any CTA is resident:
// if an all-reduce CTA lands on an idle SM first - the flow kernel is launched BEFORE the GEMMs that feed
// it - the SM would be configured for a small carve-out and no GEMM CTA could join it until the
// all-reduce CTA leaves (measured: every ready-flag wait of the flow kernel ran into its time-out).
// Asking for the maximum carve-out on the flow kernel keeps every SM GEMM-capable.
static void prefer_max_smem(const void* fn) {
  if (const char* e = getenv("AINIC_COMM_CARVEOUT")) if (atoi(e) == 0) return;   // diagnostic: leave the driver's default split
  static std::mutex mu;
  static const void* done[64];
  static int ndone = 0;
  std::lock_guard<std::mutex> lk(mu);
  for (int i = 0; i < ndone; ++i)
    if (done[i] == fn) return;
  if (cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared) != cudaSuccess)
    cudaGetLastError();
  if (ndone < 64) done[ndone++] = fn;
}

// ---------------------------------------------------------------------------------------- dispatch
typedef void (*KernFn)(const RankCtx);
typedef void (*EmulFn)(const RankCtx*);

template <int ALGO, int CODEC, int MODE>
static KernFn kern_of() {
  if constexpr (ALGO == AINIC_ALGO_RING) return allreduce_kernel<ALGO, CODEC, MODE>;
  else return allreduce_ov_kernel<ALGO, CODEC, MODE>;
}
template <int ALGO, int CODEC, int MODE>
static EmulFn emul_of() {
  if constexpr (ALGO == AINIC_ALGO_RING) return allreduce_emul_kernel<ALGO, CODEC, MODE>;
  else return allreduce_emul_ov_kernel<ALGO, CODEC, MODE>;
}
static int threads_for(int algo) { return algo == AINIC_ALGO_RING ? kCommThreads : kCommThreadsOverlap; }

template <int ALGO>
static KernFn pick_kernel(int codec, int mode) {
  switch (codec) {
    case AINIC_CODEC_F32: return kern_of<ALGO, 0, 0>();
    case AINIC_CODEC_BFP: return mode == AINIC_BFP_HW_EXACT ? kern_of<ALGO, 1, 0>() : kern_of<ALGO, 1, 1>();
    case AINIC_CODEC_MXFP8: return kern_of<ALGO, 2, 0>();
    case AINIC_CODEC_BF16: return kern_of<ALGO, 3, 0>();
  }
  return nullptr;
}
template <int ALGO>
static EmulFn pick_emul(int codec, int mode) {
  switch (codec) {
    case AINIC_CODEC_F32: return emul_of<ALGO, 0, 0>();
    case AINIC_CODEC_BFP: return mode == AINIC_BFP_HW_EXACT ? emul_of<ALGO, 1, 0>() : emul_of<ALGO, 1, 1>();
    case AINIC_CODEC_MXFP8: return emul_of<ALGO, 2, 0>();
    case AINIC_CODEC_BF16: return emul_of<ALGO, 3, 0>();
  }
  return nullptr;
}
static KernFn pick_kernel(int algo, int codec, int mode) {
  switch (algo) {
    case AINIC_ALGO_ONESHOT: return pick_kernel<AINIC_ALGO_ONESHOT>(codec, mode);
    case AINIC_ALGO_TWOSHOT: return pick_kernel<AINIC_ALGO_TWOSHOT>(codec, mode);
    case AINIC_ALGO_RING: return (codec == AINIC_CODEC_F32 || codec == AINIC_CODEC_BFP) ? pick_kernel<AINIC_ALGO_RING>(codec, mode) : nullptr;
  }
  return nullptr;
}
static EmulFn pick_emul(int algo, int codec, int mode) {
  switch (algo) {
    case AINIC_ALGO_ONESHOT: return pick_emul<AINIC_ALGO_ONESHOT>(codec, mode);
    case AINIC_ALGO_TWOSHOT: return pick_emul<AINIC_ALGO_TWOSHOT>(codec, mode);
    case AINIC_ALGO_RING: return (codec == AINIC_CODEC_F32 || codec == AINIC_CODEC_BFP) ? pick_emul<AINIC_ALGO_RING>(codec, mode) : nullptr;
  }
  return nullptr;
}

static int resolve_algo(int algo, int nranks, size_t count) {
  if (algo != AINIC_ALGO_AUTO) return algo;
  // One-shot has one barrier phase instead of two and the same wire bytes as two-shot at N = 2
  // (measured in the trainer at 2 ranks: 0.77 vs 0.88 ms/step); beyond that its (N-1)*S traffic loses
  // to two-shot's 2(N-1)/N*S except for small, latency-bound buckets.
  if (nranks <= 2) return AINIC_ALGO_ONESHOT;
  return count * sizeof(float) <= (512u << 10) ? AINIC_ALGO_ONESHOT : AINIC_ALGO_TWOSHOT;
}

// How many CTAs a call uses.  Enough warps to keep ~2 MB in flight over NVLink, few enough to leave
// most SMs to the concurrent backward GEMMs.
static int default_ctas(int algo, int nranks, size_t count) {
  const size_t frames = bfp_frames(count);
  size_t work = frames;  // frames per rank per phase
  if (algo == AINIC_ALGO_RING) work = (frames + (size_t)nranks * kSliceFrames - 1) / ((size_t)nranks * kSliceFrames);
  else work = (frames + 15) / 16;  // 8 warps per CTA, at least two frames per warp
  size_t ctas = work < 1 ? 1 : work;
  // The kernels are memory/latency bound: a 16 MiB bucket on 64 CTAs took ~70 us, the same bytes on
  // every SM ~15 us.  Run wide and short; the trainer can narrow it with AINIC_COMM_CTAS.
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms <= 0) sms = 148;
  }
  const size_t cap = (size_t)sms;
  if (ctas > cap) ctas = cap;
  return (int)ctas;
}

static int check_ring_limits(int nranks, size_t count, int ctas) {
  const size_t frames = bfp_frames(count);
  const size_t nblocks = (frames + (size_t)nranks * kSliceFrames - 1) / ((size_t)nranks * kSliceFrames);
  const size_t per_cta = (nblocks + ctas - 1) / ctas;
  return per_cta * (2 * nranks - 2) < 4096 ? AINIC_OK : AINIC_ERR_ARG;
}

static int ensure_staging_frames(ainic_comm* c, size_t frames) {
  if (frames <= c->stage_frames) return AINIC_OK;
  // The staging area only grows.  The old one cannot be handed out again right away: a slower peer may still be
  // pulling the previous call's shard from it, and nothing on this rank can know when that ends.  With one rank
  // it is freed at once (after a device synchronise); otherwise it is retired and goes back to the heap in
  // ainic_comm_release_retired(), which the caller invokes when every rank is known to be idle (barrier).
  // Every rank sees the same call sequence, so offsets stay symmetric.
  const size_t want = frames;
  if (c->stage_c) {
    if (c->nranks == 1) {
      cudaDeviceSynchronize();
      ainic_comm_free(c, c->stage_c);
      ainic_comm_free(c, c->stage_d);
    } else {
      c->retired.push_back(c->stage_c);
      c->retired.push_back(c->stage_d);
    }
    c->stage_c = c->stage_d = nullptr;
    c->stage_frames = 0;
  }
  uint8_t* sc = static_cast<uint8_t*>(ainic_comm_alloc(c, 2 * want * 1024, 4096));
  uint8_t* sd = static_cast<uint8_t*>(ainic_comm_alloc(c, 2 * want * 2048 + kRingSlack, 4096));
  if (!sc || !sd) { if (sc) ainic_comm_free(c, sc); return AINIC_ERR_ARG; }
  c->stage_c = sc;
  c->stage_d = sd;
  c->stage_frames = want;
  return AINIC_OK;
}
static int ensure_staging(ainic_comm* c, size_t count) { return ensure_staging_frames(c, bfp_frames(count)); }
// The flow kernel uses one copy of stage C (frames x <= 1024 B) and 1/nranks of the frames plus one
// rounding frame per chunk in stage D: half the frames (+ chunks) of the generic layout cover both.
static size_t flow_staging_frames(const ainic_comm* c, size_t count, int nchunks) {
  const size_t frames = bfp_frames(count);
  return (frames + 1) / 2 + (size_t)nchunks + 1;
}

extern "C" int ainic_comm_fill_ctx_(ainic_comm_t* c, RankCtx* out, const ainic_allreduce_args_t* a);

// Size the internal staging for calls of up to `count` values now (same call on every rank), so the
// first large bucket does not retire a smaller staging area (ensure_staging_frames).
extern "C" int ainic_comm_reserve(ainic_comm_t* c, size_t count) {
  if (!c || count == 0) return AINIC_ERR_ARG;
  return ensure_staging(c, count);
}
extern "C" int ainic_comm_reserve_flow(ainic_comm_t* c, size_t count, int nchunks) {
  if (!c || count == 0 || nchunks < 1) return AINIC_ERR_ARG;
  return ensure_staging_frames(c, flow_staging_frames(c, count, nchunks));
}

extern "C" int ainic_allreduce_fused(ainic_comm_t* c, const ainic_allreduce_args_t* a, void* stream) {
  if (!c || !a || !a->grad || a->count == 0) return AINIC_ERR_ARG;
  if (!c->connected) { set_error("communicator not connected"); return AINIC_ERR_ARG; }
  if (((uintptr_t)a->grad & 15) || ((uintptr_t)a->weight & 15) || ((uintptr_t)a->weight_bf16 & 7) || ((uintptr_t)a->grad_out & 15))
    return AINIC_ERR_ARG;
  const bool scattered = a->grad_scattered && a->grad_compressed && a->codec != AINIC_CODEC_F32;
  const int algo = scattered ? AINIC_ALGO_TWOSHOT : resolve_algo(a->algo, c->nranks, a->count);
  KernFn kern = pick_kernel(algo, a->codec, a->bfp_mode);
  if (!kern) { set_error("unsupported algo/codec combination"); return AINIC_ERR_UNSUPPORTED; }
  if (c->nranks > 1 && !ainic_comm_peer_ptr(c, a->grad, 0) && a->codec == AINIC_CODEC_F32) {
    set_error("fp32 (in-place) all-reduce needs the gradient bucket inside the symmetric heap");
    return AINIC_ERR_ARG;
  }
  int rc = ensure_staging(c, a->count);
  if (rc) return rc;
  int ctas = a->max_ctas > 0 ? a->max_ctas : default_ctas(algo, c->nranks, a->count);
  if (ctas > kMaxCtas) ctas = kMaxCtas;
  if (algo == AINIC_ALGO_RING && check_ring_limits(c->nranks, a->count, ctas)) { set_error("ring: too many blocks per CTA"); return AINIC_ERR_ARG; }

  RankCtx ctx{};
  rc = ainic_comm_fill_ctx_(c, &ctx, a);
  if (rc) return rc;
  // One-/two-shot run with 8-warp CTAs: 256 threads x <= 128 registers leave room for a persistent
  // GEMM CTA (192 threads, ~215 KB smem, ~15 K registers) on the same SM, so the all-reduce
  // overlaps the backward GEMMs instead of evicting them while it waits for peers.  The ring keeps
  // 16 warps (one warp per frame of an 8192-value slice).
  // (no carve-out preference here: these kernels are launched AFTER the GEMMs whose SMs they share, and with the
  // L1 shrunk to the minimum they measured 15-20 % slower on large buckets - 0.433 vs 0.362 ms at 256 MiB, 2 GPUs)
  kern<<<ctas, threads_for(algo), 0, (cudaStream_t)stream>>>(ctx);
  CUDA_TRY(cudaGetLastError());
  return AINIC_OK;
}

// Everything of a RankCtx that belongs to the communicator (peer tables, flags, staging, epoch).
// Consumes one epoch: call exactly once per fused launch, in the same order on every rank.
static int fill_ctx(ainic_comm_t* c, RankCtx* out, const ainic_allreduce_args_t* a, size_t staging_frames);
extern "C" int ainic_comm_fill_ctx_(ainic_comm_t* c, RankCtx* out, const ainic_allreduce_args_t* a) {
  if (!c || !out || !a) return AINIC_ERR_ARG;
  return fill_ctx(c, out, a, bfp_frames(a->count));
}
static int fill_ctx(ainic_comm_t* c, RankCtx* out, const ainic_allreduce_args_t* a, size_t staging_frames) {
  if (!c->connected) { set_error("communicator not connected"); return AINIC_ERR_ARG; }
  int rc = ensure_staging_frames(c, staging_frames);
  if (rc) return rc;
  RankCtx& ctx = *out;
  ctx.rank = c->rank;
  ctx.nranks = c->nranks;
  ctx.grad = a->grad;
  ctx.weight = a->weight;
  ctx.wshadow = static_cast<__nv_bfloat16*>(a->weight_bf16);
  ctx.grad_out = a->grad_out;
  for (int p = 0; p < c->nranks; ++p) {
    ctx.peer_grad[p] = p == c->rank ? a->grad : static_cast<const float*>(ainic_comm_peer_ptr(c, a->grad, p));
    ctx.stage_c[p] = c->peer_base[p] + (c->stage_c - c->heap);
    ctx.stage_d[p] = c->peer_base[p] + (c->stage_d - c->heap);
    ctx.flags[p] = reinterpret_cast<uint32_t*>(c->peer_base[p]);
  }
  ctx.counters = c->counters;
  ctx.done_slot = c->next_done_slot;
  c->next_done_slot = nullptr;
  ctx.count = a->count;
  ctx.stage_c_half = c->stage_frames * 1024;
  ctx.lr = a->lr;
  // the call number lives in device memory (epoch_dev): identical launch arguments from call to call
  ctx.epoch = 0;
  ctx.parity = 0;
  ctx.epoch_ptr = c->epoch_dev;
  ctx.timeout_ns = c->timeout_ns;
  ctx.pf_depth = pf_depth();
  ctx.precompressed = 0;
  if (a->grad_compressed && a->codec != AINIC_CODEC_F32) {
    // The GEMM epilogue already produced this rank's frames: read "stage C" from there.  The caller
    // alternates between two such buffers per bucket, which gives the same reuse guarantee as the
    // parity halves of the internal staging area.
    if (!ainic_comm_peer_ptr(c, a->grad_compressed, c->rank)) {
      set_error("grad_compressed must live in the symmetric heap");
      return AINIC_ERR_ARG;
    }
    for (int p = 0; p < c->nranks; ++p) ctx.stage_c[p] = static_cast<uint8_t*>(ainic_comm_peer_ptr(c, a->grad_compressed, p));
    ctx.parity = 0;
    ctx.stage_c_half = 0;
    ctx.precompressed = 1;
    if (a->grad_scattered) {
      // every rank's GEMM epilogue stored its frames of my shard into this buffer already
      if (a->count % kFrameVals) { set_error("scattered gradients need count % 512 == 0"); return AINIC_ERR_ARG; }
      ctx.precompressed = 2;
    }
  }
  return AINIC_OK;
}
extern "C" void ainic_set_error_(const char* msg) { set_error(msg ? msg : ""); }

// ---------------------------------------------------------------------------------------- flow (chunk-pipelined, persistent)
typedef void (*FlowFn)(const FlowCtx);
typedef void (*FlowEmulFn)(const FlowCtx*);
static FlowFn pick_flow(int codec, int mode) {
  switch (codec) {
    case AINIC_CODEC_F32: return allreduce_flow_kernel<0, 0>;
    case AINIC_CODEC_BFP: return mode == AINIC_BFP_HW_EXACT ? allreduce_flow_kernel<1, 0> : allreduce_flow_kernel<1, 1>;
    case AINIC_CODEC_MXFP8: return allreduce_flow_kernel<2, 0>;
    case AINIC_CODEC_BF16: return allreduce_flow_kernel<3, 0>;
  }
  return nullptr;
}
static FlowEmulFn pick_flow_emul(int codec, int mode) {
  switch (codec) {
    case AINIC_CODEC_F32: return allreduce_flow_emul_kernel<0, 0>;
    case AINIC_CODEC_BFP: return mode == AINIC_BFP_HW_EXACT ? allreduce_flow_emul_kernel<1, 0> : allreduce_flow_emul_kernel<1, 1>;
    case AINIC_CODEC_MXFP8: return allreduce_flow_emul_kernel<2, 0>;
    case AINIC_CODEC_BF16: return allreduce_flow_emul_kernel<3, 0>;
  }
  return nullptr;
}
// Chunk table shared by the real and the emulated entry: frame ranges + stage D slots.
static int fill_chunks(FlowCtx& fc, int nranks, size_t total, const ainic_flow_chunk_t* chunks, int nchunks) {
  if (!chunks || nchunks < 1 || nchunks > kMaxChunks) { set_error("flow: 1..32 chunks"); return AINIC_ERR_ARG; }
  size_t d_off = 0;
  for (int k = 0; k < nchunks; ++k) {
    const ainic_flow_chunk_t& ch = chunks[k];
    if (ch.count == 0 || (ch.offset % kFrameVals) || (ch.count % kFrameVals) || ch.offset + ch.count > total ||
        ch.ready_index > 32767 || ch.ready_need < 0 || ch.ready_need > 65535) {
      set_error("flow: chunk offset / count must be multiples of 512 inside the bucket");
      return AINIC_ERR_ARG;
    }
    fc.chunk[k].frame0 = (uint32_t)(ch.offset / kFrameVals);
    fc.chunk[k].frames = (uint32_t)(ch.count / kFrameVals);
    fc.chunk[k].d_off = (uint32_t)d_off;
    fc.chunk[k].ready_idx = (int16_t)(ch.ready_index < 0 ? -1 : ch.ready_index);
    fc.chunk[k].ready_need = (uint16_t)ch.ready_need;
    d_off += (fc.chunk[k].frames + nranks - 1) / nranks;
  }
  fc.nchunks = nchunks;
  return AINIC_OK;
}

extern "C" int ainic_allreduce_flow(ainic_comm_t* c, const ainic_allreduce_args_t* a, const ainic_flow_chunk_t* chunks,
                                    int nchunks, const uint32_t* local_ready, void* stream) {
  if (!c || !a || !a->grad || a->count == 0) return AINIC_ERR_ARG;
  if (!c->connected) { set_error("communicator not connected"); return AINIC_ERR_ARG; }
  if (((uintptr_t)a->grad & 15) || ((uintptr_t)a->weight & 15) || ((uintptr_t)a->weight_bf16 & 7) || ((uintptr_t)a->grad_out & 15))
    return AINIC_ERR_ARG;
  if (a->grad_scattered) { set_error("flow: scattered input is not supported"); return AINIC_ERR_UNSUPPORTED; }
  if (a->codec != AINIC_CODEC_F32 && !a->grad_compressed) {
    set_error("flow: a compressed codec needs grad_compressed (frames written by the GEMM epilogue or a codec kernel)");
    return AINIC_ERR_UNSUPPORTED;
  }
  FlowFn kern = pick_flow(a->codec, a->bfp_mode);
  if (!kern) { set_error("unsupported codec"); return AINIC_ERR_UNSUPPORTED; }
  if (c->nranks > 1 && !ainic_comm_peer_ptr(c, a->grad, 0) && a->codec == AINIC_CODEC_F32) {
    set_error("fp32 (in-place) all-reduce needs the gradient bucket inside the symmetric heap");
    return AINIC_ERR_ARG;
  }
  FlowCtx fc;
  memset(&fc, 0, sizeof(fc));
  int rc = fill_ctx(c, &fc.c, a, flow_staging_frames(c, a->count, nchunks));
  if (rc) return rc;
  rc = fill_chunks(fc, c->nranks, a->count, chunks, nchunks);
  if (rc) return rc;
  fc.local_ready = local_ready;
  // per-chunk time stamps of the last call (diagnostic; 1280 bytes inside the counters page)
  fc.trace = getenv("AINIC_FLOW_TRACE") ? reinterpret_cast<unsigned long long*>(reinterpret_cast<uint8_t*>(c->counters) + 1024) : nullptr;
  fc.c.epoch_ptr = c->epoch_dev + 1;   // flow calls count separately: their flag words (phases 4, 5) are their own
  fc.c.stage_c_half = 0;
  fc.c.parity = 0;
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms <= 0) sms = 148;
  }
  int ctas = a->max_ctas > 0 ? a->max_ctas : sms - 1;   // persistent workers beside the GEMM CTAs ...
  if (ctas > sms - 1) ctas = sms - 1;
  if (ctas < 1) ctas = 1;
  prefer_max_smem((const void*)kern);
  kern<<<ctas + 1, kCommThreadsOverlap, 0, (cudaStream_t)stream>>>(fc);   // ... plus the forwarder CTA
  CUDA_TRY(cudaGetLastError());
  return AINIC_OK;
}

extern "C" int ainic_allreduce_flow_emulate(int nranks, float* grads, float* weights, void* weights_bf16, float* grads_out,
                                            size_t count, float lr, int codec, int bfp_mode, const size_t* chunk_counts,
                                            int nchunks, int ctas_per_rank, void* stream) {
  if (nranks < 1 || nranks > kMaxRanks || !grads || count == 0 || !chunk_counts) return AINIC_ERR_ARG;
  if (((uintptr_t)grads & 15) || ((uintptr_t)weights & 15) || ((uintptr_t)grads_out & 15) || ((uintptr_t)weights_bf16 & 7) || (count % kFrameVals)) {
    set_error("flow emulate: 16-byte aligned buffers and count % 512 == 0");
    return AINIC_ERR_ARG;
  }
  FlowEmulFn kern = pick_flow_emul(codec, bfp_mode);
  if (!kern) return AINIC_ERR_UNSUPPORTED;
  ainic_flow_chunk_t chunks[kMaxChunks];
  if (nchunks < 1 || nchunks > kMaxChunks) return AINIC_ERR_ARG;
  {
    size_t off = 0;
    for (int k = 0; k < nchunks; ++k) off += chunk_counts[k];
    if (off != count) { set_error("flow emulate: chunk counts must add up to count"); return AINIC_ERR_ARG; }
    for (int k = 0; k < nchunks; ++k) {   // last chunk first, like the layers of a backward pass
      off -= chunk_counts[nchunks - 1 - k];
      chunks[k] = {off, chunk_counts[nchunks - 1 - k], k, 1};
    }
  }
  const int ctas = ctas_per_rank > 0 ? ctas_per_rank : 4;
  int dev = 0, sms = 0, per_sm = 0;
  CUDA_TRY(cudaGetDevice(&dev));
  CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kCommThreadsOverlap, 0));
  if ((long)(ctas + 1) * nranks > (long)sms * per_sm) { set_error("emulation grid exceeds co-resident capacity"); return AINIC_ERR_ARG; }
  const size_t frames = bfp_frames(count);
  const size_t c_bytes = align_up(frames * 1024, 256), d_bytes = align_up((frames + (size_t)nchunks * nranks) * 2048, 256);
  const size_t flag_bytes = kFlagWords * sizeof(uint32_t), ctr_bytes = 256;
  const size_t per_rank = c_bytes + d_bytes + flag_bytes + ctr_bytes;
  uint8_t* scratch = nullptr;
  FlowCtx* d_ctx = nullptr;
  CUDA_TRY(cudaMalloc(&scratch, per_rank * nranks + 256));
  if (cudaMalloc(&d_ctx, sizeof(FlowCtx) * nranks) != cudaSuccess) { cudaFree(scratch); return AINIC_ERR_CUDA; }
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(scratch, 0, per_rank * nranks + 256, s);
  // local ready counters: all producers "finished" (value 1) - the wait path still runs
  uint32_t* ready = reinterpret_cast<uint32_t*>(scratch + per_rank * nranks);
  if (e == cudaSuccess) e = cudaMemsetAsync(ready, 1, 256, s);   // every word 0x01010101 >= 1
  FlowCtx* h = new FlowCtx[nranks];
  memset(h, 0, sizeof(FlowCtx) * nranks);
  auto base = [&](int r) { return scratch + (size_t)r * per_rank; };
  int rc = AINIC_OK;
  for (int r = 0; r < nranks && rc == AINIC_OK; ++r) {
    RankCtx& x = h[r].c;
    x.rank = r;
    x.nranks = nranks;
    x.grad = grads + (size_t)r * count;
    x.weight = weights ? weights + (size_t)r * count : nullptr;
    x.wshadow = weights_bf16 ? static_cast<__nv_bfloat16*>(weights_bf16) + (size_t)r * count : nullptr;
    x.grad_out = grads_out ? grads_out + (size_t)r * count : nullptr;
    for (int p = 0; p < nranks; ++p) {
      x.peer_grad[p] = grads + (size_t)p * count;
      x.stage_c[p] = base(p);
      x.stage_d[p] = base(p) + c_bytes;
      x.flags[p] = reinterpret_cast<uint32_t*>(base(p) + c_bytes + d_bytes);
    }
    x.counters = reinterpret_cast<unsigned long long*>(base(r) + c_bytes + d_bytes + flag_bytes);
    x.count = count;
    x.lr = lr;
    x.epoch = 1;
    x.timeout_ns = 5ull * 1000 * 1000 * 1000;
    x.pf_depth = pf_depth();
    rc = fill_chunks(h[r], nranks, count, chunks, nchunks);
    h[r].local_ready = ready;
    // the flow kernel takes its input in wire format: encode every rank's gradient into its stage C first
    if (rc == AINIC_OK && codec != AINIC_CODEC_F32) {
      x.precompressed = 1;
      if (codec == AINIC_CODEC_BFP) rc = ainic_bfp_compress(x.grad, count, base(r), bfp_mode, s);
      else if (codec == AINIC_CODEC_MXFP8) rc = ainic_mx_compress(x.grad, count, base(r), s);
      else rc = ainic_cast_f32_bf16(x.grad, base(r), count, s);
    }
  }
  if (rc == AINIC_OK && e == cudaSuccess) e = cudaMemcpyAsync(d_ctx, h, sizeof(FlowCtx) * nranks, cudaMemcpyHostToDevice, s);
  const FlowCtx* arg0 = d_ctx;
  void* args[] = {(void*)&arg0};
  if (rc == AINIC_OK && e == cudaSuccess) e = cudaLaunchCooperativeKernel((const void*)kern, dim3(ctas + 1, nranks), dim3(kCommThreadsOverlap), args, 0, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  unsigned long long ctr[8] = {};
  if (rc == AINIC_OK && e == cudaSuccess) cudaMemcpy(ctr, h[0].c.counters, sizeof(ctr), cudaMemcpyDeviceToHost);
  delete[] h;
  cudaFree(scratch);
  cudaFree(d_ctx);
  if (rc) return rc;
  if (e != cudaSuccess) { set_error(std::string("flow emulate: ") + cudaGetErrorString(e)); return AINIC_ERR_CUDA; }
  if (ctr[CTR_TIMEOUTS]) { set_error("flow emulate: flag wait timed out"); return AINIC_ERR_TIMEOUT; }
  return AINIC_OK;
}

// ---------------------------------------------------------------------------------------- emulation
extern "C" int ainic_allreduce_emulate(int nranks, float* grads, float* weights, void* weights_bf16, float* grads_out,
                                       size_t count, float lr, int codec, int bfp_mode, int algo, int ctas_per_rank,
                                       void* stream) {
  if (nranks < 1 || nranks > kMaxRanks || !grads || count == 0) return AINIC_ERR_ARG;
  // rank r's rows start at base + r * count and are accessed with 128-bit loads / stores
  if (((uintptr_t)grads & 15) || ((uintptr_t)weights & 15) || ((uintptr_t)grads_out & 15) || ((uintptr_t)weights_bf16 & 7) ||
      (nranks > 1 && (count & 3))) {
    set_error("emulate: buffers must be 16-byte aligned and count a multiple of 4 when nranks > 1");
    return AINIC_ERR_ARG;
  }
  algo = resolve_algo(algo, nranks, count);
  EmulFn kern = pick_emul(algo, codec, bfp_mode);
  if (!kern) return AINIC_ERR_UNSUPPORTED;
  int ctas = ctas_per_rank > 0 ? ctas_per_rank : 4;
  if (algo == AINIC_ALGO_RING && check_ring_limits(nranks, count, ctas)) return AINIC_ERR_ARG;
  int dev = 0, sms = 0, per_sm = 0;
  CUDA_TRY(cudaGetDevice(&dev));
  CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads_for(algo), 0));
  if ((long)ctas * nranks > (long)sms * per_sm) { set_error("emulation grid exceeds co-resident capacity"); return AINIC_ERR_ARG; }

  const size_t frames = bfp_frames(count);
  const size_t c_bytes = 2 * frames * 1024, d_bytes = 2 * frames * 2048 + kRingSlack;
  const size_t flag_bytes = kFlagWords * sizeof(uint32_t), ctr_bytes = 64;
  const size_t per_rank = align_up(c_bytes, 256) + align_up(d_bytes, 256) + flag_bytes + ctr_bytes;
  uint8_t* scratch = nullptr;
  RankCtx* d_ctx = nullptr;
  CUDA_TRY(cudaMalloc(&scratch, per_rank * nranks));
  if (cudaMalloc(&d_ctx, sizeof(RankCtx) * nranks) != cudaSuccess) {
    cudaFree(scratch);
    set_error("emulate: cudaMalloc of the context table failed");
    return AINIC_ERR_CUDA;
  }
  cudaStream_t s = (cudaStream_t)stream;
  if (cudaMemsetAsync(scratch, 0, per_rank * nranks, s) != cudaSuccess) {
    cudaFree(scratch);
    cudaFree(d_ctx);
    set_error("emulate: clearing the scratch failed");
    return AINIC_ERR_CUDA;
  }
  RankCtx h[kMaxRanks];
  memset(h, 0, sizeof(h));
  auto base = [&](int r) { return scratch + (size_t)r * per_rank; };
  for (int r = 0; r < nranks; ++r) {
    RankCtx& x = h[r];
    x.rank = r;
    x.nranks = nranks;
    x.grad = grads + (size_t)r * count;
    x.weight = weights ? weights + (size_t)r * count : nullptr;
    x.wshadow = weights_bf16 ? static_cast<__nv_bfloat16*>(weights_bf16) + (size_t)r * count : nullptr;
    x.grad_out = grads_out ? grads_out + (size_t)r * count : nullptr;
    for (int p = 0; p < nranks; ++p) {
      x.peer_grad[p] = grads + (size_t)p * count;
      x.stage_c[p] = base(p);
      x.stage_d[p] = base(p) + align_up(c_bytes, 256);
      x.flags[p] = reinterpret_cast<uint32_t*>(base(p) + align_up(c_bytes, 256) + align_up(d_bytes, 256));
    }
    x.counters = reinterpret_cast<unsigned long long*>(base(r) + align_up(c_bytes, 256) + align_up(d_bytes, 256) + flag_bytes);
    x.done_slot = nullptr;
    x.count = count;
    x.stage_c_half = frames * 1024;
    x.lr = lr;
    x.epoch = 1;
    x.parity = 0;
    x.timeout_ns = 5ull * 1000 * 1000 * 1000;
    x.pf_depth = pf_depth();
  }
  cudaError_t e = cudaMemcpyAsync(d_ctx, h, sizeof(RankCtx) * nranks, cudaMemcpyHostToDevice, s);
  const RankCtx* arg0 = d_ctx;
  void* args[] = {(void*)&arg0};
  if (e == cudaSuccess) e = cudaLaunchCooperativeKernel((const void*)kern, dim3(ctas, nranks), dim3(threads_for(algo)), args, 0, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  unsigned long long ctr[8] = {};
  if (e == cudaSuccess) cudaMemcpy(ctr, h[0].counters, sizeof(ctr), cudaMemcpyDeviceToHost);
  cudaFree(scratch);
  cudaFree(d_ctx);
  if (e != cudaSuccess) { set_error(std::string("emulate: ") + cudaGetErrorString(e)); return AINIC_ERR_CUDA; }
  if (ctr[CTR_TIMEOUTS]) { set_error("emulate: flag wait timed out"); return AINIC_ERR_TIMEOUT; }
  return AINIC_OK;
}
