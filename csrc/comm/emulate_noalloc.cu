// ainic_allreduce_emulate_noalloc: the single-GPU N-rank all-reduce emulation for callers that are
// not allowed to allocate device memory.
//
// Who that is: a service that handed all of HBM to its own pool at start-up, so a fresh cudaMalloc
// would simply fail; and latency-critical loops that cannot afford the implicit device-wide
// synchronisation cudaFree performs (every stream of the process stalls behind it).  ainic_allreduce_emulate() is
// unusable for them because it cudaMallocs its staging area, flag words, counters and per-rank
// context table on every call and frees them afterwards.
//
// This variant keeps all of that scratch in ONE file-scope __device__ array of fixed size
// (kScratchBytes = 2 MiB, resident on every GPU that loads this library until the process exits)
// and passes the per-rank contexts to the kernel by value, so the call performs no allocation of
// any kind: no cudaMalloc, no cudaFree, no allocating library routine.  The price is a hard limit:
// at most AINIC_NOALLOC_MAX_RANKS (4) emulated ranks and AINIC_NOALLOC_MAX_COUNT (32768) values
// per rank; anything larger is rejected with AINIC_ERR_ARG and a message in errbuf.
//
// Rules for callers:
//  * The first call on each device looks up the device address of the scratch array
//    (cudaGetSymbolAddress, with that device current); the very first call in the process also
//    page-locks a 4 KiB host array the counters are copied into.  Make that first call - with the
//    algorithm / codec you will use later, so the kernel is loaded - before any stream capture
//    starts in the process.
//  * All calls share the one array and it keeps whatever the previous call left in it.  The
//    entry point re-zeroes the flag words and counters on `stream` before every launch; the
//    staging bytes need no clearing because every byte that is read was written earlier in the
//    same call.  Two calls whose GPU work overlaps (different streams, or two host threads) would
//    trample each other's flags and staging data - serialise them.
//  * Like ainic_allreduce_emulate() the call returns after the kernel has finished.  All of its
//    work (flag clearing, the launch, the counter read-back) is asynchronous work on `stream`
//    followed by one cudaStreamSynchronize(stream); nothing is issued to the legacy default
//    stream.  Measured on B200 / CUDA 12.9: a call made from a second thread while the first holds
//    a cudaStreamCaptureModeGlobal capture open still invalidates that capture (the cooperative
//    launch and the synchronisation are outside what that mode tolerates from other threads), so
//    keep such captures in thread-local or relaxed mode if this entry point runs beside them.
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>

#include <mutex>

#include "allreduce.cuh"

using namespace ainic;
using namespace ainic::comm;

namespace {

constexpr int kNaMaxRanks = AINIC_NOALLOC_MAX_RANKS;
constexpr size_t kNaMaxCount = AINIC_NOALLOC_MAX_COUNT;
constexpr size_t kNaMaxFrames = (kNaMaxCount + kFrameVals - 1) / kFrameVals;            // 64
constexpr size_t kNaStageC = 2 * kNaMaxFrames * 1024;                                     // both parity halves, widest codec (bf16)
// staging D holds the two-shot published shard (<= frames * 2048 B) or the ring inbox:
// ceil(frames / (16 N)) blocks x (2N - 2) raw slices of 32 KiB; the worst case for N <= 4 and 64 frames is N = 3.
constexpr size_t kNaStageD = 256 * 1024;
constexpr size_t kNaFlags = (size_t)kFlagWords * sizeof(uint32_t);                        // 64 KiB
constexpr size_t kNaCounters = 256;
constexpr size_t kNaPerRank = kNaStageC + kNaStageD + kNaFlags + kNaCounters;
constexpr size_t kScratchBytes = 2u << 20;
static_assert(kNaPerRank * kNaMaxRanks <= kScratchBytes, "scratch layout does not fit the static array");

// The scratch itself.  In host code this name is only a handle for cudaGetSymbolAddress - it is
// not a device address and must never be handed to a kernel directly.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_setup_mu;
unsigned char* g_scratch_dev_ptr[kMaxDevices] = {};
// Host landing area for the counter read-back.  Page-locked once (cudaHostRegister pins existing
// memory, it allocates nothing) so that the copy is a true asynchronous DMA on the caller's stream;
// if pinning is refused the copy still works through the driver's pageable path.
alignas(4096) unsigned long long g_ctr_host[4096 / sizeof(unsigned long long)];
bool g_ctr_host_tried = false;
static_assert(sizeof(g_ctr_host) >= sizeof(unsigned long long) * kNaMaxRanks * kNumCounters, "counter landing area");

struct EmulArgs {
  RankCtx ctx[kNaMaxRanks];
};
static_assert(sizeof(EmulArgs) <= 4000, "contexts must fit in kernel parameter space");
typedef void (*ByValFnT)(const EmulArgs);

template <int ALGO, int CODEC, int MODE>
__global__ void AINIC_COMM_RING_ATTR allreduce_emul_byval_kernel(const __grid_constant__ EmulArgs args) {
  run_rank<ALGO, CODEC, MODE>(args.ctx[blockIdx.y], blockIdx.x, gridDim.x);
}
template <int ALGO, int CODEC, int MODE>
__global__ void AINIC_COMM_OVERLAP_ATTR allreduce_emul_byval_ov_kernel(const __grid_constant__ EmulArgs args) {
  run_rank<ALGO, CODEC, MODE>(args.ctx[blockIdx.y], blockIdx.x, gridDim.x);
}
template <int ALGO, int CODEC, int MODE>
ByValFnT byval_of() {
  if constexpr (ALGO == AINIC_ALGO_RING) return allreduce_emul_byval_kernel<ALGO, CODEC, MODE>;
  else return allreduce_emul_byval_ov_kernel<ALGO, CODEC, MODE>;
}

typedef void (*ByValFn)(const EmulArgs);

template <int ALGO>
ByValFn pick_codec(int codec, int mode) {
  switch (codec) {
    case AINIC_CODEC_F32: return byval_of<ALGO, 0, 0>();
    case AINIC_CODEC_BFP: return mode == AINIC_BFP_HW_EXACT ? byval_of<ALGO, 1, 0>() : byval_of<ALGO, 1, 1>();
    case AINIC_CODEC_MXFP8: return byval_of<ALGO, 2, 0>();
    case AINIC_CODEC_BF16: return byval_of<ALGO, 3, 0>();
  }
  return nullptr;
}
ByValFn pick(int algo, int codec, int mode) {
  switch (algo) {
    case AINIC_ALGO_ONESHOT: return pick_codec<AINIC_ALGO_ONESHOT>(codec, mode);
    case AINIC_ALGO_TWOSHOT: return pick_codec<AINIC_ALGO_TWOSHOT>(codec, mode);
    case AINIC_ALGO_RING:
      return (codec == AINIC_CODEC_F32 || codec == AINIC_CODEC_BFP) ? pick_codec<AINIC_ALGO_RING>(codec, mode) : nullptr;
  }
  return nullptr;
}

int fail(char* errbuf, size_t n, int rc, const char* fmt, const char* detail = "") {
  if (errbuf && n) snprintf(errbuf, n, fmt, detail);
  return rc;
}

// Device address of the scratch array on the current device; resolved once per device.
unsigned char* scratch_on_current_device(char* errbuf, size_t n, int* rc) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    *rc = fail(errbuf, n, AINIC_ERR_CUDA, "no-alloc all-reduce: cudaGetDevice failed: %s", cudaGetErrorString(e));
    return nullptr;
  }
  std::lock_guard<std::mutex> lk(g_setup_mu);
  if (!g_scratch_dev_ptr[dev]) {
    unsigned char* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) {
      *rc = fail(errbuf, n, AINIC_ERR_CUDA, "no-alloc all-reduce: cudaGetSymbolAddress(scratch) failed: %s", cudaGetErrorString(e));
      return nullptr;
    }
    g_scratch_dev_ptr[dev] = p;
  }
  if (!g_ctr_host_tried) {
    g_ctr_host_tried = true;
    if (cudaHostRegister(g_ctr_host, sizeof(g_ctr_host), cudaHostRegisterPortable) != cudaSuccess) cudaGetLastError();
  }
  return g_scratch_dev_ptr[dev];
}

}  // namespace

extern "C" int ainic_allreduce_emulate_noalloc(int nranks, float* grads, float* weights, void* weights_bf16,
                                               float* grads_out, size_t count, float lr, int codec, int bfp_mode,
                                               int algo, int ctas_per_rank, void* stream, char* errbuf,
                                               size_t errbuf_len) {
  if (errbuf && errbuf_len) errbuf[0] = 0;
  if (!grads || count == 0 || nranks < 1) return fail(errbuf, errbuf_len, AINIC_ERR_ARG, "no-alloc all-reduce: bad argument%s");
  if (nranks > kNaMaxRanks) {
    if (errbuf && errbuf_len) snprintf(errbuf, errbuf_len, "no-alloc all-reduce: %d ranks requested, the static scratch holds at most %d", nranks, kNaMaxRanks);
    return AINIC_ERR_ARG;
  }
  if (count > kNaMaxCount) {
    if (errbuf && errbuf_len) snprintf(errbuf, errbuf_len, "no-alloc all-reduce: %zu values per rank requested, the static scratch holds at most %zu", count, kNaMaxCount);
    return AINIC_ERR_ARG;
  }
  // rank r's rows start at grads + r * count: every row must stay 16-byte aligned for the 128-bit accesses
  if (((uintptr_t)grads & 15) || ((uintptr_t)weights & 15) || ((uintptr_t)grads_out & 15) || ((uintptr_t)weights_bf16 & 7) ||
      (nranks > 1 && (count & 3)))
    return fail(errbuf, errbuf_len, AINIC_ERR_ARG,
                "no-alloc all-reduce: buffers must be 16-byte aligned and count a multiple of 4 when nranks > 1%s");
  if (algo == AINIC_ALGO_AUTO) algo = (nranks <= 2 || count * sizeof(float) <= (512u << 10)) ? AINIC_ALGO_ONESHOT : AINIC_ALGO_TWOSHOT;
  ByValFn kern = pick(algo, codec, bfp_mode);
  if (!kern) return fail(errbuf, errbuf_len, AINIC_ERR_UNSUPPORTED, "no-alloc all-reduce: unsupported algorithm/codec combination%s");
  const int ctas = ctas_per_rank > 0 ? ctas_per_rank : 4;
  if (ctas > kMaxCtas) return fail(errbuf, errbuf_len, AINIC_ERR_ARG, "no-alloc all-reduce: too many CTAs per rank%s");

  int rc = AINIC_OK;
  unsigned char* scratch = scratch_on_current_device(errbuf, errbuf_len, &rc);
  if (!scratch) return rc;

  int dev = 0, sms = 0, per_sm = 0;
  cudaGetDevice(&dev);
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, algo == AINIC_ALGO_RING ? kCommThreads : kCommThreadsOverlap, 0) != cudaSuccess)
    return fail(errbuf, errbuf_len, AINIC_ERR_CUDA, "no-alloc all-reduce: occupancy query failed: %s", cudaGetErrorString(cudaGetLastError()));
  if ((long)ctas * nranks > (long)sms * per_sm)
    return fail(errbuf, errbuf_len, AINIC_ERR_ARG, "no-alloc all-reduce: grid would not be co-resident%s");

  const size_t frames = bfp_frames(count);
  cudaStream_t s = (cudaStream_t)stream;
  EmulArgs args;
  memset(&args, 0, sizeof(args));
  auto base = [&](int r) { return scratch + (size_t)r * kNaPerRank; };
  for (int r = 0; r < nranks; ++r) {
    RankCtx& x = args.ctx[r];
    x.rank = r;
    x.nranks = nranks;
    x.grad = grads + (size_t)r * count;
    x.weight = weights ? weights + (size_t)r * count : nullptr;
    x.wshadow = weights_bf16 ? static_cast<__nv_bfloat16*>(weights_bf16) + (size_t)r * count : nullptr;
    x.grad_out = grads_out ? grads_out + (size_t)r * count : nullptr;
    for (int p = 0; p < nranks; ++p) {
      x.peer_grad[p] = grads + (size_t)p * count;
      x.stage_c[p] = base(p);
      x.stage_d[p] = base(p) + kNaStageC;
      x.flags[p] = reinterpret_cast<uint32_t*>(base(p) + kNaStageC + kNaStageD);
    }
    x.counters = reinterpret_cast<unsigned long long*>(base(r) + kNaStageC + kNaStageD + kNaFlags);
    x.done_slot = nullptr;
    x.count = count;
    x.stage_c_half = frames * 1024;
    x.lr = lr;
    x.epoch = 1;
    x.parity = 0;
    x.timeout_ns = 5ull * 1000 * 1000 * 1000;
    // The previous call left its epoch-1 flags and its counters behind: they must read zero again.
    if (cudaMemsetAsync(base(r) + kNaStageC + kNaStageD, 0, kNaFlags + kNaCounters, s) != cudaSuccess)
      return fail(errbuf, errbuf_len, AINIC_ERR_CUDA, "no-alloc all-reduce: clearing flags failed: %s", cudaGetErrorString(cudaGetLastError()));
  }
  void* kargs[] = {(void*)&args};
  cudaError_t e = cudaLaunchCooperativeKernel((const void*)kern, dim3(ctas, nranks), dim3(algo == AINIC_ALGO_RING ? kCommThreads : kCommThreadsOverlap), kargs, 0, s);
  // Time-out check: every rank's counter block comes back with an asynchronous copy on the caller's
  // stream, then ONE stream synchronisation.  Nothing here touches the legacy default stream (a
  // synchronous cudaMemcpy would, serialising the call against every blocking stream of the process).
  unsigned long long (*ctr)[kNumCounters] = reinterpret_cast<unsigned long long (*)[kNumCounters]>(g_ctr_host);
  for (int r = 0; r < nranks && e == cudaSuccess; ++r)
    e = cudaMemcpyAsync(ctr[r], args.ctx[r].counters, sizeof(ctr[r]), cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  if (e != cudaSuccess) return fail(errbuf, errbuf_len, AINIC_ERR_CUDA, "no-alloc all-reduce: launch failed: %s", cudaGetErrorString(e));
  for (int r = 0; r < nranks; ++r)
    if (ctr[r][CTR_TIMEOUTS]) return fail(errbuf, errbuf_len, AINIC_ERR_TIMEOUT, "no-alloc all-reduce: a flag wait timed out%s");
  return AINIC_OK;
}
