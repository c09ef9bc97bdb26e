// Flow all-reduce: ONE persistent launch walks a whole bucket chunk by chunk (in the trainer a chunk is
// a layer), each chunk going through the two-shot steps of allreduce.cuh as soon as its producer
// says it is complete:
//
//   chunk k   local ready (the weight-gradient GEMM's last CTA raised the chunk's flag)
//             -> [compress]  -> tell every peer "my chunk k is readable"          (flag phase 4)
//             -> while that message travels: pull + apply the published shards of chunk k - 1
//             -> all peers ready: reduce my 1/N of chunk k in fp32, SGD, publish it re-encoded
//             -> tell every peer "my shard of chunk k is published"               (flag phase 5)
//
// This is the reference's two-deep request pipeline (sw/mlp_mpi_example_f32.cpp:735-788: layer i's
// all-reduce runs under layer i-1's backward GEMMs) taken to its limit.  The host issues nothing
// per layer; the readiness signal is a device flag written by the GEMM epilogue ("epilogue flips the
// ready flag"), the cross-GPU flag round trips of chunk k hide behind the data movement of chunk
// k - 1, and what is left exposed after the last weight-gradient GEMM is one chunk's reduce + gather
// instead of a whole multi-layer bucket's.  The NIC of the reference starts on a buffer when the
// host rings the doorbell (hw/all_reduce.sv:210-244); here the doorbell is rung by the producer kernel.
//
// Flags: CTA b of a rank only waits for CTA b of its peers, as in allreduce.cuh.  Values are
// (call number << 8) | (k + 1), monotonically increasing over chunks and calls; the call number
// lives in device memory, so launch arguments are identical from call to call (graph replay).
// No staging parity is needed: a rank leaves call s only after every peer has published every shard,
// i.e. after every peer has finished reading this rank's stage C, and a peer's READY(0) of call s + 1
// can only be sent after it has finished reading this rank's stage D in call s.
#pragma once
#include "allreduce.cuh"

namespace ainic {
namespace comm {

constexpr int kMaxChunks = AINIC_FLOW_MAX_CHUNKS;
constexpr int kPhaseFlowReady = 4, kPhaseFlowPub = 5;

struct FlowChunk {
  uint32_t frame0;      // first frame of the chunk inside the bucket
  uint32_t frames;      // complete frames (a chunk is a whole number of frames)
  uint32_t d_off;       // slot of this chunk's first published frame in stage D
  int16_t ready_idx;    // index into local_ready, or -1: ready at launch
  uint16_t ready_need;  // the chunk is complete once local_ready[ready_idx] >= ready_need (number of producers)
};
struct FlowCtx {
  RankCtx c;                       // c.count covers the whole bucket; c.epoch_ptr = the flow call counter
  FlowChunk chunk[kMaxChunks];     // in processing order
  int nchunks;
  const uint32_t* local_ready;     // device words counting finished producers of each chunk (nullable); zeroed by the
                                   // caller before the producers of this call are launched
};

// Thread 0 polls the producer's flag; everyone leaves together.
__device__ __forceinline__ void wait_local_ready(const FlowCtx& fc, int k, WaitStats& ws) {
  const int idx = fc.chunk[k].ready_idx;
  if (fc.local_ready && idx >= 0 && threadIdx.x == 0) {
    unsigned long long t0 = 0;
    uint32_t polls = 0;
    const uint32_t need = fc.chunk[k].ready_need ? fc.chunk[k].ready_need : 1u;
    while (ptx::ld_acquire_gpu(fc.local_ready + idx) < need) {
      __nanosleep(polls > 16 ? 512 : 128);   // a GEMM CTA shares this SM: stay out of its issue slots
      if ((++polls & 0xFFu) == 0) {
        const unsigned long long now = ptx::globaltimer_ns();
        if (t0 == 0) t0 = now;
        else if (now - t0 > fc.c.timeout_ns) { atomicAdd(&ws.timeouts, 1u); break; }
      }
    }
  }
  __syncthreads();
}

__device__ __forceinline__ uint32_t flow_shard(const FlowChunk& ch, int N) { return (ch.frames + (uint32_t)N - 1u) / (uint32_t)N; }
__device__ __forceinline__ uint32_t flow_shard_lim(const FlowChunk& ch, int N, int s) {
  const uint32_t sh = flow_shard(ch, N), lo = (uint32_t)s * sh;
  return ch.frames <= lo ? 0u : (ch.frames - lo < sh ? ch.frames - lo : sh);
}
// The three data phases sit inside one loop over the chunks.  Passing the values their address
// arithmetic derives from through an empty asm keeps the compiler from hoisting that arithmetic out of
// the chunk loop and holding it in registers across the other phases.  (ptxas still parks a few
// just-loaded frame words in local memory between request and use - ~300 bytes per thread, L1-resident,
// a handful of instructions per 512-value frame.)
__device__ __forceinline__ void launder(uint32_t& first, uint32_t& step, int& lane) {
  asm volatile("" : "+r"(first), "+r"(step), "+r"(lane));
}
template <int CODEC, int MODE>
__device__ __forceinline__ void flow_compress(const RankCtx& c, const FlowChunk ch, uint32_t first, uint32_t step, int lane) {
  launder(first, step, lane);
  const int N = c.nranks;
  const uint32_t shard = flow_shard(ch, N);
  for (int s = 0; s < N; ++s) {
    const uint32_t base = ch.frame0 + (uint32_t)s * shard, lim = flow_shard_lim(ch, N, s);
#pragma unroll 1
    for (uint32_t j0 = first; j0 < lim; j0 += 2 * step) {
      if (c.pf_depth) {
        const uint32_t ja = j0 + c.pf_depth * 2 * step;
        if (ja < lim) prefetch_frame_f32(c.grad, base + ja, lane);
        if (ja + step < lim) prefetch_frame_f32(c.grad, base + ja + step, lane);
      }
      compress_pair<CODEC, MODE>(c, 0, base + j0, base + j0 + step, j0 + step < lim, lane);
    }
  }
}
template <int CODEC, int MODE>
__device__ __forceinline__ void flow_reduce(const RankCtx& c, const FlowChunk ch, uint32_t first, uint32_t step, int lane) {
  launder(first, step, lane);
  using FC = FrameCodec<CODEC, MODE>;
  const int N = c.nranks;
  const uint32_t shard = flow_shard(ch, N);
  const uint32_t base = ch.frame0 + (uint32_t)c.rank * shard, lim = flow_shard_lim(ch, N, c.rank);
  const uint32_t d_off = ch.d_off;
  auto src = [&](int p, uint32_t j) -> const uint8_t* {
    return CODEC == 0 ? reinterpret_cast<const uint8_t*>(c.peer_grad[p]) + (size_t)(base + j) * 2048
                      : c.stage_c[p] + (size_t)(base + j) * FC::kBytes;
  };
  reduce_stream<CODEC, MODE>(c, base, first, step, lim, lane, src, [&](uint32_t j, float (&acc)[16], float (&w)[16]) {
    if constexpr (CODEC == 0) {
      // in place: peers pull the reduced shard out of my gradient buffer
      if (N > 1) frame_store_f32(c.grad + (size_t)(base + j) * kFrameVals, kFrameVals, acc, lane);
    } else {
      // publish into my stage D, then apply exactly what the peers will decode
      frame_encode_rt<CODEC, MODE>(acc, c.stage_d[c.rank] + (size_t)(d_off + j) * FC::kBytes, lane);
    }
    apply_full_w(c, base + j, acc, w, lane);
  });
}
template <int CODEC, int MODE>
__device__ __forceinline__ void flow_gather(const RankCtx& c, const FlowChunk ch, uint32_t first, uint32_t step, int lane) {
  launder(first, step, lane);
  gather_stream<CODEC, MODE>(c, flow_shard(ch, c.nranks), ch.frames, first, step, lane, ch.frame0, ch.d_off);
}

template <int CODEC, int MODE>
__device__ void flow_body(const FlowCtx& fc, int cta, int nctas, WaitStats& ws) {
  const RankCtx& c = fc.c;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  const int N = c.nranks, K = fc.nchunks;
  const uint32_t first = (uint32_t)cta + (uint32_t)warp * nctas, step = (uint32_t)nctas * wpb;
  const uint32_t vbase = ws.epoch << 8;
  // One loop body for "start chunk k" and "finish chunk k - 1" (k = K only finishes).
#pragma unroll 1
  for (int k = 0; k <= K; ++k) {
    if (k < K) {
      // 1. chunk k is complete on this rank; compress it unless the producer already wrote frames
      wait_local_ready(fc, k, ws);
      if (k == 0) mark_phase(ws, 0);
      if (CODEC != 0 && !c.precompressed) flow_compress<CODEC, MODE>(c, fc.chunk[k], first, step, lane);
      if (N > 1) signal_all(c, kPhaseFlowReady, cta, vbase | (uint32_t)(k + 1));
    }
    // 2. while that flag travels: pull + apply the previous chunk's published shards
    if (N > 1 && k > 0) {
      wait_all(c, kPhaseFlowPub, cta, vbase | (uint32_t)k, ws);
      if (k == K) mark_phase(ws, 1);
      flow_gather<CODEC, MODE>(c, fc.chunk[k - 1], first, step, lane);
    }
    if (k == K) break;
    // 3. my shard of chunk k: fp32 sum in rank order, publish re-encoded, apply what peers will decode
    if (N > 1) wait_all(c, kPhaseFlowReady, cta, vbase | (uint32_t)(k + 1), ws);
    flow_reduce<CODEC, MODE>(c, fc.chunk[k], first, step, lane);
    if (N > 1) signal_all(c, kPhaseFlowPub, cta, vbase | (uint32_t)(k + 1));
  }
  if (CODEC == 0 && N > 1) {   // peers read my gradient in place: do not return (and let the next GEMM overwrite it) before they are done
    signal_all(c, kPhaseFlowReady, cta, vbase | 0xFEu);
    wait_all(c, kPhaseFlowReady, cta, vbase | 0xFEu, ws);
  }
}

template <int CODEC, int MODE>
__device__ __forceinline__ void run_flow(const FlowCtx& fc, int cta, int nctas) {
  __shared__ WaitStats ws;
  if (threadIdx.x == 0) begin_call(fc.c, ws);
  const unsigned long long t0 = ptx::globaltimer_ns();
  __syncthreads();
  flow_body<CODEC, MODE>(fc, cta, nctas, ws);
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned long long t1 = ptx::globaltimer_ns();
    const RankCtx& c = fc.c;
    if (ws.timeouts) atomicAdd(c.counters + CTR_TIMEOUTS, (unsigned long long)ws.timeouts);
    if (cta == 0) {
      atomicAdd(c.counters + CTR_BUSY_NS, t1 - t0);
      atomicAdd(c.counters + CTR_WAIT_NS, ws.wait_ns);
      atomicAdd(c.counters + CTR_LAUNCHES, 1ull);
      c.counters[CTR_LAST_NS] = t1 - t0;
      if (ws.mark[0]) atomicAdd(c.counters + CTR_TO_BARRIER0_NS, ws.mark[0] - t0);   // kernel start -> first chunk ready locally
      if (ws.mark[1]) atomicAdd(c.counters + CTR_TO_BARRIER1_NS, t1 - ws.mark[1]);   // last chunk published everywhere -> end (the exposed tail)
    }
    end_call(c, ws, nctas);
  }
}

// 256 threads x 168 registers = 43 K registers: still fits beside a persistent GEMM CTA (192 threads x <= 80).
#define AINIC_COMM_FLOW_ATTR __maxnreg__(168)
template <int CODEC, int MODE>
__global__ void AINIC_COMM_FLOW_ATTR allreduce_flow_kernel(const __grid_constant__ FlowCtx fc) {
  run_flow<CODEC, MODE>(fc, blockIdx.x, gridDim.x);
}
// Single-GPU emulation: blockIdx.y selects the rank (cooperative launch keeps all of them resident).
template <int CODEC, int MODE>
__global__ void AINIC_COMM_FLOW_ATTR allreduce_flow_emul_kernel(const FlowCtx* fcs) {
  run_flow<CODEC, MODE>(fcs[blockIdx.y], blockIdx.x, gridDim.x);
}

}  // namespace comm
}  // namespace ainic
