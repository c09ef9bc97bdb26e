// Flow all-reduce: ONE persistent launch walks a whole bucket chunk by chunk (in the trainer a chunk is
// a layer).  Every WARP is an independent pipeline with one role; nothing in the data path ever waits
// on a CTA-wide barrier:
//
//   forwarder   one warp (an extra CTA): watches the local doorbell of chunk k - the counter the
//               weight-gradient GEMM's last CTA bumps when every tile is stored - and pushes
//               "rank r's chunk k is readable" to every rank                          (flag phase 4)
//   reduce      waits until chunk k is readable on every rank, pulls its frames of this rank's 1/N of
//               the chunk from all ranks, sums in fp32 in rank order, applies SGD, publishes the sum
//               re-encoded; the LAST reduce warp of the rank to finish chunk k (local ticket) tells
//               every peer "rank r's shard of chunk k is published"                   (flag phase 5)
//   gather      waits for a peer's published shard of chunk k, pulls its frames, applies SGD
//
// This is the reference's two-deep request pipeline (sw/mlp_mpi_example_f32.cpp:735-788: layer i's
// all-reduce runs under layer i-1's backward GEMMs) taken to its limit: the host issues nothing per
// layer, the doorbell is rung by the producer kernel ("epilogue flips the ready flag"; the NIC of the
// reference starts when the host writes the start CSR, hw/all_reduce.sv:210-244), reduce of chunk
// k + 1 and gather of chunk k run at the same time in different warps, and what is left exposed after
// the last weight-gradient GEMM is one chunk's reduce + gather.
// (A first version kept all warps of a CTA in lock step - compress, signal, gather, wait, reduce,
// signal, with CTA barriers in between: every phase of every chunk paid its own pipeline ramp and
// two system-scope fences; measured 0.88 vs 0.65 ms/step at 2 GPUs against one fused call per bucket.)
//
// The input must already be in wire format: frames written by the GEMM epilogue (precompressed) or
// plain fp32 (codec 0).  Flag values are the call number, kept in device memory (identical launch
// arguments from call to call: graph replay); one flag word per (source rank, chunk), single writer.
// No staging parity is needed: a rank leaves call s only after it has pulled every peer's published
// shards, i.e. after every peer has finished reading this rank's frames, and a peer's "readable" of
// call s + 1 can only be sent after that peer has left call s.
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
  unsigned long long* trace;       // nullable: [5][kMaxChunks] %globaltimer stamps of the last call (ns since kernel start of
                                   // CTA 0 / the forwarder): doorbell seen, reduce got READY, reduce done, gather got PUB, gather done
  unsigned long long t_origin;     // unused on the host side
};
constexpr int kTraceDoorbell = 0, kTraceReduceGo = 1, kTraceReduceDone = 2, kTraceGatherGo = 3, kTraceGatherDone = 4;
__device__ __forceinline__ void flow_stamp(const FlowCtx& fc, int what, int k, bool on, int lane) {
  if (fc.trace && on && lane == 0) fc.trace[what * kMaxChunks + k] = ptx::globaltimer_ns();
}

// ---- flag words (inside the per-rank flag area of the communicator; see flag_index)
//   [phase 4][src][k]       rank src's chunk k is readable          (written by src's forwarder)
//   [phase 5][src][k]       rank src's shard of chunk k is published (written by src's last reduce warp)
//   [phase 5][src][48]      rank src has pulled everything it needs  (fp32 in-place variant only)
//   [phase 5][self][64 + k] local ticket of the reduce warps for chunk k; [64 + 48] of the gather warps
constexpr int kFlowDoneSlot = 48, kFlowTicketBase = 64;
static_assert(kMaxChunks <= kFlowDoneSlot && kFlowTicketBase + kFlowDoneSlot < kMaxCtas, "flow flag slots");

// Lanes 0 .. N-1 of a warp poll one flag each (lane p: source rank p; skip_self leaves out this rank).
__device__ __forceinline__ void warp_wait_flags(const RankCtx& c, int phase, int slot, uint32_t value, bool skip_self, int lane,
                                                WaitStats& ws) {
  if (lane < c.nranks && !(skip_self && lane == c.rank)) {
    const uint32_t* f = c.flags[c.rank] + flag_index(phase, lane, slot);
    unsigned long long t0 = 0;
    uint32_t polls = 0;
    // Poll with RELAXED loads and fence once at the end: an acquire load at system scope carries a fence
    // that invalidates the SM's L1 every time, and a thousand warps polling that way (measured: 0.95 ms per
    // step at 2 GPUs against 0.59 without flow mode) take the memory pipeline away from everybody else.
    while ((int)(ptx::ld_relaxed_sys(f) - value) < 0) {
      __nanosleep(polls > 32 ? 2048 : (polls > 4 ? 512 : 128));   // GEMM CTAs share this SM: stay out of their issue slots
      if ((++polls & 0xFFu) == 0) {
        const unsigned long long now = ptx::globaltimer_ns();
        if (t0 == 0) t0 = now;
        else if (now - t0 > c.timeout_ns) { atomicAdd(&ws.timeouts, 1u); break; }
      }
    }
    ptx::fence_acq_rel_sys();   // the one acquire: everything the flag's writer released is visible from here on
  }
  __syncwarp();
}
// Lanes 0 .. N-1 store `value` into slot [phase][this rank][slot] of every rank (lane p -> rank p).
__device__ __forceinline__ void warp_post_flags(const RankCtx& c, int phase, int slot, uint32_t value, bool skip_self, int lane) {
  if (lane < c.nranks && !(skip_self && lane == c.rank)) {
    __threadfence_system();
    ptx::st_release_sys(c.flags[lane] + flag_index(phase, c.rank, slot), value);
  }
}
// A warp has finished its part of something `total` warps of this rank work on; returns true in the
// last one (all lanes).  The ticket word resets itself.
__device__ __forceinline__ bool warp_ticket_last(const RankCtx& c, int slot, uint32_t total, int lane) {
  __syncwarp();
  uint32_t last = 0;
  if (lane == 0) {
    __threadfence();   // this warp's stores before its ticket (the last warp then fences at system scope)
    uint32_t* t = c.flags[c.rank] + flag_index(kPhaseFlowPub, c.rank, kFlowTicketBase + slot);
    if (atomicAdd(t, 1u) == total - 1u) {
      *reinterpret_cast<volatile uint32_t*>(t) = 0u;
      last = 1;
    }
  }
  return __shfl_sync(kFullMask, last, 0) != 0;
}

__device__ __forceinline__ uint32_t flow_shard(const FlowChunk& ch, int N) { return (ch.frames + (uint32_t)N - 1u) / (uint32_t)N; }
__device__ __forceinline__ uint32_t flow_shard_lim(const FlowChunk& ch, int N, int s) {
  const uint32_t sh = flow_shard(ch, N), lo = (uint32_t)s * sh;
  return ch.frames <= lo ? 0u : (ch.frames - lo < sh ? ch.frames - lo : sh);
}

// The forwarder: local doorbells -> "readable" flags on every rank (this one included).
__device__ void flow_forward(const FlowCtx& fc, WaitStats& ws, int lane) {
  const RankCtx& c = fc.c;
  for (int k = 0; k < fc.nchunks; ++k) {
    const int idx = fc.chunk[k].ready_idx;
    if (fc.local_ready && idx >= 0 && lane == 0) {
      const uint32_t need = fc.chunk[k].ready_need ? fc.chunk[k].ready_need : 1u;
      unsigned long long t0 = 0;
      uint32_t polls = 0;
      while (ptx::ld_acquire_gpu(fc.local_ready + idx) < need) {
        __nanosleep(128);
        if ((++polls & 0x3FFu) == 0) {
          const unsigned long long now = ptx::globaltimer_ns();
          if (t0 == 0) t0 = now;
          else if (now - t0 > c.timeout_ns) { atomicAdd(&ws.timeouts, 1u); break; }
        }
      }
    }
    __syncwarp();
    if (k == 0 && lane == 0) ws.mark[0] = ptx::globaltimer_ns();
    flow_stamp(fc, kTraceDoorbell, k, true, lane);
    warp_post_flags(c, kPhaseFlowReady, k, ws.epoch, false, lane);
  }
}

template <int CODEC, int MODE>
__device__ void flow_reduce_role(const FlowCtx& fc, WaitStats& ws, uint32_t first, uint32_t step, uint32_t nwarps, bool timekeeper,
                                 int lane) {
  using FC = FrameCodec<CODEC, MODE>;
  const RankCtx& c = fc.c;
  const int N = c.nranks;
#pragma unroll 1
  for (int k = 0; k < fc.nchunks; ++k) {
    const FlowChunk ch = fc.chunk[k];
    warp_wait_flags(c, kPhaseFlowReady, k, ws.epoch, false, lane, ws);
    flow_stamp(fc, kTraceReduceGo, k, timekeeper, lane);
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
    flow_stamp(fc, kTraceReduceDone, k, timekeeper, lane);
    if (N > 1 && warp_ticket_last(c, k, nwarps, lane)) warp_post_flags(c, kPhaseFlowPub, k, ws.epoch, true, lane);
  }
}

template <int CODEC, int MODE>
__device__ void flow_gather_role(const FlowCtx& fc, WaitStats& ws, uint32_t first, uint32_t step, uint32_t nwarps, bool timekeeper,
                                 int lane) {
  const RankCtx& c = fc.c;
#pragma unroll 1
  for (int k = 0; k < fc.nchunks; ++k) {
    const FlowChunk ch = fc.chunk[k];
    warp_wait_flags(c, kPhaseFlowPub, k, ws.epoch, true, lane, ws);
    if (timekeeper && k == fc.nchunks - 1 && lane == 0) ws.mark[1] = ptx::globaltimer_ns();
    flow_stamp(fc, kTraceGatherGo, k, timekeeper, lane);
    gather_stream<CODEC, MODE>(c, flow_shard(ch, c.nranks), ch.frames, first, step, lane, ch.frame0, ch.d_off);
    flow_stamp(fc, kTraceGatherDone, k, timekeeper, lane);
  }
  if (CODEC == 0) {
    // peers read my gradient in place: nobody may leave (and let the next GEMM overwrite its gradient)
    // before every rank has pulled what it needs
    if (warp_ticket_last(c, kFlowDoneSlot, nwarps, lane)) warp_post_flags(c, kPhaseFlowPub, kFlowDoneSlot, ws.epoch, true, lane);
    warp_wait_flags(c, kPhaseFlowPub, kFlowDoneSlot, ws.epoch, true, lane, ws);
  }
}

// How many of a CTA's 8 warps reduce (the rest gather): a rank reduces 1/N of the frames and gathers the rest.
__device__ __forceinline__ int flow_reduce_warps(int nranks, int wpb) {
  if (nranks <= 1) return wpb;
  const int r = (wpb + nranks / 2) / nranks;
  return r < 1 ? 1 : (r > wpb - 1 ? wpb - 1 : r);
}

// Grid: gridDim.x - 1 worker CTAs + one forwarder CTA (the last).
template <int CODEC, int MODE>
__device__ __forceinline__ void run_flow(const FlowCtx& fc, int cta, int nctas_total) {
  __shared__ WaitStats ws;
  if (threadIdx.x == 0) begin_call(fc.c, ws);
  const unsigned long long t0 = ptx::globaltimer_ns();
  __syncthreads();
  const RankCtx& c = fc.c;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  const int workers = nctas_total - 1;
  if (cta == workers) {
    if (warp == 0) flow_forward(fc, ws, lane);
  } else {
    const int nr = flow_reduce_warps(c.nranks, wpb), ng = wpb - nr;
    if (warp < nr) flow_reduce_role<CODEC, MODE>(fc, ws, (uint32_t)(cta * nr + warp), (uint32_t)(workers * nr), (uint32_t)(workers * nr),
                                                 cta == 0 && warp == 0, lane);
    else flow_gather_role<CODEC, MODE>(fc, ws, (uint32_t)(cta * ng + warp - nr), (uint32_t)(workers * ng), (uint32_t)(workers * ng),
                                       cta == 0 && warp == nr, lane);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned long long t1 = ptx::globaltimer_ns();
    if (ws.timeouts) atomicAdd(c.counters + CTR_TIMEOUTS, (unsigned long long)ws.timeouts);
    if (cta == workers) {   // forwarder: when the first chunk became readable here
      if (ws.mark[0]) atomicAdd(c.counters + CTR_TO_BARRIER0_NS, ws.mark[0] - t0);
    }
    if (cta == 0) {
      atomicAdd(c.counters + CTR_BUSY_NS, t1 - t0);
      atomicAdd(c.counters + CTR_LAUNCHES, 1ull);
      c.counters[CTR_LAST_NS] = t1 - t0;
      if (ws.mark[1]) atomicAdd(c.counters + CTR_TO_BARRIER1_NS, t1 - ws.mark[1]);   // last chunk published everywhere -> end (the exposed tail)
    }
    end_call(c, ws, nctas_total);
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
