// Fused gradient all-reduce + SGD weight update over peer memory (K10-K12).
//
// This is the B200 form of the reference's collective engine hw/all_reduce.sv (3 FSMs, 5 command
// FIFOs, 8 fp32 adders), its BFP adapter hw/bfp_adapter.sv and its weight-update engine
// hw/weight_update.sv: ONE kernel compresses the local gradient, exchanges it with the peers over
// NVLink (loads/stores on IPC-mapped peer pointers), reduces in fp32, decompresses and applies
// w = fmaf(-lr, sum, w) - no NCCL call and no separate optimizer kernel.
//
//   oneshot  every rank pulls every peer's compressed bucket and reduces all of it (latency-optimal)
//   twoshot  pull-reduce own 1/N shard -> re-encode -> every rank pulls the reduced shards
//            (bandwidth-optimal: 2(N-1)/N * S * r bytes per GPU)
//   ring     the reference schedule itself: 32 KiB slices travel r -> r-1, BFP re-quantisation on
//            every hop, owner keeps the unquantised sum (hw/all_reduce.sv:891-1086)
//
// Cross-GPU synchronisation uses monotonically increasing epoch flags pushed into the consumer's
// memory (st.release.sys) and polled locally (ld.acquire.sys); CTA b of one rank only ever waits
// on CTA b of other ranks, so no grid-wide barrier is needed and the kernel cannot deadlock against
// itself as long as every rank launches the same grid.  Every spin has a %globaltimer timeout (the
// reference's wait() spins forever and its abort input is unconnected: SURVEY D9/D10).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "../../include/ainic.h"
#include "../common/frame.cuh"

namespace ainic {
namespace comm {

constexpr int kMaxRanks = AINIC_MAX_RANKS;
constexpr int kMaxCtas = 512;
constexpr int kFlagPhases = 6;          // 0 start, 1 shard published, 2 in-place reads done, 3 ring messages,
                                        // 4 / 5 chunk ready / chunk published of the flow kernel (flow.cuh)
constexpr int kFlagWords = kFlagPhases * kMaxRanks * kMaxCtas;
constexpr int kCommThreads = 512;          // 16 warps = one 8192-value ring slice per step
constexpr int kCommThreadsOverlap = 256;   // one-/two-shot/NVLS: small enough to share an SM with a GEMM CTA
constexpr int kSliceFrames = 16;           // BUF_SIZE = 512 cache lines = 8192 fp32 (hw/all_reduce.sv:102-103)
constexpr int kNumCounters = 8;

enum { CTR_BUSY_NS = 0, CTR_WAIT_NS = 1, CTR_LAUNCHES = 2, CTR_TIMEOUTS = 3, CTR_TICKET = 4, CTR_LAST_NS = 5,
       CTR_TO_BARRIER0_NS = 6, CTR_TO_BARRIER1_NS = 7 };

struct RankCtx {
  int rank, nranks;
  float* grad;
  float* weight;
  __nv_bfloat16* wshadow;
  float* grad_out;
  const float* peer_grad[kMaxRanks];
  uint8_t* stage_c[kMaxRanks];
  uint8_t* stage_d[kMaxRanks];
  uint32_t* flags[kMaxRanks];
  unsigned long long* counters;
  volatile uint32_t* done_slot;  // host-visible completion word (nullable)
  size_t count;
  size_t stage_c_half;           // bytes of one parity half of stage_c
  float lr;
  uint32_t epoch;                // increments on every fused call (all ranks issue the same sequence)
  uint32_t parity;               // which half of stage_c this call writes (toggles per staging call)
  uint32_t* epoch_ptr;           // non-null: the call number lives in device memory (this call is *epoch_ptr + 1,
                                 // parity = its low bit; the last CTA stores it back).  Launch arguments are then
                                 // the same for every call, so a captured launch can be replayed (CUDA graphs).
  uint32_t precompressed;        // stage_c already holds this rank's encoded gradient (GEMM epilogue wrote it)
  uint32_t pf_depth;             // frames of local weights / gradients requested into L2 ahead of their loads (0 = off)
  unsigned long long timeout_ns;
};

__device__ __forceinline__ int flag_index(int phase, int src, int cta) { return (phase * kMaxRanks + src) * kMaxCtas + cta; }

// Spin until *f >= target (wrap-safe); returns false on timeout.
__device__ __forceinline__ bool spin_until(const uint32_t* f, uint32_t target, unsigned long long timeout_ns) {
  unsigned long long t0 = 0;
  uint32_t polls = 0;
  while ((int)(ptx::ld_acquire_sys(f) - target) < 0) {
    // Back off: this CTA shares its SM with a GEMM CTA whose single TMA-producer and MMA-issuer
    // threads must not lose issue slots to a tight polling loop.
    if (polls > 2) __nanosleep(polls > 64 ? 256 : 64);
    if ((++polls & 0x3FFu) == 0) {
      const unsigned long long now = ptx::globaltimer_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > timeout_ns) return false;
    }
  }
  return true;
}

struct WaitStats {
  uint32_t epoch, parity;      // this call's number / staging half (from the arguments or from device memory)
  unsigned long long wait_ns;
  unsigned long long mark[2];  // %globaltimer after the first / second cross-rank barrier (CTA 0 reports them)
  uint32_t timeouts;
};
// Thread 0, before the first __syncthreads of the kernel.
__device__ __forceinline__ void begin_call(const RankCtx& c, WaitStats& ws) {
  ws.wait_ns = 0; ws.timeouts = 0; ws.mark[0] = 0; ws.mark[1] = 0;
  if (c.epoch_ptr) {
    ws.epoch = *reinterpret_cast<volatile uint32_t*>(c.epoch_ptr) + 1u;
    ws.parity = ws.epoch & 1u;
  } else {
    ws.epoch = c.epoch;
    ws.parity = c.parity;
  }
}
// Thread 0 of every CTA, after its work is done: the last CTA to arrive publishes the call number for
// the next launch (every CTA of this launch has read it long before) and raises the done word.
__device__ __forceinline__ void end_call(const RankCtx& c, const WaitStats& ws, int nctas) {
  __threadfence_system();
  const unsigned long long ticket = atomicAdd(c.counters + CTR_TICKET, 1ull);
  if (ticket == (unsigned long long)(nctas - 1)) {
    c.counters[CTR_TICKET] = 0;
    if (c.epoch_ptr) *reinterpret_cast<volatile uint32_t*>(c.epoch_ptr) = ws.epoch;
    if (c.done_slot) {
      __threadfence_system();
      // 512'd1 into the request's done cache line (hw/all_reduce.sv:1368-1375); 2 = completed, but some
      // flag wait of this launch timed out (ainic_nic_wait maps it to AINIC_ERR_TIMEOUT)
      *c.done_slot = c.counters[CTR_TIMEOUTS] ? 2u : 1u;
    }
  }
}
__device__ __forceinline__ void mark_phase(WaitStats& ws, int i) {
  if (threadIdx.x == 0) ws.mark[i] = ptx::globaltimer_ns();
}

// All stores of this CTA issued before the call become visible to every peer that observes the flag.
__device__ __forceinline__ void signal_all(const RankCtx& c, int phase, int cta, uint32_t value) {
  __syncthreads();
  if (threadIdx.x < c.nranks && (int)threadIdx.x != c.rank) {
    __threadfence_system();
    ptx::st_release_sys(c.flags[threadIdx.x] + flag_index(phase, c.rank, cta), value);
  }
}
__device__ __forceinline__ void signal_one(const RankCtx& c, int peer, int phase, int cta, uint32_t value) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    ptx::st_release_sys(c.flags[peer] + flag_index(phase, c.rank, cta), value);
  }
}
__device__ __forceinline__ void wait_all(const RankCtx& c, int phase, int cta, uint32_t value, WaitStats& ws) {
  if (threadIdx.x < c.nranks && (int)threadIdx.x != c.rank) {
    const unsigned long long t0 = ptx::globaltimer_ns();
    if (!spin_until(c.flags[c.rank] + flag_index(phase, threadIdx.x, cta), value, c.timeout_ns)) atomicAdd(&ws.timeouts, 1u);
    if (threadIdx.x == (c.rank == 0 ? 1 : 0)) ws.wait_ns += ptx::globaltimer_ns() - t0;
  }
  __syncthreads();
}
__device__ __forceinline__ void wait_one(const RankCtx& c, int peer, int phase, int cta, uint32_t value, WaitStats& ws) {
  if (threadIdx.x == 0) {
    const unsigned long long t0 = ptx::globaltimer_ns();
    if (!spin_until(c.flags[c.rank] + flag_index(phase, peer, cta), value, c.timeout_ns)) atomicAdd(&ws.timeouts, 1u);
    ws.wait_ns += ptx::globaltimer_ns() - t0;
  }
  __syncthreads();
}

// w = fmaf(-lr, g, w) on one frame; also the bf16 shadow and the optional reduced-gradient output.
__device__ __forceinline__ void apply_frame(const RankCtx& c, size_t frame, const float (&g)[16], int lane) {
  const size_t base = frame * kFrameVals;
  if (base >= c.count) return;
  const int valid = (int)(c.count - base < (size_t)kFrameVals ? c.count - base : (size_t)kFrameVals);
  if (c.grad_out) frame_store_f32(c.grad_out + base, valid, g, lane);
  if (c.weight) {
    float w[16];
    frame_load_f32(c.weight + base, valid, w, lane);
#pragma unroll
    for (int i = 0; i < 16; ++i) w[i] = fmaf(-c.lr, g[i], w[i]);
    frame_store_f32(c.weight + base, valid, w, lane);
    if (c.wshadow) frame_store_bf16(c.wshadow + base, valid, w, lane);
  }
}

__device__ __forceinline__ void load_grad_frame(const RankCtx& c, size_t frame, float (&v)[16], int lane) {
  const size_t base = frame * kFrameVals;
  const int valid = base >= c.count ? 0 : (int)(c.count - base < (size_t)kFrameVals ? c.count - base : (size_t)kFrameVals);
  if (valid == 0) {
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = 0.0f;
  } else {
    frame_load_f32(c.grad + base, valid, v, lane);
  }
}

// acc = sum over ranks 0..N-1 (rank order, fp32 RNE adds - the 8 fadd lanes of
// hw/all_reduce.sv:1090-1101) of decode(src(p)).  Loads are issued for four peers at a time before
// any is decoded so that ~4 x frame bytes per warp are in flight over NVLink.
template <int CODEC, int MODE, class SrcFn>
__device__ __forceinline__ void reduce_frame(int N, SrcFn src, float (&acc)[16], int lane) {
  using FC = FrameCodec<CODEC, MODE>;
  float v[16];
  for (int p0 = 0; p0 < N; p0 += 4) {
    typename FC::Raw raw[4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (p0 + k < N) FC::load(src(p0 + k), raw[k], lane);
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (p0 + k < N) {  // warp-uniform
        FC::unpack(raw[k], v, lane);
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i] = (p0 + k == 0) ? v[i] : acc[i] + v[i];
      }
  }
}
// fp32 tail frame (count % 512 != 0): peers' buffers are not padded, so predicate per element.
__device__ __forceinline__ void reduce_tail_f32(const RankCtx& c, size_t f, float (&acc)[16], int lane) {
  float v[16];
  const size_t base = f * kFrameVals;
  for (int p = 0; p < c.nranks; ++p) {
    frame_load_f32(c.peer_grad[p] + base, (int)(c.count - base), v, lane);
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = p == 0 ? v[i] : acc[i] + v[i];
  }
}

// ------------------------------------------------------------------------------------------ two frames per iteration
// The per-frame work is a chain of dependent memory operations (peer load ~2-3 us over NVLink ->
// decode -> weight load -> store), and an 8-warp CTA has little thread-level parallelism to hide it
// (measured: 49 us for a 16 MiB bucket at 2 ranks, ~4x the HBM time).  The compression loop
// therefore handles TWO full frames per warp iteration; the reduce / gather loops are software
// pipelined (reduce_stream / gather_stream below).
// Only the last frame of a bucket can be partial (count % 512 != 0); it is kept out of the hot
// loops and handled once per phase by the warp that owns it, through the predicated slow path.
__device__ __forceinline__ bool frame_full(const RankCtx& c, size_t f) { return (f + 1) * kFrameVals <= c.count; }

// Where frame f of rank p is read from in a reduce phase.
template <int CODEC, int MODE>
__device__ __forceinline__ const uint8_t* reduce_src(const RankCtx& c, size_t parity_off, int p, size_t f) {
  return CODEC == 0 ? reinterpret_cast<const uint8_t*>(c.peer_grad[p]) + f * 2048
                    : c.stage_c[p] + parity_off + f * FrameCodec<CODEC, MODE>::kBytes;
}

// Compress up to two full frames of the local gradient into stage C.
template <int CODEC, int MODE>
__device__ __forceinline__ void compress_pair(const RankCtx& c, size_t parity_off, size_t f0, size_t f1, bool has1, int lane) {
  using FC = FrameCodec<CODEC, MODE>;
  uint8_t* dst = c.stage_c[c.rank] + parity_off;
  float a[16], b[16];
  frame_load_f32(c.grad + f0 * kFrameVals, kFrameVals, a, lane);
  if (has1) frame_load_f32(c.grad + f1 * kFrameVals, kFrameVals, b, lane);
  FC::encode(a, dst + f0 * FC::kBytes, lane);
  if (has1) FC::encode(b, dst + f1 * FC::kBytes, lane);
}

// ------------------------------------------------------------------------------------------ software-pipelined streams
// Measured on the pair loops above at 8 ranks (16 MiB bucket): reduce phase 38 us, gather phase 23 us
// for ~10 us of wire + HBM time.  A warp's iterations were a serial chain: peer frames (NVLink
// round trip, ~3 us) -> decode -> weight frame (HBM round trip, ~1 us) -> store -> next iteration.
// The streams below keep the next batch of peer frames AND the weight frame of the current
// iteration in flight while the current batch is consumed, so a warp pays max(remote, local)
// instead of their sum, and nothing waits between iterations.

// w = fmaf(-lr, g, w) on one FULL frame whose weights were requested earlier.
__device__ __forceinline__ void apply_full_w(const RankCtx& c, size_t f, const float (&g)[16], float (&w)[16], int lane) {
  if (c.grad_out) frame_store_f32(c.grad_out + f * kFrameVals, kFrameVals, g, lane);
  if (c.weight) {
#pragma unroll
    for (int i = 0; i < 16; ++i) w[i] = fmaf(-c.lr, g[i], w[i]);
    frame_store_f32(c.weight + f * kFrameVals, kFrameVals, w, lane);
    if (c.wshadow) frame_store_bf16(c.wshadow + f * kFrameVals, kFrameVals, w, lane);
  }
}

// L2 prefetch of one frame of local fp32 data (weights or gradients); one lane issues it.
__device__ __forceinline__ void prefetch_frame_f32(const float* base, size_t frame, int lane) {
  if (lane == 0) ptx::prefetch_l2_bulk(base + frame * kFrameVals, kFrameVals * 4);
}

// Reduce frames base + j, j = first, first + step, ... < lim over all ranks in rank order and hand
// each sum to sink(j, acc, w) together with the frame's (already loaded) weights.  Peers are read in
// batches of PB; batch k + 1 (possibly of the next frame) is requested before batch k is decoded.
// src(p, j) = where rank p's copy of frame base + j is read from.
template <int CODEC, int MODE, class Src, class Sink>
__device__ __forceinline__ void reduce_stream(const RankCtx& c, size_t base_, size_t first_, size_t step_, size_t lim_, int lane,
                                              Src src, Sink sink) {
  using FC = FrameCodec<CODEC, MODE>;
  using Raw = typename FC::Raw;
  constexpr int PB = sizeof(Raw) > 32 ? 2 : 4;
  // frame indices fit 32 bits (2^32 frames = 8 TiB of fp32): half the index registers of size_t
  const uint32_t base = (uint32_t)base_, first = (uint32_t)first_, step = (uint32_t)step_, lim = (uint32_t)lim_;
  const int N = c.nranks;
  const int nb = (N + PB - 1) / PB;
  if (first >= lim) return;
  Raw cur[PB], nxt[PB];
  uint32_t j = first;
  int b = 0;
  // the weight frames (HBM, ~1 us away) are pulled into L2 pf frames ahead of their loads
  const uint32_t pf = c.weight ? c.pf_depth * step : 0;   // prefetch distance in frames of this warp's sequence
  for (uint32_t d = first; pf && d < lim && d - first < pf; d += step) prefetch_frame_f32(c.weight, base + d, lane);
#pragma unroll
  for (int k = 0; k < PB; ++k)
    if (k < N) FC::load(src(k, j), cur[k], lane);
  float acc[16], w[16], v[16];
#pragma unroll 1
  for (;;) {
    uint32_t nj = j;
    int nbq = b + 1;
    if (nbq == nb) { nbq = 0; nj = j + step; }
    const bool more = nj < lim;
    if (b == 0 && pf && j + pf < lim) {
      prefetch_frame_f32(c.weight, base + j + pf, lane);
      // uncompressed input: this rank's own copy of the frame is local HBM as well
      if (CODEC == 0 && lane == 1) ptx::prefetch_l2_bulk(src(c.rank, j + pf), 2048);
    }
    if (more) {
#pragma unroll
      for (int k = 0; k < PB; ++k)
        if (nbq * PB + k < N) FC::load(src(nbq * PB + k, nj), nxt[k], lane);
    }
    if (b == 0 && c.weight) frame_load_f32(c.weight + (size_t)(base + j) * kFrameVals, kFrameVals, w, lane);
#pragma unroll
    for (int k = 0; k < PB; ++k)
      if (b * PB + k < N) {  // warp-uniform
        FC::unpack(cur[k], v, lane);
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i] = (b == 0 && k == 0) ? v[i] : acc[i] + v[i];
      }
    if (b == nb - 1) sink(j, acc, w);
    if (!more) break;
#pragma unroll
    for (int k = 0; k < PB; ++k) cur[k] = nxt[k];
    j = nj;
    b = nbq;
  }
}

// Two-shot gather: apply every other rank's published (reduced) shard.  A warp's (peer, frame) items
// form one flat sequence; they are fetched G at a time, group k + 1 while group k is applied, and
// the weights of item i + 1 are requested before item i is decoded.  A second cursor runs pf_depth
// items ahead of the loads and asks the L2 for the weight frames (no registers held meanwhile).
// The shards belong to a chunk of `full` complete frames that starts at bucket frame `frame0`; rank
// p's published frame j of this chunk sits in its stage D at slot d_off + j (whole bucket: 0, 0).
template <int CODEC, int MODE>
__device__ __forceinline__ void gather_stream(const RankCtx& c, size_t shard_, size_t full_, size_t first_, size_t step_, int lane,
                                              size_t frame0_ = 0, size_t d_off_ = 0) {
  using FC = FrameCodec<CODEC, MODE>;
  using Raw = typename FC::Raw;
  constexpr int G = sizeof(Raw) > 32 ? 1 : 4;
  constexpr uint32_t kNone = 0xFFFFFFFFu;
  const uint32_t shard = (uint32_t)shard_, full = (uint32_t)full_, first = (uint32_t)first_, step = (uint32_t)step_;
  const uint32_t frame0 = (uint32_t)frame0_, d_off = (uint32_t)d_off_;
  const int N = c.nranks;
  auto lim_of = [&](int d) -> uint32_t {
    const uint32_t lo = (uint32_t)((c.rank + d) % N) * shard;
    return full <= lo ? 0u : (full - lo < shard ? full - lo : shard);
  };
  // cursor over the item sequence: peers are visited staggered (rank + 1, rank + 2, ...) so the links are used evenly
  struct Cur { int dp; uint32_t j; };
  auto advance = [&](Cur& q) { q.j += step; while (q.dp < N && q.j >= lim_of(q.dp)) { ++q.dp; q.j = first; } };
  auto frame_of = [&](const Cur& q) -> uint32_t { return frame0 + (uint32_t)((c.rank + q.dp) % N) * shard + q.j; };
  Cur it{1, first};
  while (it.dp < N && it.j >= lim_of(it.dp)) ++it.dp;
  Cur ahead = it;
  const uint32_t pf = c.weight ? c.pf_depth : 0;
  for (uint32_t d = 0; d < pf && ahead.dp < N; ++d) { prefetch_frame_f32(c.weight, frame_of(ahead), lane); advance(ahead); }
  Raw cur[G], nxt[G];
  uint32_t cf[G], nf[G];
  auto fetch = [&](Raw (&r)[G], uint32_t (&fi)[G]) {
#pragma unroll
    for (int k = 0; k < G; ++k) {
      if (it.dp < N) {
        const int p = (c.rank + it.dp) % N;
        const uint32_t f = frame_of(it);
        fi[k] = f;
        FC::load(CODEC == 0 ? reinterpret_cast<const uint8_t*>(c.peer_grad[p]) + (size_t)f * 2048
                            : c.stage_d[p] + (size_t)(d_off + it.j) * FC::kBytes, r[k], lane);
        advance(it);
        if (pf && ahead.dp < N) { prefetch_frame_f32(c.weight, frame_of(ahead), lane); advance(ahead); }
      } else {
        fi[k] = kNone;
      }
    }
  };
  fetch(cur, cf);
  float wa[16], wb[16], g[16];
  if (cf[0] != kNone && c.weight) frame_load_f32(c.weight + (size_t)cf[0] * kFrameVals, kFrameVals, wa, lane);
#pragma unroll 1
  while (cf[0] != kNone) {
    fetch(nxt, nf);
#pragma unroll
    for (int k = 0; k < G; ++k) {
      if (cf[k] != kNone) {  // warp-uniform; valid items are a prefix of the group
        float (&wc)[16] = (k & 1) ? wb : wa;
        float (&wn)[16] = (k & 1) ? wa : wb;
        const uint32_t nfi = (k + 1 < G) ? cf[(k + 1) % G] : nf[0];
        if (nfi != kNone && c.weight) frame_load_f32(c.weight + (size_t)nfi * kFrameVals, kFrameVals, wn, lane);
        FC::unpack(cur[k], g, lane);
        apply_full_w(c, cf[k], g, wc, lane);
      }
    }
#pragma unroll
    for (int k = 0; k < G; ++k) { cur[k] = nxt[k]; cf[k] = nf[k]; }
    if (G & 1) {
#pragma unroll
      for (int i = 0; i < 16; ++i) wa[i] = wb[i];
    }
  }
}

// ---- slow path for the one partial frame of a bucket (never inlined into the hot loops)
template <int CODEC, int MODE>
__device__ __noinline__ void tail_compress(const RankCtx& c, size_t parity_off, size_t f, int lane) {
  float a[16];
  load_grad_frame(c, f, a, lane);
  FrameCodec<CODEC, MODE>::encode(a, c.stage_c[c.rank] + parity_off + f * FrameCodec<CODEC, MODE>::kBytes, lane);
}
// Reduce the partial frame over all ranks; publish it (two-shot: in place for fp32, re-encoded into
// stage D slot j otherwise) when `publish`; apply it.
template <int CODEC, int MODE>
__device__ __noinline__ void tail_reduce_apply(const RankCtx& c, size_t parity_off, size_t f, bool publish, size_t j, int lane) {
  using FC = FrameCodec<CODEC, MODE>;
  float acc[16];
  if (CODEC == 0) reduce_tail_f32(c, f, acc, lane);
  else reduce_frame<CODEC, MODE>(c.nranks, [&](int p) { return reduce_src<CODEC, MODE>(c, parity_off, p, f); }, acc, lane);
  if (publish) {
    if (CODEC == 0) {
      if (c.nranks > 1) frame_store_f32(c.grad + f * kFrameVals, (int)(c.count - f * kFrameVals), acc, lane);
    } else {
      uint8_t* dst = c.stage_d[c.rank] + j * FC::kBytes;
      FC::encode(acc, dst, lane);
      __syncwarp();
      frame_decode<CODEC, MODE>(dst, acc, lane);
    }
  }
  apply_frame(c, f, acc, lane);
}
template <int CODEC, int MODE>
__device__ __noinline__ void tail_gather_apply(const RankCtx& c, int p, size_t f, size_t j, int lane) {
  float g[16];
  if (CODEC == 0) frame_load_f32(c.peer_grad[p] + f * kFrameVals, (int)(c.count - f * kFrameVals), g, lane);
  else frame_decode<CODEC, MODE>(c.stage_d[p] + j * FrameCodec<CODEC, MODE>::kBytes, g, lane);
  apply_frame(c, f, g, lane);
}

// ------------------------------------------------------------------------------------------ one-shot
template <int CODEC, int MODE>
__device__ void oneshot_body(const RankCtx& c, int cta, int nctas, WaitStats& ws) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  const size_t frames = bfp_frames(c.count);
  const size_t full = c.count / kFrameVals;                       // frames [0, full) are complete
  const int N = c.nranks;
  const size_t parity_off = ws.parity ? c.stage_c_half : 0;
  const size_t first = (size_t)cta + (size_t)warp * nctas, step = (size_t)nctas * wpb;
  const bool own_tail = full < frames && (int)(full % nctas) == cta && (int)((full / nctas) % wpb) == warp;
  if (CODEC != 0 && !c.precompressed) {
#pragma unroll 1
    for (size_t f0 = first; f0 < full; f0 += 2 * step) {
      if (c.pf_depth) {   // the fp32 gradient is read exactly once: fetch it into L2 ahead of the loads
        const size_t fa = f0 + (size_t)c.pf_depth * 2 * step;
        if (fa < full) prefetch_frame_f32(c.grad, fa, lane);
        if (fa + step < full) prefetch_frame_f32(c.grad, fa + step, lane);
      }
      compress_pair<CODEC, MODE>(c, parity_off, f0, f0 + step, f0 + step < full, lane);
    }
    if (own_tail) tail_compress<CODEC, MODE>(c, parity_off, full, lane);
  }
  if (N > 1) {
    signal_all(c, 0, cta, ws.epoch);
    wait_all(c, 0, cta, ws.epoch, ws);
  } else {
    __syncthreads();
  }
  mark_phase(ws, 0);
  reduce_stream<CODEC, MODE>(c, 0, first, step, full, lane,
                             [&](int p, size_t f) { return reduce_src<CODEC, MODE>(c, parity_off, p, f); },
                             [&](size_t f, float (&acc)[16], float (&w)[16]) { apply_full_w(c, f, acc, w, lane); });
  if (own_tail) tail_reduce_apply<CODEC, MODE>(c, parity_off, full, false, 0, lane);
  if (CODEC == 0 && N > 1) {  // peers read my gradient in place: do not return before they are done
    signal_all(c, 2, cta, ws.epoch);
    wait_all(c, 2, cta, ws.epoch, ws);
  }
}

// ------------------------------------------------------------------------------------------ two-shot
template <int CODEC, int MODE>
__device__ void twoshot_body(const RankCtx& c, int cta, int nctas, WaitStats& ws) {
  using FC = FrameCodec<CODEC, MODE>;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  const size_t frames = bfp_frames(c.count);
  const size_t full = c.count / kFrameVals;
  const int N = c.nranks;
  const size_t shard = (frames + N - 1) / N;  // frames per shard
  const size_t parity_off = ws.parity ? c.stage_c_half : 0;
  const size_t first = (size_t)cta + (size_t)warp * nctas, step = (size_t)nctas * wpb;
  // the partial frame (if any): shard ts, index tj inside it; CTA b handles the same j of every shard,
  // so that a peer's CTA b finds exactly the frames it needs behind CTA b's flag
  const bool has_tail = full < frames;
  const int ts = has_tail ? (int)(full / shard) : -1;
  const size_t tj = has_tail ? full % shard : 0;
  const bool own_tail = has_tail && (int)(tj % nctas) == cta && (int)((tj / nctas) % wpb) == warp;
  // number of complete frames in shard s
  auto shard_full = [&](int s) -> size_t {
    const size_t lo = (size_t)s * shard;
    return full <= lo ? 0 : (full - lo < shard ? full - lo : shard);
  };
  // phase 0: compress the whole local gradient
  if (CODEC != 0 && !c.precompressed) {
    for (int s = 0; s < N; ++s) {
      const size_t base = (size_t)s * shard, lim = shard_full(s);
#pragma unroll 1
      for (size_t j0 = first; j0 < lim; j0 += 2 * step) {
        if (c.pf_depth) {
          const size_t ja = j0 + (size_t)c.pf_depth * 2 * step;
          if (ja < lim) prefetch_frame_f32(c.grad, base + ja, lane);
          if (ja + step < lim) prefetch_frame_f32(c.grad, base + ja + step, lane);
        }
        compress_pair<CODEC, MODE>(c, parity_off, base + j0, base + j0 + step, j0 + step < lim, lane);
      }
    }
    if (own_tail) tail_compress<CODEC, MODE>(c, parity_off, full, lane);
  }
  if (N > 1) {
    signal_all(c, 0, cta, ws.epoch);
    wait_all(c, 0, cta, ws.epoch, ws);
  } else {
    __syncthreads();
  }
  mark_phase(ws, 0);
  // phase 1: reduce my shard from every rank, publish it re-encoded, update my shard's weights
  {
    const size_t base = (size_t)c.rank * shard, lim = shard_full(c.rank);
    // scattered input (precompressed == 2): every rank's GEMM epilogue already stored its copy of my
    // shard into my inbox, slot [p][j] - the reduce phase touches local memory only
    const bool inbox = CODEC != 0 && c.precompressed == 2;
    auto src = [&](int p, size_t j) -> const uint8_t* {
      return inbox ? c.stage_c[c.rank] + ((size_t)p * shard + j) * FC::kBytes : reduce_src<CODEC, MODE>(c, parity_off, p, base + j);
    };
    reduce_stream<CODEC, MODE>(c, base, first, step, lim, lane, src, [&](size_t j, float (&acc)[16], float (&w)[16]) {
      if constexpr (CODEC == 0) {
        // in place: peers pull the reduced shard out of my gradient buffer
        if (N > 1) frame_store_f32(c.grad + (base + j) * kFrameVals, kFrameVals, acc, lane);
      } else {
        // publish into my stage D (the peers pull it in phase 2), then apply exactly what they will
        // decode (fixes SURVEY D11).  Pushing the frame into every peer's stage D instead was measured
        // at 8 ranks: the gather phase shrinks (57 -> 39 us on an 80 MiB bucket) but the reduce phase
        // grows by more (66 -> 98 us); see profiles/README.md section 3.
        frame_encode_rt<CODEC, MODE>(acc, c.stage_d[c.rank] + j * FC::kBytes, lane);
      }
      apply_full_w(c, base + j, acc, w, lane);
    });
    if (own_tail && ts == c.rank) tail_reduce_apply<CODEC, MODE>(c, parity_off, full, true, tj, lane);
  }
  if (N == 1) return;
  signal_all(c, 1, cta, ws.epoch);
  wait_all(c, 1, cta, ws.epoch, ws);
  mark_phase(ws, 1);
  // phase 2: pull every other rank's reduced shard
  gather_stream<CODEC, MODE>(c, shard, full, first, step, lane);
  if (own_tail && ts != c.rank) tail_gather_apply<CODEC, MODE>(c, ts, full, tj, lane);
  if (CODEC == 0) {
    signal_all(c, 2, cta, ws.epoch);
    wait_all(c, 2, cta, ws.epoch, ws);
  }
}

// ------------------------------------------------------------------------------------------ ring (reference-faithful)
template <int CODEC, int MODE>
__device__ void ring_body(const RankCtx& c, int cta, int nctas, WaitStats& ws) {
  using FC = FrameCodec<CODEC, MODE>;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;  // blockDim == 512: warp = frame within the slice
  const int N = c.nranks, n = c.rank;
  const size_t frames = bfp_frames(c.count);
  const size_t block_frames = (size_t)N * kSliceFrames;
  const size_t nblocks = (frames + block_frames - 1) / block_frames;
  const int down = (n + N - 1) % N, up = (n + 1) % N;
  const size_t slot_bytes = (size_t)kSliceFrames * FC::kBytes;
  const int msgs_per_block = 2 * N - 2;
  float v[16], in[16];
  if (N == 1) {
    for (size_t blk = cta; blk < nblocks; blk += nctas) {
      const size_t f = blk * block_frames + warp;
      load_grad_frame(c, f, v, lane);
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = 0.0f + v[i];
      apply_frame(c, f, v, lane);
    }
    return;
  }
  // My inbox is written by `up`; `down`'s inbox is written by me.  The start barrier guarantees
  // every rank has finished its previous fused call (stream order), i.e. all inboxes are drained -
  // the role the read credits play in the reference (hw/all_reduce.sv:468-483).
  signal_all(c, 0, cta, ws.epoch);
  wait_all(c, 0, cta, ws.epoch, ws);
  uint32_t sent = 0, recvd = 0;
  const uint32_t seq_base = ws.epoch << 12;
  for (size_t blk = cta; blk < nblocks; blk += nctas) {
    uint8_t* outbox = c.stage_d[down] + blk * msgs_per_block * slot_bytes;
    const uint8_t* inbox = c.stage_d[n] + blk * msgs_per_block * slot_bytes;
    const size_t f0 = blk * block_frames;
    // reduce-scatter: SEND_LOCAL, REDUCE, REDUCE_OUTPUT
    for (int t = 0; t < N; ++t) {
      const int s = (n + t) % N;
      const size_t f = f0 + (size_t)s * kSliceFrames + warp;
      load_grad_frame(c, f, v, lane);
      if (t == 0) {
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = 0.0f + v[i];
      } else {
        ++recvd;
        wait_one(c, up, 3, cta, seq_base + recvd, ws);
        frame_decode<CODEC, MODE>(inbox + (size_t)(t - 1) * slot_bytes + (size_t)warp * FC::kBytes, in, lane);
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = in[i] + v[i];
      }
      if (t == N - 1) apply_frame(c, f, v, lane);  // owner: un-quantised sum (hw/all_reduce.sv:984)
      FC::encode(v, outbox + (size_t)t * slot_bytes + (size_t)warp * FC::kBytes, lane);
      ++sent;
      signal_one(c, down, 3, cta, seq_base + sent);
    }
    // all-gather: FORWARD_OUTPUT, OUTPUT(_SEND)
    for (int h = 1; h < N; ++h) {
      const int s = (n + h - 1) % N;
      const size_t f = f0 + (size_t)s * kSliceFrames + warp;
      ++recvd;
      wait_one(c, up, 3, cta, seq_base + recvd, ws);
      frame_decode<CODEC, MODE>(inbox + (size_t)(N - 2 + h) * slot_bytes + (size_t)warp * FC::kBytes, in, lane);
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = in[i] + 0.0f;
      apply_frame(c, f, v, lane);
      if (h < N - 1) {
        FC::encode(v, outbox + (size_t)(N - 1 + h) * slot_bytes + (size_t)warp * FC::kBytes, lane);
        ++sent;
        signal_one(c, down, 3, cta, seq_base + sent);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------ kernel wrappers
template <int ALGO, int CODEC, int MODE>
__device__ __forceinline__ void run_rank(const RankCtx& c, int cta, int nctas) {
  __shared__ WaitStats ws;
  if (threadIdx.x == 0) begin_call(c, ws);
  const unsigned long long t0 = ptx::globaltimer_ns();
  __syncthreads();
  if (ALGO == AINIC_ALGO_ONESHOT) oneshot_body<CODEC, MODE>(c, cta, nctas, ws);
  else if (ALGO == AINIC_ALGO_TWOSHOT) twoshot_body<CODEC, MODE>(c, cta, nctas, ws);
  else ring_body<CODEC, MODE>(c, cta, nctas, ws);
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned long long t1 = ptx::globaltimer_ns();
    if (ws.timeouts) atomicAdd(c.counters + CTR_TIMEOUTS, (unsigned long long)ws.timeouts);
    if (cta == 0) {
      atomicAdd(c.counters + CTR_BUSY_NS, t1 - t0);
      atomicAdd(c.counters + CTR_WAIT_NS, ws.wait_ns);
      atomicAdd(c.counters + CTR_LAUNCHES, 1ull);
      c.counters[CTR_LAST_NS] = t1 - t0;
      if (ws.mark[0]) atomicAdd(c.counters + CTR_TO_BARRIER0_NS, ws.mark[0] - t0);
      if (ws.mark[1]) atomicAdd(c.counters + CTR_TO_BARRIER1_NS, ws.mark[1] - t0);
    }
    end_call(c, ws, nctas);
  }
}

// The context is a __grid_constant__ parameter: it stays in the constant bank and its peer tables are
// indexed there (LDC), instead of being copied to local memory for the dynamic indexing by rank.
// Ring: 16 warps (one per frame of a slice).  One-/two-shot: 8-warp CTAs capped at 144 registers,
// i.e. 36 K registers per CTA, which fits beside a persistent GEMM CTA (320 threads x 80 = 25.6 K) on one SM.
#define AINIC_COMM_RING_ATTR __launch_bounds__(kCommThreads)
#define AINIC_COMM_OVERLAP_ATTR __maxnreg__(144)
template <int ALGO>
constexpr int comm_threads() { return ALGO == AINIC_ALGO_RING ? kCommThreads : kCommThreadsOverlap; }

template <int ALGO, int CODEC, int MODE>
__global__ void AINIC_COMM_RING_ATTR allreduce_kernel(const __grid_constant__ RankCtx c) {
  run_rank<ALGO, CODEC, MODE>(c, blockIdx.x, gridDim.x);
}
template <int ALGO, int CODEC, int MODE>
__global__ void AINIC_COMM_OVERLAP_ATTR allreduce_ov_kernel(const __grid_constant__ RankCtx c) {
  run_rank<ALGO, CODEC, MODE>(c, blockIdx.x, gridDim.x);
}
template <int ALGO, int CODEC, int MODE>
__global__ void AINIC_COMM_OVERLAP_ATTR allreduce_emul_ov_kernel(const RankCtx* ctxs) {
  run_rank<ALGO, CODEC, MODE>(ctxs[blockIdx.y], blockIdx.x, gridDim.x);
}
// Single-GPU emulation: blockIdx.y selects the rank (cooperative launch keeps all of them resident).
template <int ALGO, int CODEC, int MODE>
__global__ void __launch_bounds__(kCommThreads) allreduce_emul_kernel(const RankCtx* ctxs) {
  run_rank<ALGO, CODEC, MODE>(ctxs[blockIdx.y], blockIdx.x, gridDim.x);
}

}  // namespace comm
}  // namespace ainic
