// NIC-style command interface (C1a command queues, C1e completion, C17 host driver API).
//
// Mirrors the driver of sw/mlp_mpi_example_f32.cpp:35-180:
//   all_reduce_setup(done, count, node_id)        -> ainic_nic_setup
//   all_reduce(buf, weight_out, flags, ..., &rq)  -> ainic_nic_all_reduce   (non-blocking post)
//   wait(done, &rq)                               -> ainic_nic_wait         (spin on a done slot)
//   get_all_reduce_latency / get_host_stall_cycles-> ainic_nic_latency_ns / ainic_nic_stall_ns
// Kept from the reference: flat f32 buffers, request handle {idx,count,buf,...}, EIGHT round-robin
// 64-byte done slots (slot = sequence number mod 8, hw/all_reduce.sv:1228,:1368-1375), in-order
// completion, up to 8 requests outstanding (readme.pdf p.2).  Fixed: per-call count,
// runtime lr (D6), wait() has a timeout (D10), abort works (D9), N = 1..8 (D8).
//
// Two backends behind one API:
//   GPU  (AINIC_WITH_CUDA): the post launches the fused all-reduce kernel on the caller's stream;
//        the kernel's last CTA writes the done word into pinned, device-mapped host memory.
//   fake (always available): every NIC has a worker thread that drains its command ring in order
//        and computes with the CPU oracle - "N NICs in one process", like the reference's withheld
//        3-instance RTL testbench (readme.pdf p.2-5).  This is also the timed CPU baseline.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#include "../../include/ainic.h"
#include "../oracle/oracle.h"

#ifdef AINIC_WITH_CUDA
#include <cuda_runtime.h>
extern "C" void ainic_comm_set_done_slot_(ainic_comm_t* c, volatile uint32_t* dev_visible_slot);
#endif

namespace {

using Clock = std::chrono::steady_clock;

struct Cmd {
  float* buf;
  float* weight;
  int flags;
  size_t count;
  int slot;
  uint64_t seq;
  bool inplace_sum;  // MPI-style: no weight update, result replaces buf
};

struct Fabric {  // shared by the linked fake NICs of one process
  int n = 1;
  std::mutex mu;
  std::condition_variable cv;
  int arrived = 0;
  unsigned long long generation = 0;
  std::vector<Cmd> posted;
  std::atomic<bool> aborted{false};
  bool barrier() {
    std::unique_lock<std::mutex> lk(mu);
    const unsigned long long gen = generation;
    if (++arrived == n) {
      arrived = 0;
      generation++;
      cv.notify_all();
      return true;
    }
    while (generation == gen) {
      if (cv.wait_for(lk, std::chrono::milliseconds(50)) == std::cv_status::timeout && aborted.load()) return false;
    }
    return true;
  }
};

}  // namespace

struct ainic_nic {
  int backend = 0, rank = 0, nranks = 1;
  ainic_comm_t* comm = nullptr;
  volatile uint32_t* done = nullptr;       // 8 slots x 16 words, host memory
  volatile uint32_t* done_dev = nullptr;   // device view of the same memory (GPU backend)
  size_t count = 0;
  int codec = AINIC_CODEC_BFP, bfp_mode = AINIC_BFP_CLEAN, algo = AINIC_ALGO_AUTO;
  float lr = 0.1f;
  uint64_t seq = 0;
  std::atomic<bool> aborted{false};
  // fake backend
  std::shared_ptr<Fabric> fabric;
  std::thread worker;
  std::mutex mu;
  std::condition_variable cv;
  std::deque<Cmd> ring;                    // depth <= 8
  bool stop = false;
  std::atomic<uint64_t> busy_ns{0}, stall_ns{0};
  std::vector<float> dev_weights;          // the "on-card DDR" weight copy (hw/weight_update.sv:511-664)
};

static void fake_execute(ainic_nic* nic, const Cmd& c) {
  Fabric& F = *nic->fabric;
  const int N = F.n;
  const auto t0 = Clock::now();
  {
    std::lock_guard<std::mutex> lk(F.mu);
    F.posted[nic->rank] = c;
  }
  if (!F.barrier()) return;                 // all ranks posted request seq
  const auto t1 = Clock::now();
  std::vector<float> all((size_t)N * c.count), out((size_t)N * c.count);
  for (int r = 0; r < N; ++r) memcpy(all.data() + (size_t)r * c.count, F.posted[r].buf, c.count * sizeof(float));
  if (!F.barrier()) return;                 // everyone has copied the inputs; buffers may be overwritten
  int algo = nic->algo;
  if (algo == AINIC_ALGO_AUTO) algo = (N <= 2 || c.count * 4 <= (512u << 10)) ? AINIC_ALGO_ONESHOT : AINIC_ALGO_TWOSHOT;
  if (nic->codec == AINIC_CODEC_F32 && algo != AINIC_ALGO_RING) ainic_ref_allreduce_exact(all.data(), N, c.count, out.data());
  else if (algo == AINIC_ALGO_RING) ainic_ref_allreduce_ring(all.data(), N, c.count, nic->codec == AINIC_CODEC_BFP, nic->bfp_mode, out.data());
  else if (algo == AINIC_ALGO_ONESHOT) ainic_ref_allreduce_oneshot(all.data(), N, c.count, nic->codec, nic->bfp_mode, out.data());
  else ainic_ref_allreduce_twoshot(all.data(), N, c.count, nic->codec, nic->bfp_mode, out.data());
  const float* mine = out.data() + (size_t)nic->rank * c.count;
  if (c.inplace_sum) {
    memcpy(c.buf, mine, c.count * sizeof(float));
  } else if (c.weight) {
    // flags: (re)load the initial weights into the device-resident copy (hw/weight_update.sv:624-637)
    if (c.flags || nic->dev_weights.size() != c.count) nic->dev_weights.assign(c.weight, c.weight + c.count);
    ainic_ref_sgd(c.weight, mine, c.count, nic->lr);
    nic->dev_weights.assign(c.weight, c.weight + c.count);
  }
  const auto t2 = Clock::now();
  nic->stall_ns += std::chrono::duration_cast<std::chrono::nanoseconds>(t1 - t0).count();
  nic->busy_ns += std::chrono::duration_cast<std::chrono::nanoseconds>(t2 - t0).count();
  std::atomic_thread_fence(std::memory_order_release);
  nic->done[c.slot * 16] = 1;               // 512'd1 into the done cache line (hw/all_reduce.sv:1370)
}

static void fake_worker(ainic_nic* nic) {
  for (;;) {
    Cmd c;
    {
      std::unique_lock<std::mutex> lk(nic->mu);
      nic->cv.wait(lk, [&] { return nic->stop || !nic->ring.empty(); });
      if (nic->ring.empty()) return;
      c = nic->ring.front();
    }
    fake_execute(nic, c);
    {
      std::lock_guard<std::mutex> lk(nic->mu);
      nic->ring.pop_front();
    }
    nic->cv.notify_all();
  }
}

extern "C" int ainic_nic_open(ainic_nic_t** out, ainic_comm_t* comm, int backend, int rank, int nranks) {
  if (!out || nranks < 1 || nranks > AINIC_MAX_RANKS || rank < 0 || rank >= nranks) return AINIC_ERR_ARG;
  ainic_nic* nic = new ainic_nic();
  nic->backend = backend;
  nic->rank = rank;
  nic->nranks = nranks;
  nic->comm = comm;
  if (backend == 0) {
#ifdef AINIC_WITH_CUDA
    if (!comm) { delete nic; return AINIC_ERR_ARG; }
    void* h = nullptr;
    if (cudaHostAlloc(&h, 4096, cudaHostAllocMapped) != cudaSuccess) { delete nic; return AINIC_ERR_CUDA; }
    memset(h, 0, 4096);
    void* d = nullptr;
    if (cudaHostGetDevicePointer(&d, h, 0) != cudaSuccess) { cudaFreeHost(h); delete nic; return AINIC_ERR_CUDA; }
    nic->done = static_cast<volatile uint32_t*>(h);
    nic->done_dev = static_cast<volatile uint32_t*>(d);
#else
    delete nic;
    return AINIC_ERR_UNSUPPORTED;
#endif
  } else {
    void* h = nullptr;
    if (posix_memalign(&h, 4096, 4096)) { delete nic; return AINIC_ERR_ARG; }
    memset(h, 0, 4096);
    nic->done = static_cast<volatile uint32_t*>(h);
    nic->fabric = std::make_shared<Fabric>();
    nic->fabric->n = 1;
    nic->fabric->posted.resize(1);
    nic->rank = rank;
    nic->worker = std::thread(fake_worker, nic);
  }
  *out = nic;
  return AINIC_OK;
}

extern "C" int ainic_nic_fake_link(ainic_nic_t** nics, int n) {
  if (!nics || n < 1 || n > AINIC_MAX_RANKS) return AINIC_ERR_ARG;
  auto f = std::make_shared<Fabric>();
  f->n = n;
  f->posted.resize(n);
  for (int i = 0; i < n; ++i) {
    if (!nics[i] || nics[i]->backend != 1) return AINIC_ERR_ARG;
    nics[i]->fabric = f;
    nics[i]->rank = i;
    nics[i]->nranks = n;
  }
  return AINIC_OK;
}

extern "C" int ainic_nic_close(ainic_nic_t* nic) {
  if (!nic) return AINIC_ERR_ARG;
  if (nic->backend == 1) {
    {
      std::lock_guard<std::mutex> lk(nic->mu);
      nic->stop = true;
    }
    nic->cv.notify_all();
    if (nic->worker.joinable()) nic->worker.join();
    free(const_cast<uint32_t*>(nic->done));
  } else {
#ifdef AINIC_WITH_CUDA
    cudaDeviceSynchronize();
    cudaFreeHost(const_cast<uint32_t*>(nic->done));
#endif
  }
  delete nic;
  return AINIC_OK;
}

extern "C" int ainic_nic_setup(ainic_nic_t* nic, size_t count, int codec, int bfp_mode, int algo, float lr) {
  if (!nic) return AINIC_ERR_ARG;
  nic->count = count;
  nic->codec = codec;
  nic->bfp_mode = bfp_mode;
  nic->algo = algo;
  nic->lr = lr;
  return AINIC_OK;
}

extern "C" volatile uint32_t* ainic_nic_done_buffer(ainic_nic_t* nic) { return nic ? nic->done : nullptr; }

static int post(ainic_nic* nic, float* buf, float* weight, int flags, size_t count, bool inplace, ainic_rqst_t* rq,
                void* stream) {
  if (!nic || !buf || !rq) return AINIC_ERR_ARG;
  if (count == 0) count = nic->count;      // the reference ignores the per-call count
  if (count == 0) return AINIC_ERR_ARG;
  if (nic->aborted.load()) return AINIC_ERR_ABORTED;
  const int slot = (int)(nic->seq & (AINIC_DONE_SLOTS - 1));
  // the slot's previous request (seq - 8) must have completed: at most 8 outstanding
  if (nic->seq >= AINIC_DONE_SLOTS && nic->done[slot * 16] == 0) return AINIC_ERR_BUSY;
  nic->done[slot * 16] = 0;                // done[done_idx<<4] = 0, sw/mlp_mpi_example_f32.cpp:115
  std::atomic_thread_fence(std::memory_order_seq_cst);
  rq->idx = slot;
  rq->count = (int)count;
  rq->buf = buf;
  rq->dev_buf = weight;
  rq->seq = nic->seq;
  if (nic->backend == 1) {
    Cmd c{buf, weight, flags, count, slot, nic->seq, inplace};
    {
      std::lock_guard<std::mutex> lk(nic->mu);
      nic->ring.push_back(c);
    }
    nic->cv.notify_all();
  } else {
#ifdef AINIC_WITH_CUDA
    ainic_allreduce_args_t a{};
    a.grad = buf;
    a.weight = inplace ? nullptr : weight;
    a.grad_out = inplace ? buf : nullptr;
    a.count = count;
    a.lr = nic->lr;
    a.codec = nic->codec;
    a.bfp_mode = nic->bfp_mode;
    a.algo = nic->algo;
    // fp32 one-shot reads peers' buffers while writing its own: not safe in place
    if (inplace && a.codec == AINIC_CODEC_F32 && nic->nranks > 1) a.algo = AINIC_ALGO_TWOSHOT;
    ainic_comm_set_done_slot_(nic->comm, nic->done_dev + slot * 16);
    int rc = ainic_allreduce_fused(nic->comm, &a, stream);
    if (rc) {   // nothing was launched: do not leave the slot armed for somebody else's next launch
      ainic_comm_set_done_slot_(nic->comm, nullptr);
      nic->done[slot * 16] = 1;
      return rc;
    }
#else
    return AINIC_ERR_UNSUPPORTED;
#endif
  }
  nic->seq++;
  return AINIC_OK;
}

extern "C" int ainic_nic_all_reduce(ainic_nic_t* nic, float* buf, float* weight_out, int flags, size_t count,
                                    ainic_rqst_t* rqst, void* stream) {
  return post(nic, buf, weight_out, flags, count, false, rqst, stream);
}
extern "C" int ainic_iallreduce(ainic_nic_t* nic, float* buf, size_t count, ainic_rqst_t* rqst, void* stream) {
  return post(nic, buf, nullptr, 0, count, true, rqst, stream);
}

extern "C" int ainic_nic_wait(ainic_nic_t* nic, ainic_rqst_t* rq, double timeout_s) {
  if (!nic || !rq || rq->idx < 0 || rq->idx >= AINIC_DONE_SLOTS) return AINIC_ERR_ARG;
  const auto t0 = Clock::now();
  volatile uint32_t* slot = nic->done + rq->idx * 16;
  unsigned spins = 0;
  while (*slot == 0) {                      // while (done[idx<<4] == 0), sw/mlp_mpi_example_f32.cpp:163-168
    if (nic->aborted.load()) return AINIC_ERR_ABORTED;
    if ((++spins & 0x3FF) == 0) {
      if (timeout_s > 0 && std::chrono::duration<double>(Clock::now() - t0).count() > timeout_s) return AINIC_ERR_TIMEOUT;
      std::this_thread::yield();
    }
  }
  std::atomic_thread_fence(std::memory_order_acquire);
  // the device raises 2 instead of 1 when the launch gave up waiting for a peer (spin time-out):
  // the weights may hold partial sums, the caller must not treat the request as completed
  return *slot == 2u ? AINIC_ERR_TIMEOUT : AINIC_OK;
}

extern "C" int ainic_nic_abort(ainic_nic_t* nic) {
  if (!nic) return AINIC_ERR_ARG;
  nic->aborted.store(true);
  if (nic->fabric) {
    nic->fabric->aborted.store(true);
    nic->fabric->cv.notify_all();
  }
  return AINIC_OK;
}

extern "C" uint64_t ainic_nic_latency_ns(ainic_nic_t* nic) {
  if (!nic) return 0;
  if (nic->backend == 1) return nic->busy_ns.load();
#ifdef AINIC_WITH_CUDA
  uint64_t c[8] = {};
  if (ainic_comm_counters(nic->comm, c) == AINIC_OK) return c[0];
#endif
  return 0;
}
extern "C" uint64_t ainic_nic_stall_ns(ainic_nic_t* nic) {
  if (!nic) return 0;
  if (nic->backend == 1) return nic->stall_ns.load();
#ifdef AINIC_WITH_CUDA
  uint64_t c[8] = {};
  if (ainic_comm_counters(nic->comm, c) == AINIC_OK) return c[1];
#endif
  return 0;
}
