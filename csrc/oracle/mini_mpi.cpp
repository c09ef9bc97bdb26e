#include "mini_mpi.h"

#include <string.h>

#include <chrono>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <vector>

struct mmpi_world {
  int n = 1;
  std::mutex mu;
  std::condition_variable cv;
  int arrived = 0;
  unsigned long long generation = 0;
  std::vector<const void*> slots;   // one pointer per rank, valid between two barriers
  std::vector<float*> rbufs;
  std::vector<size_t> counts;
  std::vector<float> scratch;
};

extern "C" int mmpi_size(mmpi_world_t* w) { return w->n; }

extern "C" void mmpi_barrier(mmpi_world_t* w, int) {
  std::unique_lock<std::mutex> lk(w->mu);
  const unsigned long long gen = w->generation;
  if (++w->arrived == w->n) {
    w->arrived = 0;
    w->generation++;
    w->cv.notify_all();
  } else {
    w->cv.wait(lk, [&] { return w->generation != gen; });
  }
}

extern "C" void mmpi_bcast(mmpi_world_t* w, int rank, float* buf, size_t count, int root) {
  if (rank == root) w->slots[0] = buf;
  mmpi_barrier(w, rank);
  if (rank != root) memcpy(buf, w->slots[0], count * sizeof(float));
  mmpi_barrier(w, rank);
}

extern "C" void mmpi_scatter(mmpi_world_t* w, int rank, const float* send, float* recv, size_t count, int root) {
  if (rank == root) w->slots[0] = send;
  mmpi_barrier(w, rank);
  memcpy(recv, static_cast<const float*>(w->slots[0]) + (size_t)rank * count, count * sizeof(float));
  mmpi_barrier(w, rank);
}

extern "C" void mmpi_iallreduce_sum(mmpi_world_t* w, int rank, float* buf, size_t count, mmpi_request_t* req) {
  w->rbufs[rank] = buf;
  w->counts[rank] = count;
  req->active = 1;
}

extern "C" void mmpi_wait(mmpi_world_t* w, int rank, mmpi_request_t* req) {
  if (!req->active) return;
  const size_t count = w->counts[rank];
  mmpi_barrier(w, rank);                       // every rank has posted its buffer
  if (rank == 0) w->scratch.resize(count);
  mmpi_barrier(w, rank);
  // each rank reduces a contiguous slice of the vector, in rank order -> bitwise deterministic
  const size_t lo = count * rank / w->n, hi = count * (rank + 1) / w->n;
  for (size_t i = lo; i < hi; ++i) {
    float s = w->rbufs[0][i];
    for (int r = 1; r < w->n; ++r) s += w->rbufs[r][i];
    w->scratch[i] = s;
  }
  mmpi_barrier(w, rank);
  memcpy(w->rbufs[rank], w->scratch.data(), count * sizeof(float));
  mmpi_barrier(w, rank);
  req->active = 0;
}

extern "C" double mmpi_wtime(void) {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

extern "C" int mmpi_run(int nranks, mmpi_main_fn fn, void* user) {
  mmpi_world w;
  w.n = nranks;
  w.slots.assign(nranks, nullptr);
  w.rbufs.assign(nranks, nullptr);
  w.counts.assign(nranks, 0);
  std::vector<std::thread> th;
  for (int r = 0; r < nranks; ++r) th.emplace_back([&, r] { fn(&w, r, user); });
  for (auto& t : th) t.join();
  return 0;
}
