// Gradient bucket plan shared by the GPU trainer (csrc/api/mlp.cu) and the CPU golden model
// (csrc/oracle/oracle.cpp exports it so that the plan is testable without a GPU).
//
// The reference hands the NIC one bucket per layer (sw/mlp_mpi_example_f32.cpp:754).  Here the layers
// are contiguous in memory and consecutive layers can share one fused all-reduce call.  Walking down
// from the last layer (the order the backward pass produces gradients in), a bucket is closed
//   * after schedule[i] layers when an explicit schedule is given (what is left over after the
//     schedule is used up goes one layer per bucket), or
//   * as soon as it holds bucket_bytes of fp32 gradients otherwise;
// layer 0 always closes the last bucket.  lo[l] / hi[l] = lowest / highest layer of l's bucket.
#pragma once
#include <stddef.h>

#include <vector>

namespace ainic {

inline void plan_buckets(const std::vector<size_t>& layer_bytes, size_t bucket_bytes, const std::vector<int>& schedule,
                         std::vector<int>& lo, std::vector<int>& hi) {
  const int n = (int)layer_bytes.size();
  lo.assign(n, 0);
  hi.assign(n, 0);
  int top = n - 1, in_bucket = 0;
  size_t pend = 0, sched = 0;
  for (int l = n - 1; l >= 0; --l) {
    pend += layer_bytes[l];
    ++in_bucket;
    bool close = l == 0;
    if (!schedule.empty()) {
      const int want = sched < schedule.size() ? schedule[sched] : 1;
      close = close || in_bucket >= (want > 0 ? want : 1);
    } else {
      close = close || pend >= bucket_bytes;
    }
    if (!close) continue;
    for (int k = l; k <= top; ++k) { lo[k] = l; hi[k] = top; }
    top = l - 1;
    pend = 0;
    in_bucket = 0;
    ++sched;
  }
}

}  // namespace ainic
