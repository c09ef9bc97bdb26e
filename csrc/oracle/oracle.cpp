// CPU oracle implementation.  See oracle.h for what each entry models and where in the
// reference the behaviour comes from.
#include "oracle.h"
#include "../api/bucket_plan.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "../common/bfp_codec.h"

using namespace ainic;

// ------------------------------------------------------------------------------------------
// codecs
// ------------------------------------------------------------------------------------------
extern "C" size_t ainic_ref_bfp_bytes(size_t count) { return bfp_bytes(count); }
extern "C" size_t ainic_ref_mx_bytes(size_t count) { return mx_bytes(count); }

static inline float at_or_zero(const float* x, size_t i, size_t count) { return i < count ? x[i] : 0.0f; }

extern "C" void ainic_ref_bfp_compress(const float* x, size_t count, uint8_t* out, int mode) {
  const size_t frames = bfp_frames(count);
#pragma omp parallel for schedule(static)
  for (long long f = 0; f < (long long)frames; ++f) {
    uint8_t* fr = out + (size_t)f * kBfpFrameBytes;
    for (int b = 0; b < 32; ++b) {
      const size_t base = (size_t)f * kFrameVals + (size_t)b * kBfpBlock;
      uint32_t E = 0;
      for (int j = 0; j < kBfpBlock; ++j) {
        uint32_t e = bfp_exp_field(f2u(at_or_zero(x, base + j, count)));
        E = e > E ? e : E;
      }
      for (int j = 0; j < kBfpBlock; ++j)
        fr[b * kBfpBlock + j] = (uint8_t)(int8_t)bfp_quant(at_or_zero(x, base + j, count), E, mode);
      fr[kBfpExpOff + b] = (uint8_t)E;
    }
  }
}

extern "C" void ainic_ref_bfp_decompress(const uint8_t* in, size_t count, float* out, int mode) {
  const size_t frames = bfp_frames(count);
#pragma omp parallel for schedule(static)
  for (long long f = 0; f < (long long)frames; ++f) {
    const uint8_t* fr = in + (size_t)f * kBfpFrameBytes;
    for (int b = 0; b < 32; ++b) {
      const uint32_t E = fr[kBfpExpOff + b];
      for (int j = 0; j < kBfpBlock; ++j) {
        const size_t i = (size_t)f * kFrameVals + (size_t)b * kBfpBlock + j;
        if (i < count) out[i] = bfp_dequant((int32_t)(int8_t)fr[b * kBfpBlock + j], E, mode);
      }
    }
  }
}

extern "C" void ainic_ref_mx_compress(const float* x, size_t count, uint8_t* out) {
  const size_t frames = bfp_frames(count);
#pragma omp parallel for schedule(static)
  for (long long f = 0; f < (long long)frames; ++f) {
    uint8_t* fr = out + (size_t)f * kMxFrameBytes;
    for (int b = 0; b < 16; ++b) {
      const size_t base = (size_t)f * kFrameVals + (size_t)b * kMxBlock;
      uint32_t E = 0;
      for (int j = 0; j < kMxBlock; ++j) {
        uint32_t e = bfp_exp_field(f2u(at_or_zero(x, base + j, count)));
        E = e > E ? e : E;
      }
      const uint32_t sb = mx_scale_byte(E);
      const float inv = mx_inv_scale(sb);
      for (int j = 0; j < kMxBlock; ++j)
        fr[b * kMxBlock + j] = (uint8_t)f32_to_e4m3(at_or_zero(x, base + j, count) * inv);
      fr[kMxScaleOff + b] = (uint8_t)sb;
    }
  }
}

extern "C" void ainic_ref_mx_decompress(const uint8_t* in, size_t count, float* out) {
  const size_t frames = bfp_frames(count);
#pragma omp parallel for schedule(static)
  for (long long f = 0; f < (long long)frames; ++f) {
    const uint8_t* fr = in + (size_t)f * kMxFrameBytes;
    for (int b = 0; b < 16; ++b) {
      const float s = mx_scale(fr[kMxScaleOff + b]);
      for (int j = 0; j < kMxBlock; ++j) {
        const size_t i = (size_t)f * kFrameVals + (size_t)b * kMxBlock + j;
        if (i < count) out[i] = e4m3_to_f32(fr[b * kMxBlock + j]) * s;
      }
    }
  }
}

extern "C" void ainic_ref_bfp_roundtrip(float* x, size_t count, int mode) {
  std::vector<uint8_t> tmp(bfp_bytes(count));
  ainic_ref_bfp_compress(x, count, tmp.data(), mode);
  ainic_ref_bfp_decompress(tmp.data(), count, x, mode);
}

static inline float bf16_round(float x) {
  uint32_t u = f2u(x);
  if ((u & 0x7FFFFFFFu) > 0x7F800000u) return u2f(u | 0x400000u);
  u += 0x7FFFu + ((u >> 16) & 1u);
  return u2f(u & 0xFFFF0000u);
}

// Q(x) for a whole buffer under a codec id (0 none, 1 BFP, 2 MX, 3 bf16).
static void codec_roundtrip(float* x, size_t count, int codec, int mode) {
  if (codec == 1) {
    ainic_ref_bfp_roundtrip(x, count, mode);
  } else if (codec == 2) {
    std::vector<uint8_t> tmp(mx_bytes(count));
    ainic_ref_mx_compress(x, count, tmp.data());
    ainic_ref_mx_decompress(tmp.data(), count, x);
  } else if (codec == 3) {
    for (size_t i = 0; i < count; ++i) x[i] = bf16_round(x[i]);
  }
}

// ------------------------------------------------------------------------------------------
// all-reduce models
// ------------------------------------------------------------------------------------------
extern "C" void ainic_ref_allreduce_exact(const float* grads, int nranks, size_t count, float* out) {
#pragma omp parallel for schedule(static)
  for (long long i = 0; i < (long long)count; ++i) {
    float s = grads[i];
    for (int r = 1; r < nranks; ++r) s += grads[(size_t)r * count + i];
    for (int r = 0; r < nranks; ++r) out[(size_t)r * count + i] = s;
  }
}

extern "C" void ainic_ref_allreduce_ring(const float* grads, int nranks, size_t count, int compress,
                                         int mode, float* out) {
  const int N = nranks;
  const size_t kSlice = 8192;  // BUF_SIZE=512 cache lines (hw/all_reduce.sv:102-103)
  const size_t block = (size_t)N * kSlice;
  const size_t nblocks = (count + block - 1) / block;
#pragma omp parallel for schedule(dynamic)
  for (long long b = 0; b < (long long)nblocks; ++b) {
    std::vector<float> p(kSlice), local(kSlice);
    for (int s = 0; s < N; ++s) {
      const size_t off = (size_t)b * block + (size_t)s * kSlice;
      auto load = [&](int rank, float* dst) {
        for (size_t i = 0; i < kSlice; ++i)
          dst[i] = (off + i < count) ? grads[(size_t)rank * count + off + i] : 0.0f;
      };
      auto store = [&](int rank, const float* src) {
        for (size_t i = 0; i < kSlice && off + i < count; ++i) out[(size_t)rank * count + off + i] = src[i];
      };
      // SEND_LOCAL at rank s: 0 (+) local (hw/all_reduce.sv:908)
      load(s, p.data());
      for (size_t i = 0; i < kSlice; ++i) p[i] = 0.0f + p[i];
      // REDUCE / REDUCE_OUTPUT at ranks s-1 ... s+1
      for (int k = 1; k < N; ++k) {
        const int rank = ((s - k) % N + N) % N;
        if (compress) ainic_ref_bfp_roundtrip(p.data(), kSlice, mode);
        load(rank, local.data());
        for (size_t i = 0; i < kSlice; ++i) p[i] = p[i] + local[i];
      }
      const int owner = (s + 1) % N;
      store(owner, p.data());  // un-quantised sum_out -> wu_gradient_fifo (hw/all_reduce.sv:984)
      // FORWARD_OUTPUT / OUTPUT(_SEND): owner-1, owner-2, ... each receives Q(previous) (+) 0
      for (int h = 1; h < N; ++h) {
        const int rank = ((owner - h) % N + N) % N;
        if (compress) ainic_ref_bfp_roundtrip(p.data(), kSlice, mode);
        for (size_t i = 0; i < kSlice; ++i) p[i] = p[i] + 0.0f;
        store(rank, p.data());
      }
    }
  }
}

extern "C" void ainic_ref_allreduce_oneshot(const float* grads, int nranks, size_t count, int codec,
                                            int mode, float* out) {
  std::vector<float> q((size_t)nranks * count);
  memcpy(q.data(), grads, q.size() * sizeof(float));
  for (int r = 0; r < nranks; ++r) codec_roundtrip(q.data() + (size_t)r * count, count, codec, mode);
  ainic_ref_allreduce_exact(q.data(), nranks, count, out);
}

extern "C" void ainic_ref_allreduce_twoshot(const float* grads, int nranks, size_t count, int codec,
                                            int mode, float* out) {
  // Reduce exactly like one-shot, then the owner re-encodes; because every rank (owner too)
  // applies decode(encode(sum)), the shard split does not change the values.
  ainic_ref_allreduce_oneshot(grads, nranks, count, codec, mode, out);
  codec_roundtrip(out, count, codec, mode);
  for (int r = 1; r < nranks; ++r) memcpy(out + (size_t)r * count, out, count * sizeof(float));
}

extern "C" void ainic_ref_bucket_plan(int nlayers, const size_t* layer_bytes, size_t bucket_bytes, const int* schedule,
                                      int nsched, int* lo, int* hi) {
  std::vector<size_t> bytes(layer_bytes, layer_bytes + nlayers);
  std::vector<int> sched(schedule ? schedule : nullptr, schedule ? schedule + nsched : nullptr), vlo, vhi;
  ainic::plan_buckets(bytes, bucket_bytes, sched, vlo, vhi);
  for (int l = 0; l < nlayers; ++l) { lo[l] = vlo[l]; hi[l] = vhi[l]; }
}

extern "C" void ainic_ref_sgd(float* w, const float* g, size_t count, float lr) {
#pragma omp parallel for schedule(static)
  for (long long i = 0; i < (long long)count; ++i) w[i] = fmaf(-lr, g[i], w[i]);
}

// ------------------------------------------------------------------------------------------
// MLP
// ------------------------------------------------------------------------------------------
extern "C" void ainic_ref_fc_fwd(const float* x, const float* w, const float* bias, float* y,
                                 uint8_t* mask, int mb, int k, int n, int fuse) {
  const bool use_bias = (fuse == 1 || fuse == 3), relu = (fuse == 2 || fuse == 3);
#pragma omp parallel for schedule(static)
  for (int i = 0; i < mb; ++i) {
    for (int j = 0; j < n; ++j) {
      const float* xr = x + (size_t)i * k;
      const float* wr = w + (size_t)j * k;
      float acc = 0.0f;
#pragma omp simd reduction(+ : acc)
      for (int t = 0; t < k; ++t) acc += xr[t] * wr[t];
      if (use_bias) acc += bias[j];
      if (relu) {
        if (mask) mask[(size_t)i * n + j] = acc > 0.0f;
        acc = acc > 0.0f ? acc : 0.0f;
      }
      y[(size_t)i * n + j] = acc;
    }
  }
}

extern "C" void ainic_ref_fc_bwd(const float* x, const float* w, float* dy, const uint8_t* mask,
                                 float* dx, float* dw, float* dbias, int mb, int k, int n, int fuse) {
  const bool relu = (fuse == 2 || fuse == 3);
  if (relu && mask) {
#pragma omp parallel for schedule(static)
    for (long long i = 0; i < (long long)mb * n; ++i)
      if (!mask[i]) dy[i] = 0.0f;
  }
  if (dbias) {
    for (int j = 0; j < n; ++j) {
      float s = 0.0f;
      for (int i = 0; i < mb; ++i) s += dy[(size_t)i * n + j];
      dbias[j] = s;
    }
  }
  if (dx) {
#pragma omp parallel for schedule(static)
    for (int i = 0; i < mb; ++i) {
      float* dxr = dx + (size_t)i * k;
      for (int t = 0; t < k; ++t) dxr[t] = 0.0f;
      for (int j = 0; j < n; ++j) {
        const float d = dy[(size_t)i * n + j];
        const float* wr = w + (size_t)j * k;
#pragma omp simd
        for (int t = 0; t < k; ++t) dxr[t] += d * wr[t];
      }
    }
  }
#pragma omp parallel for schedule(static)
  for (int j = 0; j < n; ++j) {
    float* dwr = dw + (size_t)j * k;
    for (int t = 0; t < k; ++t) dwr[t] = 0.0f;
    for (int i = 0; i < mb; ++i) {
      const float d = dy[(size_t)i * n + j];
      const float* xr = x + (size_t)i * k;
#pragma omp simd
      for (int t = 0; t < k; ++t) dwr[t] += d * xr[t];
    }
  }
}

extern "C" float ainic_ref_softmax_loss(const float* logits, const int32_t* labels, float* probs,
                                        float* dlogits, int mb, int n, float loss_weight) {
  double loss = 0.0;
  for (int i = 0; i < mb; ++i) {
    const float* l = logits + (size_t)i * n;
    float m = l[0];
    for (int j = 1; j < n; ++j) m = l[j] > m ? l[j] : m;
    double sum = 0.0;
    for (int j = 0; j < n; ++j) sum += exp((double)l[j] - m);
    const int lab = labels[i];
    for (int j = 0; j < n; ++j) {
      const float p = (float)(exp((double)l[j] - m) / sum);
      if (probs) probs[(size_t)i * n + j] = p;
      if (dlogits) dlogits[(size_t)i * n + j] = loss_weight * (p - (j == lab ? 1.0f : 0.0f)) / (float)mb;
    }
    loss += -((double)l[lab] - m - log(sum));
  }
  return (float)(loss / mb);
}

static inline uint64_t splitmix64(uint64_t& s) {
  uint64_t z = (s += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

extern "C" void ainic_ref_init_buf(float* buf, size_t n, uint64_t seed, int init_pos, int init_one) {
  uint64_t s = seed;
  for (size_t i = 0; i < n; ++i) {
    const double u = (double)(splitmix64(s) >> 11) * (1.0 / 9007199254740992.0);
    buf[i] = (float)(init_one ? 1.0 : (init_pos ? u : (0.05 - u / 10.0)));
  }
}

// ---------------------------------------------------------------- symmetric-heap offset allocator (csrc/comm/heap_alloc.h)
// The allocator is plain host code; these hooks let the CPU test tier drive it without a GPU.
#include "../comm/heap_alloc.h"
extern "C" void* ainic_ref_heap_new(size_t capacity) {
  auto* h = new ainic::HeapAlloc();
  h->reset(capacity);
  return h;
}
extern "C" void ainic_ref_heap_delete(void* h) { delete static_cast<ainic::HeapAlloc*>(h); }
// returns the offset, or (size_t)-1 when the request does not fit
extern "C" size_t ainic_ref_heap_alloc(void* h, size_t bytes, size_t align) { return static_cast<ainic::HeapAlloc*>(h)->alloc(bytes, align); }
extern "C" int ainic_ref_heap_free(void* h, size_t off) { return static_cast<ainic::HeapAlloc*>(h)->free(off) ? 1 : 0; }
// out[0] bytes in use, out[1] high-water offset, out[2] number of free blocks below it
extern "C" void ainic_ref_heap_stats(void* h, size_t out[3]) {
  auto* a = static_cast<ainic::HeapAlloc*>(h);
  out[0] = a->bytes_in_use();
  out[1] = a->high_water();
  out[2] = a->free_blocks();
}
