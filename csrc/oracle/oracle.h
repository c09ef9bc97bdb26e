// CPU oracle: the golden model every GPU kernel is checked against, and the timed CPU baseline.
// The reference ships no tests (its RTL testbench hw/sim is withheld, readme.pdf p.2-5), so the
// oracle restates (a) the BFP codec bit-for-bit, (b) the N-rank ring schedule of
// hw/all_reduce.sv:891-1086 with per-hop re-quantisation, (c) the exact and
// one-/two-shot reductions the GPU kernels implement, and (d) the MLP of
// sw/mlp_mpi_example_f32.cpp:700-790 (fwd, softmax loss, bwd, SGD).
#pragma once
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

// ---- codecs (frame layout: see csrc/common/bfp_codec.h) -------------------------------------
// count need not be a multiple of 512; the tail of the last frame is encoded as +0.0f.
size_t ainic_ref_bfp_bytes(size_t count);
size_t ainic_ref_mx_bytes(size_t count);
void ainic_ref_bfp_compress(const float* x, size_t count, uint8_t* out, int mode);
void ainic_ref_bfp_decompress(const uint8_t* in, size_t count, float* out, int mode);
void ainic_ref_mx_compress(const float* x, size_t count, uint8_t* out);
void ainic_ref_mx_decompress(const uint8_t* in, size_t count, float* out);
// Q(x) = decode(encode(x)) in place.
void ainic_ref_bfp_roundtrip(float* x, size_t count, int mode);

// ---- all-reduce models ------------------------------------------------------------------------
// grads: [nranks][count] row-major.  out: [nranks][count] = what each rank ends up applying.
// Exact fp32 sum in rank order 0..N-1 (what MPI_Iallreduce(SUM) of
// sw/mlp_mpi_example_f32.cpp:615 computes up to association order).
void ainic_ref_allreduce_exact(const float* grads, int nranks, size_t count, float* out);
// Reference-faithful ring: 8192-value slices, data flows r -> r-1, incoming (+) local, BFP
// re-quantisation on every hop when compress != 0, owner keeps the unquantised sum.
void ainic_ref_allreduce_ring(const float* grads, int nranks, size_t count, int compress, int mode,
                              float* out);
// codec: 0 = none (fp32), 1 = BFP, 2 = MXFP8, 3 = bf16.
// One-shot: every rank sums decode(encode(g_p)) for p = 0..N-1 in rank order.
void ainic_ref_allreduce_oneshot(const float* grads, int nranks, size_t count, int codec, int mode,
                                 float* out);
// Two-shot: shard owner sums decode(encode(g_p)) in rank order, re-encodes the sum, every rank
// (owner included) applies decode(encode(sum)).  Shards are frame aligned: shard_len =
// ceil(frames/nranks)*512 values.
void ainic_ref_allreduce_twoshot(const float* grads, int nranks, size_t count, int codec, int mode,
                                 float* out);
// w[i] = fmaf(-lr, g[i], w[i])  (hw/weight_update.sv:440-452 with a runtime lr).
void ainic_ref_sgd(float* w, const float* g, size_t count, float lr);
/* Gradient bucket plan of the GPU trainer (csrc/api/bucket_plan.h): lo[l] / hi[l] = lowest / highest layer
 * of layer l's bucket.  schedule (nullable, nsched entries) = layers per bucket from the last layer down. */
void ainic_ref_bucket_plan(int nlayers, const size_t* layer_bytes, size_t bucket_bytes, const int* schedule, int nsched,
                           int* lo, int* hi);

// ---- MLP (sw/mlp_mpi_example_f32.cpp) ---------------------------------------
// Row-major everywhere.  W[l] is [C[l+1]][C[l]] (out x in).  fuse: 0 none, 1 bias, 2 relu, 3 bias+relu.
// y[mb][n] = act(x[mb][k] . W[n][k] + b[n]); mask (may be NULL) gets 1 where y > 0.
void ainic_ref_fc_fwd(const float* x, const float* w, const float* bias, float* y, uint8_t* mask,
                      int mb, int k, int n, int fuse);
// dy is modified in place by the relu mask when fuse has relu.  dx may be NULL (first layer).
void ainic_ref_fc_bwd(const float* x, const float* w, float* dy, const uint8_t* mask, float* dx,
                      float* dw, float* dbias, int mb, int k, int n, int fuse);
// probs = softmax(logits), returns mean cross-entropy; dlogits = loss_weight*(probs-onehot)/mb.
float ainic_ref_softmax_loss(const float* logits, const int32_t* labels, float* probs,
                             float* dlogits, int mb, int n, float loss_weight);
// Deterministic init of the reference (0.05 - U[0,1)/10, sw/mlp_mpi_example_f32.cpp:182-189) on
// top of a splitmix64/xorshift generator (libxsmm's RNG is not in the snapshot).
void ainic_ref_init_buf(float* buf, size_t n, uint64_t seed, int init_pos, int init_one);

#ifdef __cplusplus
}
#endif
