/* b200-ainic public C API.
 *
 * Everything the reference exposes through its host driver functions
 * (sw/mlp_mpi_example_f32.cpp:35-180: ikl_setup / all_reduce_setup / all_reduce / wait /
 * get_all_reduce_latency / get_host_stall_cycles) plus the kernels the FPGA datapath and the
 * libxsmm_dnn calls turn into on a B200.  All entry points are extern "C", take raw device
 * pointers and a cudaStream_t passed as void*, and return AINIC_OK (0) or a negative error.
 */
#ifndef AINIC_H_
#define AINIC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AINIC_OK 0
#define AINIC_ERR_ARG (-1)
#define AINIC_ERR_CUDA (-2)
#define AINIC_ERR_TIMEOUT (-3)
#define AINIC_ERR_BUSY (-4)
#define AINIC_ERR_ABORTED (-5)
#define AINIC_ERR_UNSUPPORTED (-6)

#define AINIC_MAX_RANKS 8
#define AINIC_DONE_SLOTS 8 /* hw/all_reduce.sv:1228 3-bit done_id; sw/mlp_mpi_example_f32.cpp:115-118 */

/* gradient wire formats */
#define AINIC_CODEC_F32 0
#define AINIC_CODEC_BFP 1  /* reference BFP16x8: hw/bfp_adapter.sv */
#define AINIC_CODEC_MXFP8 2
#define AINIC_CODEC_BF16 3
/* BFP numerics */
#define AINIC_BFP_HW_EXACT 0
#define AINIC_BFP_CLEAN 1
/* all-reduce algorithms */
#define AINIC_ALGO_AUTO 0
#define AINIC_ALGO_ONESHOT 1
#define AINIC_ALGO_TWOSHOT 2
#define AINIC_ALGO_RING 3 /* reference-faithful schedule, hw/all_reduce.sv:686-1088 */
#define AINIC_ALGO_NVLS 4

const char* ainic_version(void);
const char* ainic_last_error(void);

/* ---------------------------------------------------------------- codecs (K8/K8', K14) */
size_t ainic_bfp_bytes(size_t count);
size_t ainic_mx_bytes(size_t count);
int ainic_bfp_compress(const float* x, size_t count, uint8_t* out, int mode, void* stream);
int ainic_bfp_decompress(const uint8_t* in, size_t count, float* out, int mode, void* stream);
int ainic_mx_compress(const float* x, size_t count, uint8_t* out, void* stream);
int ainic_mx_decompress(const uint8_t* in, size_t count, float* out, void* stream);

/* ---------------------------------------------------------------- fused SGD (K9, hw/weight_update.sv:440-452)
 * w = fmaf(-lr, g, w); optional bf16 shadow of the new weights for the tensor-core GEMMs. */
int ainic_sgd(float* w, const float* g, void* w_bf16 /*nullable*/, size_t count, float lr, void* stream);
int ainic_cast_f32_bf16(const float* x, void* y_bf16, size_t count, void* stream);

/* ---------------------------------------------------------------- MLP kernels (K1-K5) */
#define AINIC_GEMM_NT 0 /* C[M,N] = A[M,K] . B[N,K]^T  bf16 out  (forward)        */
#define AINIC_GEMM_NN 1 /* C[M,N] = A[M,K] . B[K,N]    bf16 out  (backward data)  */
#define AINIC_GEMM_TN 2 /* C[M,N] = A[K,M]^T . B[K,N]  fp32 out  (backward weight)*/
/* GEMM -> reduce-scatter in one kernel: the backward-weight epilogue stores every compressed frame
 * straight into the memory of the rank that will reduce it (plain stores on peer-mapped pointers,
 * tile by tile while the tensor cores work on the next tile).  A bucket of F frames is cut into
 * nranks shards of shard_frames = ceil(F / nranks); frame f belongs to rank f / shard_frames and
 * lands in that rank's inbox at slot [rank][f % shard_frames] (row pitch shard_frames frames).
 * No counterpart in the reference, whose NIC DMA-reads the finished fp32 buffer
 * (hw/all_reduce.sv:246-532); it is what sitting the codec and the wire behind the adder tree
 * becomes when the "adder tree" is a tensor-core GEMM. */
typedef struct ainic_gemm_scatter {
  void* inbox[AINIC_MAX_RANKS]; /* inbox base of every rank as mapped into this process ([rank] is local) */
  int nranks, rank;
  unsigned shard_frames;        /* frames per owner */
  unsigned frame0;              /* index of this GEMM's first frame inside the bucket */
} ainic_gemm_scatter_t;
typedef struct ainic_gemm_epilogue {
  const float* bias;      /* [N], added before activation (bf16 outputs) */
  uint8_t* mask_out;      /* [M][ldc] ReLU mask written by the forward pass */
  const uint8_t* mask_in; /* [M][ldc] ReLU mask applied by the backward-data pass */
  float alpha;            /* fp32 outputs: scale */
  int relu;
  int accumulate;         /* fp32 outputs: C += result */
  void* cgrad;            /* AINIC_GEMM_TN only, nullable: write the result as BFP/MXFP8 frames of the flat
                             [M][ldc] bucket instead of fp32 (the layout ainic_allreduce_fused reads when
                             args.grad_compressed is set); needs N % 32 == 0 and ldc % 32 == 0 */
  int codec;              /* AINIC_CODEC_BFP or AINIC_CODEC_MXFP8 when cgrad is set */
  int bfp_mode;
  const ainic_gemm_scatter_t* scatter; /* nullable; with it cgrad is ignored and the frames go to the owners' inboxes */
  uint32_t* done_flag;    /* nullable, device: incremented once (release, gpu scope) by the last CTA of the launch after
                             every output tile has been stored - the "ready" doorbell a flow all-reduce waits on */
  uint32_t* done_ticket;  /* device scratch word for done_flag (zero before the first launch; the kernel resets it) */
} ainic_gemm_epilogue_t;
/* bf16 operands, fp32 accumulation on tcgen05 tensor cores.  block_n = 0 picks the tile width,
 * max_ctas = 0 uses every SM (a smaller value leaves SMs for a concurrent all-reduce kernel). */
int ainic_gemm_bf16(int layout, const void* a, int lda, const void* b, int ldb, void* c, int ldc, int M, int N,
                    int K, const ainic_gemm_epilogue_t* epi, int block_n, int max_ctas, void* stream);

/* Diagnostics: the next `launches` launches of the 1-CTA bf16 kernel write seven time stamps per CTA
 * (%globaltimer ns: entry, after griddepcontrol.wait, first operand stage landed, last MMA issued, first
 * accumulator complete, epilogue done, exit) into consecutive [148][8] uint64 blocks of the device buffer
 * `buf`; buf = NULL switches it off.  Not thread safe; scripts/gemm_timeline.py. */
int ainic_gemm_trace(void* buf, int launches);

/* fp32 operands on the tensor cores as tf32 (tcgen05.mma.kind::tf32), fp32 result with the alpha / accumulate
 * epilogue; layouts as ainic_gemm_bf16.  The tensor-core form of the reference's all-fp32 MLP
 * (libxsmm_dnn_fc_*_exec_f32, sw/mlp_mpi_example_f32.cpp:708-709, :741-742): no bf16 rounding anywhere.
 * N % 8 == 0, K / lda / ldb / ldc % 4 == 0. */
int ainic_gemm_tf32(int layout, const float* a, int lda, const float* b, int ldb, float* c, int ldc, int M, int N, int K,
                    const ainic_gemm_epilogue_t* epi, int block_n, int max_ctas, void* stream);

/* MX block-scaled fp8 GEMM (K14): e4m3 operands with one ue8m0 scale per 32 values along K, multiplied by
 * tcgen05.mma.kind::mxf8f6f4.block_scale - the tensor core reads the scales from tensor memory.  The block
 * geometry is the MXFP8 wire codec's (and the 32-value generalisation of the reference's 16-value
 * shared-exponent blocks, hw/bfp_adapter.sv:59-76).  Scales travel in 512-byte "atoms" covering 128 rows x
 * 4 K-blocks: byte (r % 32) * 16 + (r / 32) * 4 + kb, atoms laid out [rows / 128][K / 128].
 *   ainic_mx_quantize_rows   x [rows][ld], K contiguous  -> q [rows][K] e4m3 + atoms   (operands of NT; A of NN)
 *   ainic_mx_quantize_cols   x [K][ld], K the slow dim   -> q [K][cols] e4m3 + atoms   (operands of TN; B of NN)
 * Same bytes and scales as ainic_mx_compress on the same 32-value blocks.  K % 128 == 0. */
size_t ainic_mx_sf_bytes(size_t rows, size_t K);
int ainic_mx_quantize_rows(const void* x, int x_is_bf16, int rows, int K, int ld, void* q, void* sf, void* stream);
int ainic_mx_quantize_cols(const void* x, int x_is_bf16, int K, int cols, int ld, void* q, void* sf, void* stream);
/* c = bf16 (bias / ReLU / mask epilogue) or fp32 (alpha / accumulate); block_n 128 (default) or 256. */
int ainic_gemm_mxfp8(int layout, const void* a_e4m3, int lda, const void* sfa, const void* b_e4m3, int ldb, const void* sfb,
                     void* c, int ldc, int out_fp32, int M, int N, int K, const ainic_gemm_epilogue_t* epi, int block_n,
                     int max_ctas, void* stream);

/* Row softmax + cross-entropy; dlogits = loss_weight*(p - onehot)/mb as bf16; loss_sum += sum_i CE_i.
 * (libxsmm_dnn_smax_{fwd,bwd}_exec_f32, sw/mlp_mpi_example_f32.cpp:718-728) */
int ainic_softmax_xent(const void* logits_bf16, const int32_t* labels, void* probs_bf16 /*nullable*/,
                       void* dlogits_bf16, float* loss_sum, int mb, int n, int ld, float loss_weight, void* stream);
/* fp32 logits / dlogits variant for the all-fp32 (tf32 tensor core) configuration */
int ainic_softmax_xent_f32(const float* logits, const int32_t* labels, float* dlogits, float* loss_sum, int mb, int n, int ld,
                           float loss_weight, void* stream);
/* x[i] = mask[i] ? x[i] : 0 on bf16 data (ReLU backward of the loss gradient); count % 8 == 0 */
int ainic_apply_mask_bf16(void* x_bf16, const uint8_t* mask, size_t count, void* stream);
/* dbias[n] = sum over rows of dy[mb][n] (bf16 in, fp32 out) */
int ainic_bias_grad(const void* dy_bf16, float* dbias, int mb, int n, int ld, void* stream);

/* ---------------------------------------------------------------- communicator (C1, C14, C21)
 * One communicator per process/GPU.  The symmetric heap is one device allocation per rank whose
 * IPC handle is exchanged by the caller (torch.distributed store, files, MPI ...): create ->
 * export handle -> all-gather handles out of band -> connect.  nranks == 1 needs no exchange. */
typedef struct ainic_comm ainic_comm_t;
#define AINIC_IPC_HANDLE_BYTES 64
int ainic_comm_create(ainic_comm_t** out, int rank, int nranks, size_t heap_bytes);
int ainic_comm_export(ainic_comm_t* c, void* handle_out /*64 B*/);
int ainic_comm_connect(ainic_comm_t* c, const void* all_handles /*nranks*64 B, rank order*/);
int ainic_comm_destroy(ainic_comm_t* c);
/* Allocation inside the symmetric heap (first fit, see ainic_comm_free); every rank must allocate in the same order so that
 * offsets match (the reference's flat DMA-able buffers from fpga.allocBuffer,
 * sw/mlp_mpi_example_f32.cpp:396-401). */
void* ainic_comm_alloc(ainic_comm_t* c, size_t bytes, size_t align);
/* Returns a block to the heap: same call in the same order on every rank (the allocator is deterministic, so
 * offsets stay symmetric), and no kernel anywhere may still use the block.  First fit with coalescing. */
int ainic_comm_free(ainic_comm_t* c, void* ptr);
/* Staging areas that a larger bucket outgrew are only retired on a multi-rank communicator (a peer may still be
 * reading them); this returns them to the heap.  Call on every rank behind a barrier with no fused call in flight.
 * Returns the number of blocks released. */
int ainic_comm_release_retired(ainic_comm_t* c);
/* out[0] bytes in use, out[1] high-water offset, out[2] free blocks below it, out[3] capacity */
int ainic_comm_heap_stats(ainic_comm_t* c, size_t out[4]);
void* ainic_comm_peer_ptr(ainic_comm_t* c, const void* local_ptr, int peer);
int ainic_comm_rank(const ainic_comm_t* c);
int ainic_comm_size(const ainic_comm_t* c);

typedef struct ainic_allreduce_args {
  float* grad;        /* flat fp32 gradient bucket inside the symmetric heap (sum of all ranks on return
                         only for algo RING/f32 paths; compressed paths leave it untouched) */
  float* weight;      /* flat fp32 master weights, updated in place: w -= lr * sum(g)   (nullable) */
  void* weight_bf16;  /* bf16 shadow of the updated weights (nullable) */
  float* grad_out;    /* where to write the reduced gradient when no weight update is fused (nullable) */
  size_t count;       /* values */
  float lr;
  int codec, bfp_mode, algo;
  int max_ctas;       /* 0 = default */
  void* grad_compressed; /* nullable: this rank's gradient already in `codec` frame layout inside the symmetric
                            heap (written by the GEMM epilogue); the kernel then skips its compression pass and
                            `grad` is only used for its size.  The caller alternates between two such buffers
                            from call to call on the same bucket. */
  int grad_scattered; /* nonzero: grad_compressed is this rank's INBOX, already filled by every rank's GEMM
                         epilogue (ainic_gemm_scatter_t layout, shard_frames = ceil(frames / nranks)): the
                         reduce phase reads local memory only.  Two-shot only; algo is ignored. */
} ainic_allreduce_args_t;
/* Fused compress -> (peer pull) -> fp32 reduce -> decompress -> SGD, one kernel launch on stream. */
int ainic_allreduce_fused(ainic_comm_t* c, const ainic_allreduce_args_t* args, void* stream);
/* Flow all-reduce (csrc/comm/flow.cuh): one persistent launch that walks the bucket chunk by chunk - each
 * chunk goes through compress -> reduce own 1/N -> publish -> gather as soon as its producers have
 * bumped the chunk's device-side ready counter (ainic_gemm_epilogue_t.done_flag), the flag round
 * trips of chunk k hidden behind the data movement of chunk k - 1.  In the trainer a chunk is a
 * layer, launched once per step at the start of the backward pass.  The reference's host posts one
 * request per layer and keeps two in flight (sw/mlp_mpi_example_f32.cpp:735-788); this is that
 * pipeline with the doorbell moved into the producer kernel.  Chunks are processed in array order;
 * offset and count are in values, multiples of 512; args->grad / weight / weight_bf16 /
 * grad_compressed are the bucket bases and args->count its total extent.  local_ready (device,
 * nullable) must be zeroed by the caller before this call's producers are launched. */
#define AINIC_FLOW_MAX_CHUNKS 32
typedef struct ainic_flow_chunk {
  size_t offset, count;  /* values from the bucket base; multiples of 512 */
  int ready_index;       /* index into local_ready, or -1: complete when the call is launched */
  int ready_need;        /* producers that must have finished (counter value to wait for); 0 means 1 */
} ainic_flow_chunk_t;
int ainic_allreduce_flow(ainic_comm_t* c, const ainic_allreduce_args_t* args, const ainic_flow_chunk_t* chunks, int nchunks,
                         const uint32_t* local_ready, void* stream);
/* Single-GPU emulation of the flow all-reduce (all ranks' buffers on one device, one cooperative
 * launch): chunk_counts[nchunks] cut [0, count) into consecutive chunks, processed last chunk first
 * like the layers of a backward pass. */
int ainic_allreduce_flow_emulate(int nranks, float* grads, float* weights, void* weights_bf16, float* grads_out, size_t count,
                                 float lr, int codec, int bfp_mode, const size_t* chunk_counts, int nchunks,
                                 int ctas_per_rank, void* stream);
/* With AINIC_FLOW_TRACE set, every flow call records %globaltimer stamps per chunk (processing order): out[w * 32 + k],
 * w = 0 doorbell seen by the forwarder, 1 reduce released, 2 reduce done, 3 gather released, 4 gather done (one
 * time-keeping warp each).  The software analogue of tapping the FSM states with SignalTap (hw/iochan.stp). */
int ainic_comm_flow_trace(ainic_comm_t* c, uint64_t out[160]);
/* Pre-size the communicator's staging for calls of up to `count` values (collective: same call on
 * every rank).  Optional - staging also grows on demand, but growth abandons the old area. */
int ainic_comm_reserve(ainic_comm_t* c, size_t count);
/* Same for ainic_allreduce_flow calls of up to `count` values in `nchunks` chunks (needs about half the staging). */
int ainic_comm_reserve_flow(ainic_comm_t* c, size_t count, int nchunks);
/* Device-side phase counters of the last completed fused all-reduce (ns from %globaltimer):
 * [0] total busy, [1] waiting for peers, [2] launches.  The analogue of lpbk_latency /
 * stall_* (hw/all_reduce.sv:92-98, :892-1072). */
int ainic_comm_counters(ainic_comm_t* c, uint64_t out[8]);

/* ---------------------------------------------------------------- NVLS: in-switch reduce + multicast (K13)
 * A second, VMM-backed heap whose memory is bound to an NVSwitch multicast object.  Gradient
 * buckets allocated from it can be reduced with multimem.ld_reduce / multimem.st
 * (ainic_allreduce_nvls: uncompressed fp32, count % 512 == 0).  Set-up is three collective stages
 * with a barrier (provided by the caller) between them; any failure returns
 * AINIC_ERR_UNSUPPORTED with the reason in ainic_last_error() and the P2P kernels remain usable.
 * The reference has no counterpart (its switch is plain Ethernet); the work it replaces is the
 * ring of hw/all_reduce.sv:891-1086. */
typedef struct ainic_nvls ainic_nvls_t;
int ainic_nvls_supported(void);
int ainic_nvls_stage1(ainic_nvls_t** out, int rank, int nranks, size_t bytes, const char* sock_name);
int ainic_nvls_stage2(ainic_nvls_t* h);
int ainic_nvls_stage3(ainic_nvls_t* h);
void* ainic_nvls_alloc(ainic_nvls_t* h, size_t bytes);
void* ainic_nvls_mc_ptr(ainic_nvls_t* h, const void* uc_ptr);
int ainic_nvls_destroy(ainic_nvls_t* h);
int ainic_allreduce_nvls(ainic_comm_t* c, ainic_nvls_t* h, const ainic_allreduce_args_t* args, void* stream);

/* Single-GPU emulation of an N-rank all-reduce: all ranks' buffers live on one device and one
 * cooperative launch runs every rank's CTAs (the software analogue of the withheld 3-instance RTL
 * testbench, readme.pdf p.2-5).  grads/weights: [nranks][count]. */
int ainic_allreduce_emulate(int nranks, float* grads, float* weights, void* weights_bf16, float* grads_out,
                            size_t count, float lr, int codec, int bfp_mode, int algo, int ctas_per_rank,
                            void* stream);

/* Same emulation for callers that may not allocate device memory (a service whose HBM is fully
 * claimed by its own pool, a loop that cannot take cudaFree's device-wide sync).  No counterpart in the reference: the FPGA owns its buffers statically.
 * All scratch (staging, flags, counters) lives in one 2 MiB static __device__ array per GPU and
 * the rank contexts travel by value, so the call never allocates.  Limits: nranks <=
 * AINIC_NOALLOC_MAX_RANKS, count <= AINIC_NOALLOC_MAX_COUNT values per rank; larger requests
 * return AINIC_ERR_ARG with the reason in errbuf (nullable).  The first call on a device resolves
 * the scratch address and must precede any stream capture; calls share the scratch, so calls whose
 * GPU work overlaps race - serialise them.  Returns after the kernel has finished. */
#define AINIC_NOALLOC_MAX_RANKS 4
#define AINIC_NOALLOC_MAX_COUNT 32768
int ainic_allreduce_emulate_noalloc(int nranks, float* grads, float* weights, void* weights_bf16, float* grads_out,
                                    size_t count, float lr, int codec, int bfp_mode, int algo, int ctas_per_rank,
                                    void* stream, char* errbuf, size_t errbuf_len);

/* ---------------------------------------------------------------- native data-parallel MLP trainer (C18)
 * The training application of sw/mlp_mpi_example_f32.cpp:191-881 as a native executor: forward
 * GEMMs, softmax loss, backward GEMMs on two streams and one fused all-reduce+SGD per layer on a
 * third, per-layer events instead of the reference's two-deep request ring (:735-788).
 * sizes = C[0..nlayers]; fuse: 0 none, 1 bias, 2 ReLU(mask), 3 bias+ReLU (:479-489). */
typedef struct ainic_mlp ainic_mlp_t;
int ainic_mlp_create(ainic_mlp_t** out, ainic_comm_t* comm, int nlayers, const int* sizes, int mb, int fuse, float lr,
                     float loss_weight, int codec, int bfp_mode, int algo);
/* compute_dtype: 0 = bf16 activations / weight shadows on kind::f16 tensor cores (ainic_mlp_create); 1 = the
 * reference's all-fp32 data flow on kind::tf32 tensor cores: fp32 activations, GEMMs read the fp32 master
 * weights directly, no shadow (fuse must be 0; x passed to step/forward is fp32). */
int ainic_mlp_create2(ainic_mlp_t** out, ainic_comm_t* comm, int nlayers, const int* sizes, int mb, int fuse, float lr,
                      float loss_weight, int codec, int bfp_mode, int algo, int compute_dtype);
int ainic_mlp_destroy(ainic_mlp_t* m);
/* free_heap != 0 also returns the weight / gradient arenas to the communicator's heap: only with one rank or behind a
 * barrier with every rank idle, and with the same value on every rank (ainic_mlp_destroy: free_heap = one rank). */
int ainic_mlp_destroy_ex(ainic_mlp_t* m, int free_heap);
/* what: 0 weight f32, 1 grad f32, 2 shadow bf16, 3 bias f32, 4 dbias f32, 5 act bf16, 6 dact bf16, 7 relu mask */
void* ainic_mlp_buffer(ainic_mlp_t* m, int what, int layer);
int ainic_mlp_sync_shadows(ainic_mlp_t* m, void* stream);
int ainic_mlp_forward(ainic_mlp_t* m, const void* x_bf16, void* stream);
int ainic_mlp_step(ainic_mlp_t* m, const void* x_bf16, const int32_t* labels, float* loss_out /*device, nullable*/,
                   void* stream);
int ainic_mlp_wait_updates(ainic_mlp_t* m, void* stream);
/* DETAILED_PROFILE (sw/mlp_mpi_example_f32.cpp:32, :809-815): enable = 1 starts (and clears) per-phase device
 * timing, 0 stops it, -1 only reads.  out_ms: forward FC, loss, backward+weight-gradient FC, tail until the
 * last weight update; returns the number of profiled steps.  Profiling serialises the step at its end. */
unsigned long long ainic_mlp_profile(ainic_mlp_t* m, int enable, double out_ms[4]);
unsigned long long ainic_mlp_launches(ainic_mlp_t* m);
/* Gradient bucketing: consecutive layers are contiguous in memory, so one fused all-reduce can cover
 * several.  bytes = fp32 gradient bytes to collect (from the last layer down) before a call is
 * issued; 0 = one call per layer like the reference (sw/mlp_mpi_example_f32.cpp:754).  Default: 0 on
 * one rank, 80 MiB from 2 ranks; AINIC_BUCKET_MB overrides at create time. */
int ainic_mlp_set_bucket_bytes(ainic_mlp_t* m, size_t bytes);
size_t ainic_mlp_bucket_bytes(ainic_mlp_t* m);
void* ainic_mlp_loss_ptr(ainic_mlp_t* m);

/* ---------------------------------------------------------------- NIC-style command interface (C1a, C17)
 * ainic_nic_* mirrors the driver of sw/mlp_mpi_example_f32.cpp:65-180: setup once, post up to 8
 * requests that complete in order, each signalling one of 8 round-robin 64-byte done slots that
 * the host polls.  backend: 0 = GPU (fused kernels over the communicator), 1 = host fake device
 * (CPU threads + oracle math; lets the protocol be tested without a GPU). */
typedef struct ainic_nic ainic_nic_t;
typedef struct ainic_rqst { /* fpga_rqst, sw/mlp_mpi_example_f32.cpp:41-46 */
  int idx;
  int count;
  float* buf;
  volatile float* dev_buf;
  uint64_t seq;
} ainic_rqst_t;
int ainic_nic_open(ainic_nic_t** out, ainic_comm_t* comm /*null for fake backend*/, int backend, int rank,
                   int nranks);
int ainic_nic_close(ainic_nic_t* nic);
/* all_reduce_setup(done, count, node_id): programs total length + node id, returns the done buffer. */
int ainic_nic_setup(ainic_nic_t* nic, size_t count, int codec, int bfp_mode, int algo, float lr);
volatile uint32_t* ainic_nic_done_buffer(ainic_nic_t* nic); /* 8 slots x 16 words */
/* all_reduce(buf, weight_out, flags, ...): flags != 0 = (re)load initial weights into the
 * device-resident copy first (hw/weight_update.sv:624-637). */
int ainic_nic_all_reduce(ainic_nic_t* nic, float* buf, float* weight_out, int flags, size_t count,
                         ainic_rqst_t* rqst, void* stream);
int ainic_nic_wait(ainic_nic_t* nic, ainic_rqst_t* rqst, double timeout_s /*<=0: forever*/);
int ainic_nic_abort(ainic_nic_t* nic); /* kill_syn_e0, unconnected in the reference (hw/all_reduce.sv:83) */
uint64_t ainic_nic_latency_ns(ainic_nic_t* nic);     /* get_all_reduce_latency */
uint64_t ainic_nic_stall_ns(ainic_nic_t* nic);       /* get_host_stall_cycles */
/* fake backend only: connect N fake NICs of one process into a ring. */
int ainic_nic_fake_link(ainic_nic_t** nics, int n);

/* MPI-style entry (sw/mlp_mpi_example_f32.cpp:615-620: MPI_Iallreduce(IN_PLACE, buf, count, FLOAT, SUM) + Wait) */
int ainic_iallreduce(ainic_nic_t* nic, float* buf, size_t count, ainic_rqst_t* rqst, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* AINIC_H_ */
