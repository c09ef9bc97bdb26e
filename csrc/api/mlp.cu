// Native data-parallel MLP trainer: the B200 form of the reference's training application
// sw/mlp_mpi_example_f32.cpp:191-881 (main).  One process per GPU; per iteration (the hot loop at
// :701-788):
//   forward  L tcgen05 GEMMs (+bias/ReLU/mask epilogue)              :707-710
//   loss     fused softmax + cross-entropy fwd/bwd                   :718-728
//   backward for l = L-1 .. 0:  dW_l (fp32, straight into the flat gradient bucket) on the wgrad
//            stream, dX_l on the main stream, then the fused all-reduce + SGD of bucket l on the
//            comm stream - layer l's reduction overlaps layer l-1's GEMMs like request[i%2] does in
//            the reference (:735-788), but with up to L requests in flight instead of 2.
// The next iteration's forward of layer l waits only on layer l's update event.
// Weights are flat fp32 [C[l+1]][C[l]] buffers in the symmetric heap (the reference's
// fpga.allocBuffer'd fil_libxsmm, :396-401) with a bf16 shadow for the tensor cores.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include "../../include/ainic.h"
#include "bucket_plan.h"

namespace {

struct Layer {
  int in = 0, out = 0;
  float* w = nullptr;            // fp32 master, symmetric heap
  float* g = nullptr;            // fp32 gradient bucket, symmetric heap
  __nv_bfloat16* wb = nullptr;   // bf16 shadow
  float* bias = nullptr;         // fp32 [out]
  float* dbias = nullptr;
  uint8_t* mask = nullptr;       // [mb][out] ReLU mask
  cudaEvent_t dy_ready = nullptr;   // dact[l+1] complete (main stream)
  cudaEvent_t dw_ready = nullptr;   // gradient bucket complete (wgrad stream)
  cudaEvent_t updated = nullptr;    // weights + shadow updated (comm stream)
  cudaEvent_t dx_done = nullptr;    // backward-data GEMM finished reading the shadow (main stream)
  uint8_t* cg[2] = {nullptr, nullptr};  // compressed gradient (BFP / MXFP8 frames), symmetric heap, one per step parity
  size_t slot = 0;               // values reserved in the flat arenas (count rounded up to 1024)
};

}  // namespace

struct ainic_mlp {
  ainic_comm_t* comm = nullptr;
  int nlayers = 0, mb = 0, fuse = 0;
  bool f32 = false;                  // all-fp32 data flow on kind::tf32 tensor cores (no bf16 anywhere)
  float lr = 0.1f, loss_weight = 0.1f;
  int codec = AINIC_CODEC_F32, bfp_mode = AINIC_BFP_CLEAN, algo = AINIC_ALGO_AUTO;
  int comm_ctas = 0, gemm_ctas = 0, wgrad_ctas = 0;
  std::vector<Layer> layers;
  std::vector<__nv_bfloat16*> act;   // act[0..L]   [mb][C[l]]
  std::vector<__nv_bfloat16*> dact;  // dact[1..L]  [mb][C[l]]
  float* loss_sum = nullptr;         // device scalar
  cudaStream_t wgrad_stream = nullptr, comm_stream = nullptr;
  cudaEvent_t fwd_done = nullptr;
  // DETAILED_PROFILE of the reference (sw/mlp_mpi_example_f32.cpp:32, :699-751): timestamps on the
  // main stream at step start / forward end / loss end / backward end, plus the last update.
  bool profile = false;
  cudaEvent_t prof[5] = {};
  double phase_ms[4] = {0, 0, 0, 0};   // fwd_fc, loss (fwd+bwd), bwdupd_fc, update tail; accumulated
  unsigned long long prof_steps = 0;
  bool first_step = true;
  unsigned long long launches = 0;
  unsigned long long step_idx = 0;
  bool fuse_compress = false;        // weight-gradient GEMM epilogue emits the wire format directly
  bool fuse_scatter = false;         // ... and stores every frame into its owner's inbox (GEMM -> reduce-scatter)
  // All layers' masters / gradients / shadows / compressed gradients live back to back in one
  // arena each (layer 0 first), so any run of consecutive layers is ONE flat bucket for the fused
  // all-reduce.  bucket_bytes = fp32 gradient bytes to collect (walking down from the last layer)
  // before a call is issued; 0 = one call per layer, as in the reference (:754).
  float* w_arena = nullptr;
  float* g_arena = nullptr;
  __nv_bfloat16* wb_arena = nullptr;
  uint8_t* cg_arena[2] = {nullptr, nullptr};
  size_t bucket_bytes = 0;
  std::vector<int> bucket_layers;    // optional explicit schedule: layers per bucket from the last layer down
                                     // (AINIC_BUCKET_LAYERS="4,3,2,1"); overrides bucket_bytes when non-empty
  // CUDA-graph replay of the whole iteration (one launch per step instead of ~30 kernel launches and
  // ~100 event operations).  One instantiated graph per (input, labels, loss, parity) tuple.
  struct StepGraph {
    const void* x; const int32_t* labels; float* loss_out; unsigned parity; size_t bucket_bytes;
    cudaGraphExec_t exec; unsigned long long kernels;
  };
  std::vector<StepGraph> graphs;
  bool use_graph = false;
  bool in_capture = false;
  bool events_stale = false;         // the per-layer events were last recorded inside a capture
  // Flow mode (csrc/comm/flow.cuh): ONE persistent all-reduce launch per step, started with the backward
  // pass; it takes each layer as soon as that layer's weight-gradient GEMM (and the backward-data GEMM
  // still reading the layer's bf16 shadow) have bumped the layer's device-side ready counter.
  bool flow = false;
  uint32_t* flow_ready = nullptr;    // [nlayers] producers finished this step (zeroed at the start of every backward)
  uint32_t* flow_tickets = nullptr;  // [2 * nlayers] last-CTA tickets of the dW / dX GEMMs
  cudaEvent_t bwd_start = nullptr;
  cudaEvent_t step_done = nullptr;   // end of the last graph-replayed step on the caller's stream
  cudaStream_t cap_stream = nullptr; // origin stream of captures (the caller's may be the legacy stream, which cannot capture)
};

static size_t max_bucket_count(const ainic_mlp* m);

static int fail(const char* what) {
  fprintf(stderr, "ainic_mlp: %s failed: %s\n", what, cudaGetErrorString(cudaGetLastError()));
  return AINIC_ERR_CUDA;
}
#define TRY_CUDA(x) do { if ((x) != cudaSuccess) return fail(#x); } while (0)
#define TRY(x) do { int _rc = (x); if (_rc) return _rc; } while (0)

extern "C" int ainic_mlp_create(ainic_mlp_t** out, ainic_comm_t* comm, int nlayers, const int* sizes, int mb, int fuse,
                                float lr, float loss_weight, int codec, int bfp_mode, int algo) {
  return ainic_mlp_create2(out, comm, nlayers, sizes, mb, fuse, lr, loss_weight, codec, bfp_mode, algo, 0);
}

extern "C" int ainic_mlp_create2(ainic_mlp_t** out, ainic_comm_t* comm, int nlayers, const int* sizes, int mb, int fuse,
                                 float lr, float loss_weight, int codec, int bfp_mode, int algo, int compute_dtype) {
  if (compute_dtype != 0 && compute_dtype != 1) return AINIC_ERR_ARG;
  if (compute_dtype == 1 && fuse != 0) return AINIC_ERR_UNSUPPORTED;   // fp32 outputs have no bias / ReLU epilogue
  // mb is the K extent of the weight-gradient GEMM (TMA rows of 16 bytes: a multiple of 8 bf16 values)
  if (!out || !comm || nlayers < 1 || !sizes || mb < 8 || mb % 8) return AINIC_ERR_ARG;
  for (int i = 0; i <= nlayers; ++i)
    if (sizes[i] < 8 || sizes[i] % 8) return AINIC_ERR_ARG;
  ainic_mlp* m = new ainic_mlp();
  m->comm = comm;
  m->nlayers = nlayers;
  m->mb = mb;
  m->fuse = fuse;
  m->lr = lr;
  m->loss_weight = loss_weight;
  m->codec = codec;
  m->bfp_mode = bfp_mode;
  m->algo = algo;
  m->f32 = compute_dtype == 1;
  if (const char* e = getenv("AINIC_COMM_CTAS")) m->comm_ctas = atoi(e);
  if (const char* e = getenv("AINIC_GEMM_CTAS")) m->gemm_ctas = atoi(e);
  if (const char* e = getenv("AINIC_WGRAD_CTAS")) m->wgrad_ctas = atoi(e);
  // With a block-compressed wire format the weight-gradient GEMM can write the wire format itself (the
  // reference's BFP adapter sits right behind the adder tree the same way, hw/all_reduce.sv:745-772):
  // no fp32 gradient in memory, no compression pass in the all-reduce.  Measured on the reference MLP
  // (ms/step, off -> on, one fused call per 80 MiB bucket): 2 ranks 0.651 -> 0.586, 4 ranks 0.698 ->
  // 0.657, 8 ranks 0.844 -> 0.800.  Default: on whenever there is a peer; AINIC_FUSE_COMPRESS=0/1/2 overrides.
  {
    const int n = ainic_comm_size(comm);
    int mode = n >= 2 ? 1 : 0;
    if (const char* e = getenv("AINIC_FUSE_COMPRESS")) mode = atoi(e);
    m->fuse_compress = mode >= 1 && n > 1 && !m->f32 && (codec == AINIC_CODEC_BFP || codec == AINIC_CODEC_MXFP8);
    // AINIC_FUSE_COMPRESS=2: the epilogue also performs the reduce-scatter's data movement - each frame
    // is stored into the inbox of the rank that owns it, over NVLink, while the GEMM is still running.
    // Bit-identical results (scripts/mp_check.py 3b).  The frames are staged through shared memory and
    // leave as whole row segments (epilogue_store_frames); stored straight from the tcgen05.ld register
    // layout - 16-byte pieces of 32 different frames per warp instruction - the same epilogue was 1.7x
    // slower.  Measured: about level with the compressed epilogue + pull at 2 and 4 ranks
    // (profiles/README.md section 1), not yet measured at 8 ranks, hence opt-in.
    m->fuse_scatter = m->fuse_compress && mode == 2;
  }
  m->layers.resize(nlayers);
  m->act.resize(nlayers + 1);
  m->dact.resize(nlayers + 1);
  int prio_lo = 0, prio_hi = 0;
  cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
  TRY_CUDA(cudaStreamCreateWithPriority(&m->wgrad_stream, cudaStreamNonBlocking, prio_lo));
  TRY_CUDA(cudaStreamCreateWithPriority(&m->comm_stream, cudaStreamNonBlocking, prio_hi));
  TRY_CUDA(cudaEventCreateWithFlags(&m->fwd_done, cudaEventDisableTiming));
  TRY_CUDA(cudaMalloc(&m->loss_sum, 256));
  for (int l = 0; l <= nlayers; ++l) {
    const size_t esz = m->f32 ? 4 : 2;
    TRY_CUDA(cudaMalloc(&m->act[l], (size_t)mb * sizes[l] * esz));
    if (l > 0) TRY_CUDA(cudaMalloc(&m->dact[l], (size_t)mb * sizes[l] * esz));
  }
  // Symmetric allocations, identical order on every rank: one arena per kind, a layer's slot is
  // its count rounded up to 1024 values (keeps every slot 4 KiB aligned and frame aligned; the pad
  // stays zero in gradients and weights, so a multi-layer bucket can run straight across it).
  size_t total = 0;
  bool can_fuse = m->fuse_compress;
  for (int l = 0; l < nlayers; ++l) {
    Layer& L = m->layers[l];
    L.in = sizes[l];
    L.out = sizes[l + 1];
    const size_t n = (size_t)L.in * L.out;
    L.slot = (n + 1023) & ~(size_t)1023;
    total += L.slot;
    if (n % 512 || L.in % 32) can_fuse = false;
  }
  m->fuse_compress = can_fuse;
  m->w_arena = static_cast<float*>(ainic_comm_alloc(comm, total * 4, 4096));
  m->g_arena = static_cast<float*>(ainic_comm_alloc(comm, total * 4, 4096));
  if (!m->w_arena || !m->g_arena) return AINIC_ERR_ARG;
  TRY_CUDA(cudaMemset(m->w_arena, 0, total * 4));
  TRY_CUDA(cudaMemset(m->g_arena, 0, total * 4));
  TRY_CUDA(cudaMalloc(&m->wb_arena, total * 2));
  TRY_CUDA(cudaMemset(m->wb_arena, 0, total * 2));
  const size_t frame_bytes = codec == AINIC_CODEC_BFP ? ainic_bfp_bytes(512) : ainic_mx_bytes(512);
  m->fuse_scatter = m->fuse_scatter && m->fuse_compress;
  // scatter mode: a bucket of F frames is laid out as nranks rows of ceil(F / nranks) frames, up to
  // nranks - 1 frames more than F; 8 spare frames per layer keep the next bucket's region clear
  const size_t gap = m->fuse_scatter ? AINIC_MAX_RANKS : 0;
  if (m->fuse_compress) {
    const size_t cg_frames = total / 512 + gap * nlayers;
    for (int i = 0; i < 2; ++i) {
      m->cg_arena[i] = static_cast<uint8_t*>(ainic_comm_alloc(comm, cg_frames * frame_bytes, 4096));
      if (!m->cg_arena[i]) return AINIC_ERR_ARG;
      TRY_CUDA(cudaMemset(m->cg_arena[i], 0, cg_frames * frame_bytes));
    }
  }
  {
    // default bucket size: 80 MiB as soon as there is a peer.  Each fused call pays three cross-GPU flag
    // exchanges plus the launch skew between ranks whatever its size; measured on the reference MLP
    // (ms/step, per-layer -> 80 MiB): 0.767 -> 0.700 at 2 ranks, 0.959 -> 0.784 at 8 ranks.
    const int n = ainic_comm_size(comm);
    // 2 ranks (one-shot kernel, one flag exchange per call) do better with 48 MiB - three layers of the reference
    // model per call - than with 80 (0.567 vs 0.578 ms/step); at 4 ranks 48 / 64 / 80 MiB measure 0.624 / 0.609 /
    // 0.596 and at 8 ranks every schedule with more than two calls was slower (profiles/README.md R2.1, R2.9).
    m->bucket_bytes = n == 2 ? (size_t)48 << 20 : n > 2 ? (size_t)80 << 20 : 0;
    if (const char* e = getenv("AINIC_BUCKET_MB")) m->bucket_bytes = (size_t)(atof(e) * 1048576.0);
    if (const char* e = getenv("AINIC_BUCKET_LAYERS")) {   // e.g. "4,3,2,1": big buckets first, a small exposed tail
      for (const char* p = e; *p;) {
        char* end = nullptr;
        const long v = strtol(p, &end, 10);
        if (end == p) break;
        if (v > 0) m->bucket_layers.push_back((int)v);
        p = *end ? end + 1 : end;
      }
    }
  }
  {
    // Flow mode (AINIC_FLOW=1): opt-in.  Correct at every world size (bit-identical to one call per bucket), but
    // measured slower: 2 ranks 1.03-1.08 ms/step against 0.56-0.59 with one fused call per 80 MiB bucket.  The
    // per-layer time line (AINIC_FLOW_TRACE) shows why: a reduce warp needs ~58 us for its ~7 frames of a layer
    // and ~15 us more for the ticket / flag fences, against one layer of GEMMs every ~33 us, so the kernel falls
    // behind from the second layer on (profiles/README.md, section "flow mode").
    const int n = ainic_comm_size(comm);
    (void)n;
    int want = 0;
    if (const char* e = getenv("AINIC_FLOW")) want = atoi(e);
    // the flow kernel takes wire-format input: fp32, or frames written by the weight-gradient epilogue
    const bool wire_ok = codec == AINIC_CODEC_F32 || m->fuse_compress;
    m->flow = want != 0 && nlayers <= AINIC_FLOW_MAX_CHUNKS && !m->fuse_scatter && wire_ok && !m->f32;
  }
  if (m->flow) {
    TRY(ainic_comm_reserve_flow(comm, total, nlayers));
    TRY_CUDA(cudaMalloc(&m->flow_ready, (size_t)nlayers * 4));
    TRY_CUDA(cudaMalloc(&m->flow_tickets, (size_t)nlayers * 8));
    TRY_CUDA(cudaMemset(m->flow_ready, 0, (size_t)nlayers * 4));
    TRY_CUDA(cudaMemset(m->flow_tickets, 0, (size_t)nlayers * 8));
    TRY_CUDA(cudaEventCreateWithFlags(&m->bwd_start, cudaEventDisableTiming));
  } else {
    TRY(ainic_comm_reserve(comm, max_bucket_count(m)));
  }
  // Graph replay is on by default (AINIC_MLP_GRAPH=0 switches it off).  It needs launch arguments that do
  // not change from step to step: the fused all-reduce keeps its call number (the cross-GPU flag value)
  // in device memory for that reason, and the compressed-gradient double buffer is part of the graph key.
  m->use_graph = true;
  if (const char* e = getenv("AINIC_MLP_GRAPH")) m->use_graph = atoi(e) != 0;
  TRY_CUDA(cudaEventCreateWithFlags(&m->step_done, cudaEventDisableTiming));
  TRY_CUDA(cudaStreamCreateWithFlags(&m->cap_stream, cudaStreamNonBlocking));
  size_t off = 0;
  for (int l = 0; l < nlayers; ++l) {
    Layer& L = m->layers[l];
    L.w = m->w_arena + off;
    L.g = m->g_arena + off;
    L.wb = m->wb_arena + off;
    if (m->fuse_compress) {
      L.cg[0] = m->cg_arena[0] + (off / 512 + gap * l) * frame_bytes;
      L.cg[1] = m->cg_arena[1] + (off / 512 + gap * l) * frame_bytes;
    }
    off += L.slot;
    TRY_CUDA(cudaMalloc(&L.bias, (size_t)L.out * 4));
    TRY_CUDA(cudaMalloc(&L.dbias, (size_t)L.out * 4));
    TRY_CUDA(cudaMemset(L.bias, 0, (size_t)L.out * 4));
    if (fuse >= 2) TRY_CUDA(cudaMalloc(&L.mask, (size_t)mb * L.out));
    TRY_CUDA(cudaEventCreateWithFlags(&L.dy_ready, cudaEventDisableTiming));
    TRY_CUDA(cudaEventCreateWithFlags(&L.dw_ready, cudaEventDisableTiming));
    TRY_CUDA(cudaEventCreateWithFlags(&L.updated, cudaEventDisableTiming));
    TRY_CUDA(cudaEventCreateWithFlags(&L.dx_done, cudaEventDisableTiming));
  }
  *out = m;
  return AINIC_OK;
}

// free_heap: give the weight / gradient arenas back to the communicator's symmetric heap.  Only safe when no
// peer can still be reading them, i.e. with one rank or behind a barrier (MLPTrainer.close does that); every rank
// must then make the same choice so that the heaps stay symmetric.
extern "C" int ainic_mlp_destroy_ex(ainic_mlp_t* m, int free_heap) {
  if (!m) return AINIC_ERR_ARG;
  cudaDeviceSynchronize();
  if (free_heap) {
    for (int i = 1; i >= 0; --i)
      if (m->cg_arena[i]) ainic_comm_free(m->comm, m->cg_arena[i]);
    if (m->g_arena) ainic_comm_free(m->comm, m->g_arena);
    if (m->w_arena) ainic_comm_free(m->comm, m->w_arena);
  }
  for (auto& L : m->layers) {
    cudaFree(L.bias); cudaFree(L.dbias); cudaFree(L.mask);
    cudaEventDestroy(L.dy_ready); cudaEventDestroy(L.dw_ready); cudaEventDestroy(L.updated); cudaEventDestroy(L.dx_done);
  }
  cudaFree(m->wb_arena);
  for (auto p : m->act) cudaFree(p);
  for (auto p : m->dact) cudaFree(p);
  cudaFree(m->loss_sum);
  cudaFree(m->flow_ready);
  cudaFree(m->flow_tickets);
  if (m->bwd_start) cudaEventDestroy(m->bwd_start);
  cudaEventDestroy(m->fwd_done);
  cudaEventDestroy(m->step_done);
  cudaStreamDestroy(m->cap_stream);
  for (auto& g : m->graphs) cudaGraphExecDestroy(g.exec);
  cudaStreamDestroy(m->wgrad_stream);
  cudaStreamDestroy(m->comm_stream);
  delete m;
  return AINIC_OK;
}
extern "C" int ainic_mlp_destroy(ainic_mlp_t* m) {
  return m ? ainic_mlp_destroy_ex(m, ainic_comm_size(m->comm) == 1) : AINIC_ERR_ARG;
}

// Raw buffer access for initialisation, checkpointing and tests.
// what: 0 weight fp32, 1 grad fp32, 2 shadow bf16, 3 bias fp32, 4 dbias fp32, 5 act bf16 (l in 0..L), 6 dact bf16 (1..L), 7 mask
extern "C" void* ainic_mlp_buffer(ainic_mlp_t* m, int what, int l) {
  if (!m || l < 0) return nullptr;
  if (what == 5) return l <= m->nlayers ? m->act[l] : nullptr;
  if (what == 6) return (l >= 1 && l <= m->nlayers) ? m->dact[l] : nullptr;
  if (l >= m->nlayers) return nullptr;
  Layer& L = m->layers[l];
  switch (what) {
    case 0: return L.w;
    case 1: return L.g;
    case 2: return L.wb;
    case 3: return L.bias;
    case 4: return L.dbias;
    case 7: return L.mask;
  }
  return nullptr;
}

// Refresh every bf16 shadow from the fp32 masters (after init / checkpoint load).
extern "C" int ainic_mlp_sync_shadows(ainic_mlp_t* m, void* stream) {
  if (!m) return AINIC_ERR_ARG;
  if (!m->f32)
    for (auto& L : m->layers) TRY(ainic_cast_f32_bf16(L.w, L.wb, (size_t)L.in * L.out, stream));
  m->first_step = true;
  return AINIC_OK;
}

extern "C" unsigned long long ainic_mlp_launches(ainic_mlp_t* m) { return m ? m->launches : 0; }
// Largest bucket (in values, arena pads included) the current bucket size produces.
// Bucket plan (csrc/api/bucket_plan.h): blo[l] / bhi[l] = lowest / highest layer of l's bucket.
static void plan_buckets(const ainic_mlp* m, std::vector<int>& blo, std::vector<int>& bhi) {
  std::vector<size_t> bytes(m->nlayers);
  for (int l = 0; l < m->nlayers; ++l) bytes[l] = (size_t)m->layers[l].in * m->layers[l].out * 4;
  ainic::plan_buckets(bytes, m->bucket_bytes, m->bucket_layers, blo, bhi);
}
static size_t bucket_values(const ainic_mlp* m, int lo, int hi) {   // arena pads included, last layer exact
  size_t count = 0;
  for (int k = lo; k < hi; ++k) count += m->layers[k].slot;
  return count + (size_t)m->layers[hi].in * m->layers[hi].out;
}
static size_t max_bucket_count(const ainic_mlp* m) {
  std::vector<int> blo, bhi;
  plan_buckets(m, blo, bhi);
  size_t best = 0;
  for (int l = 0; l < m->nlayers; ++l)
    if (blo[l] == l) best = std::max(best, bucket_values(m, l, bhi[l]));
  return best;
}
// Gradient bucket size in fp32 bytes (0 = one all-reduce per layer).  Takes effect at the next step;
// collective (every rank must set the same value at the same point).
extern "C" int ainic_mlp_set_bucket_bytes(ainic_mlp_t* m, size_t bytes) {
  if (!m) return AINIC_ERR_ARG;
  m->bucket_bytes = bytes;
  return ainic_comm_reserve(m->comm, max_bucket_count(m));
}
extern "C" size_t ainic_mlp_bucket_bytes(ainic_mlp_t* m) { return m ? m->bucket_bytes : 0; }
extern "C" void* ainic_mlp_loss_ptr(ainic_mlp_t* m) { return m ? m->loss_sum : nullptr; }

static int forward(ainic_mlp* m, cudaStream_t s, bool wait_updates) {
  const bool use_bias = (m->fuse == 1 || m->fuse == 3), relu = (m->fuse >= 2);
  for (int l = 0; l < m->nlayers; ++l) {
    Layer& L = m->layers[l];
    if (wait_updates && !m->first_step && !m->events_stale) TRY_CUDA(cudaStreamWaitEvent(s, L.updated, 0));
    ainic_gemm_epilogue_t epi{};
    epi.bias = use_bias ? L.bias : nullptr;
    epi.relu = relu ? 1 : 0;
    epi.mask_out = relu ? L.mask : nullptr;
    epi.alpha = 1.0f;
    if (m->f32)   // fp32 activations x fp32 master weights on kind::tf32 (libxsmm_dnn_fc_fwd_exec_f32, :708-709)
      TRY(ainic_gemm_tf32(AINIC_GEMM_NT, (const float*)m->act[l], L.in, L.w, L.in, (float*)m->act[l + 1], L.out, m->mb, L.out, L.in,
                          nullptr, 0, m->gemm_ctas, s));
    else
      TRY(ainic_gemm_bf16(AINIC_GEMM_NT, m->act[l], L.in, L.wb, L.in, m->act[l + 1], L.out, m->mb, L.out, L.in, &epi, 0,
                          m->gemm_ctas, s));
    m->launches++;
  }
  return AINIC_OK;
}

// Forward only: logits (bf16 [mb][C[L]]) stay in act[L].  x is copied/used in place as act[0].
extern "C" int ainic_mlp_forward(ainic_mlp_t* m, const void* x_bf16, void* stream) {
  if (!m || !x_bf16) return AINIC_ERR_ARG;
  cudaStream_t s = (cudaStream_t)stream;
  if (m->events_stale) TRY_CUDA(cudaStreamWaitEvent(s, m->step_done, 0));
  if (x_bf16 != m->act[0])
    TRY_CUDA(cudaMemcpyAsync(m->act[0], x_bf16, (size_t)m->mb * m->layers[0].in * (m->f32 ? 4 : 2), cudaMemcpyDeviceToDevice, s));
  return forward(m, s, true);
}

// One full training iteration.  loss_out (device, nullable) receives the mean cross-entropy of this
// rank's minibatch.  Returns after enqueueing; completion is stream/event ordered.
static int step_body(ainic_mlp* m, const void* x_bf16, const int32_t* labels, float* loss_out, cudaStream_t s) {
  const int Ln = m->nlayers;
  const bool relu = (m->fuse >= 2), use_bias = (m->fuse == 1 || m->fuse == 3);
  if (m->events_stale && !m->in_capture) {
    // the previous step was a graph replay on some stream: order after it, then behave like a first step
    TRY_CUDA(cudaStreamWaitEvent(s, m->step_done, 0));
    m->first_step = true;
    m->events_stale = false;
  }
  // act[0] is still being read by the previous step's last weight-gradient GEMM (wgrad stream is
  // in order, so layer 0's event covers all of them).
  if (!m->first_step) TRY_CUDA(cudaStreamWaitEvent(s, m->layers[0].dw_ready, 0));
  if (x_bf16 != m->act[0])
    TRY_CUDA(cudaMemcpyAsync(m->act[0], x_bf16, (size_t)m->mb * m->layers[0].in * (m->f32 ? 4 : 2), cudaMemcpyDeviceToDevice, s));
  if (m->profile) TRY_CUDA(cudaEventRecord(m->prof[0], s));
  TRY(forward(m, s, true));
  if (m->profile) TRY_CUDA(cudaEventRecord(m->prof[1], s));
  TRY_CUDA(cudaMemsetAsync(m->loss_sum, 0, 4, s));
  if (m->f32)
    TRY(ainic_softmax_xent_f32((const float*)m->act[Ln], labels, (float*)m->dact[Ln], m->loss_sum, m->mb, m->layers[Ln - 1].out,
                               m->layers[Ln - 1].out, m->loss_weight, s));
  else
    TRY(ainic_softmax_xent(m->act[Ln], labels, nullptr, m->dact[Ln], m->loss_sum, m->mb, m->layers[Ln - 1].out,
                           m->layers[Ln - 1].out, m->loss_weight, s));
  m->launches++;
  if (loss_out) TRY_CUDA(cudaMemcpyAsync(loss_out, m->loss_sum, 4, cudaMemcpyDeviceToDevice, s));
  if (relu) {  // the last layer's ReLU sits between the logits and the loss
    TRY(ainic_apply_mask_bf16(m->dact[Ln], m->layers[Ln - 1].mask, (size_t)m->mb * m->layers[Ln - 1].out, s));
    m->launches++;
  }
  if (m->profile) TRY_CUDA(cudaEventRecord(m->prof[2], s));
  std::vector<int> blo, bhi;
  plan_buckets(m, blo, bhi);
  auto bucket_count = [&](int lo, int hi) { return bucket_values(m, lo, hi); };
  const int nranks = ainic_comm_size(m->comm), rank = ainic_comm_rank(m->comm);
  if (m->flow) {
    // One persistent all-reduce for the whole backward pass, launched BEFORE the GEMMs that feed it: it
    // sits beside the GEMM CTAs (256 threads, no shared memory) and picks up layer after layer as their
    // ready counters fill.  The counters are cleared first; every producer of this step is ordered
    // after that memset (dy_ready events / stream order), the kernel itself through bwd_start.
    TRY_CUDA(cudaMemsetAsync(m->flow_ready, 0, (size_t)Ln * 4, s));
    TRY_CUDA(cudaEventRecord(m->bwd_start, s));
    TRY_CUDA(cudaStreamWaitEvent(m->comm_stream, m->bwd_start, 0));
    ainic_flow_chunk_t chunks[AINIC_FLOW_MAX_CHUNKS];
    size_t total = 0;
    for (int l = 0; l < Ln; ++l) total += m->layers[l].slot;
    for (int l = Ln - 1, k = 0; l >= 0; --l, ++k) {
      const Layer& L = m->layers[l];
      chunks[k].offset = (size_t)(L.g - m->g_arena);
      chunks[k].count = (((size_t)L.in * L.out) + 511) & ~(size_t)511;
      chunks[k].ready_index = l;
      chunks[k].ready_need = l > 0 ? 2 : 1;   // dW_l, and dX_l which reads the shadow this call rewrites
    }
    ainic_allreduce_args_t a{};
    a.grad = m->g_arena;
    a.weight = m->w_arena;
    a.weight_bf16 = m->wb_arena;
    a.count = total;
    a.lr = m->lr;
    a.codec = m->codec;
    a.bfp_mode = m->bfp_mode;
    a.max_ctas = m->comm_ctas;
    a.grad_compressed = m->fuse_compress ? m->cg_arena[m->step_idx & 1] : nullptr;
    TRY(ainic_allreduce_flow(m->comm, &a, chunks, Ln, m->flow_ready, m->comm_stream));
    m->launches++;
    for (int k = 0; k < Ln; ++k) TRY_CUDA(cudaEventRecord(m->layers[k].updated, m->comm_stream));
  }
  for (int l = Ln - 1; l >= 0; --l) {
    Layer& L = m->layers[l];
    // dact[l+1] is final here (for ReLU layers the mask was already applied by the producer below)
    TRY_CUDA(cudaEventRecord(L.dy_ready, s));
    // --- weight gradient on the wgrad stream: dW = dY^T . X -> flat fp32 bucket
    TRY_CUDA(cudaStreamWaitEvent(m->wgrad_stream, L.dy_ready, 0));
    ainic_gemm_epilogue_t we{};
    we.alpha = 1.0f;
    if (m->flow) {
      we.done_flag = m->flow_ready + l;
      we.done_ticket = m->flow_tickets + 2 * l;
    }
    uint8_t* cg = L.cg[m->step_idx & 1];   // null unless the epilogue emits the wire format
    ainic_gemm_scatter_t sc{};
    if (cg) {
      we.cgrad = cg;
      we.codec = m->codec;
      we.bfp_mode = m->bfp_mode;
      if (m->fuse_scatter) {
        // GEMM -> reduce-scatter: frame f of the bucket goes to rank f / shard, slot [rank][f % shard]
        const int lo = blo[l];
        uint8_t* inbox = m->layers[lo].cg[m->step_idx & 1];
        const size_t frames = (bucket_count(lo, bhi[l]) + 511) / 512;
        sc.nranks = nranks;
        sc.rank = rank;
        sc.shard_frames = (unsigned)((frames + nranks - 1) / nranks);
        sc.frame0 = (unsigned)((L.g - m->layers[lo].g) / 512);
        for (int r = 0; r < nranks; ++r) sc.inbox[r] = r == rank ? inbox : ainic_comm_peer_ptr(m->comm, inbox, r);
        we.scatter = &sc;
      }
    }
    if (m->f32)
      TRY(ainic_gemm_tf32(AINIC_GEMM_TN, (const float*)m->dact[l + 1], L.out, (const float*)m->act[l], L.in, L.g, L.in, L.out, L.in,
                          m->mb, nullptr, 0, m->wgrad_ctas, m->wgrad_stream));
    else
      TRY(ainic_gemm_bf16(AINIC_GEMM_TN, m->dact[l + 1], L.out, m->act[l], L.in, cg ? nullptr : L.g, L.in, L.out, L.in,
                          m->mb, &we, 0, m->wgrad_ctas, m->wgrad_stream));
    m->launches++;
    if (use_bias) {
      TRY(ainic_bias_grad(m->dact[l + 1], L.dbias, m->mb, L.out, L.out, m->wgrad_stream));
      m->launches++;
    }
    TRY_CUDA(cudaEventRecord(L.dw_ready, m->wgrad_stream));
    // --- data gradient on the main stream (critical path): dX = dY . W, masked by the previous ReLU
    if (l > 0) {
      ainic_gemm_epilogue_t de{};
      de.alpha = 1.0f;
      de.mask_in = relu ? m->layers[l - 1].mask : nullptr;
      if (m->flow) {
        de.done_flag = m->flow_ready + l;
        de.done_ticket = m->flow_tickets + 2 * l + 1;
      }
      if (m->f32)
        TRY(ainic_gemm_tf32(AINIC_GEMM_NN, (const float*)m->dact[l + 1], L.out, L.w, L.in, (float*)m->dact[l], L.in, m->mb, L.in,
                            L.out, nullptr, 0, m->gemm_ctas, s));
      else
        TRY(ainic_gemm_bf16(AINIC_GEMM_NN, m->dact[l + 1], L.out, L.wb, L.in, m->dact[l], L.in, m->mb, L.in, L.out, &de, 0,
                            m->gemm_ctas, s));
      m->launches++;
      TRY_CUDA(cudaEventRecord(L.dx_done, s));
    }
    // --- fused all-reduce + SGD on the comm stream, once the bucket that ends at layer l is full
    if (m->flow || blo[l] != l) continue;
    const int lo = l, hi = bhi[l];                      // bucket = layers lo .. hi, contiguous in the arenas
    TRY_CUDA(cudaStreamWaitEvent(m->comm_stream, L.dw_ready, 0));   // wgrad stream is in order: covers lo .. hi
    // the shadows of lo .. hi are rewritten below; the last backward-data GEMM that reads one of them
    // is layer max(lo,1)'s (main stream is in order)
    if (hi > 0) TRY_CUDA(cudaStreamWaitEvent(m->comm_stream, m->layers[lo > 0 ? lo : 1].dx_done, 0));
    const size_t count = bucket_count(lo, hi);
    ainic_allreduce_args_t a{};
    a.grad = L.g;
    a.weight = L.w;
    a.weight_bf16 = m->f32 ? nullptr : L.wb;
    a.count = count;
    a.lr = m->lr;
    a.codec = m->codec;
    a.bfp_mode = m->bfp_mode;
    a.algo = m->algo;
    a.max_ctas = m->comm_ctas;
    a.grad_compressed = cg;
    a.grad_scattered = m->fuse_scatter ? 1 : 0;
    TRY(ainic_allreduce_fused(m->comm, &a, m->comm_stream));
    m->launches++;
    for (int k = lo; k <= hi; ++k) TRY_CUDA(cudaEventRecord(m->layers[k].updated, m->comm_stream));
  }
  if (m->in_capture) {
    // join the forked streams: the graph ends when the last update has landed (every weight-gradient
    // GEMM precedes a dw_ready that some bucket's all-reduce waited for)
    for (auto& L : m->layers) TRY_CUDA(cudaStreamWaitEvent(s, L.updated, 0));
    if (m->flow) TRY_CUDA(cudaStreamWaitEvent(s, m->layers[0].dw_ready, 0));   // the wgrad stream joins here in flow mode
    return AINIC_OK;
  }
  m->first_step = false;
  m->step_idx++;
  if (m->profile) {
    // profiling mode serialises the step: backward end on the main stream, then the last update
    TRY_CUDA(cudaEventRecord(m->prof[3], s));
    for (auto& L : m->layers) TRY_CUDA(cudaStreamWaitEvent(s, L.updated, 0));
    TRY_CUDA(cudaEventRecord(m->prof[4], s));
    TRY_CUDA(cudaEventSynchronize(m->prof[4]));
    for (int i = 0; i < 4; ++i) {
      float ms = 0.f;
      cudaEventElapsedTime(&ms, m->prof[i], m->prof[i + 1]);
      m->phase_ms[i] += ms;
    }
    m->prof_steps++;
  }
  return AINIC_OK;
}

// One full training iteration (public entry).  The iteration is captured into a CUDA graph the first
// time a given (input, labels, loss) tuple is seen and replayed afterwards: the streams,
// events and kernels of step_body become graph nodes and edges, and a step costs one launch.  The
// graph of step i + 1 starts when step i's last update has landed - no overlap is lost, because the
// first thing a step needs (layer 0's weights) is the last thing the previous step produces.
extern "C" int ainic_mlp_step(ainic_mlp_t* m, const void* x_bf16, const int32_t* labels, float* loss_out, void* stream) {
  if (!m || !x_bf16 || !labels) return AINIC_ERR_ARG;
  cudaStream_t s = (cudaStream_t)stream;
  if (!m->use_graph || m->profile) return step_body(m, x_bf16, labels, loss_out, s);
  const unsigned parity = (unsigned)(m->step_idx & 1);
  ainic_mlp::StepGraph* g = nullptr;
  for (auto& e : m->graphs)
    if (e.x == x_bf16 && e.labels == labels && e.loss_out == loss_out && e.parity == parity && e.bucket_bytes == m->bucket_bytes) g = &e;
  if (!g) {
    if (m->graphs.size() >= 32) {  // callers that never reuse their buffers get no benefit: start over
      for (auto& e : m->graphs) cudaGraphExecDestroy(e.exec);
      m->graphs.clear();
    }
    if (!m->first_step && !m->events_stale) {  // a direct step is still in flight on other streams
      for (auto& L : m->layers) TRY_CUDA(cudaStreamWaitEvent(s, L.updated, 0));
    }
    const unsigned long long before = m->launches;
    const bool was_first = m->first_step;
    cudaGraph_t graph = nullptr;
    TRY_CUDA(cudaStreamBeginCapture(m->cap_stream, cudaStreamCaptureModeThreadLocal));
    m->in_capture = true;
    m->first_step = true;            // no waits on events recorded outside the capture
    const int rc = step_body(m, x_bf16, labels, loss_out, m->cap_stream);
    m->in_capture = false;
    m->first_step = was_first;
    const cudaError_t ce = cudaStreamEndCapture(m->cap_stream, &graph);
    const unsigned long long kernels = m->launches - before;
    m->launches = before;
    if (rc || ce != cudaSuccess || !graph) {
      if (graph) cudaGraphDestroy(graph);
      cudaGetLastError();
      fprintf(stderr, "ainic_mlp: graph capture failed (%s); falling back to direct launches\n", cudaGetErrorString(ce));
      m->use_graph = false;
      return step_body(m, x_bf16, labels, loss_out, s);
    }
    cudaGraphExec_t exec = nullptr;
    const cudaError_t ie = cudaGraphInstantiate(&exec, graph, 0);
    cudaGraphDestroy(graph);
    if (ie != cudaSuccess) {
      cudaGetLastError();
      m->use_graph = false;
      return step_body(m, x_bf16, labels, loss_out, s);
    }
    m->graphs.push_back({x_bf16, labels, loss_out, parity, m->bucket_bytes, exec, kernels});
    g = &m->graphs.back();
  }
  TRY_CUDA(cudaGraphLaunch(g->exec, s));
  TRY_CUDA(cudaEventRecord(m->step_done, s));
  m->launches += g->kernels;
  m->events_stale = true;
  m->first_step = false;
  m->step_idx++;
  return AINIC_OK;
}

// Per-phase device times accumulated since profiling was switched on (ms): [0] forward FC,
// [1] loss fwd+bwd, [2] backward + weight-gradient FC, [3] tail until the last weight update landed;
// returns the number of profiled steps.  enable != 0 switches the mode on (and clears), 0 off.
extern "C" unsigned long long ainic_mlp_profile(ainic_mlp_t* m, int enable, double out_ms[4]) {
  if (!m) return 0;
  const unsigned long long n = m->prof_steps;
  if (out_ms) for (int i = 0; i < 4; ++i) out_ms[i] = m->phase_ms[i];
  if (enable >= 0) {
    if (enable && !m->prof[0])
      for (auto& e : m->prof) cudaEventCreate(&e);
    m->profile = enable != 0;
    if (enable) { for (auto& p : m->phase_ms) p = 0; m->prof_steps = 0; }
  }
  return n;
}

// Block the given stream until every outstanding weight update has landed (the reference's
// wait(request[1]); wait(request[0]) at the end of an iteration, :775-787).
extern "C" int ainic_mlp_wait_updates(ainic_mlp_t* m, void* stream) {
  if (!m) return AINIC_ERR_ARG;
  if (m->first_step) return AINIC_OK;
  if (m->events_stale) {
    TRY_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, m->step_done, 0));
    return AINIC_OK;
  }
  for (auto& L : m->layers) TRY_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, L.updated, 0));
  return AINIC_OK;
}
