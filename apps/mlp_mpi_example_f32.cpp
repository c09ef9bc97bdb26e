// CPU restatement of the reference training application (sw/mlp_mpi_example_f32.cpp:191-881).
//
// Same command line:   mlp_mpi_example_f32 iters MB fuse_type type bn bk bc C1 C2 ... CN
// (defaults iters=10 MB=32 fuse=0 type=A bn=bk=bc=64, :215-223; MB is the GLOBAL minibatch, :301).
// Ranks are threads of one process over the mini-MPI shim (no MPI on this box); the number of
// ranks comes from AINIC_RANKS (default 2, BASELINE.json config 1: "world_size=2 on CPU").
//   type A  the live path of the reference (:686-790): forward, softmax, per-layer backward with
//           the gradient all-reduce + SGD handed to the NIC (here: the host fake NIC) and overlapped
//           with the next layer's backward through two in-flight requests (request[i%2]).
//   type B  the software baseline the NIC replaces (commented out in the reference, :590-680):
//           MPI_Iallreduce(IN_PLACE, SUM) + MPI_Wait + SGD on the host.
//   type F  forward only (:543-588).
// Output format follows :794-816 (GFLOP / fp time / GFLOPS / PERFDUMP line / profile breakdown).
// This binary is the oracle at application level and the timed CPU baseline.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "../csrc/oracle/mini_mpi.h"
#include "../csrc/oracle/oracle.h"
#include "../include/ainic.h"

struct Config {
  int iters = 10, global_mb = 32, fuse = 0;
  char type = 'A';
  int bn = 64, bk = 64, bc = 64;  // libxsmm blocking factors: accepted, irrelevant to row-major buffers
  std::vector<int> C;
  int nranks = 2;
  float lr = 0.1f, loss_weight = 0.1f;
  std::string codec = "bfp", algo = "auto";
};

struct Shared {
  Config cfg;
  std::vector<ainic_nic_t*> nics;
  std::vector<float> act0_global, labels_dummy;
  double seconds = 0, t_fwd = 0, t_loss = 0, t_bwd = 0, t_first_bwd = 0;
  float final_loss = 0;
  double checksum = 0;
  // CHECK=1 (the reference parses the variable but only prints it, :227-228; its per-layer L1-norm check
  // via libxsmm_matdiff sits inside #if 0, :818-826): per-layer L1 norm and a bit hash of every rank's weights
  bool check = false;
  std::vector<std::vector<double>> l1;          // [rank][layer]
  std::vector<std::vector<uint64_t>> hash;      // [rank][layer]
};

static int codec_id(const std::string& s) {
  if (s == "f32") return AINIC_CODEC_F32;
  if (s == "mxfp8") return AINIC_CODEC_MXFP8;
  if (s == "bf16") return AINIC_CODEC_BF16;
  return AINIC_CODEC_BFP;
}
static int algo_id(const std::string& s) {
  if (s == "oneshot") return AINIC_ALGO_ONESHOT;
  if (s == "twoshot") return AINIC_ALGO_TWOSHOT;
  if (s == "ring") return AINIC_ALGO_RING;
  return AINIC_ALGO_AUTO;
}

static void rank_main(mmpi_world_t* world, int rank, void* user) {
  Shared& S = *static_cast<Shared*>(user);
  const Config& cfg = S.cfg;
  const int L = (int)cfg.C.size() - 1;
  const int n = mmpi_size(world);
  const int MB = cfg.global_mb / n;  // :301
  ainic_nic_t* nic = S.nics[rank];

  std::vector<std::vector<float>> act(L + 1), dact(L + 1), W(L), dW(L), bias(L), dbias(L);
  std::vector<std::vector<uint8_t>> mask(L);
  for (int l = 0; l <= L; ++l) {
    act[l].resize((size_t)MB * cfg.C[l]);
    dact[l].resize((size_t)MB * cfg.C[l]);
  }
  for (int l = 0; l < L; ++l) {
    const size_t cnt = (size_t)cfg.C[l] * cfg.C[l + 1];
    W[l].resize(cnt);
    dW[l].resize(cnt);
    bias[l].resize(cfg.C[l + 1]);
    dbias[l].resize(cfg.C[l + 1]);
    mask[l].resize((size_t)MB * cfg.C[l + 1]);
  }
  std::vector<int32_t> labels(MB, 0);  // never initialised in the reference (:412): zero pages
  std::vector<float> probs((size_t)MB * cfg.C[L]);

  // init on rank 0, then scatter activations / broadcast weights and biases (:415-470)
  if (rank == 0) {
    S.act0_global.resize((size_t)cfg.global_mb * cfg.C[0]);
    ainic_ref_init_buf(S.act0_global.data(), S.act0_global.size(), 1, 0, 0);
    for (int l = 0; l < L; ++l) {
      ainic_ref_init_buf(W[l].data(), W[l].size(), 1000 + l, 0, 0);
      ainic_ref_init_buf(bias[l].data(), bias[l].size(), 1500 + l, 0, 0);
    }
  }
  mmpi_barrier(world, rank);
  mmpi_scatter(world, rank, S.act0_global.data(), act[0].data(), (size_t)MB * cfg.C[0], 0);
  for (int l = 0; l < L; ++l) {
    mmpi_bcast(world, rank, W[l].data(), W[l].size(), 0);
    mmpi_bcast(world, rank, bias[l].data(), bias[l].size(), 0);
  }
  ainic_nic_setup(nic, 0, codec_id(cfg.codec), AINIC_BFP_CLEAN, algo_id(cfg.algo), cfg.lr);

  double t_fwd = 0, t_loss = 0, t_bwd = 0, t_first = 0;
  float loss = 0;
  mmpi_barrier(world, rank);
  const double t0 = mmpi_wtime();
  for (int it = 0; it < cfg.iters; ++it) {
    double t = mmpi_wtime();
    for (int l = 0; l < L; ++l)
      ainic_ref_fc_fwd(act[l].data(), W[l].data(), bias[l].data(), act[l + 1].data(), mask[l].data(), MB, cfg.C[l], cfg.C[l + 1], cfg.fuse);
    t_fwd += mmpi_wtime() - t;
    if (cfg.type == 'F') continue;
    t = mmpi_wtime();
    loss = ainic_ref_softmax_loss(act[L].data(), labels.data(), probs.data(), dact[L].data(), MB, cfg.C[L], cfg.loss_weight);
    t_loss += mmpi_wtime() - t;
    ainic_rqst_t rq[2];
    mmpi_request_t mreq[2] = {{0}, {0}};
    int flags = (it == 0) ? 1 : 0;  // load the initial weights once (:700, :754-755)
    for (int l = L - 1; l >= 0; --l) {
      t = mmpi_wtime();
      ainic_ref_fc_bwd(act[l].data(), W[l].data(), dact[l + 1].data(), mask[l].data(), l > 0 ? dact[l].data() : nullptr,
                       dW[l].data(), dbias[l].data(), MB, cfg.C[l], cfg.C[l + 1], cfg.fuse);
      const double dt = mmpi_wtime() - t;
      t_bwd += dt;
      if (l == L - 1) t_first += dt;
      if (cfg.type == 'A') {
        // all_reduce(delfil[i], fil[i], flags, ..., &request[i%2]); then wait for the previous layer (:753-764)
        ainic_nic_all_reduce(nic, dW[l].data(), W[l].data(), flags, dW[l].size(), &rq[l % 2], nullptr);
        flags = 0;
        if (l < L - 1) ainic_nic_wait(nic, &rq[(l + 1) % 2], 600.0);
      } else {
        // MPI_Iallreduce(MPI_IN_PLACE, delfil[i], ..., MPI_SUM) ... MPI_Wait ... opt_exec (:615-647)
        mmpi_iallreduce_sum(world, rank, dW[l].data(), dW[l].size(), &mreq[l % 2]);
        mmpi_wait(world, rank, &mreq[l % 2]);
        ainic_ref_sgd(W[l].data(), dW[l].data(), W[l].size(), cfg.lr);
      }
    }
    if (cfg.type == 'A') ainic_nic_wait(nic, &rq[0], 600.0);  // layer 0 is the last request (:775-787)
  }
  mmpi_barrier(world, rank);
  const double t1 = mmpi_wtime();
  if (S.check) {
    for (int l = 0; l < L; ++l) {
      double n1 = 0;
      uint64_t h = 1469598103934665603ull;   // FNV-1a over the weight bits
      for (float v : W[l]) {
        n1 += fabs((double)v);
        uint32_t u;
        memcpy(&u, &v, 4);
        h = (h ^ u) * 1099511628211ull;
      }
      S.l1[rank][l] = n1;
      S.hash[rank][l] = h;
    }
    mmpi_barrier(world, rank);
  }
  if (rank == 0) {
    S.seconds = t1 - t0;
    S.t_fwd = t_fwd; S.t_loss = t_loss; S.t_bwd = t_bwd; S.t_first_bwd = t_first;
    S.final_loss = loss;
    double cs = 0;
    for (int l = 0; l < L; ++l)
      for (size_t i = 0; i < W[l].size(); i += 97) cs += W[l][i];
    S.checksum = cs;
  }
}

int main(int argc, char** argv) {
  Shared S;
  Config& c = S.cfg;
  if (argc > 1 && !strcmp(argv[1], "-h")) {
    printf("Usage: %s iters MB fuse_type type bn bk bc C1 C2 ... CN\n", argv[0]);
    printf("  type: A (NIC all-reduce, overlapped), B (MPI_Iallreduce + SGD), F (forward only); fuse_type 0-3\n");
    printf("  env: AINIC_RANKS (2), AINIC_CODEC (bfp|f32|mxfp8|bf16), AINIC_ALGO (auto|oneshot|twoshot|ring), AINIC_LR (0.1),\n");
    printf("       CHECK=1 (per-layer L1 norms and a bit comparison of every rank's weights after the run)\n");
    return 0;
  }
  int i = 1;
  if (argc > i) c.iters = atoi(argv[i++]);
  if (argc > i) c.global_mb = atoi(argv[i++]);
  if (argc > i) c.fuse = atoi(argv[i++]);
  if (argc > i) c.type = argv[i++][0];
  if (argc > i) c.bn = atoi(argv[i++]);
  if (argc > i) c.bk = atoi(argv[i++]);
  if (argc > i) c.bc = atoi(argv[i++]);
  for (; i < argc; ++i) c.C.push_back(atoi(argv[i]));
  if (c.C.size() < 2) c.C = {256, 256, 256};
  if (const char* e = getenv("AINIC_RANKS")) c.nranks = atoi(e);
  if (const char* e = getenv("AINIC_CODEC")) c.codec = e;
  if (const char* e = getenv("AINIC_ALGO")) c.algo = e;
  if (const char* e = getenv("AINIC_LR")) c.lr = (float)atof(e);
  if (c.nranks < 1 || c.nranks > AINIC_MAX_RANKS || c.global_mb % c.nranks || c.fuse < 0 || c.fuse > 3 ||
      (c.type != 'A' && c.type != 'B' && c.type != 'F')) {
    fprintf(stderr, "bad arguments (ranks 1..8 dividing MB, fuse 0..3, type A|B|F)\n");
    return 1;
  }
  const int L = (int)c.C.size() - 1;
  printf("MLP: iters %d, global MB %d (%d ranks x %d), fuse %d, type %c, blocking %d/%d/%d (ignored), layers %d\n", c.iters,
         c.global_mb, c.nranks, c.global_mb / c.nranks, c.fuse, c.type, c.bn, c.bk, c.bc, L);
  if (const char* e = getenv("CHECK")) S.check = atoi(e) != 0;
  S.l1.assign(c.nranks, std::vector<double>(L, 0.0));
  S.hash.assign(c.nranks, std::vector<uint64_t>(L, 0));
  S.nics.resize(c.nranks);
  for (int r = 0; r < c.nranks; ++r)
    if (ainic_nic_open(&S.nics[r], nullptr, 1, r, c.nranks)) return 2;
  if (ainic_nic_fake_link(S.nics.data(), c.nranks)) return 2;
  mmpi_run(c.nranks, rank_main, &S);

  double gflop = 0.0;  // :794-798
  for (int l = L - 1; l > 0; --l) gflop += 6.0 * c.global_mb * (double)c.C[l] * c.C[l + 1] * c.iters / 1e9;
  gflop += 4.0 * c.global_mb * (double)c.C[0] * c.C[1] * c.iters / 1e9;
  printf("GFLOP  = %.5g\n", gflop / c.iters);
  printf("fp time = %.5g\n", S.seconds / c.iters);
  printf("GFLOPS  = %.5g\n", gflop / S.seconds);
  printf("PERFDUMP,BP,b200-ainic-cpu,%i,%i,", c.nranks, c.global_mb / c.nranks);
  for (int l = 0; l < L; ++l) printf("%i,", c.C[l]);
  printf("%f,%f\n", S.seconds / c.iters, gflop / S.seconds);
  printf("FC time compute/loss = %.5g\n", (S.t_fwd + S.t_bwd + S.t_loss) / c.iters);
  printf("Bwdupd compute FIRST time overlaped = %.5g\n", S.t_first_bwd / c.iters);
  printf("Bwdupd compute time overlaped = %.5g\n", (S.t_bwd - S.t_first_bwd) / c.iters);
  printf("samples/s = %.5g\n", (double)c.global_mb * c.iters / S.seconds);
  printf("loss = %.6f  weight checksum = %.9g\n", S.final_loss, S.checksum);
  if (S.check) {
    bool all_same = true;
    for (int l = 0; l < L; ++l) {
      bool same = true;
      for (int r = 1; r < c.nranks; ++r) same = same && S.hash[r][l] == S.hash[0][l];
      all_same = all_same && same;
      printf("CHECK layer %d: L1 norm of weights = %.9g, replicas %s\n", l, S.l1[0][l], same ? "identical" : "DIFFER");
    }
    printf("CHECK: weight replicas %s across %d ranks\n", all_same ? "bit-identical" : "differ", c.nranks);
  }
  printf("NIC latency (rank 0, ms) = %.3f  stall = %.3f\n", ainic_nic_latency_ns(S.nics[0]) / 1e6, ainic_nic_stall_ns(S.nics[0]) / 1e6);
  for (auto* n : S.nics) ainic_nic_close(n);
  return 0;
}
