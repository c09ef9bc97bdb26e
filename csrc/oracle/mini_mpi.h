// mini-MPI: ranks are threads of one process.  The reference uses MPI only for bootstrap, data
// distribution and timing barriers (MPI_Init_thread, Comm_rank/size, Barrier, Scatter, Bcast;
// sw/mlp_mpi_example_f32.cpp:195-199, :299-300, :309, :453-470) plus the commented-out software
// baseline MPI_Iallreduce(IN_PLACE, ..., MPI_SUM) + MPI_Wait (:615-620).  No MPI exists on this box,
// so the restated application and the CPU baseline run on this shim.
#pragma once
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mmpi_world mmpi_world_t;
typedef struct mmpi_request { int active; } mmpi_request_t;
typedef void (*mmpi_main_fn)(mmpi_world_t* world, int rank, void* user);

// Runs fn on nranks threads and joins them.  Returns 0.
int mmpi_run(int nranks, mmpi_main_fn fn, void* user);
int mmpi_size(mmpi_world_t* w);
void mmpi_barrier(mmpi_world_t* w, int rank);
// root's buf (count floats) is copied into every other rank's buf.
void mmpi_bcast(mmpi_world_t* w, int rank, float* buf, size_t count, int root);
// root's send (nranks*count floats) is split; every rank receives its slice in recv.
void mmpi_scatter(mmpi_world_t* w, int rank, const float* send, float* recv, size_t count, int root);
// In-place sum over ranks in rank order (deterministic).  Blocking part happens in wait.
void mmpi_iallreduce_sum(mmpi_world_t* w, int rank, float* buf, size_t count, mmpi_request_t* req);
void mmpi_wait(mmpi_world_t* w, int rank, mmpi_request_t* req);
double mmpi_wtime(void);

#ifdef __cplusplus
}
#endif
