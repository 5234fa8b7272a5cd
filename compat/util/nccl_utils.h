// Stand-in for the reference's src/util/nccl_utils.h (:29-142) WITHOUT MPI (this image has none):
// init_mpi_nccl / init_nccl_comm bootstrap from the environment (RANK, WORLD_SIZE, B200_NCCL_ID_FILE:
// rank 0 writes the 128-byte ncclUniqueId to that file, the others poll for it), ring_exchange* forward
// to the library.  A minimal `mpi.h` replacement for the host-side MPI_Bcast/Gather calls of the ring
// driver lives in compat/mpi.h.
#ifndef NCCL_UTILS_H
#define NCCL_UTILS_H
#include <nccl.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <unistd.h>
#include "mpi.h"
#include "cuda_helper.h"
#include "../../include/b200attn.h"

#define CHECK_NCCL(call) do { ncclResult_t res__ = (call); if (res__ != ncclSuccess) { \
    fprintf(stderr, "NCCL error at %s:%d - %s\n", __FILE__, __LINE__, ncclGetErrorString(res__)); exit(1); } } while (0)

inline int init_nccl_comm(ncclComm_t* comm, int rank, int nranks, MPI_Comm = MPI_COMM_WORLD) {
    unsigned char id[128];
    if (MPI_Bcast_id(id, rank) != 0) return -1;
    return b200_ring_comm_init(comm, id, rank, nranks) == 0 ? 0 : -1;
}
inline int init_mpi_nccl(int argc, char* argv[], int* rank, int* nranks, ncclComm_t* comm) {
    MPI_Init(&argc, &argv);
    MPI_Comm_rank(MPI_COMM_WORLD, rank);
    MPI_Comm_size(MPI_COMM_WORLD, nranks);
    int ndev = 0;
    CHECK_CUDA(cudaGetDeviceCount(&ndev));
    CHECK_CUDA(cudaSetDevice(*rank % ndev));
    if (*nranks == 1) { *comm = nullptr; return 0; }
    return init_nccl_comm(comm, *rank, *nranks);
}
inline void cleanup_mpi_nccl(ncclComm_t comm) { b200_ring_comm_destroy(comm); MPI_Finalize(); }

inline void ring_exchange_kv(float* sk, float* rk, float* sv, float* rv, int size, int next_rank, int prev_rank,
                             ncclComm_t comm, cudaStream_t stream = 0) {
    if (b200_ring_exchange_kv(sk, rk, sv, rv, (size_t)size * sizeof(float), next_rank, prev_rank, comm, stream) != 0) exit(1);
}
inline void ring_exchange(float* s, float* r, int size, int next_rank, int prev_rank, ncclComm_t comm, cudaStream_t stream = 0) {
    CHECK_NCCL(ncclGroupStart());
    CHECK_NCCL(ncclSend(s, size, ncclFloat, next_rank, comm, stream));
    CHECK_NCCL(ncclRecv(r, size, ncclFloat, prev_rank, comm, stream));
    CHECK_NCCL(ncclGroupEnd());
}
#endif
