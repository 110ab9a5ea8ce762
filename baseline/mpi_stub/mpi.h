// Minimal MPI stand-in for building the reference's communicator without an MPI installation.
// Only what src/communication/mpi_nccl_communication.cu of the reference calls is declared.
// Transport: TCP sockets between the ranks of one job (see mpi_stub.cc); rank / size / rendezvous address
// come from the environment that torchrun or baseline/ref_arm.py sets (RANK, WORLD_SIZE, MASTER_ADDR,
// MASTER_PORT or MPISTUB_PORT).
#pragma once
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef long MPI_Comm;       // the reference's Python side stores it in a 64-bit ctypes integer
typedef int MPI_Datatype;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;
#define MPI_SUCCESS 0
#define MPI_COMM_WORLD ((MPI_Comm)1)
#define MPI_BYTE ((MPI_Datatype)1)
#define MPI_DATATYPE_NULL ((MPI_Datatype)0)
#define MPI_IN_PLACE ((void *)-1)
#define MPI_STATUS_IGNORE ((MPI_Status *)0)
int MPI_Init(int *argc, char ***argv);
int MPI_Finalize(void);
int MPI_Comm_rank(MPI_Comm comm, int *rank);
int MPI_Comm_size(MPI_Comm comm, int *size);
int MPI_Bcast(void *buf, int count, MPI_Datatype dt, int root, MPI_Comm comm);
int MPI_Allgather(const void *sendbuf, int sendcount, MPI_Datatype sendtype, void *recvbuf, int recvcount,
                  MPI_Datatype recvtype, MPI_Comm comm);
int MPI_Send(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm);
int MPI_Recv(void *buf, int count, MPI_Datatype dt, int source, int tag, MPI_Comm comm, MPI_Status *status);
int MPI_Barrier(MPI_Comm comm);
#ifdef __cplusplus
}
#endif
