/*
 * mpi.h — single-node MPI shim for the reference baselines (NOT a product path).
 *
 * Neither this box nor the GPU box has an MPI implementation and none can be
 * installed, but three of the four reference programs (mpi_heat2Dn.c,
 * grad1612_mpi_heat.c, grad1612_hybrid_heat.c) are MPI programs.  This header
 * plus mpishim.c implement exactly the subset of MPI those sources use
 * (SURVEY 2.4): blocking and persistent point-to-point, Cartesian topology,
 * contiguous / vector / subarray datatypes, Bcast / Barrier / Reduce /
 * Allreduce, MPI-IO write_all through a subarray file view, Wtime, Abort.
 * Ranks are processes started by the `mpiexec` of this directory and connected
 * by a full mesh of Unix socket pairs.  Numbers measured on it are labelled
 * "reference source on shim-MPI", never "OpenMPI".
 */
#ifndef MPISHIM_MPI_H_
#define MPISHIM_MPI_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef int MPI_Info;
typedef int MPI_Request;
typedef struct mpishim_file* MPI_File;
typedef long long MPI_Offset;

typedef struct MPI_Status {
  int MPI_SOURCE;
  int MPI_TAG;
  int MPI_ERROR;
  int count_bytes;
} MPI_Status;

#define MPI_SUCCESS 0
#define MPI_COMM_WORLD 0
#define MPI_COMM_NULL (-1)
#define MPI_PROC_NULL (-2)
#define MPI_ANY_SOURCE (-3)
#define MPI_ANY_TAG (-4)
#define MPI_INFO_NULL 0
#define MPI_REQUEST_NULL (-1)
#define MPI_STATUS_IGNORE ((MPI_Status*)0)
#define MPI_STATUSES_IGNORE ((MPI_Status*)0)
#define MPI_MAX_PROCESSOR_NAME 256

/* basic datatypes (ids < 16 are predefined) */
#define MPI_CHAR 1
#define MPI_INT 2
#define MPI_FLOAT 3
#define MPI_DOUBLE 4
#define MPI_BYTE 5

#define MPI_SUM 1
#define MPI_MAX 2
#define MPI_MIN 3

#define MPI_ORDER_C 0
#define MPI_MODE_CREATE 1
#define MPI_MODE_RDONLY 2
#define MPI_MODE_WRONLY 4
#define MPI_MODE_RDWR 8

int MPI_Init(int* argc, char*** argv);
int MPI_Finalize(void);
int MPI_Abort(MPI_Comm comm, int errorcode);
int MPI_Comm_size(MPI_Comm comm, int* size);
int MPI_Comm_rank(MPI_Comm comm, int* rank);
double MPI_Wtime(void);
int MPI_Get_processor_name(char* name, int* resultlen);

int MPI_Send(const void* buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm);
int MPI_Recv(void* buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm comm,
             MPI_Status* status);
int MPI_Send_init(const void* buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm,
                  MPI_Request* request);
int MPI_Recv_init(void* buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm comm,
                  MPI_Request* request);
int MPI_Startall(int count, MPI_Request* requests);
int MPI_Waitall(int count, MPI_Request* requests, MPI_Status* statuses);
int MPI_Request_free(MPI_Request* request);

int MPI_Barrier(MPI_Comm comm);
int MPI_Bcast(void* buf, int count, MPI_Datatype type, int root, MPI_Comm comm);
int MPI_Reduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype type, MPI_Op op, int root,
               MPI_Comm comm);
int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype type, MPI_Op op,
                  MPI_Comm comm);

int MPI_Cart_create(MPI_Comm comm, int ndims, const int* dims, const int* periods, int reorder,
                    MPI_Comm* newcomm);
int MPI_Cart_shift(MPI_Comm comm, int direction, int disp, int* rank_source, int* rank_dest);

int MPI_Type_contiguous(int count, MPI_Datatype oldtype, MPI_Datatype* newtype);
int MPI_Type_vector(int count, int blocklength, int stride, MPI_Datatype oldtype, MPI_Datatype* newtype);
int MPI_Type_create_subarray(int ndims, const int* sizes, const int* subsizes, const int* starts,
                             int order, MPI_Datatype oldtype, MPI_Datatype* newtype);
int MPI_Type_commit(MPI_Datatype* type);
int MPI_Type_free(MPI_Datatype* type);

int MPI_File_open(MPI_Comm comm, const char* filename, int amode, MPI_Info info, MPI_File* fh);
int MPI_File_set_view(MPI_File fh, MPI_Offset disp, MPI_Datatype etype, MPI_Datatype filetype,
                      const char* datarep, MPI_Info info);
int MPI_File_write_all(MPI_File fh, const void* buf, int count, MPI_Datatype type, MPI_Status* status);
int MPI_File_close(MPI_File* fh);

#ifdef __cplusplus
}
#endif
#endif /* MPISHIM_MPI_H_ */
