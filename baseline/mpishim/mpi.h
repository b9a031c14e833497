/*
 * mpishim: a single-node, shared-memory implementation of exactly the MPI subset used by the
 * reference programs.  HARNESS ONLY - it exists so that the UNMODIFIED
 * reference sources game_mpi.c, game_mpi_async.c, game_mpi_collective.c and game_openmp.c can be
 * built and run as CPU baselines / cross-checks on a box without an MPI installation.  The product
 * does not use MPI.  Numbers obtained with it are labelled "reference source, shim MPI".
 */
#ifndef MPISHIM_MPI_H_
#define MPISHIM_MPI_H_

#ifdef __cplusplus
extern "C" {
#endif

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Request;
typedef int MPI_File;
typedef int MPI_Info;
typedef int MPI_Op;
typedef long long MPI_Offset;
typedef struct MPI_Status {
    int MPI_SOURCE, MPI_TAG, MPI_ERROR;
} MPI_Status;

#define MPI_SUCCESS 0
#define MPI_ERR_OTHER 15
#define MPI_COMM_NULL 0
#define MPI_COMM_WORLD 1
#define MPI_DATATYPE_NULL 0
#define MPI_CHAR 1
#define MPI_INT 2
#define MPI_SUM 1
#define MPI_INFO_NULL 0
#define MPI_ANY_TAG (-1)
#define MPI_ANY_SOURCE (-2)
#define MPI_ORDER_C 0
#define MPI_STATUS_IGNORE ((MPI_Status *)0)
#define MPI_STATUSES_IGNORE ((MPI_Status *)0)
#define MPI_REQUEST_NULL (-1)
#define MPI_MODE_RDONLY 2
#define MPI_MODE_WRONLY 4
#define MPI_MODE_RDWR 8
#define MPI_MODE_CREATE 1
#define MPI_MODE_EXCL 64

int MPI_Init(int *argc, char ***argv);
int MPI_Finalize(void);
int MPI_Comm_rank(MPI_Comm comm, int *rank);
int MPI_Comm_size(MPI_Comm comm, int *size);
double MPI_Wtime(void);
int MPI_Barrier(MPI_Comm comm);

int MPI_Cart_create(MPI_Comm old, int ndims, const int dims[], const int periods[], int reorder, MPI_Comm *cart);
int MPI_Cart_coords(MPI_Comm comm, int rank, int maxdims, int coords[]);
int MPI_Cart_rank(MPI_Comm comm, const int coords[], int *rank);

int MPI_Allreduce(const void *sendbuf, void *recvbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm);
int MPI_Send(const void *buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm);
int MPI_Recv(void *buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm comm, MPI_Status *status);
int MPI_Send_init(const void *buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm, MPI_Request *req);
int MPI_Recv_init(void *buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm comm, MPI_Request *req);
int MPI_Startall(int count, MPI_Request reqs[]);
int MPI_Waitall(int count, MPI_Request reqs[], MPI_Status statuses[]);
int MPI_Wait(MPI_Request *req, MPI_Status *status);
int MPI_Request_free(MPI_Request *req);

int MPI_Type_vector(int count, int blocklength, int stride, MPI_Datatype oldtype, MPI_Datatype *newtype);
int MPI_Type_create_subarray(int ndims, const int sizes[], const int subsizes[], const int starts[], int order,
                             MPI_Datatype oldtype, MPI_Datatype *newtype);
int MPI_Type_commit(MPI_Datatype *type);
int MPI_Type_free(MPI_Datatype *type);

int MPI_File_open(MPI_Comm comm, const char *filename, int amode, MPI_Info info, MPI_File *fh);
int MPI_File_close(MPI_File *fh);
int MPI_File_delete(const char *filename, MPI_Info info);
int MPI_File_set_view(MPI_File fh, MPI_Offset disp, MPI_Datatype etype, MPI_Datatype filetype, const char *datarep,
                      MPI_Info info);
int MPI_File_read_all(MPI_File fh, void *buf, int count, MPI_Datatype type, MPI_Status *status);
int MPI_File_write_all(MPI_File fh, const void *buf, int count, MPI_Datatype type, MPI_Status *status);
int MPI_File_iread(MPI_File fh, void *buf, int count, MPI_Datatype type, MPI_Request *req);
int MPI_File_iwrite(MPI_File fh, const void *buf, int count, MPI_Datatype type, MPI_Request *req);

#ifdef __cplusplus
}
#endif
#endif
