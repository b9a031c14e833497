/*
 * mpishim.c - single-node shared-memory MPI subset (see mpi.h).  HARNESS ONLY.
 *
 * Ranks are processes forked by baseline/mpishim/mpiexec; they share one POSIX shared-memory file
 * (name in $MPISHIM_SHM) that holds a barrier, an all-reduce scratch row and one byte ring per
 * ordered rank pair.  Sends are eager (header + payload streamed into the pair's ring, blocking
 * only while the ring is full); receives drain the ring and match by tag, parking messages that
 * arrive before their receive is posted.  File views (2-D subarrays of MPI_CHAR) become pread /
 * pwrite calls per row.
 */
#define _GNU_SOURCE
#include "mpi.h"

#include <errno.h>
#include <fcntl.h>
#include <sched.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#define MAXP 64
#define RING_BYTES (1u << 20)

typedef struct {
    volatile uint64_t head; /* bytes written (producer) */
    char pad0[56];
    volatile uint64_t tail; /* bytes consumed (consumer) */
    char pad1[56];
    char data[RING_BYTES];
} Ring;

typedef struct {
    volatile int bar_count, bar_gen;
    volatile int open_result;
    int pad[13];
    volatile long long reduce[MAXP];
    Ring rings[]; /* [src * P + dst] */
} Shm;

typedef struct {
    int tag, bytes;
} MsgHeader;

typedef struct Parked {
    struct Parked *next;
    int src, tag, bytes;
    char *data;
} Parked;

typedef struct {
    int kind; /* 0 unused, 1 vector, 2 subarray */
    int count, blocklen, stride, elem;
    int sizes[2], subsizes[2], starts[2];
} TypeInfo;

typedef struct {
    int used, is_send, active, done;
    void *buf;
    int count, type, peer, tag;
} ReqInfo;

typedef struct {
    int used, fd, filetype;
} FileInfo;

static Shm *g_shm;
static int g_rank, g_size = 1;
static Parked *g_parked;
static TypeInfo g_types[256];
static int g_ntypes = 3; /* 1 = CHAR, 2 = INT */
static ReqInfo g_reqs[1024];
static FileInfo g_files[16];
static int g_cart_dims[2], g_cart_periods[2], g_cart_size, g_have_cart;

static void die(const char *msg)
{
    fprintf(stderr, "mpishim[%d]: %s\n", g_rank, msg);
    exit(EXIT_FAILURE);
}

static int elem_size(MPI_Datatype t) { return t == MPI_INT ? 4 : 1; }

/* ------------------------------------------------------------------ basics */
int MPI_Init(int *argc, char ***argv)
{
    (void)argc;
    (void)argv;
    const char *name = getenv("MPISHIM_SHM");
    if (!name) { /* not under our mpiexec: run as a single rank */
        g_size = 1;
        g_rank = 0;
        size_t bytes = sizeof(Shm) + sizeof(Ring);
        g_shm = (Shm *)calloc(1, bytes);
        return MPI_SUCCESS;
    }
    g_size = atoi(getenv("MPISHIM_SIZE"));
    g_rank = atoi(getenv("MPISHIM_RANK"));
    if (g_size < 1 || g_size > MAXP) die("bad MPISHIM_SIZE");
    int fd = shm_open(name, O_RDWR, 0600);
    if (fd < 0) die("shm_open failed");
    size_t bytes = sizeof(Shm) + sizeof(Ring) * (size_t)g_size * (size_t)g_size;
    g_shm = (Shm *)mmap(NULL, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    if (g_shm == MAP_FAILED) die("mmap failed");
    close(fd);
    return MPI_SUCCESS;
}

int MPI_Barrier(MPI_Comm comm)
{
    (void)comm;
    if (g_size == 1) return MPI_SUCCESS;
    const int gen = g_shm->bar_gen;
    if (__sync_add_and_fetch(&g_shm->bar_count, 1) == g_size) {
        g_shm->bar_count = 0;
        __sync_synchronize();
        __sync_add_and_fetch(&g_shm->bar_gen, 1);
    } else {
        int spins = 0;
        while (g_shm->bar_gen == gen)
            if (++spins > 200) sched_yield();
    }
    return MPI_SUCCESS;
}

int MPI_Finalize(void)
{
    MPI_Barrier(MPI_COMM_WORLD);
    return MPI_SUCCESS;
}

static void check_comm(MPI_Comm comm)
{
    if (comm == MPI_COMM_NULL) die("MPI_COMM_NULL passed (rank count is not a perfect square?) - aborting like MPI_ERRORS_ARE_FATAL");
}

int MPI_Comm_rank(MPI_Comm comm, int *rank)
{
    check_comm(comm);
    *rank = g_rank;
    return MPI_SUCCESS;
}

int MPI_Comm_size(MPI_Comm comm, int *size)
{
    check_comm(comm);
    *size = (comm == MPI_COMM_WORLD || !g_have_cart) ? g_size : g_cart_size;
    return MPI_SUCCESS;
}

double MPI_Wtime(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

/* -------------------------------------------------------------- topology */
int MPI_Cart_create(MPI_Comm old, int ndims, const int dims[], const int periods[], int reorder, MPI_Comm *cart)
{
    (void)reorder;
    check_comm(old);
    if (ndims != 2) die("only 2-D Cartesian communicators are supported");
    g_cart_dims[0] = dims[0];
    g_cart_dims[1] = dims[1];
    g_cart_periods[0] = periods[0];
    g_cart_periods[1] = periods[1];
    g_cart_size = dims[0] * dims[1];
    g_have_cart = 1;
    if (g_cart_size > g_size) die("Cartesian grid larger than the communicator");
    *cart = g_rank < g_cart_size ? 2 : MPI_COMM_NULL;
    return MPI_SUCCESS;
}

int MPI_Cart_coords(MPI_Comm comm, int rank, int maxdims, int coords[])
{
    check_comm(comm);
    (void)maxdims;
    coords[0] = rank / g_cart_dims[1];
    coords[1] = rank % g_cart_dims[1];
    return MPI_SUCCESS;
}

int MPI_Cart_rank(MPI_Comm comm, const int coords[], int *rank)
{
    check_comm(comm);
    int c[2];
    for (int d = 0; d < 2; ++d) {
        c[d] = coords[d];
        if (g_cart_periods[d]) c[d] = ((c[d] % g_cart_dims[d]) + g_cart_dims[d]) % g_cart_dims[d];
        else if (c[d] < 0 || c[d] >= g_cart_dims[d]) die("Cart_rank: coordinate out of range");
    }
    *rank = c[0] * g_cart_dims[1] + c[1];
    return MPI_SUCCESS;
}

/* ------------------------------------------------------------ collectives */
int MPI_Allreduce(const void *sendbuf, void *recvbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm)
{
    check_comm(comm);
    if (count != 1 || type != MPI_INT || op != MPI_SUM) die("Allreduce: only 1 x MPI_INT, MPI_SUM is supported");
    const int n = g_have_cart ? g_cart_size : g_size;
    if (g_size == 1) {
        *(int *)recvbuf = *(const int *)sendbuf;
        return MPI_SUCCESS;
    }
    g_shm->reduce[g_rank] = *(const int *)sendbuf;
    MPI_Barrier(comm);
    long long sum = 0;
    for (int i = 0; i < n; ++i) sum += g_shm->reduce[i];
    MPI_Barrier(comm);
    *(int *)recvbuf = (int)sum;
    return MPI_SUCCESS;
}

/* --------------------------------------------------------- point to point */
static Ring *ring(int src, int dst) { return &g_shm->rings[(size_t)src * (size_t)g_size + (size_t)dst]; }

static void ring_write(Ring *r, const void *src, size_t n)
{
    const char *p = (const char *)src;
    while (n) {
        uint64_t head = r->head, tail = r->tail;
        size_t space = RING_BYTES - (size_t)(head - tail);
        if (!space) {
            sched_yield();
            continue;
        }
        size_t off = (size_t)(head % RING_BYTES);
        size_t chunk = n < space ? n : space;
        if (chunk > RING_BYTES - off) chunk = RING_BYTES - off;
        memcpy(r->data + off, p, chunk);
        __sync_synchronize();
        r->head = head + chunk;
        p += chunk;
        n -= chunk;
    }
}

static void ring_read(Ring *r, void *dst, size_t n)
{
    char *p = (char *)dst;
    while (n) {
        uint64_t head = r->head, tail = r->tail;
        size_t avail = (size_t)(head - tail);
        if (!avail) {
            sched_yield();
            continue;
        }
        __sync_synchronize();
        size_t off = (size_t)(tail % RING_BYTES);
        size_t chunk = n < avail ? n : avail;
        if (chunk > RING_BYTES - off) chunk = RING_BYTES - off;
        memcpy(p, r->data + off, chunk);
        __sync_synchronize();
        r->tail = tail + chunk;
        p += chunk;
        n -= chunk;
    }
}

static int type_bytes(int count, MPI_Datatype t)
{
    if (t == MPI_CHAR || t == MPI_INT) return count * elem_size(t);
    TypeInfo *ti = &g_types[t];
    if (ti->kind != 1) die("unsupported datatype in send/recv");
    return count * ti->count * ti->blocklen * ti->elem;
}

/* gather `buf` (count x type) into contiguous bytes / scatter back */
static void pack(const void *buf, int count, MPI_Datatype t, char *out)
{
    if (t == MPI_CHAR || t == MPI_INT) {
        memcpy(out, buf, (size_t)count * (size_t)elem_size(t));
        return;
    }
    TypeInfo *ti = &g_types[t];
    const char *b = (const char *)buf;
    size_t blk = (size_t)ti->blocklen * (size_t)ti->elem, str = (size_t)ti->stride * (size_t)ti->elem;
    for (int i = 0; i < ti->count; ++i) memcpy(out + (size_t)i * blk, b + (size_t)i * str, blk);
}

static void unpack(void *buf, int count, MPI_Datatype t, const char *in)
{
    if (t == MPI_CHAR || t == MPI_INT) {
        memcpy(buf, in, (size_t)count * (size_t)elem_size(t));
        return;
    }
    TypeInfo *ti = &g_types[t];
    char *b = (char *)buf;
    size_t blk = (size_t)ti->blocklen * (size_t)ti->elem, str = (size_t)ti->stride * (size_t)ti->elem;
    for (int i = 0; i < ti->count; ++i) memcpy(b + (size_t)i * str, in + (size_t)i * blk, blk);
}

static void do_send(const void *buf, int count, MPI_Datatype t, int dest, int tag)
{
    const int bytes = type_bytes(count, t);
    MsgHeader h = {tag, bytes};
    Ring *r = ring(g_rank, dest);
    ring_write(r, &h, sizeof h);
    if (t == MPI_CHAR || t == MPI_INT) {
        ring_write(r, buf, (size_t)bytes);
    } else {
        char *tmp = (char *)malloc((size_t)bytes);
        pack(buf, count, t, tmp);
        ring_write(r, tmp, (size_t)bytes);
        free(tmp);
    }
}

static void do_recv(void *buf, int count, MPI_Datatype t, int src, int tag, MPI_Status *st)
{
    const int want = type_bytes(count, t);
    /* a message parked earlier? */
    for (Parked **pp = &g_parked; *pp; pp = &(*pp)->next) {
        Parked *p = *pp;
        if (p->src == src && (tag == MPI_ANY_TAG || p->tag == tag)) {
            if (p->bytes > want) die("message longer than the receive buffer");
            unpack(buf, count, t, p->data);
            if (st) { st->MPI_SOURCE = src; st->MPI_TAG = p->tag; st->MPI_ERROR = MPI_SUCCESS; }
            *pp = p->next;
            free(p->data);
            free(p);
            return;
        }
    }
    Ring *r = ring(src, g_rank);
    for (;;) {
        MsgHeader h;
        ring_read(r, &h, sizeof h);
        if (tag == MPI_ANY_TAG || h.tag == tag) {
            if (h.bytes > want) die("message longer than the receive buffer");
            if (t == MPI_CHAR || t == MPI_INT) {
                ring_read(r, buf, (size_t)h.bytes);
            } else {
                char *tmp = (char *)malloc((size_t)h.bytes);
                ring_read(r, tmp, (size_t)h.bytes);
                unpack(buf, count, t, tmp);
                free(tmp);
            }
            if (st) { st->MPI_SOURCE = src; st->MPI_TAG = h.tag; st->MPI_ERROR = MPI_SUCCESS; }
            return;
        }
        Parked *p = (Parked *)malloc(sizeof *p);
        p->src = src;
        p->tag = h.tag;
        p->bytes = h.bytes;
        p->data = (char *)malloc((size_t)h.bytes);
        ring_read(r, p->data, (size_t)h.bytes);
        p->next = NULL;
        Parked **pp = &g_parked;
        while (*pp) pp = &(*pp)->next;
        *pp = p;
    }
}

int MPI_Send(const void *buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm)
{
    check_comm(comm);
    do_send(buf, count, type, dest, tag);
    return MPI_SUCCESS;
}

int MPI_Recv(void *buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm comm, MPI_Status *status)
{
    check_comm(comm);
    if (source == MPI_ANY_SOURCE) die("MPI_ANY_SOURCE is not supported");
    do_recv(buf, count, type, source, tag, status);
    return MPI_SUCCESS;
}

static int new_req(void)
{
    for (int i = 0; i < (int)(sizeof g_reqs / sizeof g_reqs[0]); ++i)
        if (!g_reqs[i].used) {
            memset(&g_reqs[i], 0, sizeof g_reqs[i]);
            g_reqs[i].used = 1;
            return i;
        }
    die("out of request slots");
    return -1;
}

int MPI_Send_init(const void *buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm, MPI_Request *req)
{
    check_comm(comm);
    int i = new_req();
    g_reqs[i].is_send = 1;
    g_reqs[i].buf = (void *)buf;
    g_reqs[i].count = count;
    g_reqs[i].type = type;
    g_reqs[i].peer = dest;
    g_reqs[i].tag = tag;
    *req = i;
    return MPI_SUCCESS;
}

int MPI_Recv_init(void *buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm comm, MPI_Request *req)
{
    check_comm(comm);
    int i = new_req();
    g_reqs[i].is_send = 0;
    g_reqs[i].buf = buf;
    g_reqs[i].count = count;
    g_reqs[i].type = type;
    g_reqs[i].peer = source;
    g_reqs[i].tag = tag;
    *req = i;
    return MPI_SUCCESS;
}

int MPI_Startall(int count, MPI_Request reqs[])
{
    /* sends are eager: they complete here; receives complete in Wait */
    for (int i = 0; i < count; ++i) {
        ReqInfo *r = &g_reqs[reqs[i]];
        r->active = 1;
        r->done = 0;
        if (r->is_send) {
            do_send(r->buf, r->count, r->type, r->peer, r->tag);
            r->done = 1;
        }
    }
    return MPI_SUCCESS;
}

int MPI_Wait(MPI_Request *req, MPI_Status *status)
{
    if (*req == MPI_REQUEST_NULL) return MPI_SUCCESS;
    ReqInfo *r = &g_reqs[*req];
    if (r->active && !r->done) {
        do_recv(r->buf, r->count, r->type, r->peer, r->tag, status);
        r->done = 1;
    }
    r->active = 0;
    if (r->peer == -100) { /* completed file request: one-shot */
        r->used = 0;
        *req = MPI_REQUEST_NULL;
    }
    return MPI_SUCCESS;
}

int MPI_Waitall(int count, MPI_Request reqs[], MPI_Status statuses[])
{
    for (int i = 0; i < count; ++i) MPI_Wait(&reqs[i], statuses ? &statuses[i] : MPI_STATUS_IGNORE);
    return MPI_SUCCESS;
}

int MPI_Request_free(MPI_Request *req)
{
    if (*req >= 0) g_reqs[*req].used = 0;
    *req = MPI_REQUEST_NULL;
    return MPI_SUCCESS;
}

/* --------------------------------------------------------------- datatypes */
int MPI_Type_vector(int count, int blocklength, int stride, MPI_Datatype oldtype, MPI_Datatype *newtype)
{
    if (g_ntypes >= 256) die("out of datatype slots");
    TypeInfo *t = &g_types[g_ntypes];
    memset(t, 0, sizeof *t);
    t->kind = 1;
    t->count = count;
    t->blocklen = blocklength;
    t->stride = stride;
    t->elem = elem_size(oldtype);
    *newtype = g_ntypes++;
    return MPI_SUCCESS;
}

int MPI_Type_create_subarray(int ndims, const int sizes[], const int subsizes[], const int starts[], int order,
                             MPI_Datatype oldtype, MPI_Datatype *newtype)
{
    if (ndims != 2 || order != MPI_ORDER_C || oldtype != MPI_CHAR) die("only 2-D C-order MPI_CHAR subarrays are supported");
    if (g_ntypes >= 256) die("out of datatype slots");
    TypeInfo *t = &g_types[g_ntypes];
    memset(t, 0, sizeof *t);
    t->kind = 2;
    t->elem = 1;
    for (int d = 0; d < 2; ++d) {
        t->sizes[d] = sizes[d];
        t->subsizes[d] = subsizes[d];
        t->starts[d] = starts[d];
    }
    *newtype = g_ntypes++;
    return MPI_SUCCESS;
}

int MPI_Type_commit(MPI_Datatype *type)
{
    (void)type;
    return MPI_SUCCESS;
}

int MPI_Type_free(MPI_Datatype *type)
{
    *type = MPI_DATATYPE_NULL; /* slots are not recycled; the programs create a handful */
    return MPI_SUCCESS;
}

/* ------------------------------------------------------------------ files */
int MPI_File_open(MPI_Comm comm, const char *filename, int amode, MPI_Info info, MPI_File *fh)
{
    (void)info;
    check_comm(comm);
    int slot = -1;
    for (int i = 1; i < 16; ++i)
        if (!g_files[i].used) { slot = i; break; }
    if (slot < 0) die("out of file slots");
    int flags = (amode & MPI_MODE_RDWR) ? O_RDWR : (amode & MPI_MODE_WRONLY) ? O_WRONLY : O_RDONLY;
    int fd = -1, result = MPI_SUCCESS;
    if (amode & MPI_MODE_CREATE) {
        /* collective create: rank 0 creates (honouring EXCL), everybody learns the outcome */
        if (g_rank == 0) {
            fd = open(filename, flags | O_CREAT | ((amode & MPI_MODE_EXCL) ? O_EXCL : 0), 0644);
            g_shm->open_result = fd < 0 ? MPI_ERR_OTHER : MPI_SUCCESS;
        }
        MPI_Barrier(comm);
        result = g_shm->open_result;
        if (result == MPI_SUCCESS && g_rank != 0) fd = open(filename, flags);
        MPI_Barrier(comm);
    } else {
        MPI_Barrier(comm);
        fd = open(filename, flags);
    }
    if (result != MPI_SUCCESS || fd < 0) return MPI_ERR_OTHER;
    g_files[slot].used = 1;
    g_files[slot].fd = fd;
    g_files[slot].filetype = 0;
    *fh = slot;
    return MPI_SUCCESS;
}

int MPI_File_close(MPI_File *fh)
{
    close(g_files[*fh].fd);
    g_files[*fh].used = 0;
    *fh = 0;
    MPI_Barrier(MPI_COMM_WORLD);
    return MPI_SUCCESS;
}

int MPI_File_delete(const char *filename, MPI_Info info)
{
    (void)info;
    return unlink(filename) == 0 ? MPI_SUCCESS : MPI_ERR_OTHER;
}

int MPI_File_set_view(MPI_File fh, MPI_Offset disp, MPI_Datatype etype, MPI_Datatype filetype, const char *datarep,
                      MPI_Info info)
{
    (void)datarep;
    (void)info;
    if (disp != 0 || etype != MPI_CHAR || g_types[filetype].kind != 2) die("File_set_view: only (0, MPI_CHAR, 2-D subarray) is supported");
    g_files[fh].filetype = filetype;
    return MPI_SUCCESS;
}

static void file_io(MPI_File fh, char *buf, int count, int writing)
{
    TypeInfo *t = &g_types[g_files[fh].filetype];
    const long long W = t->sizes[1], w = t->subsizes[1];
    long long done = 0;
    for (int row = 0; row < t->subsizes[0] && done < count; ++row) {
        long long n = w < count - done ? w : count - done;
        long long off = (long long)(t->starts[0] + row) * W + t->starts[1];
        long long pos = 0;
        while (pos < n) {
            ssize_t r = writing ? pwrite(g_files[fh].fd, buf + done + pos, (size_t)(n - pos), (off_t)(off + pos))
                                : pread(g_files[fh].fd, buf + done + pos, (size_t)(n - pos), (off_t)(off + pos));
            if (r < 0 && errno == EINTR) continue;
            if (r <= 0) die(writing ? "file write failed" : "file read failed / short file");
            pos += r;
        }
        done += n;
    }
}

int MPI_File_read_all(MPI_File fh, void *buf, int count, MPI_Datatype type, MPI_Status *status)
{
    (void)status;
    if (type != MPI_CHAR) die("File_read_all: MPI_CHAR only");
    file_io(fh, (char *)buf, count, 0);
    MPI_Barrier(MPI_COMM_WORLD);
    return MPI_SUCCESS;
}

int MPI_File_write_all(MPI_File fh, const void *buf, int count, MPI_Datatype type, MPI_Status *status)
{
    (void)status;
    if (type != MPI_CHAR) die("File_write_all: MPI_CHAR only");
    file_io(fh, (char *)buf, count, 1);
    MPI_Barrier(MPI_COMM_WORLD);
    return MPI_SUCCESS;
}

static int file_request(void)
{
    int i = new_req();
    g_reqs[i].peer = -100;
    g_reqs[i].active = 1;
    g_reqs[i].done = 1;
    return i;
}

int MPI_File_iread(MPI_File fh, void *buf, int count, MPI_Datatype type, MPI_Request *req)
{
    if (type != MPI_CHAR) die("File_iread: MPI_CHAR only");
    file_io(fh, (char *)buf, count, 0); /* completes immediately; the reference waits right away (A:194-198) */
    *req = file_request();
    return MPI_SUCCESS;
}

int MPI_File_iwrite(MPI_File fh, const void *buf, int count, MPI_Datatype type, MPI_Request *req)
{
    if (type != MPI_CHAR) die("File_iwrite: MPI_CHAR only");
    file_io(fh, (char *)buf, count, 1);
    *req = file_request();
    return MPI_SUCCESS;
}
