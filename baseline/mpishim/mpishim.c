/*
 * mpishim.c — the subset of MPI used by the reference programs, for one node.
 * See mpi.h.  Transport: one Unix stream socket pair per pair of ranks, created
 * by ./mpiexec before it forks the ranks; messages are {header, payload} and
 * are matched by (source, tag, communicator) in arrival order.  Sends are
 * "eager with progress": a send that would block keeps draining incoming
 * traffic into the unexpected-message queues, so the reference's
 * send-before-receive patterns (mpi_heat2Dn.c:179-192) cannot deadlock.
 */
#define _GNU_SOURCE
#include "mpi.h"

#include <errno.h>
#include <fcntl.h>
#include <poll.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

/* ------------------------------------------------------------------ state */
typedef struct msg {
  int src, tag, comm, nbytes;
  char* data;
  struct msg* next;
} msg_t;

typedef struct {
  int fd;
  /* reassembly of the message currently arriving */
  int have_hdr;
  int hdr_got;
  int hdr[4];
  char* body;
  int body_got;
  /* complete, not yet received messages (FIFO) */
  msg_t *head, *tail;
} peer_t;

static int g_rank = -1, g_size = 0, g_inited = 0;
static peer_t* g_peer = NULL;

/* communicators: 0 = world; others = Cartesian copies of world */
#define MAX_COMMS 16
typedef struct { int used, ndims, dims[3], periods[3]; } comm_t;
static comm_t g_comm[MAX_COMMS];

/* datatypes */
typedef struct { long off, len; } blk_t;
typedef struct { int used, nblk; blk_t* blk; long extent, size; } dtype_t;
#define MAX_TYPES 256
static dtype_t g_type[MAX_TYPES];

/* persistent requests */
typedef struct {
  int used, is_send, active;
  void* buf; int count; MPI_Datatype type; int peer, tag, comm;
} req_t;
#define MAX_REQS 1024
static req_t g_req[MAX_REQS];

struct mpishim_file { int fd; MPI_Offset disp; MPI_Datatype filetype; MPI_Comm comm; };

static void die(const char* what) {
  fprintf(stderr, "[mpishim rank %d] fatal: %s (%s)\n", g_rank, what, strerror(errno));
  exit(70);
}

/* ---------------------------------------------------------------- datatypes */
static void basic_type(int id, long sz) {
  g_type[id].used = 1; g_type[id].nblk = 1;
  g_type[id].blk = (blk_t*)malloc(sizeof(blk_t));
  g_type[id].blk[0].off = 0; g_type[id].blk[0].len = sz;
  g_type[id].extent = sz; g_type[id].size = sz;
}
static dtype_t* T(MPI_Datatype t) {
  if (t <= 0 || t >= MAX_TYPES || !g_type[t].used) { errno = EINVAL; die("bad datatype"); }
  return &g_type[t];
}
static int new_type(void) {
  for (int i = 16; i < MAX_TYPES; ++i) if (!g_type[i].used) { memset(&g_type[i], 0, sizeof g_type[i]); g_type[i].used = 1; return i; }
  errno = ENOMEM; die("too many datatypes"); return -1;
}
static void push_blk(dtype_t* d, long off, long len) {
  if (d->nblk && d->blk[d->nblk - 1].off + d->blk[d->nblk - 1].len == off) { d->blk[d->nblk - 1].len += len; return; }
  d->blk = (blk_t*)realloc(d->blk, sizeof(blk_t) * (size_t)(d->nblk + 1));
  d->blk[d->nblk].off = off; d->blk[d->nblk].len = len; d->nblk++;
}
static void append_type(dtype_t* d, const dtype_t* old, long base) {
  for (int b = 0; b < old->nblk; ++b) push_blk(d, base + old->blk[b].off, old->blk[b].len);
}

int MPI_Type_contiguous(int count, MPI_Datatype oldtype, MPI_Datatype* newtype) {
  dtype_t old = *T(oldtype); int id = new_type(); dtype_t* d = &g_type[id];
  for (int i = 0; i < count; ++i) append_type(d, &old, (long)i * old.extent);
  d->extent = (long)count * old.extent; d->size = (long)count * old.size;
  *newtype = id; return MPI_SUCCESS;
}
int MPI_Type_vector(int count, int blocklength, int stride, MPI_Datatype oldtype, MPI_Datatype* newtype) {
  dtype_t old = *T(oldtype); int id = new_type(); dtype_t* d = &g_type[id];
  for (int i = 0; i < count; ++i)
    for (int b = 0; b < blocklength; ++b) append_type(d, &old, ((long)i * stride + b) * old.extent);
  d->extent = count > 0 ? (((long)(count - 1)) * stride + blocklength) * old.extent : 0;
  d->size = (long)count * blocklength * old.size;
  *newtype = id; return MPI_SUCCESS;
}
int MPI_Type_create_subarray(int ndims, const int* sizes, const int* subsizes, const int* starts,
                             int order, MPI_Datatype oldtype, MPI_Datatype* newtype) {
  if (order != MPI_ORDER_C || ndims < 1 || ndims > 3) { errno = EINVAL; die("subarray: only C order, 1-3 dims"); }
  dtype_t old = *T(oldtype); int id = new_type(); dtype_t* d = &g_type[id];
  long total = 1, nsub = 1;
  for (int k = 0; k < ndims; ++k) { total *= sizes[k]; nsub *= subsizes[k]; }
  int idx[3] = {0, 0, 0};
  for (long n = 0; n < nsub; ++n) {
    long lin = 0;
    for (int k = 0; k < ndims; ++k) lin = lin * sizes[k] + (starts[k] + idx[k]);
    append_type(d, &old, lin * old.extent);
    for (int k = ndims - 1; k >= 0; --k) { if (++idx[k] < subsizes[k]) break; idx[k] = 0; }
  }
  d->extent = total * old.extent; d->size = nsub * old.size;
  *newtype = id; return MPI_SUCCESS;
}
int MPI_Type_commit(MPI_Datatype* type) { (void)T(*type); return MPI_SUCCESS; }
int MPI_Type_free(MPI_Datatype* type) {
  if (*type >= 16 && *type < MAX_TYPES && g_type[*type].used) { free(g_type[*type].blk); memset(&g_type[*type], 0, sizeof(dtype_t)); }
  *type = 0; return MPI_SUCCESS;
}

static long packed_size(int count, MPI_Datatype t) { return (long)count * T(t)->size; }
static void pack(char* dst, const void* src, int count, MPI_Datatype t) {
  const dtype_t* d = T(t); const char* s = (const char*)src;
  for (int c = 0; c < count; ++c)
    for (int b = 0; b < d->nblk; ++b) { memcpy(dst, s + (long)c * d->extent + d->blk[b].off, (size_t)d->blk[b].len); dst += d->blk[b].len; }
}
static void unpack(void* dst, const char* src, int count, MPI_Datatype t, long avail) {
  const dtype_t* d = T(t); char* o = (char*)dst;
  for (int c = 0; c < count; ++c)
    for (int b = 0; b < d->nblk; ++b) {
      long n = d->blk[b].len; if (n > avail) n = avail; if (n <= 0) return;
      memcpy(o + (long)c * d->extent + d->blk[b].off, src, (size_t)n); src += n; avail -= n;
    }
}

/* ---------------------------------------------------------------- transport */
static void enqueue(peer_t* p, msg_t* m) { m->next = NULL; if (p->tail) p->tail->next = m; else p->head = m; p->tail = m; }

/* Read whatever is available from peer q (non-blocking). */
static void drain(int q) {
  peer_t* p = &g_peer[q];
  if (p->fd < 0) return;
  for (;;) {
    if (!p->have_hdr) {
      ssize_t n = read(p->fd, (char*)p->hdr + p->hdr_got, sizeof p->hdr - (size_t)p->hdr_got);
      if (n < 0) { if (errno == EAGAIN || errno == EWOULDBLOCK) return; if (errno == EINTR) continue; die("read header"); }
      if (n == 0) return; /* peer closed */
      p->hdr_got += (int)n;
      if (p->hdr_got < (int)sizeof p->hdr) continue;
      p->have_hdr = 1; p->body_got = 0;
      p->body = (char*)malloc((size_t)(p->hdr[3] > 0 ? p->hdr[3] : 1));
    }
    while (p->body_got < p->hdr[3]) {
      ssize_t n = read(p->fd, p->body + p->body_got, (size_t)(p->hdr[3] - p->body_got));
      if (n < 0) { if (errno == EAGAIN || errno == EWOULDBLOCK) return; if (errno == EINTR) continue; die("read body"); }
      if (n == 0) return;
      p->body_got += (int)n;
    }
    msg_t* m = (msg_t*)malloc(sizeof(msg_t));
    m->src = p->hdr[0]; m->tag = p->hdr[1]; m->comm = p->hdr[2]; m->nbytes = p->hdr[3]; m->data = p->body;
    enqueue(p, m);
    p->have_hdr = 0; p->hdr_got = 0; p->body = NULL;
  }
}
static void drain_all(void) { for (int q = 0; q < g_size; ++q) if (q != g_rank) drain(q); }

/* Block until something can happen: any peer readable, or `wfd` writable. */
static void wait_io(int wfd) {
  struct pollfd pf[1024]; int n = 0;
  for (int q = 0; q < g_size && n < 1023; ++q) if (q != g_rank && g_peer[q].fd >= 0) { pf[n].fd = g_peer[q].fd; pf[n].events = POLLIN; pf[n].revents = 0; ++n; }
  if (wfd >= 0) { pf[n].fd = wfd; pf[n].events = POLLOUT; pf[n].revents = 0; ++n; }
  if (poll(pf, (nfds_t)n, 1000) < 0 && errno != EINTR) die("poll");
}

static void write_all(int dest, const char* p, long n) {
  int fd = g_peer[dest].fd;
  while (n > 0) {
    ssize_t w = write(fd, p, (size_t)n);
    if (w < 0) {
      if (errno == EINTR) continue;
      if (errno == EAGAIN || errno == EWOULDBLOCK) { drain_all(); wait_io(fd); continue; }
      die("write");
    }
    p += w; n -= w;
  }
}

static void send_raw(const char* data, long nbytes, int dest, int tag, int comm) {
  if (dest == MPI_PROC_NULL) return;
  if (dest < 0 || dest >= g_size) { errno = EINVAL; die("send: bad destination"); }
  if (dest == g_rank) {
    msg_t* m = (msg_t*)malloc(sizeof(msg_t));
    m->src = g_rank; m->tag = tag; m->comm = comm; m->nbytes = (int)nbytes;
    m->data = (char*)malloc((size_t)(nbytes > 0 ? nbytes : 1)); memcpy(m->data, data, (size_t)nbytes);
    enqueue(&g_peer[g_rank], m); return;
  }
  int hdr[4] = {g_rank, tag, comm, (int)nbytes};
  write_all(dest, (const char*)hdr, sizeof hdr);
  write_all(dest, data, nbytes);
}

static msg_t* take_match(int source, int tag, int comm) {
  peer_t* p = &g_peer[source];
  msg_t* prev = NULL;
  for (msg_t* m = p->head; m; prev = m, m = m->next) {
    if (m->comm == comm && (tag == MPI_ANY_TAG || m->tag == tag)) {
      if (prev) prev->next = m->next; else p->head = m->next;
      if (p->tail == m) p->tail = prev;
      return m;
    }
  }
  return NULL;
}
static msg_t* recv_raw(int source, int tag, int comm) {
  for (;;) {
    if (source == MPI_ANY_SOURCE) {
      for (int q = 0; q < g_size; ++q) { msg_t* m = take_match(q, tag, comm); if (m) return m; }
    } else {
      msg_t* m = take_match(source, tag, comm); if (m) return m;
    }
    drain_all();
    if (source == MPI_ANY_SOURCE) {
      int found = 0;
      for (int q = 0; q < g_size && !found; ++q) for (msg_t* m = g_peer[q].head; m; m = m->next) if (m->comm == comm && (tag == MPI_ANY_TAG || m->tag == tag)) { found = 1; break; }
      if (found) continue;
    } else {
      int found = 0;
      for (msg_t* m = g_peer[source].head; m; m = m->next) if (m->comm == comm && (tag == MPI_ANY_TAG || m->tag == tag)) { found = 1; break; }
      if (found) continue;
    }
    wait_io(-1);
  }
}

/* ----------------------------------------------------------------- basics */
int MPI_Init(int* argc, char*** argv) {
  (void)argc; (void)argv;
  const char* r = getenv("MPISHIM_RANK"); const char* s = getenv("MPISHIM_SIZE"); const char* f = getenv("MPISHIM_FDS");
  if (!r || !s) { g_rank = 0; g_size = 1; }  /* run without mpiexec: a world of one */
  else { g_rank = atoi(r); g_size = atoi(s); }
  g_peer = (peer_t*)calloc((size_t)g_size, sizeof(peer_t));
  for (int q = 0; q < g_size; ++q) g_peer[q].fd = -1;
  if (f && g_size > 1) {
    char* copy = strdup(f); char* save = NULL; int q = 0;
    for (char* tok = strtok_r(copy, ",", &save); tok && q < g_size; tok = strtok_r(NULL, ",", &save), ++q) {
      int fd = atoi(tok); g_peer[q].fd = fd;
      if (fd >= 0) { int fl = fcntl(fd, F_GETFL, 0); if (fl < 0 || fcntl(fd, F_SETFL, fl | O_NONBLOCK) < 0) die("fcntl"); }
    }
    free(copy);
  }
  basic_type(MPI_CHAR, 1); basic_type(MPI_BYTE, 1); basic_type(MPI_INT, sizeof(int));
  basic_type(MPI_FLOAT, sizeof(float)); basic_type(MPI_DOUBLE, sizeof(double));
  memset(g_comm, 0, sizeof g_comm); g_comm[0].used = 1;
  g_inited = 1;
  return MPI_SUCCESS;
}
int MPI_Comm_size(MPI_Comm comm, int* size) { (void)comm; *size = g_size; return MPI_SUCCESS; }
int MPI_Comm_rank(MPI_Comm comm, int* rank) { (void)comm; *rank = g_rank; return MPI_SUCCESS; }
double MPI_Wtime(void) { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec; }
int MPI_Get_processor_name(char* name, int* resultlen) {
  if (gethostname(name, MPI_MAX_PROCESSOR_NAME - 1) != 0) strcpy(name, "localhost");
  name[MPI_MAX_PROCESSOR_NAME - 1] = 0; *resultlen = (int)strlen(name); return MPI_SUCCESS;
}
int MPI_Abort(MPI_Comm comm, int errorcode) {
  (void)comm; fflush(stdout);
  fprintf(stderr, "[mpishim rank %d] MPI_Abort(%d)\n", g_rank, errorcode);
  _exit(errorcode ? errorcode : 1);  /* mpiexec kills the other ranks */
}
int MPI_Finalize(void) {
  MPI_Barrier(MPI_COMM_WORLD);
  fflush(stdout);
  for (int q = 0; q < g_size; ++q) if (g_peer[q].fd >= 0) close(g_peer[q].fd);
  return MPI_SUCCESS;
}

/* ----------------------------------------------------------- point to point */
int MPI_Send(const void* buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm) {
  if (dest == MPI_PROC_NULL) return MPI_SUCCESS;
  long n = packed_size(count, type);
  const dtype_t* d = T(type);
  if (d->nblk == 1 && d->blk[0].off == 0 && d->extent == d->size) { send_raw((const char*)buf, n, dest, tag, comm); return MPI_SUCCESS; }
  char* tmp = (char*)malloc((size_t)(n > 0 ? n : 1)); pack(tmp, buf, count, type);
  send_raw(tmp, n, dest, tag, comm); free(tmp);
  return MPI_SUCCESS;
}
int MPI_Recv(void* buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm comm, MPI_Status* status) {
  if (source == MPI_PROC_NULL) { if (status) { status->MPI_SOURCE = MPI_PROC_NULL; status->MPI_TAG = MPI_ANY_TAG; status->count_bytes = 0; } return MPI_SUCCESS; }
  msg_t* m = recv_raw(source, tag, comm);
  unpack(buf, m->data, count, type, m->nbytes);
  if (status) { status->MPI_SOURCE = m->src; status->MPI_TAG = m->tag; status->MPI_ERROR = 0; status->count_bytes = m->nbytes; }
  free(m->data); free(m);
  return MPI_SUCCESS;
}

static int new_req(void) {
  for (int i = 0; i < MAX_REQS; ++i) if (!g_req[i].used) { memset(&g_req[i], 0, sizeof g_req[i]); g_req[i].used = 1; return i; }
  errno = ENOMEM; die("too many requests"); return -1;
}
int MPI_Send_init(const void* buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm, MPI_Request* request) {
  int r = new_req(); g_req[r].is_send = 1; g_req[r].buf = (void*)buf; g_req[r].count = count; g_req[r].type = type;
  g_req[r].peer = dest; g_req[r].tag = tag; g_req[r].comm = comm; *request = r; return MPI_SUCCESS;
}
int MPI_Recv_init(void* buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm comm, MPI_Request* request) {
  int r = new_req(); g_req[r].is_send = 0; g_req[r].buf = buf; g_req[r].count = count; g_req[r].type = type;
  g_req[r].peer = source; g_req[r].tag = tag; g_req[r].comm = comm; *request = r; return MPI_SUCCESS;
}
int MPI_Startall(int count, MPI_Request* requests) {
  for (int i = 0; i < count; ++i) {
    req_t* q = &g_req[requests[i]];
    q->active = 1;
    /* sends go out at once (eager): the matching receive may be posted later */
    if (q->is_send) MPI_Send(q->buf, q->count, q->type, q->peer, q->tag, q->comm);
  }
  return MPI_SUCCESS;
}
int MPI_Waitall(int count, MPI_Request* requests, MPI_Status* statuses) {
  (void)statuses;
  for (int i = 0; i < count; ++i) {
    if (requests[i] == MPI_REQUEST_NULL) continue;
    req_t* q = &g_req[requests[i]];
    if (!q->active) continue;            /* never started: completes immediately */
    if (!q->is_send) MPI_Recv(q->buf, q->count, q->type, q->peer, q->tag, q->comm, MPI_STATUS_IGNORE);
    q->active = 0;
  }
  return MPI_SUCCESS;
}
int MPI_Request_free(MPI_Request* request) { if (*request >= 0 && *request < MAX_REQS) g_req[*request].used = 0; *request = MPI_REQUEST_NULL; return MPI_SUCCESS; }

/* -------------------------------------------------------------- collectives */
#define TAG_BARRIER (-1001)
#define TAG_BCAST (-1002)
#define TAG_REDUCE (-1003)
int MPI_Barrier(MPI_Comm comm) {
  char c = 0;
  if (g_size == 1) return MPI_SUCCESS;
  if (g_rank == 0) {
    for (int q = 1; q < g_size; ++q) { msg_t* m = recv_raw(q, TAG_BARRIER, comm); free(m->data); free(m); }
    for (int q = 1; q < g_size; ++q) send_raw(&c, 1, q, TAG_BARRIER, comm);
  } else {
    send_raw(&c, 1, 0, TAG_BARRIER, comm);
    msg_t* m = recv_raw(0, TAG_BARRIER, comm); free(m->data); free(m);
  }
  return MPI_SUCCESS;
}
int MPI_Bcast(void* buf, int count, MPI_Datatype type, int root, MPI_Comm comm) {
  if (g_size == 1) return MPI_SUCCESS;
  if (g_rank == root) { for (int q = 0; q < g_size; ++q) if (q != root) MPI_Send(buf, count, type, q, TAG_BCAST, comm); }
  else MPI_Recv(buf, count, type, root, TAG_BCAST, comm, MPI_STATUS_IGNORE);
  return MPI_SUCCESS;
}
static void combine(void* acc, const void* in, int count, MPI_Datatype type, MPI_Op op) {
  for (int i = 0; i < count; ++i) {
    if (type == MPI_FLOAT) { float* a = (float*)acc; const float* b = (const float*)in; a[i] = op == MPI_SUM ? a[i] + b[i] : op == MPI_MAX ? (a[i] > b[i] ? a[i] : b[i]) : (a[i] < b[i] ? a[i] : b[i]); }
    else if (type == MPI_DOUBLE) { double* a = (double*)acc; const double* b = (const double*)in; a[i] = op == MPI_SUM ? a[i] + b[i] : op == MPI_MAX ? (a[i] > b[i] ? a[i] : b[i]) : (a[i] < b[i] ? a[i] : b[i]); }
    else if (type == MPI_INT) { int* a = (int*)acc; const int* b = (const int*)in; a[i] = op == MPI_SUM ? a[i] + b[i] : op == MPI_MAX ? (a[i] > b[i] ? a[i] : b[i]) : (a[i] < b[i] ? a[i] : b[i]); }
    else { errno = EINVAL; die("reduce: unsupported datatype"); }
  }
}
int MPI_Reduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype type, MPI_Op op, int root, MPI_Comm comm) {
  long n = packed_size(count, type);
  if (g_rank != root) { MPI_Send(sendbuf, count, type, root, TAG_REDUCE, comm); return MPI_SUCCESS; }
  /* combine in rank order so that every run gives the same bits */
  char* tmp = (char*)malloc((size_t)(n > 0 ? n : 1));
  int first = 1;
  for (int q = 0; q < g_size; ++q) {
    const void* contrib;
    if (q == root) contrib = sendbuf; else { MPI_Recv(tmp, count, type, q, TAG_REDUCE, comm, MPI_STATUS_IGNORE); contrib = tmp; }
    if (first) { memmove(recvbuf, contrib, (size_t)n); first = 0; } else combine(recvbuf, contrib, count, type, op);
  }
  free(tmp);
  return MPI_SUCCESS;
}
int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm) {
  long n = packed_size(count, type);
  char* mine = (char*)malloc((size_t)(n > 0 ? n : 1)); memcpy(mine, sendbuf, (size_t)n);
  MPI_Reduce(mine, recvbuf, count, type, op, 0, comm);
  MPI_Bcast(recvbuf, count, type, 0, comm);
  free(mine);
  return MPI_SUCCESS;
}

/* ----------------------------------------------------------------- topology */
int MPI_Cart_create(MPI_Comm comm, int ndims, const int* dims, const int* periods, int reorder, MPI_Comm* newcomm) {
  (void)comm; (void)reorder;  /* ranks are never reordered: one node, uniform links */
  if (ndims < 1 || ndims > 3) { errno = EINVAL; die("Cart_create: 1-3 dims"); }
  long prod = 1; for (int k = 0; k < ndims; ++k) prod *= dims[k];
  if (prod > g_size) { errno = EINVAL; die("Cart_create: grid larger than the communicator"); }
  int id = -1; for (int i = 1; i < MAX_COMMS; ++i) if (!g_comm[i].used) { id = i; break; }
  if (id < 0) { errno = ENOMEM; die("too many communicators"); }
  g_comm[id].used = 1; g_comm[id].ndims = ndims;
  for (int k = 0; k < ndims; ++k) { g_comm[id].dims[k] = dims[k]; g_comm[id].periods[k] = periods[k]; }
  *newcomm = (g_rank < prod) ? id : MPI_COMM_NULL;
  return MPI_SUCCESS;
}
int MPI_Cart_shift(MPI_Comm comm, int direction, int disp, int* rank_source, int* rank_dest) {
  comm_t* c = &g_comm[comm];
  int coords[3], r = g_rank;
  for (int k = c->ndims - 1; k >= 0; --k) { coords[k] = r % c->dims[k]; r /= c->dims[k]; }  /* row-major */
  int out[2];
  for (int s = 0; s < 2; ++s) {
    int cc[3]; memcpy(cc, coords, sizeof cc);
    cc[direction] += (s == 0 ? -disp : disp);
    int ok = 1;
    if (cc[direction] < 0 || cc[direction] >= c->dims[direction]) {
      if (c->periods[direction]) cc[direction] = ((cc[direction] % c->dims[direction]) + c->dims[direction]) % c->dims[direction];
      else ok = 0;
    }
    if (!ok) { out[s] = MPI_PROC_NULL; continue; }
    int rank = 0; for (int k = 0; k < c->ndims; ++k) rank = rank * c->dims[k] + cc[k];
    out[s] = rank;
  }
  *rank_source = out[0]; *rank_dest = out[1];
  return MPI_SUCCESS;
}

/* ------------------------------------------------------------------- MPI-IO */
int MPI_File_open(MPI_Comm comm, const char* filename, int amode, MPI_Info info, MPI_File* fh) {
  (void)info;
  int fl = 0;
  if (amode & MPI_MODE_RDWR) fl |= O_RDWR; else if (amode & MPI_MODE_WRONLY) fl |= O_WRONLY; else fl |= O_RDONLY;
  if (amode & MPI_MODE_CREATE) fl |= O_CREAT;   /* no O_TRUNC: MPI does not truncate either */
  int fd = open(filename, fl, 0644);
  if (fd < 0) die("MPI_File_open");
  struct mpishim_file* f = (struct mpishim_file*)calloc(1, sizeof *f);
  f->fd = fd; f->disp = 0; f->filetype = MPI_BYTE; f->comm = comm;
  *fh = f;
  MPI_Barrier(comm);
  return MPI_SUCCESS;
}
int MPI_File_set_view(MPI_File fh, MPI_Offset disp, MPI_Datatype etype, MPI_Datatype filetype, const char* datarep, MPI_Info info) {
  (void)etype; (void)datarep; (void)info;
  fh->disp = disp; fh->filetype = filetype;
  return MPI_SUCCESS;
}
int MPI_File_write_all(MPI_File fh, const void* buf, int count, MPI_Datatype type, MPI_Status* status) {
  (void)status;
  long n = packed_size(count, type);
  char* tmp = (char*)malloc((size_t)(n > 0 ? n : 1)); pack(tmp, buf, count, type);
  const dtype_t* ft = T(fh->filetype);
  if (ft->nblk == 1 && ft->blk[0].off == 0 && ft->size == ft->extent) {
    /* contiguous view (the default byte view included): one linear write */
    long w = 0;
    while (w < n) { ssize_t k = pwrite(fh->fd, tmp + w, (size_t)(n - w), (off_t)(fh->disp + w)); if (k < 0) { if (errno == EINTR) continue; die("pwrite"); } w += k; }
    free(tmp);
    MPI_Barrier(fh->comm);
    return MPI_SUCCESS;
  }
  /* the packed stream fills the holes of the file type, tile after tile */
  long done = 0, tile = 0;
  while (done < n) {
    for (int b = 0; b < ft->nblk && done < n; ++b) {
      long len = ft->blk[b].len; if (len > n - done) len = n - done;
      off_t off = (off_t)(fh->disp + tile * ft->extent + ft->blk[b].off);
      long w = 0;
      while (w < len) { ssize_t k = pwrite(fh->fd, tmp + done + w, (size_t)(len - w), off + w); if (k < 0) { if (errno == EINTR) continue; die("pwrite"); } w += k; }
      done += len;
    }
    ++tile;
  }
  free(tmp);
  MPI_Barrier(fh->comm);
  return MPI_SUCCESS;
}
int MPI_File_close(MPI_File* fh) {
  MPI_Comm c = (*fh)->comm;
  if (close((*fh)->fd) != 0) die("MPI_File_close");
  free(*fh); *fh = NULL;
  MPI_Barrier(c);
  return MPI_SUCCESS;
}
