/* hypre_b200/host/mpishim/mpishim.c -- the message-passing library behind mpi.h in this
 * directory (see there for scope and for who builds it).  One process per rank, started by mpirun_shim,
 * which hands every rank one connected Unix stream socket per peer (environment SHIM_RANK,
 * SHIM_SIZE, SHIM_FDS).  Ranks started by something else (torchrun, a job script) meet through
 * Unix sockets in a directory instead: SHIM_DIR=<dir> with SHIM_RANK/SHIM_SIZE or torchrun's
 * RANK/WORLD_SIZE.  Without those variables the library runs as a single rank.
 *
 * Progress model: all sockets are non-blocking; every blocking call (Recv, Wait*, Probe, the
 * collectives, Finalize) drives one poll()-based engine that drains incoming bytes into
 * complete messages and pushes queued outgoing bytes.  Sends are eager and buffered: the
 * payload is copied into an output queue and the send request completes at once.  Incoming
 * messages match posted receives in posting order, else wait in an arrival-ordered queue
 * (MPI's non-overtaking rule holds because each pair of ranks shares one FIFO stream).
 */
#define _GNU_SOURCE
#include "mpi.h"
#include <errno.h>
#include <fcntl.h>
#include <poll.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <time.h>
#include <unistd.h>

/* ------------------------------------------------------------------ state */
typedef struct msg
{
   int ctx, src, tag, bytes;
   char *data;
   struct msg *next;
} msg_t;

typedef struct
{
   int ctx, src, tag, bytes;
} hdr_t;

typedef struct outbuf
{
   char *data;
   size_t len, pos;
   struct outbuf *next;
} outbuf_t;

typedef struct
{
   int fd;
   /* input state machine */
   hdr_t hdr;
   size_t hdr_got;
   char *payload;
   size_t pay_got;
   /* output queue */
   outbuf_t *out_head, *out_tail;
} peer_t;

enum { REQ_FREE = 0, REQ_SEND = 1, REQ_RECV = 2 };
typedef struct
{
   int kind, done, active, persistent, posted;
   void *buf;
   int count;
   MPI_Datatype dt;
   int ctx, peer, tag;       /* peer: comm rank (source for receives, dest for sends) */
   MPI_Comm comm;
   MPI_Status st;
} req_t;

typedef struct
{
   int used, ctx, size, rank;
   int *world;               /* world rank of every member */
} comm_t;

typedef struct
{
   int used, n;
   int *world;
} group_t;

typedef struct
{
   long disp, len;
} blk_t;
typedef struct
{
   int used, nblk;
   blk_t *blk;               /* flattened byte blocks of ONE element */
   long size, extent;
} dtype_t;

#define MAX_COMMS  1024
#define MAX_GROUPS 1024
#define MAX_REQS   65536
#define MAX_TYPES  4096
#define MAX_OPS    64
#define FIRST_DERIVED 32
#define FIRST_USER_OP 100
#define COLL_TAG   0x7f000000

static int g_init = 0, g_size = 1, g_rank = 0;
static peer_t *g_peer;
static msg_t *g_unexp_head, *g_unexp_tail;
static req_t *g_req;
static int g_posted[MAX_REQS], g_nposted;       /* posted receive requests in posting order */
static comm_t g_comm[MAX_COMMS];
static group_t g_group[MAX_GROUPS];
static dtype_t g_type[MAX_TYPES];
static MPI_User_function *g_op[MAX_OPS];
static int g_next_ctx = 4;

static void die(const char *fmt, ...)
{
   va_list ap;
   va_start(ap, fmt);
   fprintf(stderr, "[mpishim rank %d] ", g_rank);
   vfprintf(stderr, fmt, ap);
   fprintf(stderr, "\n");
   va_end(ap);
   fflush(stderr);
   _exit(86);
}

/* ------------------------------------------------------------------ datatypes */
static long basic_size(MPI_Datatype dt)
{
   switch (dt)
   {
      case 1: case 2: case 15: return 1;
      case 3: return sizeof(short);
      case 4: case 7: return sizeof(int);
      case 5: case 8: return sizeof(long);
      case 6: return sizeof(long long);
      case 9: return sizeof(float);
      case 10: return sizeof(double);
      case 11: return sizeof(long double);
      case 12: return 2 * sizeof(float);
      case 13: return 2 * sizeof(double);
      case 14: return 2 * sizeof(long double);
      default: return -1;
   }
}
static long type_size(MPI_Datatype dt)
{
   if (dt < FIRST_DERIVED) { long s = basic_size(dt); if (s < 0) { die("unknown datatype %d", dt); } return s; }
   if (dt >= MAX_TYPES || !g_type[dt].used) { die("invalid datatype %d", dt); }
   return g_type[dt].size;
}
static long type_extent(MPI_Datatype dt)
{
   if (dt < FIRST_DERIVED) { return type_size(dt); }
   return g_type[dt].extent;
}
/* contiguous bytes of (buf, count, dt) -> out (malloc'd by caller, count*size bytes) */
static void pack(const void *buf, int count, MPI_Datatype dt, char *out)
{
   if (dt < FIRST_DERIVED) { memcpy(out, buf, (size_t)count * (size_t)type_size(dt)); return; }
   const dtype_t *t = &g_type[dt];
   for (int e = 0; e < count; e++)
   {
      for (int b = 0; b < t->nblk; b++)
      {
         memcpy(out, (const char *)buf + (long)e * t->extent + t->blk[b].disp, (size_t)t->blk[b].len);
         out += t->blk[b].len;
      }
   }
}
static void unpack(const char *in, long bytes, void *buf, int count, MPI_Datatype dt)
{
   if (dt < FIRST_DERIVED) { memcpy(buf, in, (size_t)bytes); return; }
   const dtype_t *t = &g_type[dt];
   long left = bytes;
   for (int e = 0; e < count && left > 0; e++)
   {
      for (int b = 0; b < t->nblk && left > 0; b++)
      {
         long l = t->blk[b].len < left ? t->blk[b].len : left;
         memcpy((char *)buf + (long)e * t->extent + t->blk[b].disp, in, (size_t)l);
         in += l; left -= l;
      }
   }
}
static MPI_Datatype new_type(int nblk)
{
   for (int i = FIRST_DERIVED; i < MAX_TYPES; i++)
   {
      if (!g_type[i].used)
      {
         g_type[i].used = 1; g_type[i].nblk = nblk;
         g_type[i].blk = (blk_t *)calloc((size_t)(nblk > 0 ? nblk : 1), sizeof(blk_t));
         g_type[i].size = 0; g_type[i].extent = 0;
         return i;
      }
   }
   die("out of datatype slots");
   return -1;
}
/* append the blocks of `count` consecutive elements of oldtype at byte displacement disp */
static int type_nblk(MPI_Datatype dt) { return dt < FIRST_DERIVED ? 1 : g_type[dt].nblk; }
static void append_blocks(dtype_t *t, int *pos, MPI_Datatype oldtype, int count, long disp)
{
   for (int e = 0; e < count; e++)
   {
      long base = disp + (long)e * type_extent(oldtype);
      if (oldtype < FIRST_DERIVED) { t->blk[*pos].disp = base; t->blk[*pos].len = type_size(oldtype); (*pos)++; }
      else
      {
         for (int b = 0; b < g_type[oldtype].nblk; b++)
         {
            t->blk[*pos].disp = base + g_type[oldtype].blk[b].disp; t->blk[*pos].len = g_type[oldtype].blk[b].len; (*pos)++;
         }
      }
   }
}
static void finish_type(dtype_t *t, long extent_hint)
{
   long lo = 0, hi = 0, size = 0;
   for (int b = 0; b < t->nblk; b++)
   {
      if (b == 0 || t->blk[b].disp < lo) { lo = t->blk[b].disp; }
      if (b == 0 || t->blk[b].disp + t->blk[b].len > hi) { hi = t->blk[b].disp + t->blk[b].len; }
      size += t->blk[b].len;
   }
   t->size = size;
   t->extent = extent_hint > 0 ? extent_hint : hi - (lo < 0 ? lo : 0);
}

/* ------------------------------------------------------------------ communicators */
static comm_t *get_comm(MPI_Comm c)
{
   if (c < 0 || c >= MAX_COMMS || !g_comm[c].used) { die("invalid communicator %d", c); }
   return &g_comm[c];
}
static MPI_Comm new_comm(int ctx, int size, int rank, const int *world)
{
   for (int i = 2; i < MAX_COMMS; i++)
   {
      if (!g_comm[i].used)
      {
         g_comm[i].used = 1; g_comm[i].ctx = ctx; g_comm[i].size = size; g_comm[i].rank = rank;
         g_comm[i].world = (int *)malloc(sizeof(int) * (size_t)size);
         memcpy(g_comm[i].world, world, sizeof(int) * (size_t)size);
         return i;
      }
   }
   die("out of communicator slots");
   return MPI_COMM_NULL;
}

/* ------------------------------------------------------------------ matching */
static int matches(int ctx, int src, int tag, const msg_t *m)
{
   return m->ctx == ctx && (src == MPI_ANY_SOURCE || src == m->src) && (tag == MPI_ANY_TAG || tag == m->tag);
}
static void complete_recv(req_t *r, msg_t *m)
{
   long cap = (long)r->count * type_size(r->dt);
   if (m->bytes > cap) { die("message of %d bytes truncated by a receive of %ld bytes (src %d tag %d)", m->bytes, cap, m->src, m->tag); }
   unpack(m->data, m->bytes, r->buf, r->count, r->dt);
   r->st.MPI_SOURCE = m->src; r->st.MPI_TAG = m->tag; r->st.MPI_ERROR = MPI_SUCCESS; r->st.shim_bytes = m->bytes;
   r->done = 1; r->posted = 0;
   free(m->data); free(m);
}
/* a complete message arrived (or was sent to self): posted receives first, in posting order */
static void deliver(msg_t *m)
{
   for (int i = 0; i < g_nposted; i++)
   {
      req_t *r = &g_req[g_posted[i]];
      if (matches(r->ctx, r->peer, r->tag, m))
      {
         memmove(&g_posted[i], &g_posted[i + 1], sizeof(int) * (size_t)(g_nposted - i - 1));
         g_nposted--;
         complete_recv(r, m);
         return;
      }
   }
   m->next = NULL;
   if (g_unexp_tail) { g_unexp_tail->next = m; } else { g_unexp_head = m; }
   g_unexp_tail = m;
}
static msg_t *find_unexpected(int ctx, int src, int tag, int remove)
{
   msg_t *prev = NULL;
   for (msg_t *m = g_unexp_head; m; prev = m, m = m->next)
   {
      if (matches(ctx, src, tag, m))
      {
         if (remove)
         {
            if (prev) { prev->next = m->next; } else { g_unexp_head = m->next; }
            if (g_unexp_tail == m) { g_unexp_tail = prev; }
         }
         return m;
      }
   }
   return NULL;
}

/* ------------------------------------------------------------------ progress engine */
static void pump_in(peer_t *p)
{
   for (;;)
   {
      if (p->hdr_got < sizeof(hdr_t))
      {
         ssize_t n = read(p->fd, (char *)&p->hdr + p->hdr_got, sizeof(hdr_t) - p->hdr_got);
         if (n == 0) { close(p->fd); p->fd = -1; return; }   /* peer closed: it has finalized */
         if (n < 0) { if (errno == EAGAIN || errno == EWOULDBLOCK || errno == EINTR) { return; } die("read: %s", strerror(errno)); }
         p->hdr_got += (size_t)n;
         if (p->hdr_got < sizeof(hdr_t)) { return; }
         p->payload = (char *)malloc((size_t)(p->hdr.bytes > 0 ? p->hdr.bytes : 1));
         p->pay_got = 0;
      }
      while (p->pay_got < (size_t)p->hdr.bytes)
      {
         ssize_t n = read(p->fd, p->payload + p->pay_got, (size_t)p->hdr.bytes - p->pay_got);
         if (n == 0) { die("peer closed in the middle of a message"); }
         if (n < 0) { if (errno == EAGAIN || errno == EWOULDBLOCK || errno == EINTR) { return; } die("read: %s", strerror(errno)); }
         p->pay_got += (size_t)n;
      }
      msg_t *m = (msg_t *)malloc(sizeof(msg_t));
      m->ctx = p->hdr.ctx; m->src = p->hdr.src; m->tag = p->hdr.tag; m->bytes = p->hdr.bytes; m->data = p->payload; m->next = NULL;
      p->payload = NULL; p->hdr_got = 0; p->pay_got = 0;
      deliver(m);
   }
}
static void pump_out(peer_t *p)
{
   while (p->out_head)
   {
      outbuf_t *o = p->out_head;
      ssize_t n = write(p->fd, o->data + o->pos, o->len - o->pos);
      if (n < 0) { if (errno == EAGAIN || errno == EWOULDBLOCK || errno == EINTR) { return; } die("write: %s", strerror(errno)); }
      o->pos += (size_t)n;
      if (o->pos < o->len) { return; }
      p->out_head = o->next;
      if (!p->out_head) { p->out_tail = NULL; }
      free(o->data); free(o);
   }
}
/* one round: wait up to timeout_ms for socket activity, then move bytes both ways */
static void progress(int timeout_ms)
{
   if (g_size == 1) { return; }
   struct pollfd pf[g_size];
   int np = 0, who[g_size];
   for (int i = 0; i < g_size; i++)
   {
      if (i == g_rank || g_peer[i].fd < 0) { continue; }
      pf[np].fd = g_peer[i].fd; pf[np].events = POLLIN | (g_peer[i].out_head ? POLLOUT : 0); pf[np].revents = 0;
      who[np++] = i;
   }
   int rc = poll(pf, (nfds_t)np, timeout_ms);
   if (rc < 0) { if (errno == EINTR) { return; } die("poll: %s", strerror(errno)); }
   for (int k = 0; k < np; k++)
   {
      if (pf[k].revents & (POLLIN | POLLHUP)) { pump_in(&g_peer[who[k]]); }
      if (pf[k].revents & POLLOUT) { pump_out(&g_peer[who[k]]); }
   }
}
static int pending_output(void)
{
   for (int i = 0; i < g_size; i++) { if (i != g_rank && g_peer[i].out_head) { return 1; } }
   return 0;
}

/* eager send of packed bytes to world rank `dest_world` */
static void post_message(int dest_world, int ctx, int src_rank, int tag, char *data, int bytes)
{
   if (dest_world == g_rank)
   {
      msg_t *m = (msg_t *)malloc(sizeof(msg_t));
      m->ctx = ctx; m->src = src_rank; m->tag = tag; m->bytes = bytes; m->data = data; m->next = NULL;
      deliver(m);
      return;
   }
   peer_t *p = &g_peer[dest_world];
   outbuf_t *o = (outbuf_t *)malloc(sizeof(outbuf_t));
   o->len = sizeof(hdr_t) + (size_t)bytes; o->pos = 0; o->next = NULL;
   o->data = (char *)malloc(o->len);
   hdr_t h = { ctx, src_rank, tag, bytes };
   memcpy(o->data, &h, sizeof(h));
   if (bytes > 0) { memcpy(o->data + sizeof(h), data, (size_t)bytes); }
   free(data);
   if (p->out_tail) { p->out_tail->next = o; } else { p->out_head = o; }
   p->out_tail = o;
   pump_out(p);
}

/* ------------------------------------------------------------------ requests */
static MPI_Request new_req(void)
{
   static int hint = 0;
   for (int k = 0; k < MAX_REQS; k++)
   {
      int i = (hint + k) % MAX_REQS;
      if (g_req[i].kind == REQ_FREE) { memset(&g_req[i], 0, sizeof(req_t)); hint = i + 1; return i; }
   }
   die("out of request slots");
   return MPI_REQUEST_NULL;
}
static void start_send(req_t *r)
{
   comm_t *c = get_comm(r->comm);
   r->active = 1; r->done = 1;
   if (r->peer == MPI_PROC_NULL) { return; }
   if (r->peer < 0 || r->peer >= c->size) { die("send to rank %d of a communicator of size %d", r->peer, c->size); }
   long bytes = (long)r->count * type_size(r->dt);
   char *data = (char *)malloc((size_t)(bytes > 0 ? bytes : 1));
   pack(r->buf, r->count, r->dt, data);
   post_message(c->world[r->peer], r->ctx, c->rank, r->tag, data, (int)bytes);
}
static void start_recv(req_t *r, int self_index)
{
   r->active = 1; r->done = 0;
   if (r->peer == MPI_PROC_NULL)
   {
      r->done = 1; r->st.MPI_SOURCE = MPI_PROC_NULL; r->st.MPI_TAG = MPI_ANY_TAG; r->st.shim_bytes = 0;
      return;
   }
   msg_t *m = find_unexpected(r->ctx, r->peer, r->tag, 1);
   if (m) { complete_recv(r, m); return; }
   if (g_nposted >= MAX_REQS) { die("too many posted receives"); }
   r->posted = 1;
   g_posted[g_nposted++] = self_index;
}
static void release_req(MPI_Request *q, MPI_Status *st)
{
   req_t *r = &g_req[*q];
   if (st && r->kind == REQ_RECV) { *st = r->st; }
   if (st && r->kind == REQ_SEND) { st->MPI_SOURCE = MPI_ANY_SOURCE; st->MPI_TAG = MPI_ANY_TAG; st->MPI_ERROR = MPI_SUCCESS; st->shim_bytes = 0; }
   if (r->persistent) { r->active = 0; }
   else { r->kind = REQ_FREE; *q = MPI_REQUEST_NULL; }
}
static int req_inactive(MPI_Request q)
{
   return q == MPI_REQUEST_NULL || (g_req[q].persistent && !g_req[q].active);
}

/* ------------------------------------------------------------------ init / finalize */
static void set_nonblocking(int fd, int peer)
{
   int fl = fcntl(fd, F_GETFL, 0);
   if (fl < 0 || fcntl(fd, F_SETFL, fl | O_NONBLOCK) < 0) { die("fcntl on the socket to rank %d: %s", peer, strerror(errno)); }
}
static void full_io(int fd, void *buf, size_t n, int out)
{
   char *p = (char *)buf;
   while (n > 0)
   {
      ssize_t k = out ? write(fd, p, n) : read(fd, p, n);
      if (k < 0 && errno == EINTR) { continue; }
      if (k <= 0) { die("rendezvous: short %s (%s)", out ? "write" : "read", strerror(errno)); }
      p += k; n -= (size_t)k;
   }
}
/* Ranks that were not started by mpirun_shim meet in a directory: rank r listens on <dir>/r<r>.sock, connects
 * to every lower rank (retrying until it is up, <= 120 s) and accepts one connection from every higher rank;
 * the connecting side sends its rank number.  The result is the same full mesh of stream sockets. */
static void rendezvous(const char *dir)
{
   struct sockaddr_un a;
   memset(&a, 0, sizeof(a));
   a.sun_family = AF_UNIX;
   if (snprintf(a.sun_path, sizeof(a.sun_path), "%s/r%d.sock", dir, g_rank) >= (int)sizeof(a.sun_path)) { die("SHIM_DIR too long"); }
   int ls = socket(AF_UNIX, SOCK_STREAM, 0);
   if (ls < 0) { die("socket: %s", strerror(errno)); }
   unlink(a.sun_path);
   if (bind(ls, (struct sockaddr *)&a, sizeof(a)) < 0 || listen(ls, g_size) < 0) { die("bind/listen %s: %s", a.sun_path, strerror(errno)); }
   for (int j = 0; j < g_rank; j++)
   {
      struct sockaddr_un b;
      memset(&b, 0, sizeof(b));
      b.sun_family = AF_UNIX;
      snprintf(b.sun_path, sizeof(b.sun_path), "%s/r%d.sock", dir, j);
      int fd = -1;
      for (int attempt = 0; attempt < 12000; attempt++)
      {
         fd = socket(AF_UNIX, SOCK_STREAM, 0);
         if (fd < 0) { die("socket: %s", strerror(errno)); }
         if (connect(fd, (struct sockaddr *)&b, sizeof(b)) == 0) { break; }
         close(fd); fd = -1;
         struct timespec ts = { 0, 10 * 1000 * 1000 };
         nanosleep(&ts, NULL);
      }
      if (fd < 0) { die("rendezvous: rank %d did not appear in %s", j, dir); }
      int me = g_rank;
      full_io(fd, &me, sizeof(me), 1);
      g_peer[j].fd = fd;
   }
   for (int k = g_rank + 1; k < g_size; k++)
   {
      int fd;
      do { fd = accept(ls, NULL, NULL); } while (fd < 0 && errno == EINTR);
      if (fd < 0) { die("accept: %s", strerror(errno)); }
      int who = -1;
      full_io(fd, &who, sizeof(who), 0);
      if (who <= g_rank || who >= g_size || g_peer[who].fd >= 0) { die("rendezvous: unexpected rank %d", who); }
      g_peer[who].fd = fd;
   }
   close(ls);
   unlink(a.sun_path);
   for (int i = 0; i < g_size; i++) { if (i != g_rank) { set_nonblocking(g_peer[i].fd, i); } }
}

int MPI_Init(int *argc, char ***argv)
{
   (void)argc; (void)argv;
   if (g_init) { return MPI_SUCCESS; }
   g_init = 1;
   const char *es = getenv("SHIM_SIZE"), *er = getenv("SHIM_RANK"), *ef = getenv("SHIM_FDS");
   const char *ed = getenv("SHIM_DIR");
   if (!ef && ed)
   {
      if (!es) { es = getenv("WORLD_SIZE"); }
      if (!er) { er = getenv("RANK"); }
   }
   if (es && er && (ef || ed)) { g_size = atoi(es); g_rank = atoi(er); }
   if (g_size < 1 || g_rank < 0 || g_rank >= g_size) { die("bad rank %d of %d", g_rank, g_size); }
   g_peer = (peer_t *)calloc((size_t)g_size, sizeof(peer_t));
   g_req = (req_t *)calloc(MAX_REQS, sizeof(req_t));
   for (int i = 0; i < g_size; i++) { g_peer[i].fd = -1; }
   if (g_size > 1 && !ef) { rendezvous(ed); }
   else if (g_size > 1)
   {
      char *copy = strdup(ef), *save = NULL;
      int i = 0;
      for (char *tok = strtok_r(copy, ",", &save); tok && i < g_size; tok = strtok_r(NULL, ",", &save), i++)
      {
         g_peer[i].fd = atoi(tok);
         if (i != g_rank)
         {
            int fl = fcntl(g_peer[i].fd, F_GETFL, 0);
            if (fl < 0 || fcntl(g_peer[i].fd, F_SETFL, fl | O_NONBLOCK) < 0) { die("fcntl on the socket to rank %d: %s", i, strerror(errno)); }
         }
      }
      free(copy);
      if (i != g_size) { die("SHIM_FDS lists %d sockets for %d ranks", i, g_size); }
   }
   int *world = (int *)malloc(sizeof(int) * (size_t)g_size);
   for (int i = 0; i < g_size; i++) { world[i] = i; }
   g_comm[0].used = 1; g_comm[0].ctx = 0; g_comm[0].size = g_size; g_comm[0].rank = g_rank; g_comm[0].world = world;
   int *self = (int *)malloc(sizeof(int)); self[0] = g_rank;
   g_comm[1].used = 1; g_comm[1].ctx = 2; g_comm[1].size = 1; g_comm[1].rank = 0; g_comm[1].world = self;
   return MPI_SUCCESS;
}
int MPI_Initialized(int *flag) { *flag = g_init; return MPI_SUCCESS; }
int MPI_Finalize(void)
{
   if (!g_init) { return MPI_SUCCESS; }
   MPI_Barrier(MPI_COMM_WORLD);
   while (pending_output()) { progress(100); }
   for (int i = 0; i < g_size; i++) { if (i != g_rank && g_peer[i].fd >= 0) { close(g_peer[i].fd); g_peer[i].fd = -1; } }
   return MPI_SUCCESS;
}
int MPI_Abort(MPI_Comm comm, int errorcode)
{
   (void)comm;
   fprintf(stderr, "[mpishim rank %d] MPI_Abort(%d)\n", g_rank, errorcode);
   fflush(NULL);
   _exit(errorcode ? errorcode : 1);
}
double MPI_Wtime(void)
{
   struct timespec ts;
   clock_gettime(CLOCK_MONOTONIC, &ts);
   return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}
double MPI_Wtick(void) { return 1e-9; }

/* ------------------------------------------------------------------ point to point */
static MPI_Request make_req(int kind, const void *buf, int count, MPI_Datatype dt, int peer, int tag, MPI_Comm comm, int persistent)
{
   comm_t *c = get_comm(comm);
   MPI_Request q = new_req();
   req_t *r = &g_req[q];
   r->kind = kind; r->buf = (void *)buf; r->count = count; r->dt = dt; r->peer = peer; r->tag = tag; r->comm = comm;
   r->ctx = c->ctx; r->persistent = persistent;
   return q;
}
int MPI_Isend(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request *req)
{
   *req = make_req(REQ_SEND, buf, count, dt, dest, tag, comm, 0);
   start_send(&g_req[*req]);
   return MPI_SUCCESS;
}
int MPI_Irsend(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request *req)
{
   return MPI_Isend(buf, count, dt, dest, tag, comm, req);
}
int MPI_Irecv(void *buf, int count, MPI_Datatype dt, int source, int tag, MPI_Comm comm, MPI_Request *req)
{
   *req = make_req(REQ_RECV, buf, count, dt, source, tag, comm, 0);
   start_recv(&g_req[*req], *req);
   return MPI_SUCCESS;
}
int MPI_Send_init(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request *req)
{
   *req = make_req(REQ_SEND, buf, count, dt, dest, tag, comm, 1);
   return MPI_SUCCESS;
}
int MPI_Recv_init(void *buf, int count, MPI_Datatype dt, int source, int tag, MPI_Comm comm, MPI_Request *req)
{
   *req = make_req(REQ_RECV, buf, count, dt, source, tag, comm, 1);
   return MPI_SUCCESS;
}
int MPI_Startall(int count, MPI_Request *reqs)
{
   for (int i = 0; i < count; i++)
   {
      if (reqs[i] == MPI_REQUEST_NULL) { continue; }
      req_t *r = &g_req[reqs[i]];
      if (r->kind == REQ_SEND) { start_send(r); } else { start_recv(r, reqs[i]); }
   }
   return MPI_SUCCESS;
}
int MPI_Request_free(MPI_Request *req)
{
   if (*req == MPI_REQUEST_NULL) { return MPI_SUCCESS; }
   req_t *r = &g_req[*req];
   if (r->posted)
   {
      for (int i = 0; i < g_nposted; i++)
      {
         if (g_posted[i] == *req) { memmove(&g_posted[i], &g_posted[i + 1], sizeof(int) * (size_t)(g_nposted - i - 1)); g_nposted--; break; }
      }
   }
   r->kind = REQ_FREE;
   *req = MPI_REQUEST_NULL;
   return MPI_SUCCESS;
}
int MPI_Wait(MPI_Request *req, MPI_Status *status)
{
   if (req_inactive(*req)) { return MPI_SUCCESS; }
   while (!g_req[*req].done) { progress(1000); }
   release_req(req, status);
   return MPI_SUCCESS;
}
int MPI_Waitall(int count, MPI_Request *reqs, MPI_Status *statuses)
{
   for (int i = 0; i < count; i++) { MPI_Wait(&reqs[i], statuses ? &statuses[i] : MPI_STATUS_IGNORE); }
   return MPI_SUCCESS;
}
int MPI_Waitany(int count, MPI_Request *reqs, int *index, MPI_Status *status)
{
   for (;;)
   {
      int live = 0;
      for (int i = 0; i < count; i++)
      {
         if (req_inactive(reqs[i])) { continue; }
         live = 1;
         if (g_req[reqs[i]].done) { *index = i; release_req(&reqs[i], status); return MPI_SUCCESS; }
      }
      if (!live) { *index = MPI_UNDEFINED; return MPI_SUCCESS; }
      progress(1000);
   }
}
int MPI_Test(MPI_Request *req, int *flag, MPI_Status *status)
{
   if (req_inactive(*req)) { *flag = 1; return MPI_SUCCESS; }
   progress(0);
   *flag = g_req[*req].done;
   if (*flag) { release_req(req, status); }
   return MPI_SUCCESS;
}
int MPI_Testall(int count, MPI_Request *reqs, int *flag, MPI_Status *statuses)
{
   progress(0);
   for (int i = 0; i < count; i++) { if (!req_inactive(reqs[i]) && !g_req[reqs[i]].done) { *flag = 0; return MPI_SUCCESS; } }
   *flag = 1;
   for (int i = 0; i < count; i++) { if (!req_inactive(reqs[i])) { release_req(&reqs[i], statuses ? &statuses[i] : MPI_STATUS_IGNORE); } }
   return MPI_SUCCESS;
}
int MPI_Send(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm)
{
   MPI_Request q;
   MPI_Isend(buf, count, dt, dest, tag, comm, &q);
   return MPI_Wait(&q, MPI_STATUS_IGNORE);
}
int MPI_Recv(void *buf, int count, MPI_Datatype dt, int source, int tag, MPI_Comm comm, MPI_Status *status)
{
   MPI_Request q;
   MPI_Irecv(buf, count, dt, source, tag, comm, &q);
   return MPI_Wait(&q, status);
}
int MPI_Iprobe(int source, int tag, MPI_Comm comm, int *flag, MPI_Status *status)
{
   comm_t *c = get_comm(comm);
   progress(0);
   msg_t *m = find_unexpected(c->ctx, source, tag, 0);
   *flag = m != NULL;
   if (m && status) { status->MPI_SOURCE = m->src; status->MPI_TAG = m->tag; status->MPI_ERROR = MPI_SUCCESS; status->shim_bytes = m->bytes; }
   return MPI_SUCCESS;
}
int MPI_Probe(int source, int tag, MPI_Comm comm, MPI_Status *status)
{
   int flag = 0;
   for (;;)
   {
      MPI_Iprobe(source, tag, comm, &flag, status);
      if (flag) { return MPI_SUCCESS; }
      progress(1000);
   }
}
int MPI_Get_count(const MPI_Status *status, MPI_Datatype dt, int *count)
{
   long sz = type_size(dt);
   *count = sz > 0 ? (int)(status->shim_bytes / sz) : 0;
   return MPI_SUCCESS;
}

/* ------------------------------------------------------------------ collectives (linear, deterministic) */
/* internal point-to-point on the collective context of the communicator (ctx + 1) */
static void coll_send(const void *buf, long bytes, int dest, int tag, comm_t *c)
{
   char *data = (char *)malloc((size_t)(bytes > 0 ? bytes : 1));
   if (bytes > 0) { memcpy(data, buf, (size_t)bytes); }
   post_message(c->world[dest], c->ctx + 1, c->rank, COLL_TAG + tag, data, (int)bytes);
}
static void coll_recv(void *buf, long bytes, int src, int tag, comm_t *c)
{
   msg_t *m;
   while (!(m = find_unexpected(c->ctx + 1, src, COLL_TAG + tag, 1))) { progress(1000); }
   if (m->bytes != bytes) { die("collective size mismatch: got %d bytes from rank %d, expected %ld", m->bytes, src, bytes); }
   if (bytes > 0) { memcpy(buf, m->data, (size_t)bytes); }
   free(m->data); free(m);
}

#define REDUCE_LOOP(T)                                                                   \
   {                                                                                     \
      const T *a = (const T *)in; T *b = (T *)inout;                                     \
      for (int i = 0; i < count; i++)                                                    \
      {                                                                                  \
         switch (op)                                                                     \
         {                                                                               \
            case 1: b[i] = a[i] + b[i]; break;                                           \
            case 2: b[i] = a[i] < b[i] ? a[i] : b[i]; break;                             \
            case 3: b[i] = a[i] > b[i] ? a[i] : b[i]; break;                             \
            case 4: b[i] = (T)((a[i] != 0) || (b[i] != 0)); break;                       \
            case 5: b[i] = (T)((a[i] != 0) && (b[i] != 0)); break;                       \
            case 8: b[i] = a[i] * b[i]; break;                                           \
            default: die("reduction operator %d not defined for this datatype", op);     \
         }                                                                               \
      }                                                                                  \
   }
#define REDUCE_LOOP_BITS(T)                                                              \
   {                                                                                     \
      const T *a = (const T *)in; T *b = (T *)inout;                                     \
      for (int i = 0; i < count; i++)                                                    \
      {                                                                                  \
         if (op == 6) { b[i] = a[i] | b[i]; } else if (op == 7) { b[i] = a[i] & b[i]; }  \
      }                                                                                  \
   }
/* inout = in (op) inout, `in` being the contribution of the lower ranks */
static void apply_op(const void *in, void *inout, int count, MPI_Datatype dt, MPI_Op op)
{
   if (op >= FIRST_USER_OP)
   {
      if (op - FIRST_USER_OP >= MAX_OPS || !g_op[op - FIRST_USER_OP]) { die("invalid user operator %d", op); }
      g_op[op - FIRST_USER_OP]((void *)in, inout, &count, &dt);
      return;
   }
   if (op == 6 || op == 7)
   {
      switch (dt)
      {
         case 1: case 2: case 15: REDUCE_LOOP_BITS(unsigned char) return;
         case 4: case 7: REDUCE_LOOP_BITS(unsigned int) return;
         case 5: case 8: REDUCE_LOOP_BITS(unsigned long) return;
         case 6: REDUCE_LOOP_BITS(unsigned long long) return;
         default: die("bitwise reduction on datatype %d", dt);
      }
   }
   switch (dt)
   {
      case 1: REDUCE_LOOP(char) break;
      case 2: case 15: REDUCE_LOOP(unsigned char) break;
      case 3: REDUCE_LOOP(short) break;
      case 4: REDUCE_LOOP(int) break;
      case 5: REDUCE_LOOP(long) break;
      case 6: REDUCE_LOOP(long long) break;
      case 7: REDUCE_LOOP(unsigned int) break;
      case 8: REDUCE_LOOP(unsigned long) break;
      case 9: REDUCE_LOOP(float) break;
      case 10: REDUCE_LOOP(double) break;
      case 11: REDUCE_LOOP(long double) break;
      default: die("reduction on datatype %d", dt);
   }
}

int MPI_Barrier(MPI_Comm comm)
{
   comm_t *c = get_comm(comm);
   char z = 0;
   if (c->size == 1) { return MPI_SUCCESS; }
   if (c->rank == 0)
   {
      for (int r = 1; r < c->size; r++) { coll_recv(&z, 1, r, 1, c); }
      for (int r = 1; r < c->size; r++) { coll_send(&z, 1, r, 2, c); }
   }
   else { coll_send(&z, 1, 0, 1, c); coll_recv(&z, 1, 0, 2, c); }
   return MPI_SUCCESS;
}
int MPI_Bcast(void *buf, int count, MPI_Datatype dt, int root, MPI_Comm comm)
{
   comm_t *c = get_comm(comm);
   if (c->size == 1) { return MPI_SUCCESS; }
   long bytes = (long)count * type_size(dt);
   char *tmp = (char *)malloc((size_t)(bytes > 0 ? bytes : 1));
   if (c->rank == root)
   {
      pack(buf, count, dt, tmp);
      for (int r = 0; r < c->size; r++) { if (r != root) { coll_send(tmp, bytes, r, 3, c); } }
   }
   else { coll_recv(tmp, bytes, root, 3, c); unpack(tmp, bytes, buf, count, dt); }
   free(tmp);
   return MPI_SUCCESS;
}
int MPI_Reduce(const void *sbuf, void *rbuf, int count, MPI_Datatype dt, MPI_Op op, int root, MPI_Comm comm)
{
   comm_t *c = get_comm(comm);
   long bytes = (long)count * type_size(dt);
   if (dt >= FIRST_DERIVED && g_type[dt].nblk != 1) { die("reduction on a non-contiguous datatype"); }
   const void *mine = (sbuf == MPI_IN_PLACE) ? rbuf : sbuf;
   if (c->rank != root) { coll_send(mine, bytes, root, 4, c); return MPI_SUCCESS; }
   /* acc = c_0 op c_1 op ... op c_{n-1}, combined in rank order */
   char *acc = (char *)malloc((size_t)(bytes > 0 ? bytes : 1)), *nxt = (char *)malloc((size_t)(bytes > 0 ? bytes : 1));
   for (int r = 0; r < c->size; r++)
   {
      char *dst = r == 0 ? acc : nxt;
      if (r == root) { memcpy(dst, mine, (size_t)bytes); } else { coll_recv(dst, bytes, r, 4, c); }
      if (r > 0) { apply_op(acc, nxt, count, dt, op); char *t = acc; acc = nxt; nxt = t; }
   }
   memcpy(rbuf, acc, (size_t)bytes);
   free(acc); free(nxt);
   return MPI_SUCCESS;
}
int MPI_Allreduce(const void *sbuf, void *rbuf, int count, MPI_Datatype dt, MPI_Op op, MPI_Comm comm)
{
   MPI_Reduce(sbuf, rbuf, count, dt, op, 0, comm);
   return MPI_Bcast(rbuf, count, dt, 0, comm);
}
int MPI_Scan(const void *sbuf, void *rbuf, int count, MPI_Datatype dt, MPI_Op op, MPI_Comm comm)
{
   comm_t *c = get_comm(comm);
   long bytes = (long)count * type_size(dt);
   const void *mine = (sbuf == MPI_IN_PLACE) ? rbuf : sbuf;
   char *own = (char *)malloc((size_t)(bytes > 0 ? bytes : 1));
   memcpy(own, mine, (size_t)bytes);
   if (c->rank > 0)
   {
      char *pre = (char *)malloc((size_t)(bytes > 0 ? bytes : 1));
      coll_recv(pre, bytes, c->rank - 1, 5, c);
      apply_op(pre, own, count, dt, op);
      free(pre);
   }
   if (c->rank + 1 < c->size) { coll_send(own, bytes, c->rank + 1, 5, c); }
   memcpy(rbuf, own, (size_t)bytes);
   free(own);
   return MPI_SUCCESS;
}
int MPI_Gatherv(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, const int *rcounts, const int *displs,
                MPI_Datatype rdt, int root, MPI_Comm comm)
{
   comm_t *c = get_comm(comm);
   long sbytes = (long)scount * type_size(sdt);
   if (c->rank != root)
   {
      char *tmp = (char *)malloc((size_t)(sbytes > 0 ? sbytes : 1));
      pack(sbuf, scount, sdt, tmp);
      coll_send(tmp, sbytes, root, 6, c);
      free(tmp);
      return MPI_SUCCESS;
   }
   long ext = type_extent(rdt), rsz = type_size(rdt);
   for (int r = 0; r < c->size; r++)
   {
      long bytes = (long)rcounts[r] * rsz;
      char *dst = (char *)rbuf + (long)displs[r] * ext;
      if (r == root)
      {
         if (sbuf != MPI_IN_PLACE)
         {
            char *tmp = (char *)malloc((size_t)(sbytes > 0 ? sbytes : 1));
            pack(sbuf, scount, sdt, tmp); unpack(tmp, sbytes, dst, rcounts[r], rdt); free(tmp);
         }
      }
      else
      {
         char *tmp = (char *)malloc((size_t)(bytes > 0 ? bytes : 1));
         coll_recv(tmp, bytes, r, 6, c); unpack(tmp, bytes, dst, rcounts[r], rdt); free(tmp);
      }
   }
   return MPI_SUCCESS;
}
int MPI_Gather(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, int rcount, MPI_Datatype rdt, int root, MPI_Comm comm)
{
   comm_t *c = get_comm(comm);
   int *cnt = (int *)malloc(sizeof(int) * (size_t)c->size), *dsp = (int *)malloc(sizeof(int) * (size_t)c->size);
   for (int r = 0; r < c->size; r++) { cnt[r] = rcount; dsp[r] = r * rcount; }
   int rc = MPI_Gatherv(sbuf, scount, sdt, rbuf, cnt, dsp, rdt, root, comm);
   free(cnt); free(dsp);
   return rc;
}
int MPI_Scatterv(const void *sbuf, const int *scounts, const int *displs, MPI_Datatype sdt, void *rbuf, int rcount,
                 MPI_Datatype rdt, int root, MPI_Comm comm)
{
   comm_t *c = get_comm(comm);
   long rbytes = (long)rcount * type_size(rdt);
   if (c->rank != root)
   {
      char *tmp = (char *)malloc((size_t)(rbytes > 0 ? rbytes : 1));
      coll_recv(tmp, rbytes, root, 7, c); unpack(tmp, rbytes, rbuf, rcount, rdt); free(tmp);
      return MPI_SUCCESS;
   }
   long ext = type_extent(sdt), ssz = type_size(sdt);
   for (int r = 0; r < c->size; r++)
   {
      long bytes = (long)scounts[r] * ssz;
      const char *src = (const char *)sbuf + (long)displs[r] * ext;
      char *tmp = (char *)malloc((size_t)(bytes > 0 ? bytes : 1));
      pack(src, scounts[r], sdt, tmp);
      if (r == root) { if (rbuf != MPI_IN_PLACE) { unpack(tmp, bytes, rbuf, rcount, rdt); } }
      else { coll_send(tmp, bytes, r, 7, c); }
      free(tmp);
   }
   return MPI_SUCCESS;
}
int MPI_Scatter(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, int rcount, MPI_Datatype rdt, int root, MPI_Comm comm)
{
   comm_t *c = get_comm(comm);
   int *cnt = (int *)malloc(sizeof(int) * (size_t)c->size), *dsp = (int *)malloc(sizeof(int) * (size_t)c->size);
   for (int r = 0; r < c->size; r++) { cnt[r] = scount; dsp[r] = r * scount; }
   int rc = MPI_Scatterv(sbuf, cnt, dsp, sdt, rbuf, rcount, rdt, root, comm);
   free(cnt); free(dsp);
   return rc;
}
int MPI_Allgatherv(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, const int *rcounts, const int *displs,
                   MPI_Datatype rdt, MPI_Comm comm)
{
   comm_t *c = get_comm(comm);
   /* in place: my block already sits in rbuf */
   const void *mine = sbuf; int mcount = scount; MPI_Datatype mdt = sdt;
   if (sbuf == MPI_IN_PLACE) { mine = (const char *)rbuf + (long)displs[c->rank] * type_extent(rdt); mcount = rcounts[c->rank]; mdt = rdt; }
   MPI_Gatherv(mine, mcount, mdt, rbuf, rcounts, displs, rdt, 0, comm);
   for (int r = 0; r < c->size; r++) { MPI_Bcast((char *)rbuf + (long)displs[r] * type_extent(rdt), rcounts[r], rdt, 0, comm); }
   return MPI_SUCCESS;
}
int MPI_Allgather(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, int rcount, MPI_Datatype rdt, MPI_Comm comm)
{
   comm_t *c = get_comm(comm);
   int *cnt = (int *)malloc(sizeof(int) * (size_t)c->size), *dsp = (int *)malloc(sizeof(int) * (size_t)c->size);
   for (int r = 0; r < c->size; r++) { cnt[r] = rcount; dsp[r] = r * rcount; }
   int rc = MPI_Allgatherv(sbuf, scount, sdt, rbuf, cnt, dsp, rdt, comm);
   free(cnt); free(dsp);
   return rc;
}
int MPI_Alltoall(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, int rcount, MPI_Datatype rdt, MPI_Comm comm)
{
   comm_t *c = get_comm(comm);
   long sbytes = (long)scount * type_size(sdt), rbytes = (long)rcount * type_size(rdt);
   for (int r = 0; r < c->size; r++)
   {
      char *tmp = (char *)malloc((size_t)(sbytes > 0 ? sbytes : 1));
      pack((const char *)sbuf + (long)r * scount * type_extent(sdt), scount, sdt, tmp);
      coll_send(tmp, sbytes, r, 8, c);
      free(tmp);
   }
   for (int r = 0; r < c->size; r++)
   {
      char *tmp = (char *)malloc((size_t)(rbytes > 0 ? rbytes : 1));
      coll_recv(tmp, rbytes, r, 8, c);
      unpack(tmp, rbytes, (char *)rbuf + (long)r * rcount * type_extent(rdt), rcount, rdt);
      free(tmp);
   }
   return MPI_SUCCESS;
}

/* ------------------------------------------------------------------ communicators and groups */
int MPI_Comm_size(MPI_Comm comm, int *size) { *size = get_comm(comm)->size; return MPI_SUCCESS; }
int MPI_Comm_rank(MPI_Comm comm, int *rank) { *rank = get_comm(comm)->rank; return MPI_SUCCESS; }
MPI_Comm MPI_Comm_f2c(MPI_Fint comm) { return (MPI_Comm)comm; }

/* all members of `comm` agree on a fresh range of `n` context pairs; returns its first id */
static int agree_ctx(MPI_Comm comm, int n)
{
   int mine = g_next_ctx, base = 0;
   MPI_Allreduce(&mine, &base, 1, MPI_INT, MPI_MAX, comm);
   g_next_ctx = base + 2 * n;
   return base;
}
int MPI_Comm_split(MPI_Comm comm, int color, int key, MPI_Comm *newcomm)
{
   comm_t *c = get_comm(comm);
   int n = c->size;
   int *ck = (int *)malloc(sizeof(int) * 2 * (size_t)n), mine[2] = { color, key };
   MPI_Allgather(mine, 2, MPI_INT, ck, 2, MPI_INT, comm);
   /* distinct colours in ascending order -> one context pair each */
   int *colors = (int *)malloc(sizeof(int) * (size_t)n), ncol = 0;
   for (int r = 0; r < n; r++)
   {
      int col = ck[2 * r], seen = 0;
      if (col == MPI_UNDEFINED) { continue; }
      for (int k = 0; k < ncol; k++) { if (colors[k] == col) { seen = 1; } }
      if (!seen) { colors[ncol++] = col; }
   }
   for (int a = 0; a < ncol; a++) { for (int b = a + 1; b < ncol; b++) { if (colors[b] < colors[a]) { int t = colors[a]; colors[a] = colors[b]; colors[b] = t; } } }
   int base = agree_ctx(comm, ncol > 0 ? ncol : 1);
   if (color == MPI_UNDEFINED) { *newcomm = MPI_COMM_NULL; free(ck); free(colors); return MPI_SUCCESS; }
   int ci = 0;
   while (colors[ci] != color) { ci++; }
   /* members of my colour ordered by (key, rank in the parent) */
   int *memb = (int *)malloc(sizeof(int) * (size_t)n), nm = 0;
   for (int r = 0; r < n; r++) { if (ck[2 * r] == color) { memb[nm++] = r; } }
   for (int a = 0; a < nm; a++)
   {
      for (int b = a + 1; b < nm; b++)
      {
         int ka = ck[2 * memb[a] + 1], kb = ck[2 * memb[b] + 1];
         if (kb < ka || (kb == ka && memb[b] < memb[a])) { int t = memb[a]; memb[a] = memb[b]; memb[b] = t; }
      }
   }
   int *world = (int *)malloc(sizeof(int) * (size_t)nm), myrank = -1;
   for (int k = 0; k < nm; k++) { world[k] = c->world[memb[k]]; if (memb[k] == c->rank) { myrank = k; } }
   *newcomm = new_comm(base + 2 * ci, nm, myrank, world);
   free(world); free(memb); free(ck); free(colors);
   return MPI_SUCCESS;
}
int MPI_Comm_dup(MPI_Comm comm, MPI_Comm *newcomm)
{
   comm_t *c = get_comm(comm);
   int base = agree_ctx(comm, 1);
   *newcomm = new_comm(base, c->size, c->rank, c->world);
   return MPI_SUCCESS;
}
int MPI_Comm_split_type(MPI_Comm comm, int split_type, int key, MPI_Info info, MPI_Comm *newcomm)
{
   (void)split_type; (void)info;
   return MPI_Comm_split(comm, 0, key, newcomm);      /* every rank lives on this one node */
}
int MPI_Comm_free(MPI_Comm *comm)
{
   if (*comm >= 2 && *comm < MAX_COMMS && g_comm[*comm].used) { free(g_comm[*comm].world); g_comm[*comm].used = 0; }
   *comm = MPI_COMM_NULL;
   return MPI_SUCCESS;
}
static MPI_Group new_group(int n, const int *world)
{
   for (int i = 1; i < MAX_GROUPS; i++)
   {
      if (!g_group[i].used)
      {
         g_group[i].used = 1; g_group[i].n = n;
         g_group[i].world = (int *)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
         memcpy(g_group[i].world, world, sizeof(int) * (size_t)n);
         return i;
      }
   }
   die("out of group slots");
   return MPI_GROUP_NULL;
}
int MPI_Comm_group(MPI_Comm comm, MPI_Group *group)
{
   comm_t *c = get_comm(comm);
   *group = new_group(c->size, c->world);
   return MPI_SUCCESS;
}
int MPI_Group_incl(MPI_Group group, int n, const int *ranks, MPI_Group *newgroup)
{
   if (group < 1 || group >= MAX_GROUPS || !g_group[group].used) { die("invalid group %d", group); }
   int *world = (int *)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
   for (int k = 0; k < n; k++)
   {
      if (ranks[k] < 0 || ranks[k] >= g_group[group].n) { die("MPI_Group_incl: rank %d outside the group", ranks[k]); }
      world[k] = g_group[group].world[ranks[k]];
   }
   *newgroup = new_group(n, world);
   free(world);
   return MPI_SUCCESS;
}
int MPI_Group_free(MPI_Group *group)
{
   if (*group >= 1 && *group < MAX_GROUPS && g_group[*group].used) { free(g_group[*group].world); g_group[*group].used = 0; }
   *group = MPI_GROUP_NULL;
   return MPI_SUCCESS;
}
/* collective over `comm`; ranks outside the group receive MPI_COMM_NULL */
int MPI_Comm_create(MPI_Comm comm, MPI_Group group, MPI_Comm *newcomm)
{
   if (group < 1 || group >= MAX_GROUPS || !g_group[group].used) { die("invalid group %d", group); }
   int base = agree_ctx(comm, 1), myrank = -1;
   group_t *g = &g_group[group];
   for (int k = 0; k < g->n; k++) { if (g->world[k] == g_rank) { myrank = k; } }
   *newcomm = myrank < 0 ? MPI_COMM_NULL : new_comm(base, g->n, myrank, g->world);
   return MPI_SUCCESS;
}
int MPI_Info_create(MPI_Info *info) { *info = 1; return MPI_SUCCESS; }
int MPI_Info_free(MPI_Info *info) { *info = MPI_INFO_NULL; return MPI_SUCCESS; }

/* ------------------------------------------------------------------ operators and datatypes */
int MPI_Op_create(MPI_User_function *fn, int commute, MPI_Op *op)
{
   (void)commute;
   for (int i = 0; i < MAX_OPS; i++) { if (!g_op[i]) { g_op[i] = fn; *op = FIRST_USER_OP + i; return MPI_SUCCESS; } }
   die("out of operator slots");
   return MPI_ERR_OTHER;
}
int MPI_Op_free(MPI_Op *op)
{
   if (*op >= FIRST_USER_OP && *op - FIRST_USER_OP < MAX_OPS) { g_op[*op - FIRST_USER_OP] = NULL; }
   *op = MPI_OP_NULL;
   return MPI_SUCCESS;
}
int MPI_Type_contiguous(int count, MPI_Datatype oldtype, MPI_Datatype *newtype)
{
   MPI_Datatype t = new_type(count * type_nblk(oldtype));
   int pos = 0;
   append_blocks(&g_type[t], &pos, oldtype, count, 0);
   finish_type(&g_type[t], (long)count * type_extent(oldtype));
   *newtype = t;
   return MPI_SUCCESS;
}
int MPI_Type_create_hvector(int count, int blocklength, MPI_Aint stride, MPI_Datatype oldtype, MPI_Datatype *newtype)
{
   MPI_Datatype t = new_type(count * blocklength * type_nblk(oldtype));
   int pos = 0;
   for (int k = 0; k < count; k++) { append_blocks(&g_type[t], &pos, oldtype, blocklength, (long)k * (long)stride); }
   finish_type(&g_type[t], count > 0 ? (long)(count - 1) * (long)stride + (long)blocklength * type_extent(oldtype) : 0);
   *newtype = t;
   return MPI_SUCCESS;
}
int MPI_Type_hvector(int count, int blocklength, MPI_Aint stride, MPI_Datatype oldtype, MPI_Datatype *newtype)
{
   return MPI_Type_create_hvector(count, blocklength, stride, oldtype, newtype);
}
int MPI_Type_vector(int count, int blocklength, int stride, MPI_Datatype oldtype, MPI_Datatype *newtype)
{
   return MPI_Type_create_hvector(count, blocklength, (MPI_Aint)stride * type_extent(oldtype), oldtype, newtype);
}
int MPI_Type_create_struct(int count, const int *blocklens, const MPI_Aint *displs, const MPI_Datatype *types, MPI_Datatype *newtype)
{
   int nblk = 0;
   for (int k = 0; k < count; k++) { nblk += blocklens[k] * type_nblk(types[k]); }
   MPI_Datatype t = new_type(nblk);
   int pos = 0;
   for (int k = 0; k < count; k++) { append_blocks(&g_type[t], &pos, types[k], blocklens[k], (long)displs[k]); }
   finish_type(&g_type[t], 0);
   *newtype = t;
   return MPI_SUCCESS;
}
int MPI_Type_struct(int count, int *blocklens, MPI_Aint *displs, MPI_Datatype *types, MPI_Datatype *newtype)
{
   return MPI_Type_create_struct(count, blocklens, displs, types, newtype);
}
int MPI_Type_commit(MPI_Datatype *dt) { (void)dt; return MPI_SUCCESS; }
int MPI_Type_free(MPI_Datatype *dt)
{
   if (*dt >= FIRST_DERIVED && *dt < MAX_TYPES && g_type[*dt].used) { free(g_type[*dt].blk); g_type[*dt].used = 0; }
   *dt = MPI_DATATYPE_NULL;
   return MPI_SUCCESS;
}
int MPI_Get_address(const void *location, MPI_Aint *address) { *address = (MPI_Aint)location; return MPI_SUCCESS; }
int MPI_Address(void *location, MPI_Aint *address) { *address = (MPI_Aint)location; return MPI_SUCCESS; }
