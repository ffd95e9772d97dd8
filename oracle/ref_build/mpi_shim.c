/* mpi_shim.c -- single-node stand-in for the nine MPI calls of the reference (see shim_mpi/mpi.h).
 *
 * MPI_Init maps one anonymous MAP_SHARED region and forks R - 1 children (R = $COMD_SHIM_NRANKS, default 1); the
 * parent is rank 0 and reaps the children in MPI_Finalize.  Point-to-point: every rank owns a mailbox of SLOTS
 * message slots; MPI_Sendrecv first posts its message into a free slot of the destination (stamped with a
 * per-(source, destination) sequence number, so two messages between the same pair keep their order), then waits for
 * the next message of `source` in its own mailbox.  Posting never waits on the receiver's progress beyond a free
 * slot, so the ring-shaped exchanges of haloExchange.c:1493-1522 cannot deadlock.  Collectives: every rank writes
 * its slot, barrier, every rank reduces all slots in rank order (deterministic), barrier.
 * Test infrastructure only -- never linked into the product. */
#define _GNU_SOURCE
#include "mpi.h"
#include <sched.h>
#include <stdatomic.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>

#define MAX_RANKS 256
#define SLOTS 8
#define COLL_BYTES 8192
#define BCAST_BYTES (1 << 20)

typedef struct { atomic_int state; int src; unsigned seq; int len; } SlotHdr;      /* state: 0 free, 1 being filled, 2 full */
typedef struct
{
   atomic_int arrived, generation;
   atomic_int failed;
   unsigned char coll[MAX_RANKS][COLL_BYTES];
   unsigned char bcast[BCAST_BYTES];
   SlotHdr hdr[MAX_RANKS][SLOTS];
} Shared;

static Shared* S = NULL;
static unsigned char* g_data = NULL;          /* [rank][slot][slotBytes] */
static size_t g_slotBytes = 0;
static int g_rank = 0, g_n = 1;
static pid_t g_kids[MAX_RANKS];
static unsigned g_seqOut[MAX_RANKS], g_seqIn[MAX_RANKS];

static inline void relax(int* spins) { if (++*spins > 200) { sched_yield(); *spins = 0; } }
static void die(const char* msg) { fprintf(stderr, "mpi_shim (rank %d): %s\n", g_rank, msg); if (S) atomic_store(&S->failed, 1); _exit(3); }

static void barrier(void)
{
   if (g_n == 1) return;
   const int gen = atomic_load(&S->generation);
   if (atomic_fetch_add(&S->arrived, 1) + 1 == g_n) { atomic_store(&S->arrived, 0); atomic_store(&S->generation, gen + 1); }
   else { int sp = 0; while (atomic_load(&S->generation) == gen) { relax(&sp); if (atomic_load(&S->failed)) _exit(3); } }
}

int MPI_Init(int* argc, char*** argv)
{
   (void)argc; (void)argv;
   const char* e = getenv("COMD_SHIM_NRANKS");
   g_n = e ? atoi(e) : 1;
   if (g_n < 1 || g_n > MAX_RANKS) { fprintf(stderr, "mpi_shim: COMD_SHIM_NRANKS must be 1..%d\n", MAX_RANKS); exit(2); }
   const char* sb = getenv("COMD_SHIM_SLOT_MB");
   g_slotBytes = (size_t)(sb ? atoi(sb) : 8) << 20;
   S = mmap(NULL, sizeof(Shared), PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
   g_data = mmap(NULL, (size_t)g_n * SLOTS * g_slotBytes, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
   if (S == MAP_FAILED || g_data == MAP_FAILED) { perror("mpi_shim: mmap"); exit(2); }
   memset(S, 0, sizeof(Shared));
   fflush(NULL);
   g_rank = 0;
   for (int r = 1; r < g_n; ++r)
   {
      pid_t p = fork();
      if (p < 0) { perror("mpi_shim: fork"); exit(2); }
      if (p == 0) { g_rank = r; break; }
      g_kids[r] = p;
   }
   barrier();
   return MPI_SUCCESS;
}

int MPI_Finalize(void)
{
   barrier();
   fflush(NULL);
   if (g_rank != 0) _exit(0);                 /* children end here: rank 0 alone returns to main and reports */
   for (int r = 1; r < g_n; ++r) { int st; waitpid(g_kids[r], &st, 0); }
   return MPI_SUCCESS;
}

int MPI_Comm_rank(MPI_Comm c, int* rank) { (void)c; *rank = g_rank; return MPI_SUCCESS; }
int MPI_Comm_size(MPI_Comm c, int* size) { (void)c; *size = g_n; return MPI_SUCCESS; }
int MPI_Barrier(MPI_Comm c) { (void)c; barrier(); return MPI_SUCCESS; }

static size_t typeSize(MPI_Datatype t)
{
   switch (t) { case MPI_BYTE: case MPI_CHAR: return 1; case MPI_INT: case MPI_FLOAT: return 4; case MPI_DOUBLE: return 8; case MPI_DOUBLE_INT: return 16; }
   die("unknown datatype"); return 0;
}

int MPI_Sendrecv(const void* sendbuf, int sendcount, MPI_Datatype sendtype, int dest, int sendtag,
                 void* recvbuf, int recvcount, MPI_Datatype recvtype, int source, int recvtag, MPI_Comm c, MPI_Status* status)
{
   (void)sendtag; (void)recvtag; (void)c;
   const size_t sbytes = (size_t)sendcount * typeSize(sendtype), rcap = (size_t)recvcount * typeSize(recvtype);
   if (sbytes > g_slotBytes) die("message larger than a mailbox slot (COMD_SHIM_SLOT_MB)");
   if (g_n == 1) { if (sbytes > rcap) die("receive buffer too small"); memcpy(recvbuf, sendbuf, sbytes); if (status) status->count = (int)sbytes; return MPI_SUCCESS; }
   /* post */
   int sp = 0, slot = -1;
   while (slot < 0)
   {
      for (int k = 0; k < SLOTS && slot < 0; ++k)
      {
         int zero = 0;
         if (atomic_compare_exchange_strong(&S->hdr[dest][k].state, &zero, 1)) slot = k;
      }
      if (slot < 0) { relax(&sp); if (atomic_load(&S->failed)) _exit(3); }
   }
   SlotHdr* h = &S->hdr[dest][slot];
   memcpy(g_data + ((size_t)dest * SLOTS + slot) * g_slotBytes, sendbuf, sbytes);
   h->src = g_rank; h->seq = g_seqOut[dest]++; h->len = (int)sbytes;
   atomic_store(&h->state, 2);
   /* receive the next message of `source` */
   const unsigned want = g_seqIn[source]++;
   sp = 0;
   for (;;)
   {
      for (int k = 0; k < SLOTS; ++k)
      {
         SlotHdr* q = &S->hdr[g_rank][k];
         if (atomic_load(&q->state) == 2 && q->src == source && q->seq == want)
         {
            if ((size_t)q->len > rcap) die("receive buffer too small");
            memcpy(recvbuf, g_data + ((size_t)g_rank * SLOTS + k) * g_slotBytes, (size_t)q->len);
            if (status) { status->count = q->len; status->MPI_SOURCE = source; status->MPI_TAG = 0; }
            atomic_store(&q->state, 0);
            return MPI_SUCCESS;
         }
      }
      relax(&sp);
      if (atomic_load(&S->failed)) _exit(3);
   }
}

int MPI_Get_count(const MPI_Status* status, MPI_Datatype t, int* count) { *count = (int)((size_t)status->count / typeSize(t)); return MPI_SUCCESS; }

typedef struct { double val; int rank; } DoubleInt;

int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype t, MPI_Op op, MPI_Comm c)
{
   (void)c;
   const size_t bytes = (size_t)count * typeSize(t);
   if (bytes > COLL_BYTES) die("Allreduce message larger than the collective slot");
   if (g_n == 1) { memmove(recvbuf, sendbuf, bytes); return MPI_SUCCESS; }
   memcpy(S->coll[g_rank], sendbuf, bytes);
   barrier();
   for (int i = 0; i < count; ++i)
   {
      if (t == MPI_INT)
      {
         int acc = ((int*)S->coll[0])[i];
         for (int r = 1; r < g_n; ++r) { const int v = ((int*)S->coll[r])[i]; if (op == MPI_SUM) acc += v; else if (op == MPI_MAX) { if (v > acc) acc = v; } else die("op on MPI_INT"); }
         ((int*)recvbuf)[i] = acc;
      }
      else if (t == MPI_DOUBLE)
      {
         double acc = ((double*)S->coll[0])[i];
         for (int r = 1; r < g_n; ++r) { const double v = ((double*)S->coll[r])[i]; if (op == MPI_SUM) acc += v; else if (op == MPI_MAX) { if (v > acc) acc = v; } else die("op on MPI_DOUBLE"); }
         ((double*)recvbuf)[i] = acc;
      }
      else if (t == MPI_FLOAT)
      {
         float acc = ((float*)S->coll[0])[i];
         for (int r = 1; r < g_n; ++r) { const float v = ((float*)S->coll[r])[i]; if (op == MPI_SUM) acc += v; else if (op == MPI_MAX) { if (v > acc) acc = v; } else die("op on MPI_FLOAT"); }
         ((float*)recvbuf)[i] = acc;
      }
      else if (t == MPI_DOUBLE_INT)
      {
         DoubleInt acc = ((DoubleInt*)S->coll[0])[i];
         for (int r = 1; r < g_n; ++r)
         {
            const DoubleInt v = ((DoubleInt*)S->coll[r])[i];
            if (op == MPI_MINLOC) { if (v.val < acc.val) acc = v; } else if (op == MPI_MAXLOC) { if (v.val > acc.val) acc = v; } else die("op on MPI_DOUBLE_INT");
         }
         ((DoubleInt*)recvbuf)[i] = acc;
      }
      else die("Allreduce datatype");
   }
   barrier();
   return MPI_SUCCESS;
}

int MPI_Bcast(void* buf, int count, MPI_Datatype t, int root, MPI_Comm c)
{
   (void)c;
   if (g_n == 1) return MPI_SUCCESS;
   const size_t bytes = (size_t)count * typeSize(t);
   for (size_t off = 0; off < bytes; off += BCAST_BYTES)
   {
      const size_t chunk = bytes - off < BCAST_BYTES ? bytes - off : BCAST_BYTES;
      if (g_rank == root) memcpy(S->bcast, (char*)buf + off, chunk);
      barrier();
      if (g_rank != root) memcpy((char*)buf + off, S->bcast, chunk);
      barrier();
   }
   return MPI_SUCCESS;
}
