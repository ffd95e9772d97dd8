/* parallel.c -- the wrappers of parallel.h:17-63 without MPI (the image has none).
 *
 * One process per GPU on ONE node.  Rank and size come from the launcher's environment
 * (RANK / WORLD_SIZE / LOCAL_RANK as set by torchrun --no-python or tools/comd-launch;
 * COMD_RANK / COMD_NRANKS also accepted).  The small host collectives CoMD needs (sums of a few
 * ints or reals, min/max-loc of the timers, one broadcast, barriers) run over a
 * POSIX shared-memory segment: every rank writes its slot, a sense-reversing barrier, every rank
 * reads all slots and reduces them in rank order -- deterministic, no device involved.  The atom
 * and F'(rho) halos do NOT come through here: they stay on the GPUs (NCCL, comd_comm.cu); this
 * layer only hands the NCCL id from rank 0 to the others.
 */
#define _GNU_SOURCE
#include "comd_host.h"
#include <fcntl.h>
#include <sched.h>
#include <stdatomic.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <time.h>
#include <unistd.h>

#define MAX_RANKS 64
#define SLOT_BYTES 1024
#define SHM_MAGIC 0x434f4d44

typedef struct
{
   atomic_int magic, arrived, generation, attached;
   unsigned char slot[MAX_RANKS][SLOT_BYTES];
} Shared;

static int g_rank = 0, g_nRanks = 1, g_localRank = 0;
static Shared* g_shm = NULL;
static char g_shmName[128] = "";

static int envInt(const char* a, const char* b, int dflt)
{
   const char* v = getenv(a);
   if (!v && b) v = getenv(b);
   return v ? atoi(v) : dflt;
}

int getNRanks(void) { return g_nRanks; }
int getMyRank(void) { return g_rank; }
int printRank(void) { return g_rank == 0; }
int localDeviceId(void) { return g_localRank; }

void barrierParallel(void)
{
   if (g_nRanks == 1) return;
   const int gen = atomic_load(&g_shm->generation);
   if (atomic_fetch_add(&g_shm->arrived, 1) + 1 == g_nRanks)
   {
      atomic_store(&g_shm->arrived, 0);
      atomic_store(&g_shm->generation, gen + 1);
   }
   else
      while (atomic_load(&g_shm->generation) == gen) sched_yield();
}

void initParallel(int* argc, char*** argv)
{
   (void)argc; (void)argv;
   g_rank = envInt("COMD_RANK", "RANK", 0);
   g_nRanks = envInt("COMD_NRANKS", "WORLD_SIZE", 1);
   g_localRank = envInt("COMD_LOCAL_RANK", "LOCAL_RANK", g_rank);
   if (g_nRanks > MAX_RANKS) { fprintf(stderr, "initParallel: at most %d ranks\n", MAX_RANKS); exit(-1); }
   if (g_nRanks == 1) return;
   const char* job = getenv("COMD_JOB");
   if (!job) job = getenv("MASTER_PORT");
   if (!job) { fprintf(stderr, "initParallel: COMD_JOB or MASTER_PORT must name the job for %d ranks\n", g_nRanks); exit(-1); }
   /* A per-launch nonce in the name: a segment left behind by a crashed run with the same MASTER_PORT and uid
    * still carries a valid magic, and a non-root rank that opened it before rank 0 had re-created it would wait
    * on stale counters for ever.  All ranks are children of one launcher (comd-launch, torchrun), so its pid
    * tells launches apart; COMD_JOB_NONCE overrides it for launchers that interpose a shell per rank. */
   const char* nonceEnv = getenv("COMD_JOB_NONCE");
   const long nonce = nonceEnv ? atol(nonceEnv) : (long)getppid();
   snprintf(g_shmName, sizeof g_shmName, "/comd_b200_%s_%d_%ld", job, (int)getuid(), nonce);
   int fd = -1;
   if (g_rank == 0)
   {
      shm_unlink(g_shmName);
      fd = shm_open(g_shmName, O_CREAT | O_EXCL | O_RDWR, 0600);
      if (fd < 0 || ftruncate(fd, sizeof(Shared)) != 0) { perror("initParallel: shm_open"); exit(-1); }
   }
   else
   {
      for (int tries = 0; fd < 0; ++tries)
      {
         fd = shm_open(g_shmName, O_RDWR, 0600);
         if (fd < 0) { if (tries > 60000) { perror("initParallel: waiting for rank 0"); exit(-1); } usleep(1000); }
      }
   }
   struct timespec nap = { 0, 1000000 };
   for (;;)
   {  /* the segment has its full size only after rank 0's ftruncate */
      off_t len = lseek(fd, 0, SEEK_END);
      if (len >= (off_t)sizeof(Shared)) break;
      nanosleep(&nap, NULL);
   }
   g_shm = mmap(NULL, sizeof(Shared), PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
   if (g_shm == MAP_FAILED) { perror("initParallel: mmap"); exit(-1); }
   close(fd);
   if (g_rank == 0)
   {
      atomic_store(&g_shm->arrived, 0);
      atomic_store(&g_shm->generation, 0);
      atomic_store(&g_shm->attached, 0);
      atomic_store(&g_shm->magic, SHM_MAGIC);
   }
   else
      while (atomic_load(&g_shm->magic) != SHM_MAGIC) nanosleep(&nap, NULL);
   atomic_fetch_add(&g_shm->attached, 1);
   /* nobody enters the first barrier before everybody is looking at THIS segment */
   while (atomic_load(&g_shm->attached) < g_nRanks) nanosleep(&nap, NULL);
   barrierParallel();
}

void destroyParallel(void)
{
   if (g_nRanks == 1 || !g_shm) return;
   barrierParallel();
   if (g_rank == 0) { atomic_store(&g_shm->magic, 0); shm_unlink(g_shmName); }
   munmap(g_shm, sizeof(Shared));
   g_shm = NULL;
}

void attachParallelDevice(ComdDev* dev) { (void)dev; }

void timestampBarrier(const char* msg)
{
   barrierParallel();
   if (!printRank()) return;
   time_t t = time(NULL);
   char* stamp = ctime(&t);
   fprintf(screenOut, "%s: %s", stamp ? strtok(stamp, "\n") : "", msg);
   fflush(screenOut);
}

/* every rank contributes `bytes`; afterwards all[r] holds rank r's contribution */
static void allGather(const void* mine, int bytes, unsigned char all[][SLOT_BYTES])
{
   if (bytes > SLOT_BYTES) { fprintf(stderr, "parallel: message of %d bytes exceeds the %d-byte slot\n", bytes, SLOT_BYTES); exit(-1); }
   memcpy(g_shm->slot[g_rank], mine, bytes);
   barrierParallel();
   for (int r = 0; r < g_nRanks; ++r) memcpy(all[r], g_shm->slot[r], bytes);
   barrierParallel();
}

#define REDUCE(T, send, recv, count, OP) do { \
   if (g_nRanks == 1) { for (int i_ = 0; i_ < (count); ++i_) (recv)[i_] = (send)[i_]; break; } \
   static unsigned char all_[MAX_RANKS][SLOT_BYTES]; \
   allGather((send), (count) * (int)sizeof(T), all_); \
   for (int i_ = 0; i_ < (count); ++i_) { \
      T acc_ = ((T*)all_[0])[i_]; \
      for (int r_ = 1; r_ < g_nRanks; ++r_) { const T v_ = ((T*)all_[r_])[i_]; OP; } \
      (recv)[i_] = acc_; } } while (0)

void addIntParallel(int* s, int* r, int n)          { REDUCE(int, s, r, n, acc_ += v_); }
void addRealParallel(real_t* s, real_t* r, int n)   { REDUCE(real_t, s, r, n, acc_ += v_); }
void addDoubleParallel(double* s, double* r, int n) { REDUCE(double, s, r, n, acc_ += v_); }
void maxIntParallel(int* s, int* r, int n)          { REDUCE(int, s, r, n, if (v_ > acc_) acc_ = v_); }
void minRankDoubleParallel(RankReduceData* s, RankReduceData* r, int n) { REDUCE(RankReduceData, s, r, n, if (v_.val < acc_.val) acc_ = v_); }
void maxRankDoubleParallel(RankReduceData* s, RankReduceData* r, int n) { REDUCE(RankReduceData, s, r, n, if (v_.val > acc_.val) acc_ = v_); }

void bcastParallel(void* buf, int len, int root)
{
   if (g_nRanks == 1) return;
   for (int off = 0; off < len; off += SLOT_BYTES)
   {
      const int chunk = len - off < SLOT_BYTES ? len - off : SLOT_BYTES;
      if (g_rank == root) memcpy(g_shm->slot[root], (char*)buf + off, chunk);
      barrierParallel();
      if (g_rank != root) memcpy((char*)buf + off, g_shm->slot[root], chunk);
      barrierParallel();
   }
}

/* NCCL bootstrap: rank 0 makes the id, everyone gets it */
int parallelCommId(unsigned char id[COMD_COMM_ID_BYTES])
{
   int rc = COMD_OK;
   if (g_rank == 0) rc = comd_comm_unique_id(id);
   bcastParallel(&rc, sizeof rc, 0);
   bcastParallel(id, COMD_COMM_ID_BYTES, 0);
   return rc;
}
