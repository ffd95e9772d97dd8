// comd_comm.cu -- the exchange step of the halo phases: self exchange for an axis with a
// single rank (what the serial reference does with memcpy, parallel.c:112-117) and
// ncclSend/ncclRecv over NVLink between GPUs otherwise (replaces MPI_Sendrecv,
// haloExchange.c:1514-1515, and the libmp layer of comm.cc).  Buffers never leave HBM.
//
// NCCL is bound at run time (dlopen) so that the same library works inside a torch
// process (which already carries its own libnccl.so.2) and in the plain-C driver.
#include "comd_dev.cuh"
#include <dlfcn.h>
#include <cstdio>
#include <cstring>

// ---- minimal NCCL surface (nccl.h 2.x ABI) ------------------------------------------
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { ncclInt32 = 2, ncclInt64 = 4, ncclFloat32 = 7, ncclFloat64 = 8 };   // ncclDataType_t values
enum { ncclSum = 0 };

struct NcclApi
{
   void* h;
   ncclResult_t (*GetUniqueId)(ncclUniqueId*);
   ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
   ncclResult_t (*CommDestroy)(ncclComm_t);
   ncclResult_t (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t);
   ncclResult_t (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t);
   ncclResult_t (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t);
   ncclResult_t (*GroupStart)(void);
   ncclResult_t (*GroupEnd)(void);
   const char* (*GetErrorString)(ncclResult_t);
};
static NcclApi g_nccl = {};

static int ncclLoad()
{
   if (g_nccl.h) return COMD_OK;
   const char* names[] = { "libnccl.so.2", "libnccl.so", nullptr };
   void* h = nullptr;
   for (int i = 0; names[i] && !h; ++i) h = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
   if (!h) { comdSetError("NCCL: cannot dlopen libnccl.so.2 (%s)", dlerror()); return COMD_E_COMM; }
#define SYM(field, name) *(void**)(&g_nccl.field) = dlsym(h, name); if (!g_nccl.field) { comdSetError("NCCL: missing symbol %s", name); return COMD_E_COMM; }
   SYM(GetUniqueId, "ncclGetUniqueId") SYM(CommInitRank, "ncclCommInitRank") SYM(CommDestroy, "ncclCommDestroy")
   SYM(Send, "ncclSend") SYM(Recv, "ncclRecv") SYM(AllReduce, "ncclAllReduce")
   SYM(GroupStart, "ncclGroupStart") SYM(GroupEnd, "ncclGroupEnd") SYM(GetErrorString, "ncclGetErrorString")
#undef SYM
   g_nccl.h = h;
   return COMD_OK;
}
#define NCCL_TRY(call) do { ncclResult_t r_ = (call); if (r_ != 0) { \
   comdSetError("%s:%d: %s -> %s", __FILE__, __LINE__, #call, g_nccl.GetErrorString(r_)); return COMD_E_COMM; } } while (0)

extern "C" int comd_comm_unique_id(unsigned char id[COMD_COMM_ID_BYTES])
{
   int rc = ncclLoad(); if (rc) return rc;
   ncclUniqueId u;
   NCCL_TRY(g_nccl.GetUniqueId(&u));
   memcpy(id, &u, COMD_COMM_ID_BYTES);
   return COMD_OK;
}

extern "C" int comd_comm_init(ComdDev* d, const unsigned char id[COMD_COMM_ID_BYTES], int rank, int nRanks)
{
   if (!d || !id) return COMD_E_ARG;
   if (rank != d->rank || nRanks != d->nRanks)
   { comdSetError("comd_comm_init: rank %d/%d does not match the geometry's %d/%d", rank, nRanks, d->rank, d->nRanks); return COMD_E_ARG; }
   int rc = ncclLoad(); if (rc) return rc;
   CUDA_TRY(cudaSetDevice(d->device));
   ncclUniqueId u; memcpy(&u, id, COMD_COMM_ID_BYTES);
   ncclComm_t c;
   NCCL_TRY(g_nccl.CommInitRank(&c, nRanks, u, rank));
   d->ncclComm = c;
   return COMD_OK;
}

extern "C" int comd_comm_allreduce_sum(ComdDev* d, double* inout, int n)
{
   if (!d || !inout || n < 0 || n > 64) return COMD_E_ARG;
   if (d->nRanks == 1 || !d->ncclComm) return COMD_OK;
   CUDA_TRY(cudaSetDevice(d->device));
   double* tmp = d->redPartial;       // scratch, > 64 doubles
   CUDA_TRY(cudaMemcpyAsync(tmp, inout, n * sizeof(double), cudaMemcpyHostToDevice, d->stream));
   NCCL_TRY(g_nccl.AllReduce(tmp, tmp, n, ncclFloat64, ncclSum, (ncclComm_t)d->ncclComm, d->stream));
   CUDA_TRY(cudaMemcpyAsync(inout, tmp, n * sizeof(double), cudaMemcpyDeviceToHost, d->stream));
   CUDA_TRY(cudaStreamSynchronize(d->stream));
   return COMD_OK;
}

extern "C" int comd_comm_allreduce_sum_int(ComdDev* d, long* inout, int n)
{
   if (!d || !inout || n < 0 || n > 64) return COMD_E_ARG;
   if (d->nRanks == 1 || !d->ncclComm) return COMD_OK;
   CUDA_TRY(cudaSetDevice(d->device));
   long* tmp = (long*)d->redPartial;
   CUDA_TRY(cudaMemcpyAsync(tmp, inout, n * sizeof(long), cudaMemcpyHostToDevice, d->stream));
   NCCL_TRY(g_nccl.AllReduce(tmp, tmp, n, ncclInt64, ncclSum, (ncclComm_t)d->ncclComm, d->stream));
   CUDA_TRY(cudaMemcpyAsync(inout, tmp, n * sizeof(long), cudaMemcpyDeviceToHost, d->stream));
   CUDA_TRY(cudaStreamSynchronize(d->stream));
   return COMD_OK;
}

extern "C" int comd_comm_barrier(ComdDev* d)
{
   long one = 1;
   return comd_comm_allreduce_sum_int(d, &one, 1);
}

// ---- self exchange: what left across one face comes back across the opposite one ----
__global__ void k_self_counts(int* cnt, int sendBase, int recvBase, int fM)
{
   cnt[recvBase + fM] = cnt[sendBase + fM + 1];       // across my - face arrives what the (same) rank sent across its + face
   cnt[recvBase + fM + 1] = cnt[sendBase + fM];
}

template <typename T> static inline void swapPtr(T*& a, T*& b) { T* t = a; a = b; b = t; }

// Generic two-message phase: sendBuf[0] -> rank nbrM, sendBuf[1] -> rank nbrP;
// recvBuf[0] <- nbrM (its + face message), recvBuf[1] <- nbrP (its - face message).
// `cache`: 12 host ints (6 send counts, 6 receive counts) of this message family.  When `useCache`
// is set the counts are the ones of the previous exchange (list mode between rebuilds: same atoms,
// same order), so neither the count messages nor the host synchronisation are needed.
static int exchangePhase(ComdDev* d, int ax, real_t* sendBuf[2], real_t* recvBuf[2], int sendBase, int recvBase, int doublesPerItem,
                         int* cache = nullptr, bool useCache = false)
{
   const int fM = 2 * ax, fP = fM + 1;
   cudaStream_t s = d->xs;
   if (d->nbrRank[fM] == d->rank)
   {
      k_self_counts<<<1, 1, 0, s>>>(d->cnt, sendBase, recvBase, fM); LAUNCH_CHECK(d);
      swapPtr(recvBuf[0], sendBuf[1]);
      swapPtr(recvBuf[1], sendBuf[0]);
      return COMD_OK;
   }
   if (!d->ncclComm) { comdSetError("halo exchange across ranks needs comd_comm_init first"); return COMD_E_STATE; }
   ncclComm_t comm = (ncclComm_t)d->ncclComm;
   const int nbrM = d->nbrRank[fM], nbrP = d->nbrRank[fP];
   if (cache && useCache)
   {
      const long nSM = cache[fM], nSP = cache[fP], nRM = cache[6 + fM], nRP = cache[6 + fP];
      NCCL_TRY(g_nccl.GroupStart());
      if (nSP) NCCL_TRY(g_nccl.Send(sendBuf[1], nSP * doublesPerItem, COMD_NCCL_REAL, nbrP, comm, s));
      if (nSM) NCCL_TRY(g_nccl.Send(sendBuf[0], nSM * doublesPerItem, COMD_NCCL_REAL, nbrM, comm, s));
      if (nRM) NCCL_TRY(g_nccl.Recv(recvBuf[0], nRM * doublesPerItem, COMD_NCCL_REAL, nbrM, comm, s));
      if (nRP) NCCL_TRY(g_nccl.Recv(recvBuf[1], nRP * doublesPerItem, COMD_NCCL_REAL, nbrP, comm, s));
      NCCL_TRY(g_nccl.GroupEnd());
      return COMD_OK;
   }
   // 1. counts travel first (the receiver must size its ncclRecv; the reference leans on
   //    MPI_Get_count, parallel.c:109).  Order matters when nbrM == nbrP (two ranks on the
   //    axis): sends to one peer match its receives in issue order.
   NCCL_TRY(g_nccl.GroupStart());
   NCCL_TRY(g_nccl.Send(d->cnt + sendBase + fP, 1, ncclInt32, nbrP, comm, s));
   NCCL_TRY(g_nccl.Send(d->cnt + sendBase + fM, 1, ncclInt32, nbrM, comm, s));
   NCCL_TRY(g_nccl.Recv(d->cnt + recvBase + fM, 1, ncclInt32, nbrM, comm, s));
   NCCL_TRY(g_nccl.Recv(d->cnt + recvBase + fP, 1, ncclInt32, nbrP, comm, s));
   NCCL_TRY(g_nccl.GroupEnd());
   CUDA_TRY(cudaMemcpyAsync(d->hCnt, d->cnt, C_COUNT * sizeof(int), cudaMemcpyDeviceToHost, s));
   CUDA_TRY(cudaStreamSynchronize(s));
   const long nSM = d->hCnt[sendBase + fM], nSP = d->hCnt[sendBase + fP];
   const long nRM = d->hCnt[recvBase + fM], nRP = d->hCnt[recvBase + fP];
   if (nRM > d->msgCap || nRP > d->msgCap || nSM > d->msgCap || nSP > d->msgCap)
   { comdSetError("halo message of %ld items exceeds capacity %ld", nRM > nRP ? nRM : nRP, d->msgCap); return COMD_E_CAPACITY; }
   if (cache) { cache[fM] = (int)nSM; cache[fP] = (int)nSP; cache[6 + fM] = (int)nRM; cache[6 + fP] = (int)nRP; }
   // 2. payload
   NCCL_TRY(g_nccl.GroupStart());
   if (nSP) NCCL_TRY(g_nccl.Send(sendBuf[1], nSP * doublesPerItem, COMD_NCCL_REAL, nbrP, comm, s));
   if (nSM) NCCL_TRY(g_nccl.Send(sendBuf[0], nSM * doublesPerItem, COMD_NCCL_REAL, nbrM, comm, s));
   if (nRM) NCCL_TRY(g_nccl.Recv(recvBuf[0], nRM * doublesPerItem, COMD_NCCL_REAL, nbrM, comm, s));
   if (nRP) NCCL_TRY(g_nccl.Recv(recvBuf[1], nRP * doublesPerItem, COMD_NCCL_REAL, nbrP, comm, s));
   NCCL_TRY(g_nccl.GroupEnd());
   return COMD_OK;
}

int comdExchangeAtomsPhase(ComdDev* d, int ax)
{
   d->fpCntValid = 0;                                        // a full atom exchange changes the ghost population
   return exchangePhase(d, ax, d->sendBuf, d->recvBuf, C_SEND0, C_RECV0, COMD_MSG_REALS, d->hAtomCnt, false);
}

int comdExchangeFprimePhase(ComdDev* d, int ax)
{
   // list mode: the arrays do not move between rebuilds, so the F' message sizes of the first exchange
   // after a rebuild hold until the next one
   const bool cached = d->useNL && (d->fpCntValid & (1 << ax));
   int rc = exchangePhase(d, ax, d->fpSend, d->fpRecv, C_FPSEND0, C_FPGOT0, 1, d->hFpCnt, cached);
   if (rc == COMD_OK && d->useNL) d->fpCntValid |= 1 << ax;
   return rc;
}

// ghost refresh between neighbour-list rebuilds: positions of the atoms of the F' halo's cell runs (neighbor.cu),
// so the message sizes are the F' halo's and share its count cache (the in-place update of unloadAtomsBuffer's
// hash-table branch, haloExchange.c:1673-1683)
int comdExchangeRefreshPhase(ComdDev* d, int ax)
{
   const bool cached = (d->fpCntValid & (1 << ax)) != 0;
   int rc = exchangePhase(d, ax, d->sendBuf, d->recvBuf, C_FPSEND0, C_FPGOT0, 3, d->hFpCnt, cached);
   if (rc == COMD_OK) d->fpCntValid |= 1 << ax;
   return rc;
}

// "if one process has to rebuild its neighbour list, all have to" (neighborList.c:236-244)
int comdAllreduceMaxFlag(ComdDev* d, int* flag)
{
   if (d->nRanks == 1 || !d->ncclComm) return COMD_OK;
   long v = *flag ? 1 : 0;
   int rc = comd_comm_allreduce_sum_int(d, &v, 1);
   *flag = v > 0;
   return rc;
}

// the same on a device-resident int32 flag, in place, no host round trip (sum of 0/1 flags)
int comdAllreduceFlagDevice(ComdDev* d, int* devFlag)
{
   if (d->nRanks == 1 || !d->ncclComm) return COMD_OK;
   NCCL_TRY(g_nccl.AllReduce(devFlag, devFlag, 1, ncclInt32, ncclSum, (ncclComm_t)d->ncclComm, d->stream));
   return COMD_OK;
}
