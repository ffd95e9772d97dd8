// comd_dev.cuh -- internal device-side context shared by the sm_100a translation units.
//
// Data layout in HBM (see DESIGN.md "Data layout"):
//   * atoms are dense structure-of-arrays, sorted by (cell, gid); no MAXATOMS padding
//     (the reference pads every cell to MAXATOMS slots, linkCells.c:111-116).
//   * cells live on a PADDED grid (nx+2)(ny+2)(nz+2), x fastest, halo cells included,
//     so the three x-adjacent cells of every stencil row are ONE contiguous atom run:
//     the 27-cell stencil is 9 contiguous runs [cellStart[c-1], cellStart[c+2]).
//   * the reference's own box numbering (getBoxFromTuple, linkCells.c:299-346) is
//     recovered by refBoxOfCell() whenever ids cross the C ABI.
#pragma once
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>
#include <stdint.h>
#include "../../include/comd_b200.h"

// Function attributes (dynamic shared-memory limit, carve-out) are per device; the launchers cache what they have set in
// static tables indexed by ComdDev::device.  comd_dev_create refuses ids beyond the table.
#define COMD_MAX_DEVICES 64

// ---- precision (mytype.h:8-20; the reference's Makefile:73-77 switches -DCOMD_DOUBLE / -DCOMD_SINGLE) --------------
// real_t is the type of every per-atom quantity, table and message on the device and across the C ABI (comd_real).
// R(x) writes a literal of that type: in the double build it is the literal itself, so the FP64 code is unchanged.
#ifdef COMD_SINGLE
typedef float real_t; typedef float2 real2; typedef float4 real4;
#define R(x) x##f
#define make_real2 make_float2
#define make_real4 make_float4
#define COMD_NCCL_REAL 7            // ncclFloat32
#define COMD_MSG_REALS 8            // AtomMsg on the wire: gid, species, r[3], p[3] = 32 B (haloExchange.h:32-38 with float real_t)
__device__ __forceinline__ real_t mulRn(real_t a, real_t b) { return __fmul_rn(a, b); }
__device__ __forceinline__ real_t addRn(real_t a, real_t b) { return __fadd_rn(a, b); }
__device__ __forceinline__ int realFloorInt(real_t x) { return __float2int_rd(x); }
__device__ __forceinline__ long long realBits(real_t x) { return (long long)__float_as_int(x); }
// one 16-byte load of a real4 record (LDG.E.128); P = PTX cache-policy qualifiers
#define COMD_LDG_REAL4(P, v, p) asm volatile("ld.global" P ".v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p))
#else
typedef double real_t; typedef double2 real2; typedef double4 real4;
#define R(x) x
#define make_real2 make_double2
#define make_real4 make_double4
#define COMD_NCCL_REAL 8            // ncclFloat64
#define COMD_MSG_REALS 7            // AtomMsg on the wire: {gid|species}, r[3], p[3] = 56 B (haloExchange.h:32-38)
__device__ __forceinline__ real_t mulRn(real_t a, real_t b) { return __dmul_rn(a, b); }
__device__ __forceinline__ real_t addRn(real_t a, real_t b) { return __dadd_rn(a, b); }
__device__ __forceinline__ int realFloorInt(real_t x) { return __double2int_rd(x); }
__device__ __forceinline__ long long realBits(real_t x) { return __double_as_longlong(x); }
// one 32-byte load of a real4 record (LDG.E.256 on sm_100a); P = PTX cache-policy qualifiers
#define COMD_LDG_REAL4(P, v, p) asm volatile("ld.global" P ".v4.f64 {%0, %1, %2, %3}, [%4];" : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w) : "l"(p))
#endif
// identity of an atom inside a message: FP64 packs {gid, species} into one 8-byte slot, FP32 uses one 4-byte slot each
__device__ __forceinline__ void msgPutId(real_t* o, int gid, int spec)
{
#ifdef COMD_SINGLE
   o[0] = __int_as_float(gid); o[1] = __int_as_float(spec);
#else
   int2 id = make_int2(gid, spec); o[0] = *reinterpret_cast<double*>(&id);
#endif
}
__device__ __forceinline__ void msgGetId(const real_t* in, int& gid, int& spec)
{
#ifdef COMD_SINGLE
   gid = __float_as_int(in[0]); spec = __float_as_int(in[1]);
#else
   double w = in[0]; int2 id = *reinterpret_cast<int2*>(&w); gid = id.x; spec = id.y;
#endif
}
#define COMD_MSG_R0 (COMD_MSG_REALS - 6)   // first position slot of a message record
#define COMD_MSG_DOUBLES COMD_MSG_REALS
#define COMD_MAX_CELL_ATOMS 256     // sanity bound used by the in-cell rank sort

// device counters (ints)
enum {
   C_NATOMS = 0,   // atoms in the sorted arrays (local + ghost)
   C_NLOCAL,       // atoms in local cells
   C_NSRC,         // size of the source pool while redistributing
   C_NSHELL,       // entries in the shell list
   C_SEND0,        // 6 send counts (face order x-,x+,y-,y+,z-,z+ as haloExchange.h:20-22)
   C_RECV0 = C_SEND0 + 6,
   C_FPSEND0 = C_RECV0 + 6,   // F' message sizes
   C_FPRECV0 = C_FPSEND0 + 6,
   C_FPGOT0 = C_FPRECV0 + 6,  // F' elements actually received per face
   C_ERR = C_FPGOT0 + 6,      // sticky error bits
   C_NLREBUILD,               // neighbour-list: rebuild required flag
   C_NLPAIRS,                 // neighbour-list: stored entries
   C_NLMAX,                   // neighbour-list: max per atom
   C_COUNT = 64
};
// error bits
#define ERRB_ESCAPE    1   // an atom left the halo shell in one step (moved > 1 cell)
#define ERRB_CAPACITY  2   // atom or message capacity exceeded
#define ERRB_CELLFULL  4   // a cell holds more than COMD_MAX_CELL_ATOMS atoms
#define ERRB_FPMISMATCH 8  // F' halo: message length != ghost population
#define ERRB_NLFULL    16  // neighbour list row overflow
#define ERRB_FIXED     64  // a pair contribution outside the fixed-point range of the in-block accumulators (force_n3.cuh)
#define ERRB_TILE      32  // a force tile's stencil rows exceed the shared-memory staging capacity
#define COMD_PAD       16  // slack elements behind every per-atom array (aligned TMA over-read)

// cell type bits
#define CT_LOCAL  1
#define CT_SHELL  2                      // within two planes of any face (may have to travel)
#define CT_FACE(f) (4 << (f))            // belongs to the 2-plane send set of face f

struct DevGeom
{
   int n[3];            // local cells per axis
   int P[3];            // padded
   int nCells;          // padded total
   real_t lMin[3], lMax[3], inv[3];
};

struct DevTable { int n; real_t x0, invDx; const real_t* v; /* entry 0; valid [-1,n+1] */ };

struct AtomArrays
{
   real_t *r[3], *p[3];
   int *gid, *spec, *cell;
   float4 *rf;                  // FP32 shadow of r relative to localMin (filter stage of the force tiles)
   real4 *rq;                   // gather copy {x, y, z, F'(rho)}: one 32-byte sector per neighbour visit (list kernels)
};

struct ComdDev
{
   int device;
   cudaStream_t stream;
   DevGeom g;
   ComdGeometry hostGeom;
   real_t shift[6][3];          // PBC shift added when packing for face f (haloExchange.c:310-323)
   int nbrRank[6];
   int rank, nRanks;
   long cap;                    // atom capacity of every per-atom array
   AtomArrays A, B;             // current (sorted) and staging
   real_t *f[3], *e;            // forces, per-atom potential energy (eV)
   int *cellStart;              // nCells+1
   int *cellCount, *cellFill;   // nCells
   unsigned char* cellType;     // nCells
   int *cellKey;                // new cell per source atom (cap)
   int *shell;                  // shell list (cap)
   int *tmpSrc, *tmpGid, *tmpCell;  // scatter staging (cap)
   int *scanBlock;              // block sums for the scan
   int *cnt;                    // C_COUNT device counters
   int *hCnt;                   // pinned mirror
   int approxAtoms;             // host-side estimate for grid sizing only
   int approxHasGhosts;         // approxAtoms counts ghosts too (set from the device counter) or only locals (set from an upload)
   // messages
   long msgCap;                 // atoms per face message
   real_t *sendBuf[2], *recvBuf[2];
   real_t *fpSend[2], *fpRecv[2];
   int *runOff;                 // F' run offsets scratch
   // potential
   int potKind;                 // 0 none, 1 LJ, 2 EAM
   double mass;
   double ljSigma, ljEps, cutoff;
   DevTable tPhi, tRho, tF;
   real_t *tabMem;
   real2 *tPRmem, *tPR;                // interleaved {phi,rho} when both tables share one grid (entry 0), else NULL
   real2 *tW[2];                        // two copies of the interleaved table, the second shifted by one entry: the 64-byte
                                         // window [i-1, i+2] of any lookup is 32-byte aligned in one of them (two 256-bit loads)
   real2 *tDR;                           // pass 3: {g1, g2-g1} of rho per interval (one 128-bit load), entry 0
   // list path: one 32-byte record per interval, so every table lookup is ONE sector (LDG.E.256)
   real4 *rvPhi, *rvRho;                 // {T[i], g1, T[i+1]+T[i-1]-2T[i], g2-g1}: value and derivative of one table
   real4 *r3;                            // {g1, g2-g1} of phi and of rho side by side (same grid only, else NULL)
   int wantEnergy;                       // this force evaluation also produces per-atom potential energies
   int eValid;                           // d->e holds the energies of the current configuration
   int optEnergyEveryStep;               // comd_dev_timestep: energies at every step (1) or only at the last one (0, default)
   int optNlThreads;                     // CTA size of the list force kernels
   int optNlBlock[3];                    // cells per work block of the list force kernels (0 = auto)
   real_t *rhobar, *dfEmbed;
   // neighbour list
   int useNL; double skin;
   int *nlList; int *nlCount; int nlStride; int nlMax; real_t* nlLastR[3]; int nlBuilds; int nlValid;
   long nlRows;                 // rows allocated in nlList
   int optNlN3;                          // pairs inside a work block evaluated once for both atoms (force_n3.cuh): 1 yes, 0 no, -1 (default) EAM yes / LJ no
   int optNlLanes, nlLanes;              // lanes per atom of the once-per-pair kernels (option nl_lanes: 4 default, 8) / of the current list
   int nlN3Shape[3];                     // block shape the current list was built for (0 = full list)
   int optNlChunksPerCta;                // target work chunks per CTA of the list force kernels (0 = default 32)
   int *nlWork; unsigned nlWorkSeq;      // work counters of the list force kernels (one per launch, ring of 64)
   int *nlCountHit;             // per atom: list prefix that held every in-cutoff entry in the last pass 1
   int nlRefreshes;             // steps served by a ghost refresh instead of a rebuild
   unsigned long long* nlStat;  // device: [0] stored pairs, [1] max per atom
   int runLayoutValid;          // d->runOff describes the cell runs of the current sort (halo_runs.cuh; list mode keeps it between rebuilds)
   // reductions
   double *redPartial; double *redOut; double *hRed;
   // stats
   cudaEvent_t evForce[2 * 128]; int nForceEv; cudaEvent_t evT0, evT1; double lastForceMs; long nLaunches;   // force-time ring (128 calls) + user timer
   // e2e staging
   real_t* hStage; long hStageDoubles;
   int* rowTab;                          // local-row table of the device-side compaction
   real_t* e2eU; int* e2eUgid;           // landing area of the uploaded inputs (9 slabs of cap doubles + gids)
   real_t* e2eOut[2]; int* e2eOutGid[2]; // compacted results (10 slabs each) of this call and of the previous one
   int e2eCur;                           // which of the two holds what the caller was handed last
   int e2eLastM;                         // ... and how many atoms that was
   cudaStream_t copyStream; cudaEvent_t evCopy;   // results that are final early leave on their own stream (comd_dev_step_host)
   cudaStream_t upStream; cudaEvent_t evUpR, evUpP, evUp;   // uploads of a speculative call (see comd_dev_step_host)
   int* e2eFlag; int* hE2eFlag;          // "inputs differ from what the device holds" (device, pinned host)
   int e2eOrderValid;                    // rowTab describes the current sort and the caller holds the atoms in that order
   int e2ePristine;                      // nothing but comd_dev_step_host has touched the atoms since its last return
   int e2eSpecArmed;                     // speculation allowed (cleared for good by the first mismatch)
   long e2eResumes;                      // comd_dev_step_host calls served without re-sorting / rebuilding
   // tuning knobs (comd_dev_set_option)
   int optForceVariant;         // 0 = thread-per-atom stencil loop (v1), 1 = tiled TMA/filter/process engine
   int optTileTx;               // target cells per tile along x
   int optTileHits;             // link-cell EAM: pass 1 writes the step's hit list, the list kernel does the pair forces (default 1)
   int posDirty;                // positions moved since the FP32 shadow was written
   int tileSmemSet[4];
   int hAtomCnt[12], hFpCnt[12], fpCntValid;   // host copies of the last message sizes (send 6 | recv 6); F' cache validity per axis
   // nccl
   void* nccl; void* ncclComm;
   // halo / compute overlap (list path, more than one rank): the exchange phases run on `commStream`
   // while the force kernels work on the blocks of cells that see no ghost (force_nl.cuh)
   cudaStream_t commStream;      // high priority
   cudaStream_t xs;              // stream the exchange code enqueues on: `stream`, or `commStream` inside an overlapped section
   cudaEvent_t evRefresh, evPass1, evFp, evStepBegin;
   cudaStream_t sideStream;      // boundary-shell force launches of an overlapped pass (force_n3.cuh)
   cudaEvent_t evSideBegin, evSideDone;
   int specStep;                 // inside a speculatively enqueued step (comd_dev_timestep): no host sync on the rebuild flag
   long nlSpecRedone;            // speculatively enqueued steps that had to be redone behind a rebuild
   int optNlSpeculate;           // 1 (default): list-mode steps are enqueued without waiting for the rebuild verdict
   cudaEvent_t evFlag;           // the flag's copy to the host has landed
   // CUDA-graph replay of the steady-state list-mode step (comd_dev_timestep): captured once per list build
   cudaGraphExec_t stepGraph; long stepGraphGen; double stepGraphDt; int stepGraphOpts;
   long stepGraphLaunches, stepGraphRefreshes, nGraphSteps;
   int capturing;                // the compute stream is in capture mode: no host-visible event records, no timing ring
   int optGraph;                 // -1 auto (small systems on one rank), 0 off, 1 on
   int optNlHilbert;             // work blocks of the list kernels handed out along a Hilbert curve (CoMD's -H)
   int* nlOrder[3]; int nlOrderSig[3][7];   // order tables per phase set (all / interior / boundary) and what each was built for, nl_blocks.cuh
   int optOverlapMode;           // 1: one-shot interior CTAs on all SMs + boundary shell on sideStream; 0: round-1 scheme
   int optOverlap;               // 1 (default): overlap where it applies
   int optOverlapSms;            // SMs an overlapped interior launch may fill (0 = default)
   int checkFused;               // the drift kernel already evaluated the list's rebuild criterion for the current positions
   int inTimestep;               // inside comd_dev_timestep: redistribute is always followed by the force call
   int refreshInFlight;          // ghost refresh of this step enqueued on commStream, evRefresh recorded behind it
   // per-phase timing of comd_dev_timestep (option "phase_timing"): marks 0..6 per step, first 32 steps of a call
   int optPhaseTiming; cudaEvent_t evPh[32][8]; float phaseSample[6][1024]; long phaseSteps; int phStep;
};
#define PH_MARK(d, id) do { if ((d)->optPhaseTiming && (d)->phStep >= 0 && (d)->phStep < 32) cudaEventRecord((d)->evPh[(d)->phStep][id], (d)->stream); } while (0)

// NVTX ranges around the phases of a step, as the reference brackets its halo phases (PUSH_RANGE / POP_RANGE,
// haloExchange.c:60-90,1503-1521): visible in Nsight Systems / ncu --nvtx, one pointer check when no tool is attached.
struct NvtxRange { explicit NvtxRange(const char* name) { nvtxRangePushA(name); } ~NvtxRange() { nvtxRangePop(); } };
#define NVTX_RANGE(name) NvtxRange nvtxRange_##__LINE__(name)

// ---- geometry helpers -----------------------------------------------------------
__host__ __device__ inline int cellIndex(const DevGeom& g, int ix, int iy, int iz)   // unpadded coords in [-1, n]
{
   return (ix + 1) + g.P[0] * ((iy + 1) + g.P[1] * (iz + 1));
}

// Owner cell of a position: the reference's getBoxFromCoord (linkCells.c:448-480) on the
// padded grid.  (x - lMin) * inv is a subtract followed by a multiply, which nvcc cannot
// contract into an FMA, so the floor() argument is bit-identical to the host's.
__device__ inline int axisCell(real_t x, real_t lMin, real_t lMax, real_t inv, int n, int& err)
{
   int i = (int)floor((x - lMin) * inv);
   if (x < lMax) { if (i == n) i = n - 1; }
   else i = n;
   if (i < -1) { i = -1; err = 1; }
   if (i > n) { i = n; err = 1; }
   return i;
}

__device__ inline int cellOfPos(const DevGeom& g, real_t x, real_t y, real_t z, int& err)
{
   int ix = axisCell(x, g.lMin[0], g.lMax[0], g.inv[0], g.n[0], err);
   int iy = axisCell(y, g.lMin[1], g.lMax[1], g.inv[1], g.n[1], err);
   int iz = axisCell(z, g.lMin[2], g.lMax[2], g.inv[2], g.n[2], err);
   return cellIndex(g, ix, iy, iz);
}

// reference box id (getBoxFromTuple, linkCells.c:299-346) of a padded cell
__host__ __device__ inline int refBoxOfTuple(const int* n, int ix, int iy, int iz)
{
   const int nl = n[0] * n[1] * n[2];
   if (iz == n[2]) return nl + 2*n[2]*n[1] + 2*n[2]*(n[0]+2) + (n[0]+2)*(n[1]+2) + (n[0]+2)*(iy+1) + (ix+1);
   if (iz == -1)   return nl + 2*n[2]*n[1] + 2*n[2]*(n[0]+2) + (n[0]+2)*(iy+1) + (ix+1);
   if (iy == n[1]) return nl + 2*n[2]*n[1] + n[2]*(n[0]+2) + (n[0]+2)*iz + (ix+1);
   if (iy == -1)   return nl + 2*n[2]*n[1] + iz*(n[0]+2) + (ix+1);
   if (ix == n[0]) return nl + n[1]*n[2] + iz*n[1] + iy;
   if (ix == -1)   return nl + iz*n[1] + iy;
   return ix + n[0]*iy + n[0]*n[1]*iz;
}
__host__ __device__ inline int refBoxOfCell(const DevGeom& g, int c)
{
   int ix = c % g.P[0] - 1, iy = (c / g.P[0]) % g.P[1] - 1, iz = c / (g.P[0] * g.P[1]) - 1;
   return refBoxOfTuple(g.n, ix, iy, iz);
}

// ---- error plumbing -------------------------------------------------------------
void comdSetError(const char* fmt, ...);
#define CUDA_TRY(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { \
   comdSetError("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); return COMD_E_CUDA; } } while (0)
#define LAUNCH_CHECK(dev) do { (dev)->nLaunches++; cudaError_t e_ = cudaGetLastError(); if (e_ != cudaSuccess) { \
   comdSetError("%s:%d: kernel launch -> %s", __FILE__, __LINE__, cudaGetErrorString(e_)); return COMD_E_CUDA; } } while (0)

static inline int gridFor(long n, int block, int maxBlocks = 148 * 16)
{
   long b = (n + block - 1) / block;
   if (b < 1) b = 1;
   if (b > maxBlocks) b = maxBlocks;
   return (int)b;
}

// once-per-pair list kernels for this potential?  LJ pair terms are cheaper than the scatter, so "auto" keeps its full list
static inline bool comdNlN3(const ComdDev* d) { return d->useNL && (d->optNlN3 < 0 ? d->potKind == 2 : d->optNlN3 != 0); }

// implemented across the translation units
int comdForceLJ(ComdDev* d);
int comdRefreshShadow(ComdDev* d);
int comdForceEAM(ComdDev* d);
int comdFprimeHalo(ComdDev* d);
int comdHaloRunLayout(ComdDev* d, cudaStream_t s);      // force_eam.cu: run starts/offsets/sizes of the six one-plane halo messages
int comdNeighborBuildIfNeeded(ComdDev* d);
int comdNeighborAllocate(ComdDev* d);
int comdCheckErrors(ComdDev* d, const char* where);
int comdExchangeAtomsPhase(ComdDev* d, int axis);       // comm.cu: moves send buffers to recv buffers (self or NCCL)
int comdExchangeFprimePhase(ComdDev* d, int axis);
int comdExchangeRefreshPhase(ComdDev* d, int axis);     // 3 doubles per atom of the F' halo's cell runs
int comdNeighborRefreshOrRebuild(ComdDev* d, int* didRefresh);   // neighbor.cu: the redistributeAtomsGpuNL decision
int comdNeighborAfterSort(ComdDev* d);                  // neighbor.cu: cell-run layout of the new sort (ghost refresh, F' halo)
int comdAllreduceMaxFlag(ComdDev* d, int* flag);
int comdAllreduceFlagDevice(ComdDev* d, int* devFlag); // comm.cu: in-place sum of a device-resident 0/1 flag over the ranks        // comm.cu: OR of a host flag over the ranks
