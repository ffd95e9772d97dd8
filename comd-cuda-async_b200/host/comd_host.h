/* comd_host.h -- the plain-C host layer: CoMD's own API surface in front of the B200 device
 * layer (include/comd_b200.h).  Type and function names, argument meaning and printed output
 * follow the reference so that CoMD.c-style drivers keep working:
 *
 *   BasePotential / LjPotential / EamPotential / InterpolationObject   CoMDTypes.h:42-53,140-201
 *   SimFlat, SpeciesData, Validate                                     CoMDTypes.h:55-135
 *   Domain, initDecomposition, processorNum                            decomposition.h, decomposition.c:18-66
 *   LinkCell, initLinkCells, getBoxFromTuple, getBoxFromCoord          linkCells.h, linkCells.c:122-182,299-346,448-480
 *   Atoms, createFccLattice, setTemperature, randomDisplacements       initAtoms.h, initAtoms.c:26-216
 *   lcg61, gasdev, mkSeed                                              random.c:21-75
 *   Command, parseCommandLine, printCmdYaml                            mycommand.h:11-42, mycommand.c:187-320
 *   timestep, computeForce, advanceVelocity, advancePosition,
 *   redistributeAtoms, kineticEnergy                                   timestep.h:10-16, timestep.c:48-220
 *   HaloExchange, initAtomHaloExchange, initForceHaloExchange,
 *   haloExchange, destroyHaloExchange                                  haloExchange.h:63-159
 *   parallel wrappers                                                  parallel.h:17-63
 *   performance timers, YAML output                                    performanceTimers.h, yamlOutput.h
 *
 * What differs, on purpose: SimFlat carries one opaque ComdDev* instead of two SimGpu structs of
 * raw device pointers; atoms live on the device between calls (host arrays are refreshed on
 * demand, syncAtomsFromDevice); there is no MAXATOMS; MPI is replaced by NCCL + a file/env
 * bootstrap because the image has no MPI.  There is no CPU force path here.
 */
#ifndef COMD_HOST_H
#define COMD_HOST_H

#include <stdint.h>
#include <stdio.h>
#include "../../include/comd_b200.h"

/* mytype.h:8-24: one switch for the whole host layer.  -DCOMD_SINGLE builds CoMD-b200-f32 / libcomd_host_f32.so against
 * libcomd_b200_f32.so (comd_real = float in include/comd_b200.h follows the same macro). */
#ifdef COMD_SINGLE
typedef float real_t;
#define FMT1 "%g"
#define EMT1 "%e"
#else
typedef double real_t;
#define FMT1 "%lg"
#define EMT1 "%le"
#endif
typedef real_t real3[3];

/* ---- constants.h:28-39 ---- */
#define amuInKilograms   1.660538921e-27
#define fsInSeconds      1.0e-15
#define AngsInMeters     1.0e-10
#define eVInJoules       1.602176565e-19
#define amuToInternalMass (amuInKilograms * AngsInMeters * AngsInMeters / (fsInSeconds * fsInSeconds * eVInJoules))
#define kB_eV            8.6173324e-5
#define hartreeToEv      27.21138505
#define bohrToAngs       0.52917721092

/* ---- decomposition.h ---- */
typedef struct DomainSt
{
   int procGrid[3];
   int procCoord[3];
   real3 globalMin, globalMax, globalExtent;
   real3 localMin, localMax, localExtent;
} Domain;
Domain* initDecomposition(int xproc, int yproc, int zproc, real3 globalExtent);
int processorNum(Domain* domain, int dix, int diy, int diz);

/* ---- linkCells.h: geometry and numbering; populations come back from the device ---- */
typedef struct LinkCellSt
{
   int gridSize[3];
   int nLocalBoxes, nHaloBoxes, nTotalBoxes;
   int useHilbert;              /* recorded for the device layer; box numbering stays getBoxFromTuple's */
   real3 localMin, localMax, boxSize, invBoxSize;
   int* nAtoms;                 /* nTotalBoxes; refreshed by sumAtoms() from the device */
} LinkCell;
LinkCell* initLinkCells(const Domain* domain, real_t cutoff, int useHilbert);
void destroyLinkCells(LinkCell** boxes);
int getBoxFromTuple(LinkCell* boxes, int x, int y, int z);
int getBoxFromCoord(LinkCell* boxes, const real_t rr[3]);
int maxOccupancy(LinkCell* boxes);

/* ---- initAtoms.h: host mirror of the local atoms, dense SoA in the reference's box order ---- */
typedef struct AtomsSt
{
   int nLocal, nGlobal;
   int capacity;
   int* gid;
   int* iSpecies;
   real_t* r[3];
   real_t* p[3];
   real_t* f[3];
   real_t* U;
} Atoms;
struct SimFlatSt;
Atoms* initAtoms(LinkCell* boxes, real_t skinDistance);
void destroyAtoms(Atoms* atoms);
void createFccLattice(int nx, int ny, int nz, real_t lat, struct SimFlatSt* s);
void setTemperature(struct SimFlatSt* s, real_t temperature);
void setVcm(struct SimFlatSt* s, real_t vcm[3]);
void randomDisplacements(struct SimFlatSt* s, real_t delta);

/* ---- random.h ---- */
double lcg61(uint64_t* seed);
real_t gasdev(uint64_t* seed);
uint64_t mkSeed(uint32_t id, uint32_t callSite);

/* ---- CoMDTypes.h ---- */
typedef struct BasePotentialSt
{
   real_t cutoff, mass, lat;
   char latticeType[8];
   char name[3];
   int atomicNo;
   int  (*force)(struct SimFlatSt* s);
   void (*print)(FILE* file, struct BasePotentialSt* pot);
   void (*destroy)(struct BasePotentialSt** pot);
} BasePotential;

typedef struct LjPotentialSt
{
   real_t cutoff, mass, lat;
   char latticeType[8];
   char name[3];
   int atomicNo;
   int  (*force)(struct SimFlatSt* s);
   void (*print)(FILE* file, BasePotential* pot);
   void (*destroy)(BasePotential** pot);
   real_t sigma, epsilon;
} LjPotential;

typedef struct InterpolationObjectSt
{
   int n;
   real_t x0, invDx;
   real_t* values;              /* entry 0 of an array valid on [-1, n+1] (eam.c:496-519) */
} InterpolationObject;

struct HaloExchangeSt;
typedef struct EamPotentialSt
{
   real_t cutoff, mass, lat;
   char latticeType[8];
   char name[3];
   int atomicNo;
   int  (*force)(struct SimFlatSt* s);
   void (*print)(FILE* file, BasePotential* pot);
   void (*destroy)(BasePotential** pot);
   InterpolationObject* phi;
   InterpolationObject* rho;
   InterpolationObject* f;
   struct HaloExchangeSt* forceExchange;
} EamPotential;

typedef struct SpeciesDataSt { char name[3]; int atomicNo; real_t mass; } SpeciesData;
typedef struct ValidateSt { double eTot0; int nAtoms0; } Validate;

/* method ids, defines.h:11-17 */
enum { THREAD_ATOM = 0, WARP_ATOM = 1, CTA_CELL = 2, THREAD_ATOM_NL = 3, WARP_ATOM_NL = 4, CPU_NL = 5 };

typedef struct SimFlatSt
{
   int nSteps, printRate;
   double dt;
   Domain* domain;
   LinkCell* boxes;
   Atoms* atoms;
   SpeciesData* species;
   real_t ePotential, eKinetic;
   BasePotential* pot;
   struct HaloExchangeSt* atomExchange;
   ComdDev* dev;                /* replaces SimGpu host/gpu (CoMDTypes.h:97-98) */
   int method;
   int gpuAsync, gpuProfile;
   int ljInterpolation, spline, usePairlist;   /* accepted, recorded, not acted on (DESIGN.md "out of scope") */
   int useHilbert;              /* -H: the device hands its work blocks to the SMs along a Hilbert curve (list path) */
   real_t skinDistance;
   int hostAtomsStale;          /* device state is newer than sim->atoms */
} SimFlat;

BasePotential* initLjPot(void);
BasePotential* initEamPot(const char* dir, const char* file, const char* type);
void eamBcastPotential(EamPotential* pot);
void interpolate(InterpolationObject* table, real_t r, real_t* f, real_t* df);

/* ---- haloExchange.h:63-145: the exchange itself runs on the device; the struct keeps the call sites and the members a
 * CoMD.c-style caller may touch.  The message buffers live in HBM (the reference's sendBufM/P, recvBufM/P and their _Async
 * twins have no host counterpart here), so loadBuffer / unloadBuffer do not move data: loadBuffer reports the bytes the
 * device packed for `face` in the last exchange (the return value haloExchange.h:84 documents), unloadBuffer is a no-op. ---- */
struct HashTableSt;
typedef struct HaloExchangeSt
{
   int nbrRank[6];
   int bufCapacity;
   int  (*loadBuffer)(void* parms, void* data, int face, char* buf);
   void (*unloadBuffer)(void* parms, void* data, int face, int bufSize, char* buf);
   void (*destroy)(void* parms);
   void* parms;
   struct HashTableSt* hashTable;   /* always NULL: the device keys ghosts by (cell, gid) order, not by a gid hash */
   int type;                    /* 0 atoms, 1 F'(rho)  (haloExchange.h:111) */
} HaloExchange;
HaloExchange* initAtomHaloExchange(Domain* domain, LinkCell* boxes);
HaloExchange* initForceHaloExchange(Domain* domain, LinkCell* boxes, int useGPU);
void haloExchange(HaloExchange* haloExchange, void* data);
void destroyHaloExchange(HaloExchange** haloExchange);

/* ---- timestep.h ---- */
double timestep(SimFlat* s, int n, real_t dt);
void computeForce(SimFlat* s);
void advanceVelocity(SimFlat* s, real_t dt);
void advancePosition(SimFlat* s, real_t dt);
void redistributeAtoms(SimFlat* s);
void kineticEnergy(SimFlat* s);
void kineticEnergyGpu(SimFlat* s);
void syncAtomsFromDevice(SimFlat* s);      /* GetLocalAtomsFromGpu, gpu_utility.c:602 */

/* ---- mycommand.h ---- */
typedef struct CommandSt
{
   char potDir[1024], potName[1024], potType[1024];
   int doeam;
   int nx, ny, nz;
   int xproc, yproc, zproc;
   int nSteps, printRate;
   double dt, lat, temperature, initialDelta, relativeSkinDistance;
   int doHilbert;
   char method[1024];
   int gpuAsync, gpuProfile, ljInterpolation, spline, usePairlist;
} Command;
Command parseCommandLine(int argc, char** argv);
void printCmdYaml(FILE* file, Command* cmd);

/* ---- parallel.h ---- */
typedef struct RankReduceDataSt { double val; int rank; } RankReduceData;
int getNRanks(void);
int getMyRank(void);
int printRank(void);
void timestampBarrier(const char* msg);
void initParallel(int* argc, char*** argv);
void destroyParallel(void);
void barrierParallel(void);
void attachParallelDevice(ComdDev* dev);   /* reductions ride on the device's NCCL communicator */
int  localDeviceId(void);
void addIntParallel(int* sendBuf, int* recvBuf, int count);
void addRealParallel(real_t* sendBuf, real_t* recvBuf, int count);
void addDoubleParallel(double* sendBuf, double* recvBuf, int count);
void maxIntParallel(int* sendBuf, int* recvBuf, int count);
void minRankDoubleParallel(RankReduceData* sendBuf, RankReduceData* recvBuf, int count);
void maxRankDoubleParallel(RankReduceData* sendBuf, RankReduceData* recvBuf, int count);
void bcastParallel(void* buf, int len, int root);
int  parallelCommId(unsigned char id[COMD_COMM_ID_BYTES]);   /* rank 0 creates, the others read */

/* ---- performanceTimers.h ---- */
enum TimerHandle { totalTimer, loopTimer, timestepTimer, positionTimer, velocityTimer, neighborListTimer, redistributeTimer,
                   atomHaloTimer, computeForceTimer, eamHaloTimer, commHaloTimer, commReduceTimer, numberOfTimers };
void profileStart(enum TimerHandle handle);
void profileStop(enum TimerHandle handle);
double getElapsedTime(enum TimerHandle handle);
void printPerformanceResults(int nGlobalAtoms, int printRate);
void printPerformanceResultsYaml(FILE* file);
void setRooflineInfo(double flopPerAtomStep, double fp64PeakTflopsPerGpu, int nRanks);
#define startTimer(h) profileStart(h)
#define stopTimer(h)  profileStop(h)

/* ---- yamlOutput.h ---- */
extern FILE* yamlFile;
#define screenOut stdout
void yamlBegin(void);
void yamlEnd(void);
void yamlAppInfo(FILE* file);
void printSeparator(FILE* file);

/* ---- CoMD.c ---- */
SimFlat* initSimulation(Command cmd);
void destroySimulation(SimFlat** ps);
void sumAtoms(SimFlat* s);
void printThings(SimFlat* s, int iStep, double elapsedTime);
void printSimulationDataYaml(FILE* file, SimFlat* s);
void sanityChecks(Command cmd, double cutoff, double latticeConst, char latticeType[8]);
Validate* initValidate(SimFlat* s);
void validateResult(const Validate* val, SimFlat* sim);

void comdDie(const char* where, int rc);    /* device error -> message + exit, like gpu_utility.h:71-84 */
#define DEV_TRY(call) do { int rc_ = (call); if (rc_ != COMD_OK) comdDie(#call, rc_); } while (0)

#endif
