/* ref_bind_b200.c -- the maintainer's patch of INTEGRATION.md section 2, as code: the device entry points the
 * reference's OWN host C calls (gpu_kernels.h:13-92), implemented on top of include/comd_b200.h.
 *
 * oracle/ref_build/Makefile links the reference's unmodified host sources (CoMD.c, timestep.c, eam.c, ljForce.c,
 * haloExchange.c, linkCells.c, ... compiled where they lie under /root/reference/src-mpi) with THIS file and
 * -lcomd_b200 instead of ref_stubs.c -> oracle/_ref/CoMD-ref-b200.  Running it with the reference's GPU methods
 * (-m thread_atom | warp_atom | cta_cell -> link-cell kernels, -m thread_atom_nl | warp_atom_nl -> list kernels) drives
 * libcomd_b200.so through the reference's own control flow: timestep() -> advanceVelocityGpu / advancePositionGpu ->
 * redistributeAtomsGpu[NL] -> computeForce -> kineticEnergyGpu, command line, print-out and YAML all the reference's.
 * tests/test_gpu_driver.py compares its trace with the reference's CPU path.
 *
 * What maps to what (left: what the reference's host code calls; right: the comd_b200.h call):
 *   updateLinkCellsGpu(sim)                      first call: comd_dev_create + set_mass + set_neighbor_list + upload_atoms
 *                                                (AllocateGpu / CopyDataToGpu of gpu_utility.c have run on the shimmed
 *                                                cuda runtime and hold nothing we need); every call: comd_dev_redistribute
 *   haloExchange(atomExchange / forceExchange)   loadBuffer / unloadBuffer of both exchanges re-plugged with no-ops (0 bytes):
 *                                                the halos travelled inside comd_dev_redistribute / comd_dev_compute_force
 *   compactCellsGpu, unloadAtomsBufferToGpu,     no-ops (not reached once the exchanges are re-plugged)
 *     buildAtomListGpu, sortAtomsGpu
 *   neighborListUpdateRequiredGpu                1: updateLinkCellsGpu is then reached every step, and comd_dev_redistribute
 *                                                takes the rebuild-or-refresh decision itself (timestep.c:278-354)
 *   buildNeighborListGpu, updateNeighborsGpu,    no-ops (the list is built by comd_dev_compute_force when it is stale)
 *     emptyNeighborListGpu, hash-table calls
 *   ljForceGpu / eamForce1Gpu                    first call: comd_dev_set_lj / comd_dev_set_eam; comd_dev_compute_force
 *   eamForce2Gpu, eamForce3Gpu,                  no-ops: passes 2 and 3 and the F'(rho) halo ran inside comd_dev_compute_force
 *     load/unloadForceBuffer*Gpu
 *   computeEnergy(sim, eLocal)                   comd_dev_energy
 *   advanceVelocityGpu / advancePositionGpu      comd_dev_advance_velocity / comd_dev_advance_position
 * The Async / Comm / KI twins belong to the GPUDirect-Async layer (out of scope): reaching one aborts with a message.
 *
 * Test infrastructure: nothing in the product links or includes this file.  No reference source is copied: the
 * reference's headers are included from where they lie, for its struct definitions and prototypes. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "CoMDTypes.h"
#include "gpu_kernels.h"
#include "linkCells.h"
#include "decomposition.h"
#include "initAtoms.h"
#include "parallel.h"
#include "defines.h"

#include "../../include/comd_b200.h"

static ComdDev* g_dev = NULL;
static SimFlat* g_sim = NULL;
static int g_potSet = 0;

#define DEV(call) do { int rc_ = (call); if (rc_ != COMD_OK) { \
   fprintf(stderr, "rank %d: %s -> %d: %s\n", getMyRank(), #call, rc_, comd_last_error()); exit(-1); } } while (0)

/* The halo messages never reach the host: atoms and F'(rho) travel inside comd_dev_redistribute / comd_dev_compute_force.  The
 * reference's HaloExchange is a plug-in interface (loadBuffer / unloadBuffer function pointers, haloExchange.h:84,98); these two
 * are plugged in for both of its exchanges, so its own loadAtomsBuffer / loadForceBuffer are not entered.  (They could not be
 * entered for long anyway: outside the libmp path nothing resets sizeMsgIndexM/P, and loadAtomsBuffer writes
 * sizeMsgM[sizeMsgIndexM++] past its three entries from the second step on -- haloExchange.c:491-495,512-513,1622.) */
static int loadNothing(void* parms, void* data, int face, char* buf) { (void)parms; (void)data; (void)face; (void)buf; return 0; }
static void unloadNothing(void* parms, void* data, int face, int bufSize, char* buf) { (void)parms; (void)data; (void)face; (void)bufSize; (void)buf; }

static void createDevice(SimFlat* sim)
{
   sim->atomExchange->loadBuffer = loadNothing; sim->atomExchange->unloadBuffer = unloadNothing;
   ComdGeometry g;
   for (int a = 0; a < 3; ++a)
   {
      g.gridSize[a] = sim->boxes->gridSize[a];
      g.localMin[a] = sim->boxes->localMin[a];     g.localMax[a] = sim->boxes->localMax[a];
      g.boxSize[a] = sim->boxes->boxSize[a];       g.invBoxSize[a] = sim->boxes->invBoxSize[a];
      g.globalExtent[a] = sim->domain->globalExtent[a];
      g.procGrid[a] = sim->domain->procGrid[a];    g.procCoord[a] = sim->domain->procCoord[a];
   }
   int nLocal = 0;
   for (int b = 0; b < sim->boxes->nLocalBoxes; ++b) nLocal += sim->boxes->nAtoms[b];
   double shell = 1.0;
   for (int a = 0; a < 3; ++a) shell *= (g.gridSize[a] + 2.0) / g.gridSize[a];
   const long cap = (long)(nLocal * (2.0 * shell - 1.0) * 1.15) + 65536;
   if (getNRanks() > 1) { fprintf(stderr, "CoMD-ref-b200: this binding test is single-rank (the host layer of the product carries the NCCL bootstrap)\n"); exit(-1); }
   DEV(comd_dev_create(&g_dev, 0, &g, cap));
   DEV(comd_dev_set_mass(g_dev, sim->pot->mass));
   const int useNL = sim->method == THREAD_ATOM_NL || sim->method == WARP_ATOM_NL;
   DEV(comd_dev_set_neighbor_list(g_dev, useNL, sim->skinDistance));
   /* the reference's host mirror: MAXATOMS slots per link cell (linkCells.c:111-116) -> dense arrays */
   int* gid = malloc(nLocal * sizeof(int)); int* spec = malloc(nLocal * sizeof(int));
   real_t* v[6];
   for (int k = 0; k < 6; ++k) v[k] = malloc(nLocal * sizeof(real_t));
   int n = 0;
   const Atoms* at = sim->atoms;
   for (int b = 0; b < sim->boxes->nLocalBoxes; ++b)
      for (int i = b * MAXATOMS; i < b * MAXATOMS + sim->boxes->nAtoms[b]; ++i, ++n)
      {
         gid[n] = at->gid[i]; spec[n] = at->iSpecies[i];
         v[0][n] = at->r.x[i]; v[1][n] = at->r.y[i]; v[2][n] = at->r.z[i];
         v[3][n] = at->p.x[i]; v[4][n] = at->p.y[i]; v[5][n] = at->p.z[i];
      }
   /* comd_dev_upload_atoms ends with a redistribute and needs no potential; the potential follows at the first force call */
   DEV(comd_dev_upload_atoms(g_dev, n, gid, spec, v[0], v[1], v[2], v[3], v[4], v[5]));
   free(gid); free(spec);
   for (int k = 0; k < 6; ++k) free(v[k]);
   g_sim = sim;
}

/* ---- redistributeAtomsGpu / redistributeAtomsGpuNL (timestep.c:222-354) ---- */
void updateLinkCellsGpu(SimFlat* sim)
{
   if (!g_dev) { createDevice(sim); return; }           /* the upload has just redistributed */
   DEV(comd_dev_redistribute(g_dev));
}
int compactCellsGpu(char* work_d, int nCells, int* d_cellList, SimGpu sim_gpu, int* d_cellOffsets, int* d_workScan, real3_old shift, cudaStream_t stream)
{
   (void)work_d; (void)nCells; (void)d_cellList; (void)sim_gpu; (void)d_cellOffsets; (void)d_workScan; (void)shift; (void)stream;
   return 0;                                            /* nothing for the host to carry: the halo stayed on the device */
}
void unloadAtomsBufferToGpu(char* buf, int nBuf, SimFlat* s, char* gpu_buf, cudaStream_t stream) { (void)buf; (void)nBuf; (void)s; (void)gpu_buf; (void)stream; }
void buildAtomListGpu(SimFlat* sim, cudaStream_t stream) { (void)sim; (void)stream; }
void sortAtomsGpu(SimFlat* sim, cudaStream_t stream) { (void)sim; (void)stream; }
int neighborListUpdateRequiredGpu(SimGpu* sim) { (void)sim; return 1; }
int pairlistUpdateRequiredGpu(SimGpu* sim) { (void)sim; return 1; }
void buildNeighborListGpu(SimGpu* sim, int method, int boundaryFlag) { (void)sim; (void)method; (void)boundaryFlag; }
void emptyNeighborListGpu(SimGpu* sim, int boundaryFlag) { (void)sim; (void)boundaryFlag; }
void updateNeighborsGpu(SimGpu sim, int* temp) { (void)sim; (void)temp; }
void initHashTableGpu(HashTableGpu* hashTable, int nMaxEntries) { (void)hashTable; (void)nMaxEntries; }
void emptyHashTableGpu(HashTableGpu* hashTable) { (void)hashTable; }
void getAtomMsgSoAPtr(char* const buffer, AtomMsgSoA* atomMsg, int n) { (void)buffer; (void)n; memset(atomMsg, 0, sizeof *atomMsg); }

/* ---- computeForce (ljForce.c:141, eam.c:196-264) ---- */
static void forceCall(void)
{
   if (!g_dev) { fprintf(stderr, "CoMD-ref-b200: force call before the first redistributeAtoms\n"); exit(-1); }
   DEV(comd_dev_compute_force(g_dev));
}
void ljForceGpu(SimGpu* sim, int interpolation, int num_cells, int* cells_list, real_t plcutoff, int method)
{
   (void)sim; (void)interpolation; (void)num_cells; (void)cells_list; (void)plcutoff; (void)method;
   if (!g_potSet)
   {
      const LjPotential* lj = (const LjPotential*)g_sim->pot;
      DEV(comd_dev_set_lj(g_dev, lj->sigma, lj->epsilon, lj->cutoff));
      g_potSet = 1;
   }
   forceCall();
}
void eamForce1Gpu(SimGpu sim, int method, int spline)
{
   (void)sim; (void)method; (void)spline;
   if (!g_potSet)
   {
      const EamPotential* e = (const EamPotential*)g_sim->pot;
      ComdTable t[3] = { { e->phi->n, e->phi->x0, e->phi->invDx, e->phi->values },
                         { e->rho->n, e->rho->x0, e->rho->invDx, e->rho->values },
                         { e->f->n, e->f->x0, e->f->invDx, e->f->values } };
      DEV(comd_dev_set_eam(g_dev, e->cutoff, &t[0], &t[1], &t[2]));
      e->forceExchange->loadBuffer = loadNothing; e->forceExchange->unloadBuffer = unloadNothing;
      g_potSet = 1;
   }
   forceCall();
}
void eamForce2Gpu(SimGpu sim, int method, int spline) { (void)sim; (void)method; (void)spline; }
void eamForce3Gpu(SimGpu sim, int method, int spline) { (void)sim; (void)method; (void)spline; }
void loadForceBufferFromGpu(char* buf, int* nbuf, int nCells, int* cellList, int* natoms_buf, int* partial_sums, SimFlat* s, char* gpu_buf, cudaStream_t stream)
{
   (void)buf; (void)nCells; (void)cellList; (void)natoms_buf; (void)partial_sums; (void)s; (void)gpu_buf; (void)stream;
   *nbuf = 0;                                           /* F'(rho) of the ghosts was exchanged between pass 2 and pass 3 on the device */
}
void unloadForceBufferToGpu(char* buf, int nBuf, int nCells, int* cellList, int* natoms_buf, int* partial_sums, SimFlat* s, char* gpu_buf, cudaStream_t stream)
{
   (void)buf; (void)nBuf; (void)nCells; (void)cellList; (void)natoms_buf; (void)partial_sums; (void)s; (void)gpu_buf; (void)stream;
}
void unloadForceScanCells(int nCells, int* cellList, int* natoms_buf, int* partial_sums, SimFlat* s, cudaStream_t stream)
{
   (void)nCells; (void)cellList; (void)natoms_buf; (void)partial_sums; (void)s; (void)stream;
}

/* ---- timestep.c:109-195 ---- */
void advanceVelocityGpu(SimGpu sim, real_t dt) { (void)sim; DEV(comd_dev_advance_velocity(g_dev, dt)); }
void advancePositionGpu(SimGpu* sim, real_t dt) { (void)sim; DEV(comd_dev_advance_position(g_dev, dt)); }
void computeEnergy(SimFlat* sim, real_t* eLocal)
{
   (void)sim;
   double e[2];
   DEV(comd_dev_energy(g_dev, e));
   eLocal[0] = e[0]; eLocal[1] = e[1];
}

/* ---- GPUDirect-Async / libmp layer: out of scope (BASELINE.json north_star) ---- */
#define ASYNC_ONLY(ret, name, args) ret name args { fprintf(stderr, "CoMD-ref-b200: %s belongs to the reference's Async/comm layer, not bound here (run without -a / comm)\n", #name); exit(-1); }
ASYNC_ONLY(void, updateNeighborsGpuAsync, (SimGpu sim, int* temp, int nCells, int* cellList, cudaStream_t stream))
ASYNC_ONLY(void, eamForce1GpuAsync, (SimGpu sim, AtomListGpu atoms_list, int num_cells, int* cells_list, int method, cudaStream_t stream, int spline))
ASYNC_ONLY(void, eamForce2GpuAsync, (SimGpu sim, AtomListGpu atoms_list, int num_cells, int* cells_list, int method, cudaStream_t stream, int spline))
ASYNC_ONLY(void, eamForce3GpuAsync, (SimGpu sim, AtomListGpu atoms_list, int num_cells, int* cells_list, int method, cudaStream_t stream, int spline))
ASYNC_ONLY(int, loadAtomsBufferFromGpu_Async, (char* sendBuf, int* sendSize, char* d_compactAtoms, int nCells, int* d_cellList, SimGpu sim_gpu, int* d_cellOffsets, int* d_workScan, real3_old shift, cudaStream_t stream, int type))
ASYNC_ONLY(int, loadAtomsBufferFromGpu_Comm, (char* sendBuf, int* sendSize, char* d_compactAtoms, int nCells, int* d_cellList, SimGpu sim_gpu, int* d_cellOffsets, int* d_workScan, real3_old shift, cudaStream_t stream, int type))
ASYNC_ONLY(int, unloadAtomsBufferToGpu_Async, (char* buf, int sendSize, SimFlat* sim, char* gpu_buf, cudaStream_t stream, int typeMem))
ASYNC_ONLY(int, unloadAtomsBufferToGpu_Comm, (char* buf, int sendSize, SimFlat* sim, char* gpu_buf, cudaStream_t stream, int typeMem))
ASYNC_ONLY(void, loadForceBufferFromGpu_Async, (char* buf, int bufSize, int nCells, int* cellList, int* natoms_buf, int* partial_sums, SimFlat* s, char* gpu_buf, cudaStream_t stream))
ASYNC_ONLY(void, loadForceBufferFromGpu_Comm, (char* buf, int bufSize, int nCells, int* cellList, int* natoms_buf, int* partial_sums, SimFlat* s, char* gpu_buf, cudaStream_t stream))
ASYNC_ONLY(void, unloadForceBufferToGpu_Async, (char* buf, int bufSize, int nCells, int* cellList, int* natoms_buf, int* partial_sums, SimFlat* s, char* gpu_buf, cudaStream_t stream, int type))
ASYNC_ONLY(void, unloadForceBufferToGpu_Comm, (char* buf, int bufSize, int nCells, int* cellList, int* natoms_buf, int* partial_sums, SimFlat* s, char* gpu_buf, cudaStream_t stream, int type))

