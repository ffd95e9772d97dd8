/* timestep.c -- the step recipe of timestep.c:48-105 of the reference on top of the device layer,
 * plus the call sites of haloExchange.h kept as thin shims (the exchanges run inside the device
 * calls: atoms inside comd_dev_redistribute, F'(rho) inside comd_dev_compute_force).
 *
 *   advanceVelocity(dt/2); advancePosition(dt); redistributeAtoms; [neighbour list]; computeForce;
 *   advanceVelocity(dt/2)                                       x nSteps, then kineticEnergy.
 *
 * With --gpuAsync 1 the whole interval is handed to comd_dev_timestep in one call (fused
 * kick+drift, no host synchronisation between kernels): the same arithmetic, bit for bit
 * (tests/test_gpu_parity.py::test_split_calls_equal_fused_timestep), but the per-phase timers
 * then only see the interval as a whole.
 */
#include "comd_host.h"
#include <stdlib.h>

void comdDie(const char* where, int rc)
{
   fprintf(stderr, "rank %d: %s failed (%d): %s\n", getMyRank(), where, rc, comd_last_error());
   exit(-1);
}

void advanceVelocity(SimFlat* s, real_t dt)
{
   DEV_TRY(comd_dev_advance_velocity(s->dev, dt));
   s->hostAtomsStale = 1;
}

void advancePosition(SimFlat* s, real_t dt)
{
   DEV_TRY(comd_dev_advance_position(s->dev, dt));
   s->hostAtomsStale = 1;
}

void redistributeAtoms(SimFlat* s)
{
   /* re-bin, atom halo incl. migration, canonical sort; with a neighbour-list method the device
    * decides between a full re-bin and an in-place ghost refresh (redistributeAtomsGpuNL) */
   startTimer(atomHaloTimer);
   haloExchange(s->atomExchange, s);
   stopTimer(atomHaloTimer);
   s->hostAtomsStale = 1;
}

void computeForce(SimFlat* s)
{
   s->pot->force(s);
}

void kineticEnergyGpu(SimFlat* s)
{
   double e[2];
   DEV_TRY(comd_dev_energy(s->dev, e));                /* device sums are double in either precision */
   real_t eLocal[2] = { (real_t)e[0], (real_t)e[1] }, sum[2];     /* timestep.c:124-128 reduces real_t */
   startTimer(commReduceTimer);
   addRealParallel(eLocal, sum, 2);
   stopTimer(commReduceTimer);
   s->ePotential = sum[0];
   s->eKinetic = sum[1];
}
void kineticEnergy(SimFlat* s) { kineticEnergyGpu(s); }

double timestep(SimFlat* s, int nSteps, real_t dt)
{
   if (s->gpuAsync)
   {
      DEV_TRY(comd_dev_timestep(s->dev, nSteps, dt));
      s->hostAtomsStale = 1;
   }
   else
      for (int ii = 0; ii < nSteps; ++ii)
      {
         startTimer(velocityTimer); advanceVelocity(s, 0.5 * dt); stopTimer(velocityTimer);
         startTimer(positionTimer); advancePosition(s, dt);       stopTimer(positionTimer);
         startTimer(redistributeTimer); redistributeAtoms(s);     stopTimer(redistributeTimer);
         startTimer(computeForceTimer); computeForce(s);          stopTimer(computeForceTimer);
         startTimer(velocityTimer); advanceVelocity(s, 0.5 * dt); stopTimer(velocityTimer);
      }
   kineticEnergyGpu(s);
   return s->ePotential;
}

/* GetLocalAtomsFromGpu (gpu_utility.c:602): bring the local atoms back, in device order */
extern void comdHostReserveAtoms(Atoms* a, int n);
void syncAtomsFromDevice(SimFlat* s)
{
   if (!s->hostAtomsStale) return;
   Atoms* a = s->atoms;
   int n = 0;
   DEV_TRY(comd_dev_num_local(s->dev, &n));
   comdHostReserveAtoms(a, n);
   DEV_TRY(comd_dev_download_atoms(s->dev, &n, a->capacity, a->gid, a->iSpecies, NULL, a->r[0], a->r[1], a->r[2],
                                   a->p[0], a->p[1], a->p[2], a->f[0], a->f[1], a->f[2], a->U));
   a->nLocal = n;
   s->hostAtomsStale = 0;
}

/* ---- haloExchange.h shims ---- */
/* loadBuffer / unloadBuffer (haloExchange.h:84,98): `data` is the SimFlat, as for haloExchange() */
static int loadDeviceBuffer(void* parms, void* data, int face, char* buf)
{
   (void)buf;
   const HaloExchange* h = parms;
   SimFlat* s = data;
   long sent = 0, got = 0;
   if (!s || face < 0 || face > 5) return 0;
   DEV_TRY(comd_dev_halo_bytes(s->dev, h->type, face, &sent, &got));
   return (int)sent;
}
static void unloadDeviceBuffer(void* parms, void* data, int face, int bufSize, char* buf)
{
   (void)parms; (void)data; (void)face; (void)bufSize; (void)buf;   /* unpacked on the device by the exchange itself */
}

static HaloExchange* makeExchange(Domain* dom, int kind)
{
   HaloExchange* h = calloc(1, sizeof *h);
   h->type = kind;
   h->loadBuffer = loadDeviceBuffer; h->unloadBuffer = unloadDeviceBuffer;
   h->parms = h;                                  /* the sub-class data of this exchange is the exchange itself */
   /* face order x-, x+, y-, y+, z-, z+ (haloExchange.h:20-22) */
   h->nbrRank[0] = processorNum(dom, -1, 0, 0); h->nbrRank[1] = processorNum(dom, +1, 0, 0);
   h->nbrRank[2] = processorNum(dom, 0, -1, 0); h->nbrRank[3] = processorNum(dom, 0, +1, 0);
   h->nbrRank[4] = processorNum(dom, 0, 0, -1); h->nbrRank[5] = processorNum(dom, 0, 0, +1);
   return h;
}

HaloExchange* initAtomHaloExchange(Domain* dom, LinkCell* boxes)
{
   HaloExchange* h = makeExchange(dom, 0);
   const int* g = boxes->gridSize;
   int maxFace = (g[0] + 2) * (g[1] + 2);
   if ((g[1] + 2) * (g[2] + 2) > maxFace) maxFace = (g[1] + 2) * (g[2] + 2);
   if ((g[0] + 2) * (g[2] + 2) > maxFace) maxFace = (g[0] + 2) * (g[2] + 2);
   h->bufCapacity = maxFace * 2;               /* cells per message; device buffers are sized by the device layer */
   return h;
}

HaloExchange* initForceHaloExchange(Domain* dom, LinkCell* boxes, int useGPU)
{
   (void)boxes; (void)useGPU;
   return makeExchange(dom, 1);
}

void haloExchange(HaloExchange* h, void* data)
{
   SimFlat* s = data;
   if (h && h->type == 1) return;              /* F'(rho): exchanged between pass 2 and pass 3 inside comd_dev_compute_force */
   DEV_TRY(comd_dev_redistribute(s->dev));
}

void destroyHaloExchange(HaloExchange** h)
{
   if (!h || !*h) return;
   free(*h);
   *h = NULL;
}
