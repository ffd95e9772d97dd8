/* Driver around the UNMODIFIED reference CoMD.c (compiled where it lies under
 * $(REF)/src-mpi): renames its main() and exposes the file-static
 * initSimulation()/SimFlat to the test-suite through a flat C API (ctypes).
 * Used to (1) pin oracle/comd_oracle.c against the reference, (2) generate
 * tests/golden/ fixtures, (3) time the reference CPU path for bench.py.
 * Test infrastructure only: never linked into the product. */
#define main comd_reference_main
#include "CoMD.c"
#undef main
#include <getopt.h>
#include <stdint.h>
#include "random.h"

void ref_get_tuple(LinkCell* boxes, int iBox, int* t);

static SimFlat* g_sim = NULL;
static LinkCell* g_altBoxes = NULL;

/* Build a simulation exactly as the reference's main() does up to the time loop
 * (CoMD.c:86-134): parseCommandLine + initSimulation (incl. first
 * redistribute, neighbour-list build and force call). */
int ref_init(int argc, char** argv)
{
   optind = 0;
   Command cmd = parseCommandLine(argc, argv);
   g_sim = initSimulation(cmd);
   sumAtoms(g_sim);
   return g_sim->atoms->nLocal;
}

void ref_destroy(void)
{
   /* destroySimulation frees structures some of which the shimmed CUDA path
    * double-registers; the oracle process is short-lived, so just drop it. */
   g_sim = NULL;
}

int ref_n_local(void) { sumAtoms(g_sim); return g_sim->atoms->nLocal; }
int ref_n_global(void) { return g_sim->atoms->nGlobal; }
int ref_n_local_boxes(void) { return g_sim->boxes->nLocalBoxes; }
int ref_n_total_boxes(void) { return g_sim->boxes->nTotalBoxes; }
int ref_maxatoms(void) { return MAXATOMS; }

/* geometry: out[0..2]=gridSize, dbl[0..2]=localMin, [3..5]=localMax, [6..8]=boxSize, [9..11]=invBoxSize */
void ref_geometry(int* grid, double* dbl)
{
   LinkCell* b = g_sim->boxes;
   for (int i = 0; i < 3; ++i)
   {
      grid[i] = b->gridSize[i];
      dbl[i] = b->localMin[i]; dbl[3 + i] = b->localMax[i];
      dbl[6 + i] = b->boxSize[i]; dbl[9 + i] = b->invBoxSize[i];
   }
}

void ref_potential(double* cutoff, double* mass, double* lat)
{
   *cutoff = g_sim->pot->cutoff; *mass = g_sim->pot->mass; *lat = g_sim->pot->lat;
}
BasePotential* ref_pot_handle(void) { return g_sim->pot; }

/* Local atoms in the reference's own traversal order (box-major, slot order).
 * r,p,f are n x 3 row-major; any pointer may be NULL. */
int ref_get_atoms(int* gid, int* box, double* r, double* p, double* f, double* U)
{
   int n = 0;
   LinkCell* b = g_sim->boxes; Atoms* a = g_sim->atoms;
   for (int iBox = 0; iBox < b->nLocalBoxes; ++iBox)
      for (int iOff = MAXATOMS * iBox, ii = 0; ii < b->nAtoms[iBox]; ++ii, ++iOff, ++n)
      {
         if (gid) gid[n] = a->gid[iOff];
         if (box) box[n] = iBox;
         if (r) { r[3*n] = a->r.x[iOff]; r[3*n+1] = a->r.y[iOff]; r[3*n+2] = a->r.z[iOff]; }
         if (p) { p[3*n] = a->p.x[iOff]; p[3*n+1] = a->p.y[iOff]; p[3*n+2] = a->p.z[iOff]; }
         if (f) { f[3*n] = a->f.x[iOff]; f[3*n+1] = a->f.y[iOff]; f[3*n+2] = a->f.z[iOff]; }
         if (U) U[n] = a->U[iOff];
      }
   return n;
}

/* Every atom incl. halo cells (ghosts), same order; returns count. */
int ref_get_all_atoms(int* gid, int* box, double* r)
{
   int n = 0;
   LinkCell* b = g_sim->boxes; Atoms* a = g_sim->atoms;
   for (int iBox = 0; iBox < b->nTotalBoxes; ++iBox)
      for (int iOff = MAXATOMS * iBox, ii = 0; ii < b->nAtoms[iBox]; ++ii, ++iOff, ++n)
      {
         if (gid) gid[n] = a->gid[iOff];
         if (box) box[n] = iBox;
         if (r) { r[3*n] = a->r.x[iOff]; r[3*n+1] = a->r.y[iOff]; r[3*n+2] = a->r.z[iOff]; }
      }
   return n;
}

void ref_box_counts(int* nAtoms) { for (int i = 0; i < g_sim->boxes->nTotalBoxes; ++i) nAtoms[i] = g_sim->boxes->nAtoms[i]; }

/* timestep() of the reference (timestep.c:48-100); returns via ref_energies. */
void ref_timestep(int nSteps) { timestep(g_sim, nSteps, g_sim->dt); }
void ref_compute_force(void) { computeForce(g_sim); kineticEnergy(g_sim); }
void ref_energies(double* ePot, double* eKin) { *ePot = g_sim->ePotential; *eKin = g_sim->eKinetic; }

/* The reference's ownership rule on an independent LinkCell geometry built with
 * an explicit cell-size cutoff (initLinkCells, linkCells.c:122; getBoxFromCoord,
 * linkCells.c:448-480).  Needed because -m cpu_nl sizes its own cells on
 * cutoff+skin while the device path sizes them on the cutoff. */
void ref_alt_linkcells(double cellCutoff)
{
   g_altBoxes = initLinkCells(g_sim->domain, cellCutoff, 0);
}
void ref_alt_geometry(int* grid, double* dbl)
{
   LinkCell* b = g_altBoxes;
   for (int i = 0; i < 3; ++i)
   {
      grid[i] = b->gridSize[i];
      dbl[i] = b->localMin[i]; dbl[3 + i] = b->localMax[i];
      dbl[6 + i] = b->boxSize[i]; dbl[9 + i] = b->invBoxSize[i];
   }
   grid[3] = b->nLocalBoxes; grid[4] = b->nTotalBoxes;
}
void ref_alt_box_from_coord(int n, const double* r, int* box, int* tuple /* 3n or NULL */)
{
   for (int i = 0; i < n; ++i)
   {
      real_t rr[3] = { r[3*i], r[3*i+1], r[3*i+2] };
      box[i] = getBoxFromCoord(g_altBoxes, rr);
      if (tuple) ref_get_tuple(g_altBoxes, box[i], tuple + 3*i);
   }
}
int ref_alt_box_from_tuple(int ix, int iy, int iz) { return getBoxFromTuple(g_altBoxes, ix, iy, iz); }

/* RNG known answers (random.c:21-75). */
unsigned long long ref_mk_seed(unsigned id, unsigned callSite) { return mkSeed(id, callSite); }
double ref_lcg61(unsigned long long* seed) { uint64_t s = *seed; double v = lcg61(&s); *seed = s; return v; }
double ref_gasdev(unsigned long long* seed) { uint64_t s = *seed; double v = gasdev(&s); *seed = s; return v; }

#ifdef REF_STANDALONE_MAIN
int main(int argc, char** argv) { return comd_reference_main(argc, argv); }
#endif
