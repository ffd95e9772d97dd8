/* comd_oracle.c -- CPU restatement of CoMD's short-range force path (the parity ORACLE).
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (comd-cuda-async_b200/, include/)
 * may import, link or call this file; only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py do, and only as the checker.
 *
 * What it restates (all citations are into /root/reference/src-mpi):
 *   - link-cell geometry, halo-cell numbering and the ownership rule   linkCells.c:122-182, 299-346, 448-480, 497-568
 *   - putAtomInBox / moveAtom / updateLinkCellsCpu / emptyHaloCells    linkCells.c:228-256, 352-405, 483-487
 *   - lcg61 / mkSeed / gasdev                                          random.c:21-75
 *   - createFccLattice / setTemperature / randomDisplacements          initAtoms.c:81-216
 *   - funcfl / setfl readers, table padding, interpolate               eam.c:496-519, 557-579, 680-872
 *   - half neighbour list with skin and its rebuild criterion          neighborList.c:101-247
 *   - ljForceCpuNL                                                     ljForce.c:146-265
 *   - eamForceCpuNL (three loops + F' halo exchange)                   eam.c:266-419
 *   - atom halo exchange (3 serial axis phases, PBC shift, FIFO of
 *     remembered ghost slots) and F' exchange, single rank / periodic  haloExchange.c:198-345, 1493-1860; hashTable.c:35-94
 *   - redistributeAtomsCpuNL, advanceVelocity/PositionCpu,
 *     kineticEnergy, timestep                                          timestep.c:48-197, 356-405
 *
 * Parity status: PINNED.  tests/test_oracle.py checks this file against the
 * reference itself (oracle/_ref/libcomd_ref.so, built from the reference's own sources
 * by oracle/ref_build/Makefile) bit-for-bit: cell ids, RNG streams, initial r and p,
 * tables, per-atom f and U at step 0 and the energy trace over time steps; and against
 * the reference's documented known answers (CoMD.c:896-900).
 *
 * The arithmetic deliberately keeps the reference's operation order (no FMA
 * contraction: build with -ffp-contract=off) so that results are bit-identical.
 * Single rank only: the single-rank periodic run is the ground truth that every
 * multi-GPU decomposition has to reproduce.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>

#ifndef ORC_MAXATOMS
#define ORC_MAXATOMS 64       /* MAXATOMS, Makefile:16 */
#endif
#define ORC_MAXNBR 1024       /* neighborList.c:58 */
#define ORC_VECW 4            /* VECTOR_WIDTH, defines.h:64 */

/* constants.h:28-39 */
static const double orcAmuToInternalMass = 1.660538921e-27 * 1.0e-10 * 1.0e-10 / (1.0e-15 * 1.0e-15 * 1.602176565e-19);
static const double orcKbEv = 8.6173324e-5;
static const double orcHartreeToEv = 27.21138505;
static const double orcBohrToAngs = 0.52917721092;

typedef struct { int n; double x0, invDx; double* v; /* v[-1..n+1] valid */ } OrcTable;

typedef struct OrcSim
{
   /* domain: one rank owns the whole periodic box (decomposition.c:18-50 with procGrid 1,1,1) */
   double gExt[3], lMin[3], lMax[3];
   /* link cells */
   int grid[3]; double boxSize[3], invBox[3];
   int nLocalBoxes, nHaloBoxes, nTotalBoxes;
   int* nAtoms;
   /* atom slabs, nTotalBoxes*MAXATOMS+1 entries; the last one is the atom at infinity (initAtoms.c:26-62) */
   int slab;
   int *gid, *species, *lid;
   double *r[3], *p[3], *f[3], *U;
   int nLocal, nGlobal;
   /* potential */
   int isEam; double cutoff, mass, lat;
   double sigma, epsilon;
   OrcTable phi, rho, fem;
   double *rhobar, *dfEmbed;
   /* neighbour list (neighborList.c:53-75) */
   double skin, skinHalf2;
   int nMaxLocal; int* list; int* nNbr; double* lastR[3];
   int nlUpdateRequired, nlForceRebuild, nlLinkCellsRequired, nStepsSinceBuild, nBuilds;
   /* FIFO of ghost slots (hashTable.c) */
   int* fifo; int fifoPut, fifoGet, fifoCap;
   /* exchange cell lists */
   int* atomCells[6]; int nAtomCells[6];
   int* fSend[6]; int* fRecv[6]; int nFCells[6];
   double pbc[6][3];
   double ePot, eKin, dt;
} OrcSim;

/* ---------------------------------------------------------------- RNG (random.c) */
double orc_lcg61(uint64_t* seed)
{
   static const double toDouble = 1.0 / UINT64_C(2305843009213693951);
   *seed *= UINT64_C(437799614237992725);
   *seed %= UINT64_C(2305843009213693951);
   return *seed * toDouble;
}
uint64_t orc_mk_seed(uint32_t id, uint32_t callSite)
{
   uint32_t a = id * UINT32_C(2654435761);
   uint32_t b = (id + callSite) * UINT32_C(2654435761);
   uint64_t s = (UINT64_C(0x100000000) * a) + b;
   for (unsigned k = 0; k < 10; ++k) orc_lcg61(&s);
   return s;
}
double orc_gasdev(uint64_t* seed)
{
   double rsq, v1, v2;
   do {
      v1 = 2.0 * orc_lcg61(seed) - 1.0;
      v2 = 2.0 * orc_lcg61(seed) - 1.0;
      rsq = v1 * v1 + v2 * v2;
   } while (rsq >= 1.0 || rsq == 0.0);
   return v2 * sqrt(-2.0 * log(rsq) / rsq);
}

/* ---------------------------------------------------------------- link cells (linkCells.c) */
static void cellsInit(OrcSim* s, double cellCutoff)
{
   for (int i = 0; i < 3; ++i)
   {
      double ext = s->gExt[i] / 1;            /* decomposition.c:44: localExtent = globalExtent/procGrid, procGrid 1 */
      s->grid[i] = (int)(ext / cellCutoff);
      s->boxSize[i] = ext / ((double)s->grid[i]);
      s->invBox[i] = 1.0 / s->boxSize[i];
   }
   const int* g = s->grid;
   s->nLocalBoxes = g[0] * g[1] * g[2];
   s->nHaloBoxes = 2 * ((g[0] + 2) * (g[1] + g[2] + 2) + (g[1] * g[2]));
   s->nTotalBoxes = s->nLocalBoxes + s->nHaloBoxes;
}

int orc_box_from_tuple(const OrcSim* s, int ix, int iy, int iz)
{
   const int* g = s->grid;
   const int nl = s->nLocalBoxes;
   if (iz == g[2])  return nl + 2*g[2]*g[1] + 2*g[2]*(g[0]+2) + (g[0]+2)*(g[1]+2) + (g[0]+2)*(iy+1) + (ix+1);
   if (iz == -1)    return nl + 2*g[2]*g[1] + 2*g[2]*(g[0]+2) + (g[0]+2)*(iy+1) + (ix+1);
   if (iy == g[1])  return nl + 2*g[2]*g[1] + g[2]*(g[0]+2) + (g[0]+2)*iz + (ix+1);
   if (iy == -1)    return nl + 2*g[2]*g[1] + iz*(g[0]+2) + (ix+1);
   if (ix == g[0])  return nl + g[1]*g[2] + iz*g[1] + iy;
   if (ix == -1)    return nl + iz*g[1] + iy;
   return ix + g[0]*iy + g[0]*g[1]*iz;
}

void orc_tuple_from_box(const OrcSim* s, int iBox, int* t)
{
   const int* g = s->grid;
   int ix, iy, iz;
   if (iBox < s->nLocalBoxes)
   {
      ix = iBox % g[0]; iy = (iBox / g[0]) % g[1]; iz = iBox / (g[0]*g[1]);
   }
   else
   {
      int k = iBox - s->nLocalBoxes;
      if (k < 2*g[1]*g[2])
      {
         if (k < g[1]*g[2]) ix = 0; else { k -= g[1]*g[2]; ix = g[0] + 1; }
         iy = 1 + k % g[1]; iz = 1 + k / g[1];
      }
      else if (k < 2*g[2]*(g[1] + g[0] + 2))
      {
         k -= 2*g[2]*g[1];
         if (k < (g[0]+2)*g[2]) iy = 0; else { k -= (g[0]+2)*g[2]; iy = g[1] + 1; }
         ix = k % (g[0]+2); iz = 1 + k / (g[0]+2);
      }
      else
      {
         k -= 2*g[2]*(g[1] + g[0] + 2);
         if (k < (g[0]+2)*(g[1]+2)) iz = 0; else { k -= (g[0]+2)*(g[1]+2); iz = g[2] + 1; }
         ix = k % (g[0]+2); iy = k / (g[0]+2);
      }
      --ix; --iy; --iz;
   }
   t[0] = ix; t[1] = iy; t[2] = iz;
}

/* ownership rule, linkCells.c:448-480 */
void orc_tuple_from_coord(const OrcSim* s, const double rr[3], int* t)
{
   for (int a = 0; a < 3; ++a)
   {
      int i = (int)(floor((rr[a] - s->lMin[a]) * s->invBox[a]));
      if (rr[a] < s->lMax[a]) { if (i == s->grid[a]) i = s->grid[a] - 1; }
      else i = s->grid[a];
      t[a] = i;
   }
}
int orc_box_from_coord(const OrcSim* s, const double rr[3])
{
   int t[3];
   orc_tuple_from_coord(s, rr, t);
   return orc_box_from_tuple(s, t[0], t[1], t[2]);
}

static int neighborBoxes(const OrcSim* s, int iBox, int* nbr)
{
   int t[3], n = 0;
   orc_tuple_from_box(s, iBox, t);
   for (int i = t[0]-1; i <= t[0]+1; ++i)
      for (int j = t[1]-1; j <= t[1]+1; ++j)
         for (int k = t[2]-1; k <= t[2]+1; ++k)
            nbr[n++] = orc_box_from_tuple(s, i, j, k);
   return n;
}

static int putAtom(OrcSim* s, int gid, int type, double x, double y, double z, double px, double py, double pz)
{
   double rr[3] = { x, y, z };
   int iBox = orc_box_from_coord(s, rr);
   int iOff = iBox * ORC_MAXATOMS + s->nAtoms[iBox];
   if (s->nAtoms[iBox] >= ORC_MAXATOMS) { fprintf(stderr, "comd_oracle: link cell %d overflows MAXATOMS=%d\n", iBox, ORC_MAXATOMS); abort(); }
   if (iBox < s->nLocalBoxes) s->nLocal++;
   s->nAtoms[iBox]++;
   s->gid[iOff] = gid; s->species[iOff] = type;
   s->r[0][iOff] = x; s->r[1][iOff] = y; s->r[2][iOff] = z;
   s->p[0][iOff] = px; s->p[1][iOff] = py; s->p[2][iOff] = pz;
   return iOff;
}

static void copyAtom(OrcSim* s, int iAtom, int iBox, int jAtom, int jBox)
{
   const int i = ORC_MAXATOMS * iBox + iAtom, j = ORC_MAXATOMS * jBox + jAtom;
   s->gid[j] = s->gid[i]; s->species[j] = s->species[i];
   for (int a = 0; a < 3; ++a) { s->r[a][j] = s->r[a][i]; s->p[a][j] = s->p[a][i]; s->f[a][j] = s->f[a][i]; }
   s->U[j] = s->U[i];
}

static void moveAtom(OrcSim* s, int iId, int iBox, int jBox)
{
   int nj = s->nAtoms[jBox];
   if (nj >= ORC_MAXATOMS) { fprintf(stderr, "comd_oracle: link cell %d overflows MAXATOMS\n", jBox); abort(); }
   copyAtom(s, iId, iBox, nj, jBox);
   s->nAtoms[jBox]++;
   s->nAtoms[iBox]--;
   int ni = s->nAtoms[iBox];
   if (ni) copyAtom(s, ni, iBox, iId, iBox);
   if (jBox > s->nLocalBoxes) --s->nLocal;    /* sic: '>' as in linkCells.c:365 */
}

static void emptyHaloCells(OrcSim* s)
{
   for (int i = s->nLocalBoxes; i < s->nTotalBoxes; ++i) s->nAtoms[i] = 0;
}

static void updateLinkCells(OrcSim* s)
{
   emptyHaloCells(s);
   for (int iBox = 0; iBox < s->nLocalBoxes; ++iBox)
   {
      int iOff = iBox * ORC_MAXATOMS, ii = 0;
      while (ii < s->nAtoms[iBox])
      {
         double rr[3] = { s->r[0][iOff+ii], s->r[1][iOff+ii], s->r[2][iOff+ii] };
         int jBox = orc_box_from_coord(s, rr);
         if (jBox != iBox) moveAtom(s, ii, iBox, jBox); else ++ii;
      }
   }
}

static void updateLocalIds(OrcSim* s)
{
   int c = 0;
   for (int iBox = 0; iBox < s->nLocalBoxes; ++iBox)
      for (int iOff = ORC_MAXATOMS*iBox, ii = 0; ii < s->nAtoms[iBox]; ++ii, ++iOff) s->lid[iOff] = c++;
}

/* ---------------------------------------------------------------- tables (eam.c:496-519, 557-579) */
static void tableInit(OrcTable* t, int n, double x0, double dx, const double* data)
{
   double* buf = (double*)calloc(n + 3, sizeof(double));
   t->v = buf + 1; t->n = n; t->invDx = 1.0 / dx; t->x0 = x0;
   for (int i = 0; i < n; ++i) t->v[i] = data[i];
   t->v[-1] = t->v[0];
   t->v[n+1] = t->v[n] = t->v[n-1];
}

static inline void interp(const OrcTable* t, double r, double* f, double* df)
{
   const double* tt = t->v;
   if (r < t->x0) r = t->x0;
   r = (r - t->x0) * (t->invDx);
   int ii = (int)floor(r);
   if (ii > t->n) { ii = t->n; r = t->n / t->invDx; }
   r = r - floor(r);
   double g1 = tt[ii+1] - tt[ii-1];
   double g2 = tt[ii+2] - tt[ii];
   *f = tt[ii] + 0.5 * r * (g1 + r * (tt[ii+1] + tt[ii-1] - 2.0 * tt[ii]));
   *df = 0.5 * (g1 + r * (g2 - g1)) * t->invDx;
}

void orc_interpolate(const OrcSim* s, int which, int n, const double* x, double* f, double* df)
{
   const OrcTable* t = which == 0 ? &s->phi : (which == 1 ? &s->rho : &s->fem);
   for (int i = 0; i < n; ++i) interp(t, x[i], f + i, df + i);
}
int orc_table_n(const OrcSim* s, int which) { return (which == 0 ? s->phi : (which == 1 ? s->rho : s->fem)).n; }
void orc_table(const OrcSim* s, int which, double* x0, double* invDx, double* padded)
{
   const OrcTable* t = which == 0 ? &s->phi : (which == 1 ? &s->rho : &s->fem);
   *x0 = t->x0; *invDx = t->invDx;
   for (int i = -1; i < t->n + 2; ++i) padded[i+1] = t->v[i];
}

/* eamReadFuncfl, eam.c:802-872 */
static int readFuncfl(OrcSim* s, const char* path)
{
   char line[4096];
   FILE* fp = fopen(path, "r");
   if (!fp) return -1;
   fgets(line, sizeof line, fp);                       /* comment / element name */
   int z, nRho, nR; double mass, lat, dRho, dR, cut; char lt[64];
   fgets(line, sizeof line, fp); sscanf(line, "%d %le %le %63s", &z, &mass, &lat, lt);
   s->lat = lat; s->mass = mass * orcAmuToInternalMass;
   fgets(line, sizeof line, fp); sscanf(line, "%d %le %d %le %le", &nRho, &dRho, &nR, &dR, &cut);
   s->cutoff = cut;
   int nb = nRho > nR ? nRho : nR;
   double* buf = (double*)malloc(nb * sizeof(double));
   for (int i = 0; i < nRho; ++i) fscanf(fp, "%lg", buf + i);
   tableInit(&s->fem, nRho, 0.0, dRho, buf);
   for (int i = 0; i < nR; ++i) fscanf(fp, "%lg", buf + i);
   for (int i = 1; i < nR; ++i)
   {
      double r = 0.0 + i * dR;
      buf[i] *= buf[i] / r;
      buf[i] *= orcHartreeToEv * orcBohrToAngs;
   }
   buf[0] = buf[1] + (buf[1] - buf[2]);
   tableInit(&s->phi, nR, 0.0, dR, buf);
   for (int i = 0; i < nR; ++i) fscanf(fp, "%lg", buf + i);
   tableInit(&s->rho, nR, 0.0, dR, buf);
   free(buf); fclose(fp);
   return 0;
}

/* eamReadSetfl, eam.c:680-757 */
static int readSetfl(OrcSim* s, const char* path)
{
   char line[4096];
   FILE* fp = fopen(path, "r");
   if (!fp) return -1;
   for (int i = 0; i < 3; ++i) fgets(line, sizeof line, fp);
   int nElems; fgets(line, sizeof line, fp); sscanf(line, "%d", &nElems);
   if (nElems != 1) { fclose(fp); return -2; }
   int nRho, nR, z; double dRho, dR, cut, mass, lat; char lt[64];
   fgets(line, sizeof line, fp); sscanf(line, "%d %le %d %le %le", &nRho, &dRho, &nR, &dR, &cut);
   s->cutoff = cut;
   fgets(line, sizeof line, fp); sscanf(line, "%d %le %le %63s", &z, &mass, &lat, lt);
   s->lat = lat; s->mass = mass * orcAmuToInternalMass;
   int nb = nRho > nR ? nRho : nR;
   double* buf = (double*)malloc(nb * sizeof(double));
   for (int i = 0; i < nRho; ++i) fscanf(fp, "%lg", buf + i);
   tableInit(&s->fem, nRho, 0.0, dRho, buf);
   for (int i = 0; i < nR; ++i) fscanf(fp, "%lg", buf + i);
   tableInit(&s->rho, nR, 0.0, dR, buf);
   for (int i = 0; i < nR; ++i) fscanf(fp, "%lg", buf + i);
   for (int i = 1; i < nR; ++i) { double r = 0.0 + i * dR; buf[i] /= r; }
   buf[0] = buf[1] + (buf[1] - buf[2]);
   tableInit(&s->phi, nR, 0.0, dR, buf);
   free(buf); fclose(fp);
   return 0;
}

/* ---------------------------------------------------------------- exchange lists (haloExchange.c:1543-1567, 1712-1801) */
static int* mkList(const OrcSim* s, int xb, int xe, int yb, int ye, int zb, int ze, int* nOut)
{
   int n = (xe-xb)*(ye-yb)*(ze-zb), c = 0;
   int* l = (int*)malloc(n * sizeof(int));
   for (int ix = xb; ix < xe; ++ix) for (int iy = yb; iy < ye; ++iy) for (int iz = zb; iz < ze; ++iz)
      l[c++] = orc_box_from_tuple(s, ix, iy, iz);
   *nOut = n;
   return l;
}

static void exchangeInit(OrcSim* s)
{
   const int nx = s->grid[0], ny = s->grid[1], nz = s->grid[2];
   for (int face = 0; face < 6; ++face)
   {
      int b[3] = { -1, -1, -1 }, e[3] = { nx+1, ny+1, nz+1 };
      int ax = face / 2;
      if (face % 2 == 0) e[ax] = b[ax] + 2; else b[ax] = e[ax] - 2;
      s->atomCells[face] = mkList(s, b[0], e[0], b[1], e[1], b[2], e[2], &s->nAtomCells[face]);
      for (int a = 0; a < 3; ++a) s->pbc[face][a] = 0.0;
      /* single rank: procCoord == 0 == procGrid-1 on every axis (haloExchange.c:310-323) */
      s->pbc[face][ax] = (face % 2 == 0) ? +1.0 : -1.0;
   }
   /* force exchange send / recv planes */
   int sb[6][6] = {
      { 0, 1,      0, ny,     0, nz }, { nx-1, nx,  0, ny,     0, nz },
      { -1, nx+1,  0, 1,      0, nz }, { -1, nx+1,  ny-1, ny,  0, nz },
      { -1, nx+1,  -1, ny+1,  0, 1  }, { -1, nx+1,  -1, ny+1,  nz-1, nz } };
   int rb[6][6] = {
      { -1, 0,     0, ny,     0, nz }, { nx, nx+1,  0, ny,     0, nz },
      { -1, nx+1,  -1, 0,     0, nz }, { -1, nx+1,  ny, ny+1,  0, nz },
      { -1, nx+1,  -1, ny+1,  -1, 0 }, { -1, nx+1,  -1, ny+1,  nz, nz+1 } };
   for (int face = 0; face < 6; ++face)
   {
      int n1, n2;
      s->fSend[face] = mkList(s, sb[face][0], sb[face][1], sb[face][2], sb[face][3], sb[face][4], sb[face][5], &n1);
      s->fRecv[face] = mkList(s, rb[face][0], rb[face][1], rb[face][2], rb[face][3], rb[face][4], rb[face][5], &n2);
      s->nFCells[face] = n1;
   }
   s->fifoCap = (s->nTotalBoxes - s->nLocalBoxes) * ORC_MAXATOMS * 2;
   s->fifo = (int*)malloc(s->fifoCap * sizeof(int));
   s->fifoPut = s->fifoGet = 0;
}

typedef struct { int gid, type; double rx, ry, rz, px, py, pz; } OrcAtomMsg;

static int nlUpdateRequired(OrcSim* s);

static int loadAtoms(OrcSim* s, int face, OrcAtomMsg* buf)
{
   double sh[3];
   for (int a = 0; a < 3; ++a) sh[a] = s->pbc[face][a] * s->gExt[a];
   int n = 0;
   for (int ic = 0; ic < s->nAtomCells[face]; ++ic)
   {
      int iBox = s->atomCells[face][ic], iOff = iBox * ORC_MAXATOMS;
      for (int ii = iOff; ii < iOff + s->nAtoms[iBox]; ++ii, ++n)
      {
         buf[n].gid = s->gid[ii]; buf[n].type = s->species[ii];
         buf[n].rx = s->r[0][ii] + sh[0]; buf[n].ry = s->r[1][ii] + sh[1]; buf[n].rz = s->r[2][ii] + sh[2];
         buf[n].px = s->p[0][ii]; buf[n].py = s->p[1][ii]; buf[n].pz = s->p[2][ii];
      }
   }
   return n;
}

static void unloadAtoms(OrcSim* s, int n, const OrcAtomMsg* buf)
{
   const int rebuild = nlUpdateRequired(s);
   for (int i = 0; i < n; ++i)
   {
      if (rebuild)
      {
         int iOff = putAtom(s, buf[i].gid, buf[i].type, buf[i].rx, buf[i].ry, buf[i].rz, buf[i].px, buf[i].py, buf[i].pz);
         if (iOff >= ORC_MAXATOMS * s->nLocalBoxes) s->fifo[s->fifoPut++] = iOff;
         else s->nlLinkCellsRequired = 1;
      }
      else
      {
         int iOff = s->fifo[s->fifoGet++];
         s->nAtoms[iOff / ORC_MAXATOMS]++;
         s->gid[iOff] = buf[i].gid; s->species[iOff] = buf[i].type;
         s->r[0][iOff] = buf[i].rx; s->r[1][iOff] = buf[i].ry; s->r[2][iOff] = buf[i].rz;
         s->p[0][iOff] = buf[i].px; s->p[1][iOff] = buf[i].py; s->p[2][iOff] = buf[i].pz;
      }
   }
}

static void atomHaloExchange(OrcSim* s)
{
   int cap = 0;
   for (int f = 0; f < 6; ++f) if (s->nAtomCells[f] > cap) cap = s->nAtomCells[f];
   OrcAtomMsg* bM = (OrcAtomMsg*)malloc((size_t)cap * ORC_MAXATOMS * sizeof(OrcAtomMsg));
   OrcAtomMsg* bP = (OrcAtomMsg*)malloc((size_t)cap * ORC_MAXATOMS * sizeof(OrcAtomMsg));
   for (int ax = 0; ax < 3; ++ax)
   {
      int nM = loadAtoms(s, 2*ax, bM);
      int nP = loadAtoms(s, 2*ax + 1, bP);
      /* self exchange (parallel.c:112-117): what went out of the - face arrives as
       * the + face's receive and vice versa; unload order M then P (haloExchange.c:1518-1519) */
      unloadAtoms(s, nP, bP);    /* recvBufM <- sendBufP */
      unloadAtoms(s, nM, bM);    /* recvBufP <- sendBufM */
   }
   free(bM); free(bP);
}

static void forceHaloExchange(OrcSim* s)
{
   int cap = 0;
   for (int f = 0; f < 6; ++f) if (s->nFCells[f] > cap) cap = s->nFCells[f];
   double* bM = (double*)malloc((size_t)cap * ORC_MAXATOMS * sizeof(double));
   double* bP = (double*)malloc((size_t)cap * ORC_MAXATOMS * sizeof(double));
   for (int ax = 0; ax < 3; ++ax)
   {
      double* sb[2] = { bM, bP };
      for (int d = 0; d < 2; ++d)
      {
         int face = 2*ax + d, n = 0;
         for (int ic = 0; ic < s->nFCells[face]; ++ic)
         {
            int iBox = s->fSend[face][ic], iOff = iBox * ORC_MAXATOMS;
            for (int ii = iOff; ii < iOff + s->nAtoms[iBox]; ++ii) sb[d][n++] = s->dfEmbed[ii];
         }
      }
      /* unload faceM from what the - neighbour sent across its + face, then faceP */
      for (int d = 0; d < 2; ++d)
      {
         int face = 2*ax + d, n = 0;
         const double* rb = sb[1 - d];
         for (int ic = 0; ic < s->nFCells[face]; ++ic)
         {
            int iBox = s->fRecv[face][ic], iOff = iBox * ORC_MAXATOMS;
            for (int ii = iOff; ii < iOff + s->nAtoms[iBox]; ++ii) s->dfEmbed[ii] = rb[n++];
         }
      }
   }
   free(bM); free(bP);
}

/* ---------------------------------------------------------------- neighbour list (neighborList.c) */
static int nlUpdateRequired(OrcSim* s)
{
   if (s->nlForceRebuild == 1) s->nlUpdateRequired = 1;
   else if (s->nlUpdateRequired == -1)
   {
      for (int iBox = 0; iBox < s->nLocalBoxes && s->nlUpdateRequired != 1; ++iBox)
         for (int iOff = ORC_MAXATOMS*iBox, ii = 0; ii < s->nAtoms[iBox] && s->nlUpdateRequired != 1; ++ii, ++iOff)
         {
            int l = s->lid[iOff];
            double dx = s->lastR[0][l] - s->r[0][iOff];
            double dy = s->lastR[1][l] - s->r[1][iOff];
            double dz = s->lastR[2][l] - s->r[2][iOff];
            if ((dx*dx + dy*dy + dz*dz) > s->skinHalf2) s->nlUpdateRequired = 1;
         }
      s->nlUpdateRequired = s->nlUpdateRequired > 0 ? 1 : 0;
   }
   return s->nlUpdateRequired;
}

static void buildNeighborList(OrcSim* s)
{
   if (s->nlUpdateRequired == -1) nlUpdateRequired(s);
   if (s->nlUpdateRequired != 1) { s->nStepsSinceBuild++; return; }
   for (int i = 0; i < s->nMaxLocal; ++i) s->nNbr[i] = 0;
   const double rc = s->cutoff + s->skin, rc2 = rc * rc;
   int nbr[27];
   for (int iBox = 0; iBox < s->nLocalBoxes; ++iBox)
   {
      const int nI = s->nAtoms[iBox];
      const int nNb = neighborBoxes(s, iBox, nbr);
      for (int jt = 0; jt < nNb; ++jt)
      {
         const int jBox = nbr[jt];
         if (jBox < iBox) continue;
         const int nJ = s->nAtoms[jBox];
         for (int iOff = ORC_MAXATOMS*iBox, ii = 0; ii < nI; ++ii, ++iOff)
         {
            const int l = s->lid[iOff];
            int* il = s->list + (size_t)ORC_MAXNBR * l;
            int nn = s->nNbr[l];
            for (int a = 0; a < 3; ++a) s->lastR[a][l] = s->r[a][iOff];
            for (int jOff = ORC_MAXATOMS*jBox, ij = 0; ij < nJ; ++ij, ++jOff)
            {
               if (iBox == jBox && ij <= ii) continue;
               double dx = s->r[0][iOff] - s->r[0][jOff];
               double dy = s->r[1][iOff] - s->r[1][jOff];
               double dz = s->r[2][iOff] - s->r[2][jOff];
               double r2 = dx*dx + dy*dy + dz*dz;
               if (r2 > rc2) continue;
               if (nn >= ORC_MAXNBR) { fprintf(stderr, "comd_oracle: neighbour list overflow\n"); abort(); }
               il[nn++] = jOff;
            }
            s->nNbr[l] = nn;
         }
      }
      for (int iOff = ORC_MAXATOMS*iBox, ii = 0; ii < nI; ++ii, ++iOff)
      {
         const int l = s->lid[iOff];
         int* il = s->list + (size_t)ORC_MAXNBR * l;
         int nn = s->nNbr[l];
         int pad = ORC_VECW - (nn % ORC_VECW);
         for (int k = 0; k < pad; ++k) il[nn++] = ORC_MAXATOMS * s->nTotalBoxes;
         s->nNbr[l] = nn;
      }
   }
   s->nlForceRebuild = 0; s->nStepsSinceBuild = 1; s->nlUpdateRequired = 0; s->nBuilds++;
}

/* ---------------------------------------------------------------- forces */
static void zeroForces(OrcSim* s)
{
   const int n = s->nTotalBoxes * ORC_MAXATOMS;
   for (int i = 0; i < n; ++i) { s->U[i] = 0.0; s->f[0][i] = 0.0; s->f[1][i] = 0.0; s->f[2][i] = 0.0; }
}

/* ljForceCpuNL, ljForce.c:146-265 (POT_SHIFT 1.0, ljForce.c:83) */
static void ljForce(OrcSim* s)
{
   const double sigma = s->sigma, epsilon = s->epsilon, rCut = s->cutoff, rCut2 = rCut * rCut;
   zeroForces(s);
   const double s6 = sigma*sigma*sigma*sigma*sigma*sigma;
   const double rCut6 = s6 / (rCut2*rCut2*rCut2);
   const double eShift = 1.0 * rCut6 * (rCut6 - 1.0);
   double *rx = s->r[0], *ry = s->r[1], *rz = s->r[2], *fx = s->f[0], *fy = s->f[1], *fz = s->f[2], *U = s->U;
   for (int iBox = 0; iBox < s->nLocalBoxes; ++iBox)
      for (int iOff = ORC_MAXATOMS*iBox, ii = 0; ii < s->nAtoms[iBox]; ++ii, ++iOff)
      {
         const int l = s->lid[iOff];
         const int* il = s->list + (size_t)ORC_MAXNBR * l;
         const int nn = s->nNbr[l];
         const double irx = rx[iOff], iry = ry[iOff], irz = rz[iOff];
         double ifx = 0.0, ify = 0.0, ifz = 0.0, iU = 0.0;
         for (int k = 0; k < nn; ++k)
         {
            const int jOff = il[k];
            double drx = irx - rx[jOff], dry = iry - ry[jOff], drz = irz - rz[jOff];
            double r2 = drx*drx + dry*dry + drz*drz;
            double fr = 0.0, eLocal = 0.0;
            if (r2 <= rCut2)
            {
               r2 = 1.0 / r2;
               double r6 = (s6*r2) * (r2*r2);
               eLocal = r6 * (r6 - 1.0) - eShift;
               iU += eLocal;
               fr = -4.0*epsilon*r6*r2*(12.0*r6 - 6.0);
               ifx -= drx*fr; ify -= dry*fr; ifz -= drz*fr;
            }
            U[jOff] += eLocal;
            fx[jOff] += drx*fr; fy[jOff] += dry*fr; fz[jOff] += drz*fr;
         }
         U[iOff] += iU; fx[iOff] += ifx; fy[iOff] += ify; fz[iOff] += ifz;
      }
   double ePot = 0.0;
   for (int iBox = 0; iBox < s->nLocalBoxes; ++iBox)
      for (int iOff = ORC_MAXATOMS*iBox, ii = 0; ii < s->nAtoms[iBox]; ++ii, ++iOff)
      {
         U[iOff] *= 0.5;
         ePot += U[iOff];
      }
   s->ePot = ePot * 4.0 * epsilon;
}

/* eamForceCpuNL, eam.c:266-419 */
static void eamForce(OrcSim* s)
{
   const double rCut2 = s->cutoff * s->cutoff;
   const int nAll = s->nTotalBoxes * ORC_MAXATOMS;
   double etot = 0.0;
   zeroForces(s);
   memset(s->dfEmbed, 0, nAll * sizeof(double));
   memset(s->rhobar, 0, nAll * sizeof(double));
   double *rx = s->r[0], *ry = s->r[1], *rz = s->r[2], *fx = s->f[0], *fy = s->f[1], *fz = s->f[2];
   for (int iBox = 0; iBox < s->nLocalBoxes; ++iBox)
      for (int iOff = ORC_MAXATOMS*iBox, ii = 0; ii < s->nAtoms[iBox]; ++ii, ++iOff)
      {
         const int l = s->lid[iOff];
         const int* il = s->list + (size_t)ORC_MAXNBR * l;
         const int nn = s->nNbr[l];
         for (int k = 0; k < nn; ++k)
         {
            const int jOff = il[k];
            double r2 = 0.0, dr[3];
            dr[0] = rx[iOff] - rx[jOff]; dr[1] = ry[iOff] - ry[jOff]; dr[2] = rz[iOff] - rz[jOff];
            r2 += dr[0]*dr[0] + dr[1]*dr[1] + dr[2]*dr[2];
            if (r2 > rCut2) continue;
            double r = sqrt(r2);
            double phiT, dPhi, rhoT, dRho;
            interp(&s->phi, r, &phiT, &dPhi);
            interp(&s->rho, r, &rhoT, &dRho);
            fx[iOff] -= dPhi*dr[0]/r; fy[iOff] -= dPhi*dr[1]/r; fz[iOff] -= dPhi*dr[2]/r;
            fx[jOff] += dPhi*dr[0]/r; fy[jOff] += dPhi*dr[1]/r; fz[jOff] += dPhi*dr[2]/r;
            if (jOff / ORC_MAXATOMS < s->nLocalBoxes) etot += phiT; else etot += 0.5*phiT;
            s->U[iOff] += 0.5*phiT; s->U[jOff] += 0.5*phiT;
            s->rhobar[iOff] += rhoT; s->rhobar[jOff] += rhoT;
         }
      }
   for (int iBox = 0; iBox < s->nLocalBoxes; ++iBox)
      for (int iOff = ORC_MAXATOMS*iBox, ii = 0; ii < s->nAtoms[iBox]; ++ii, ++iOff)
      {
         double fE, dfE;
         interp(&s->fem, s->rhobar[iOff], &fE, &dfE);
         s->dfEmbed[iOff] = dfE;
         etot += fE;
         s->U[iOff] += fE;
      }
   forceHaloExchange(s);
   for (int iBox = 0; iBox < s->nLocalBoxes; ++iBox)
      for (int iOff = ORC_MAXATOMS*iBox, ii = 0; ii < s->nAtoms[iBox]; ++ii, ++iOff)
      {
         const int l = s->lid[iOff];
         const int* il = s->list + (size_t)ORC_MAXNBR * l;
         const int nn = s->nNbr[l];
         for (int k = 0; k < nn; ++k)
         {
            const int jOff = il[k];
            double r2 = 0.0, dr[3];
            dr[0] = rx[iOff] - rx[jOff]; dr[1] = ry[iOff] - ry[jOff]; dr[2] = rz[iOff] - rz[jOff];
            r2 += dr[0]*dr[0] + dr[1]*dr[1] + dr[2]*dr[2];
            if (r2 >= rCut2) continue;
            double r = sqrt(r2);
            double rhoT, dRho;
            interp(&s->rho, r, &rhoT, &dRho);
            const double w = (s->dfEmbed[iOff] + s->dfEmbed[jOff]);
            fx[iOff] -= w*dRho*dr[0]/r; fy[iOff] -= w*dRho*dr[1]/r; fz[iOff] -= w*dRho*dr[2]/r;
            fx[jOff] += w*dRho*dr[0]/r; fy[jOff] += w*dRho*dr[1]/r; fz[jOff] += w*dRho*dr[2]/r;
         }
      }
   s->ePot = etot;
}

static void computeForce(OrcSim* s) { if (s->isEam) eamForce(s); else ljForce(s); }

/* ---------------------------------------------------------------- integrator (timestep.c) */
static void kineticEnergy(OrcSim* s)
{
   double eK = 0.0;
   for (int iBox = 0; iBox < s->nLocalBoxes; ++iBox)
      for (int iOff = ORC_MAXATOMS*iBox, ii = 0; ii < s->nAtoms[iBox]; ++ii, ++iOff)
      {
         double invMass = 0.5 / s->mass;
         eK += (s->p[0][iOff]*s->p[0][iOff] + s->p[1][iOff]*s->p[1][iOff] + s->p[2][iOff]*s->p[2][iOff]) * invMass;
      }
   s->eKin = eK;
}

static void advanceVelocity(OrcSim* s, double dt)
{
   for (int iBox = 0; iBox < s->nLocalBoxes; ++iBox)
      for (int iOff = ORC_MAXATOMS*iBox, ii = 0; ii < s->nAtoms[iBox]; ++ii, ++iOff)
         for (int a = 0; a < 3; ++a) s->p[a][iOff] += dt * s->f[a][iOff];
}

static void advancePosition(OrcSim* s, double dt)
{
   for (int iBox = 0; iBox < s->nLocalBoxes; ++iBox)
      for (int iOff = ORC_MAXATOMS*iBox, ii = 0; ii < s->nAtoms[iBox]; ++ii, ++iOff)
      {
         double invMass = 1.0 / s->mass;
         for (int a = 0; a < 3; ++a) s->r[a][iOff] += dt * s->p[a][iOff] * invMass;
      }
   s->nlUpdateRequired = -1;
}

/* redistributeAtomsCpuNL, timestep.c:356-405 */
static void redistribute(OrcSim* s)
{
   int rebuild = nlUpdateRequired(s);
   emptyHaloCells(s);
   if (rebuild) { s->fifoPut = 0; updateLinkCells(s); }
   s->fifoGet = 0;
   atomHaloExchange(s);
   if (s->nlLinkCellsRequired > 0)
   {
      s->nlForceRebuild = 1;
      updateLocalIds(s);
      emptyHaloCells(s);
      updateLinkCells(s);
      s->fifoPut = 0;
      s->fifoGet = 0;
      atomHaloExchange(s);
      s->nlLinkCellsRequired = 0;
   }
   if (rebuild) updateLocalIds(s);
}

void orc_timestep(OrcSim* s, int nSteps)
{
   const double dt = s->dt;
   for (int i = 0; i < nSteps; ++i)
   {
      advanceVelocity(s, 0.5*dt);
      advancePosition(s, dt);
      redistribute(s);
      buildNeighborList(s);
      computeForce(s);
      advanceVelocity(s, 0.5*dt);
   }
   kineticEnergy(s);
}

/* ---------------------------------------------------------------- set-up (CoMD.c:200-327, initAtoms.c) */
static void allocAtoms(OrcSim* s)
{
   s->slab = ORC_MAXATOMS * s->nTotalBoxes + 1;
   const int n = s->slab;
   s->nAtoms = (int*)calloc(s->nTotalBoxes, sizeof(int));
   s->gid = (int*)calloc(n, sizeof(int)); s->species = (int*)calloc(n, sizeof(int));
   s->lid = (int*)calloc(ORC_MAXATOMS * s->nLocalBoxes, sizeof(int));
   for (int a = 0; a < 3; ++a)
   {
      s->r[a] = (double*)calloc(n, sizeof(double)); s->p[a] = (double*)calloc(n, sizeof(double));
      s->f[a] = (double*)calloc(n, sizeof(double));
      s->r[a][n-1] = FLT_MAX;
   }
   s->U = (double*)calloc(n, sizeof(double));
   s->rhobar = (double*)calloc(n, sizeof(double)); s->dfEmbed = (double*)calloc(n, sizeof(double));
   s->nMaxLocal = ORC_MAXATOMS * s->nLocalBoxes;
   s->list = (int*)malloc((size_t)s->nMaxLocal * ORC_MAXNBR * sizeof(int));
   s->nNbr = (int*)calloc(s->nMaxLocal, sizeof(int));
   for (int a = 0; a < 3; ++a) s->lastR[a] = (double*)calloc(s->nMaxLocal, sizeof(double));
   s->nlUpdateRequired = 1; s->nlForceRebuild = 1; s->nlLinkCellsRequired = 0; s->nStepsSinceBuild = 0; s->nBuilds = 0;
}

static void createFcc(OrcSim* s, int nx, int ny, int nz, double lat)
{
   static const double basis[4][3] = { {0.25,0.25,0.25}, {0.25,0.75,0.75}, {0.75,0.25,0.75}, {0.75,0.75,0.25} };
   int b[3], e[3];
   for (int a = 0; a < 3; ++a) { b[a] = floor(s->lMin[a] / lat); e[a] = ceil(s->lMax[a] / lat); }
   for (int ix = b[0]; ix < e[0]; ++ix) for (int iy = b[1]; iy < e[1]; ++iy) for (int iz = b[2]; iz < e[2]; ++iz)
      for (int ib = 0; ib < 4; ++ib)
      {
         double rx = (ix + basis[ib][0]) * lat, ry = (iy + basis[ib][1]) * lat, rz = (iz + basis[ib][2]) * lat;
         if (rx < s->lMin[0] || rx >= s->lMax[0]) continue;
         if (ry < s->lMin[1] || ry >= s->lMax[1]) continue;
         if (rz < s->lMin[2] || rz >= s->lMax[2]) continue;
         int id = ib + 4 * (iz + nz * (iy + ny * ix));
         putAtom(s, id, 0, rx, ry, rz, 0.0, 0.0, 0.0);
      }
   s->nGlobal = s->nLocal;
}

static void setTemperature(OrcSim* s, double T)
{
   const double mass = s->mass;
   for (int iBox = 0; iBox < s->nLocalBoxes; ++iBox)
      for (int iOff = ORC_MAXATOMS*iBox, ii = 0; ii < s->nAtoms[iBox]; ++ii, ++iOff)
      {
         double sigma = sqrt(orcKbEv * T / mass);
         uint64_t seed = orc_mk_seed(s->gid[iOff], 123);
         s->p[0][iOff] = mass * sigma * orc_gasdev(&seed);
         s->p[1][iOff] = mass * sigma * orc_gasdev(&seed);
         s->p[2][iOff] = mass * sigma * orc_gasdev(&seed);
      }
   if (T == 0.0) return;
   /* setVcm(0): computeVcm sums p and mass in traversal order */
   double v[4] = { 0., 0., 0., 0. };
   for (int iBox = 0; iBox < s->nLocalBoxes; ++iBox)
      for (int iOff = ORC_MAXATOMS*iBox, ii = 0; ii < s->nAtoms[iBox]; ++ii, ++iOff)
      {
         v[0] += s->p[0][iOff]; v[1] += s->p[1][iOff]; v[2] += s->p[2][iOff]; v[3] += mass;
      }
   double sh[3] = { 0. - v[0]/v[3], 0. - v[1]/v[3], 0. - v[2]/v[3] };
   for (int iBox = 0; iBox < s->nLocalBoxes; ++iBox)
      for (int iOff = ORC_MAXATOMS*iBox, ii = 0; ii < s->nAtoms[iBox]; ++ii, ++iOff)
         for (int a = 0; a < 3; ++a) s->p[a][iOff] += mass * sh[a];
   kineticEnergy(s);
   double temp = (s->eKin / s->nGlobal) / orcKbEv / 1.5;
   double scale = sqrt(T / temp);
   for (int iBox = 0; iBox < s->nLocalBoxes; ++iBox)
      for (int iOff = ORC_MAXATOMS*iBox, ii = 0; ii < s->nAtoms[iBox]; ++ii, ++iOff)
         for (int a = 0; a < 3; ++a) s->p[a][iOff] *= scale;
   kineticEnergy(s);
}

static void randomDisplacements(OrcSim* s, double delta)
{
   for (int iBox = 0; iBox < s->nLocalBoxes; ++iBox)
      for (int iOff = ORC_MAXATOMS*iBox, ii = 0; ii < s->nAtoms[iBox]; ++ii, ++iOff)
      {
         uint64_t seed = orc_mk_seed(s->gid[iOff], 457);
         s->r[0][iOff] += (2.0 * orc_lcg61(&seed) - 1.0) * delta;
         s->r[1][iOff] += (2.0 * orc_lcg61(&seed) - 1.0) * delta;
         s->r[2][iOff] += (2.0 * orc_lcg61(&seed) - 1.0) * delta;
      }
}

/* potKind: 0 = LJ (ljCutMult*sigma cutoff; BASELINE uses 2.5, the fork ships 5.0, ljForce.c:114),
 *          1 = EAM funcfl, 2 = EAM setfl (potPath = file).
 * cellCutoff <= 0 selects the reference's cutoff+skin (CoMD.c:257-268).
 * Mirrors initSimulation (CoMD.c:200-327) incl. the first redistribute, list build and force call. */
OrcSim* orc_create(int potKind, const char* potPath, double ljCutMult,
                   int nx, int ny, int nz, double latIn, double temperature, double delta,
                   double relSkin, double dt, int doInitialForce)
{
   OrcSim* s = (OrcSim*)calloc(1, sizeof(OrcSim));
   s->isEam = potKind != 0;
   if (potKind == 0)
   {
      s->sigma = 2.315; s->epsilon = 0.167; s->mass = 63.55 * orcAmuToInternalMass;
      s->lat = 3.615; s->cutoff = ljCutMult * s->sigma;
   }
   else
   {
      int rc = potKind == 1 ? readFuncfl(s, potPath) : readSetfl(s, potPath);
      if (rc != 0) { free(s); return NULL; }
   }
   const double lat = latIn < 0.0 ? s->lat : latIn;
   s->gExt[0] = nx * lat; s->gExt[1] = ny * lat; s->gExt[2] = nz * lat;
   for (int a = 0; a < 3; ++a) { s->lMin[a] = 0.0; s->lMax[a] = 0.0 + 1 * (s->gExt[a] / 1); }
   s->skin = s->cutoff * relSkin;
   s->skinHalf2 = (s->skin / 2.0) * (s->skin / 2.0);
   s->dt = dt;
   cellsInit(s, s->cutoff + s->skin);
   allocAtoms(s);
   createFcc(s, nx, ny, nz, lat);
   setTemperature(s, temperature);
   randomDisplacements(s, delta);
   exchangeInit(s);
   if (doInitialForce)
   {
      redistribute(s);
      buildNeighborList(s);
      computeForce(s);
      kineticEnergy(s);
   }
   return s;
}

void orc_destroy(OrcSim* s)
{
   if (!s) return;
   free(s->nAtoms); free(s->gid); free(s->species); free(s->lid);
   for (int a = 0; a < 3; ++a) { free(s->r[a]); free(s->p[a]); free(s->f[a]); free(s->lastR[a]); }
   free(s->U); free(s->rhobar); free(s->dfEmbed); free(s->list); free(s->nNbr); free(s->fifo);
   for (int f = 0; f < 6; ++f) { free(s->atomCells[f]); free(s->fSend[f]); free(s->fRecv[f]); }
   if (s->isEam) { free(s->phi.v - 1); free(s->rho.v - 1); free(s->fem.v - 1); }
   free(s);
}

/* Replace the state by caller-supplied local atoms (n x 3 row-major r and p), re-bin them in
 * the given order, rebuild ghosts + list and recompute forces: lets the tests hand the oracle
 * exactly the positions the device holds. */
void orc_set_atoms(OrcSim* s, int n, const int* gid, const double* r, const double* p)
{
   for (int i = 0; i < s->nTotalBoxes; ++i) s->nAtoms[i] = 0;
   s->nLocal = 0;
   for (int i = 0; i < n; ++i)
      putAtom(s, gid[i], 0, r[3*i], r[3*i+1], r[3*i+2], p ? p[3*i] : 0.0, p ? p[3*i+1] : 0.0, p ? p[3*i+2] : 0.0);
   s->nGlobal = s->nLocal;
   s->nlForceRebuild = 1; s->nlUpdateRequired = 1; s->nlLinkCellsRequired = 0;
   updateLocalIds(s);
   redistribute(s);
   buildNeighborList(s);
   computeForce(s);
   kineticEnergy(s);
}

int orc_n_local(const OrcSim* s) { int n = 0; for (int i = 0; i < s->nLocalBoxes; ++i) n += s->nAtoms[i]; return n; }
int orc_n_builds(const OrcSim* s) { return s->nBuilds; }
void orc_energies(const OrcSim* s, double* ePot, double* eKin) { *ePot = s->ePot; *eKin = s->eKin; }
void orc_potential(const OrcSim* s, double* cutoff, double* mass, double* lat) { *cutoff = s->cutoff; *mass = s->mass; *lat = s->lat; }
void orc_geometry(const OrcSim* s, int* grid, double* dbl)
{
   for (int i = 0; i < 3; ++i)
   {
      grid[i] = s->grid[i];
      dbl[i] = s->lMin[i]; dbl[3+i] = s->lMax[i]; dbl[6+i] = s->boxSize[i]; dbl[9+i] = s->invBox[i];
   }
   grid[3] = s->nLocalBoxes; grid[4] = s->nTotalBoxes;
}

int orc_get_atoms(const OrcSim* s, int* gid, int* box, double* r, double* p, double* f, double* U)
{
   int n = 0;
   for (int iBox = 0; iBox < s->nLocalBoxes; ++iBox)
      for (int iOff = ORC_MAXATOMS*iBox, ii = 0; ii < s->nAtoms[iBox]; ++ii, ++iOff, ++n)
      {
         if (gid) gid[n] = s->gid[iOff];
         if (box) box[n] = iBox;
         for (int a = 0; a < 3; ++a)
         {
            if (r) r[3*n+a] = s->r[a][iOff];
            if (p) p[3*n+a] = s->p[a][iOff];
            if (f) f[3*n+a] = s->f[a][iOff];
         }
         if (U) U[n] = s->U[iOff];
      }
   return n;
}

int orc_get_all_atoms(const OrcSim* s, int* gid, int* box, double* r)
{
   int n = 0;
   for (int iBox = 0; iBox < s->nTotalBoxes; ++iBox)
      for (int iOff = ORC_MAXATOMS*iBox, ii = 0; ii < s->nAtoms[iBox]; ++ii, ++iOff, ++n)
      {
         if (gid) gid[n] = s->gid[iOff];
         if (box) box[n] = iBox;
         if (r) for (int a = 0; a < 3; ++a) r[3*n+a] = s->r[a][iOff];
      }
   return n;
}

/* A second, independent cell geometry (cells sized on an explicit cutoff, as the device
 * path does) for the bit-exact binning tests: same rule, different grid. */
OrcSim* orc_cells_only(const double gExt[3], double cellCutoff)
{
   OrcSim* s = (OrcSim*)calloc(1, sizeof(OrcSim));
   for (int a = 0; a < 3; ++a) { s->gExt[a] = gExt[a]; s->lMin[a] = 0.0; s->lMax[a] = 0.0 + 1 * (gExt[a] / 1); }
   cellsInit(s, cellCutoff);
   return s;
}
void orc_bin(const OrcSim* s, int n, const double* r, int* box, int* tuple)
{
   for (int i = 0; i < n; ++i)
   {
      int t[3];
      orc_tuple_from_coord(s, r + 3*i, t);
      box[i] = orc_box_from_tuple(s, t[0], t[1], t[2]);
      if (tuple) { tuple[3*i] = t[0]; tuple[3*i+1] = t[1]; tuple[3*i+2] = t[2]; }
   }
}
