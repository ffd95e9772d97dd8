/* atoms.c -- deterministic initial state on the host: FCC lattice, gid-seeded Maxwell-Boltzmann
 * momenta rescaled to exactly T, gid-seeded random displacements.
 *
 * Restates initAtoms.c:26-216 and random.c:21-75 of the reference.  The reference walks its
 * padded per-box slabs (box 0..nLocalBoxes-1, atoms in the order putAtomInBox appended them);
 * here the host atoms are one dense SoA list kept in that very order (a stable counting sort by
 * owner box after the lattice is generated), so every sum below adds the same numbers in the
 * same order and the initial r and p are bit-identical to the reference's on the same libm.
 */
#include "comd_host.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ---- random.c ---- */
#define LCG_MOD UINT64_C(2305843009213693951)       /* 2^61 - 1 */
double lcg61(uint64_t* seed)
{
   static const double toUnit = 1.0 / LCG_MOD;
   *seed *= UINT64_C(437799614237992725);
   *seed %= LCG_MOD;
   return *seed * toUnit;
}

real_t gasdev(uint64_t* seed)
{
   real_t v1, v2, rsq;
   for (;;)
   {  /* polar Box-Muller, second variate discarded */
      v1 = 2.0 * lcg61(seed) - 1.0;
      v2 = 2.0 * lcg61(seed) - 1.0;
      rsq = v1 * v1 + v2 * v2;
      if (rsq < 1.0 && rsq != 0.0) break;
   }
   return v2 * sqrt(-2.0 * log(rsq) / rsq);
}

uint64_t mkSeed(uint32_t id, uint32_t callSite)
{
   const uint32_t hi = id * UINT32_C(2654435761);
   const uint32_t lo = (id + callSite) * UINT32_C(2654435761);
   uint64_t s = (UINT64_C(0x100000000) * hi) + lo;
   for (int k = 0; k < 10; ++k) lcg61(&s);
   return s;
}

/* ---- storage ---- */
static void reserveAtoms(Atoms* a, int n)
{
   if (n <= a->capacity) return;
   int cap = a->capacity ? a->capacity : 1024;
   while (cap < n) cap += cap / 2 + 1024;
   a->gid = realloc(a->gid, cap * sizeof(int));
   a->iSpecies = realloc(a->iSpecies, cap * sizeof(int));
   a->U = realloc(a->U, cap * sizeof(real_t));
   for (int k = 0; k < 3; ++k)
   {
      a->r[k] = realloc(a->r[k], cap * sizeof(real_t));
      a->p[k] = realloc(a->p[k], cap * sizeof(real_t));
      a->f[k] = realloc(a->f[k], cap * sizeof(real_t));
   }
   a->capacity = cap;
}

Atoms* initAtoms(LinkCell* boxes, real_t skinDistance)
{
   (void)boxes; (void)skinDistance;          /* no MAXATOMS slabs, no host neighbour list */
   Atoms* a = calloc(1, sizeof *a);
   return a;
}

void destroyAtoms(Atoms* a)
{
   if (!a) return;
   free(a->gid); free(a->iSpecies); free(a->U);
   for (int k = 0; k < 3; ++k) { free(a->r[k]); free(a->p[k]); free(a->f[k]); }
   free(a);
}

void comdHostReserveAtoms(Atoms* a, int n) { reserveAtoms(a, n); }

/* stable counting sort of the host atoms by owner box; fills boxes->nAtoms */
static void orderByBox(SimFlat* s)
{
   Atoms* a = s->atoms;
   LinkCell* lc = s->boxes;
   const int n = a->nLocal;
   int* box = malloc((n > 0 ? n : 1) * sizeof(int));
   int* start = calloc(lc->nTotalBoxes + 1, sizeof(int));
   for (int i = 0; i < n; ++i)
   {
      const real_t rr[3] = { a->r[0][i], a->r[1][i], a->r[2][i] };
      box[i] = getBoxFromCoord(lc, rr);
      start[box[i] + 1]++;
   }
   for (int b = 0; b < lc->nTotalBoxes; ++b) { lc->nAtoms[b] = start[b + 1]; start[b + 1] += start[b]; }
   int* dst = malloc((n > 0 ? n : 1) * sizeof(int));
   for (int i = 0; i < n; ++i) dst[i] = start[box[i]]++;
   real_t* tmp = malloc((n > 0 ? n : 1) * sizeof(real_t));
   int* tmpI = (int*)tmp;
   #define PERMUTE(arr, T, buf) do { for (int i = 0; i < n; ++i) (buf)[dst[i]] = (arr)[i]; memcpy((arr), (buf), n * sizeof(T)); } while (0)
   PERMUTE(a->gid, int, tmpI);
   PERMUTE(a->iSpecies, int, tmpI);
   for (int k = 0; k < 3; ++k) { PERMUTE(a->r[k], real_t, tmp); PERMUTE(a->p[k], real_t, tmp); }
   #undef PERMUTE
   free(tmp); free(dst); free(start); free(box);
}

void createFccLattice(int nx, int ny, int nz, real_t lat, SimFlat* s)
{
   (void)nx;
   static const real_t basis[4][3] = { { 0.25, 0.25, 0.25 }, { 0.25, 0.75, 0.75 }, { 0.75, 0.25, 0.75 }, { 0.75, 0.75, 0.25 } };
   const real_t* lo = s->domain->localMin;
   const real_t* hi = s->domain->localMax;
   int begin[3], end[3];
   for (int a = 0; a < 3; ++a) { begin[a] = floor(lo[a] / lat); end[a] = ceil(hi[a] / lat); }
   Atoms* at = s->atoms;
   at->nLocal = 0;
   for (int ix = begin[0]; ix < end[0]; ++ix)
      for (int iy = begin[1]; iy < end[1]; ++iy)
         for (int iz = begin[2]; iz < end[2]; ++iz)
            for (int ib = 0; ib < 4; ++ib)
            {
               const real_t x = (ix + basis[ib][0]) * lat, y = (iy + basis[ib][1]) * lat, z = (iz + basis[ib][2]) * lat;
               if (x < lo[0] || x >= hi[0] || y < lo[1] || y >= hi[1] || z < lo[2] || z >= hi[2]) continue;
               reserveAtoms(at, at->nLocal + 1);
               const int i = at->nLocal++;
               at->gid[i] = ib + 4 * (iz + nz * (iy + ny * ix));
               at->iSpecies[i] = 0;
               at->r[0][i] = x; at->r[1][i] = y; at->r[2][i] = z;
               for (int k = 0; k < 3; ++k) { at->p[k][i] = 0.0; at->f[k][i] = 0.0; }
               at->U[i] = 0.0;
            }
   orderByBox(s);
   profileStart(commReduceTimer);
   addIntParallel(&at->nLocal, &at->nGlobal, 1);
   profileStop(commReduceTimer);
   if (at->nGlobal != 4 * nx * ny * nz)
   {
      if (printRank()) fprintf(stderr, "createFccLattice: %d atoms instead of %d\n", at->nGlobal, 4 * nx * ny * nz);
      exit(4);
   }
}

static void hostKineticEnergy(SimFlat* s)
{
   const Atoms* a = s->atoms;
   real_t loc[2] = { s->ePotential, 0 }, sum[2];
   for (int i = 0; i < a->nLocal; ++i)
   {
      const real_t invMass = 0.5 / s->species[a->iSpecies[i]].mass;
      loc[1] += (a->p[0][i] * a->p[0][i] + a->p[1][i] * a->p[1][i] + a->p[2][i] * a->p[2][i]) * invMass;
   }
   profileStart(commReduceTimer);
   addRealParallel(loc, sum, 2);
   profileStop(commReduceTimer);
   s->ePotential = sum[0];
   s->eKinetic = sum[1];
}

void setVcm(SimFlat* s, real_t newVcm[3])
{
   Atoms* a = s->atoms;
   real_t loc[4] = { 0., 0., 0., 0. }, sum[4];
   for (int i = 0; i < a->nLocal; ++i)
   {
      loc[0] += a->p[0][i]; loc[1] += a->p[1][i]; loc[2] += a->p[2][i];
      loc[3] += s->species[a->iSpecies[i]].mass;
   }
   profileStart(commReduceTimer);
   addRealParallel(loc, sum, 4);
   profileStop(commReduceTimer);
   real_t shift[3];
   for (int k = 0; k < 3; ++k) shift[k] = newVcm[k] - sum[k] / sum[3];
   for (int i = 0; i < a->nLocal; ++i)
   {
      const real_t mass = s->species[a->iSpecies[i]].mass;
      for (int k = 0; k < 3; ++k) a->p[k][i] += mass * shift[k];
   }
}

void setTemperature(SimFlat* s, real_t temperature)
{
   Atoms* a = s->atoms;
   for (int i = 0; i < a->nLocal; ++i)
   {
      const real_t mass = s->species[a->iSpecies[i]].mass;
      const real_t sigma = sqrt(kB_eV * temperature / mass);
      uint64_t seed = mkSeed(a->gid[i], 123);
      a->p[0][i] = mass * sigma * gasdev(&seed);
      a->p[1][i] = mass * sigma * gasdev(&seed);
      a->p[2][i] = mass * sigma * gasdev(&seed);
   }
   if (temperature == 0.0) return;
   real_t zero[3] = { 0., 0., 0. };
   setVcm(s, zero);
   hostKineticEnergy(s);
   const real_t temp = (s->eKinetic / a->nGlobal) / kB_eV / 1.5;
   const real_t scale = sqrt(temperature / temp);
   for (int i = 0; i < a->nLocal; ++i)
      for (int k = 0; k < 3; ++k) a->p[k][i] *= scale;
   hostKineticEnergy(s);
}

void randomDisplacements(SimFlat* s, real_t delta)
{
   Atoms* a = s->atoms;
   for (int i = 0; i < a->nLocal; ++i)
   {
      uint64_t seed = mkSeed(a->gid[i], 457);
      a->r[0][i] += (2.0 * lcg61(&seed) - 1.0) * delta;
      a->r[1][i] += (2.0 * lcg61(&seed) - 1.0) * delta;
      a->r[2][i] += (2.0 * lcg61(&seed) - 1.0) * delta;
   }
}
