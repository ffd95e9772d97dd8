/* geometry.c -- rank layout and link-cell geometry on the host.
 *
 * Same arithmetic, operation for operation, as the reference's initDecomposition
 * (decomposition.c:18-50), processorNum (:57-66), initLinkCells (linkCells.c:122-182),
 * getBoxFromTuple (:299-346) and getBoxFromCoord (:448-480): boxSize and invBoxSize computed
 * here are handed to the device verbatim (ComdGeometry), so the owner cell of an atom is the same
 * bit pattern on the CPU reference, on this host and on the GPU.
 */
#include "comd_host.h"
#include <math.h>
#include <stdlib.h>

Domain* initDecomposition(int xproc, int yproc, int zproc, real3 globalExtent)
{
   Domain* dom = calloc(1, sizeof *dom);
   const int grid[3] = { xproc, yproc, zproc };
   int rank = getMyRank();
   /* rank -> (i,j,k): x fastest */
   dom->procCoord[0] = rank % xproc;
   rank /= xproc;
   dom->procCoord[1] = rank % yproc;
   dom->procCoord[2] = rank / yproc;
   for (int a = 0; a < 3; ++a)
   {
      dom->procGrid[a] = grid[a];
      dom->globalMin[a] = 0;
      dom->globalMax[a] = globalExtent[a];
      dom->globalExtent[a] = dom->globalMax[a] - dom->globalMin[a];
      dom->localExtent[a] = dom->globalExtent[a] / dom->procGrid[a];
      dom->localMin[a] = dom->globalMin[a] + dom->procCoord[a] * dom->localExtent[a];
      dom->localMax[a] = dom->globalMin[a] + (dom->procCoord[a] + 1) * dom->localExtent[a];
   }
   return dom;
}

int processorNum(Domain* dom, int dix, int diy, int diz)
{
   const int d[3] = { dix, diy, diz };
   int c[3];
   for (int a = 0; a < 3; ++a) c[a] = (dom->procCoord[a] + d[a] + dom->procGrid[a]) % dom->procGrid[a];
   return c[0] + dom->procGrid[0] * (c[1] + dom->procGrid[1] * c[2]);
}

LinkCell* initLinkCells(const Domain* dom, real_t cutoff, int useHilbert)
{
   LinkCell* lc = calloc(1, sizeof *lc);
   lc->useHilbert = useHilbert;   /* -H: acted on by the device layer (option nl_hilbert), for any grid size -- the reference needs powers of two (linkCells.c:151-155) */
   for (int a = 0; a < 3; ++a)
   {
      lc->localMin[a] = dom->localMin[a];
      lc->localMax[a] = dom->localMax[a];
      lc->gridSize[a] = dom->localExtent[a] / cutoff;
      lc->boxSize[a] = dom->localExtent[a] / ((real_t)lc->gridSize[a]);
      lc->invBoxSize[a] = 1.0 / lc->boxSize[a];
   }
   const int* g = lc->gridSize;
   lc->nLocalBoxes = g[0] * g[1] * g[2];
   lc->nHaloBoxes = 2 * ((g[0] + 2) * (g[1] + g[2] + 2) + g[1] * g[2]);
   lc->nTotalBoxes = lc->nLocalBoxes + lc->nHaloBoxes;
   lc->nAtoms = calloc(lc->nTotalBoxes, sizeof(int));
   if (g[0] < 2 || g[1] < 2 || g[2] < 2)
   {
      if (printRank()) fprintf(stderr, "initLinkCells: fewer than 2 link cells on an axis (%d,%d,%d)\n", g[0], g[1], g[2]);
      exit(3);
   }
   return lc;
}

void destroyLinkCells(LinkCell** boxes)
{
   if (!boxes || !*boxes) return;
   free((*boxes)->nAtoms);
   free(*boxes);
   *boxes = NULL;
}

/* Local cells first (x fastest), then the halo planes: -x, +x (ny*nz each), -y, +y (nz rows of
 * nx+2), -z, +z (whole (nx+2)(ny+2) planes) -- the numbering of linkCells.c:299-346. */
int getBoxFromTuple(LinkCell* lc, int ix, int iy, int iz)
{
   const int nx = lc->gridSize[0], ny = lc->gridSize[1], nz = lc->gridSize[2];
   const int afterX = lc->nLocalBoxes + 2 * nz * ny;
   const int afterY = afterX + 2 * nz * (nx + 2);
   if (iz == nz) return afterY + (nx + 2) * (ny + 2) + (nx + 2) * (iy + 1) + (ix + 1);
   if (iz == -1) return afterY + (nx + 2) * (iy + 1) + (ix + 1);
   if (iy == ny) return afterX + nz * (nx + 2) + (nx + 2) * iz + (ix + 1);
   if (iy == -1) return afterX + iz * (nx + 2) + (ix + 1);
   if (ix == nx) return lc->nLocalBoxes + ny * nz + iz * ny + iy;
   if (ix == -1) return lc->nLocalBoxes + iz * ny + iy;
   return ix + nx * iy + nx * ny * iz;
}

int getBoxFromCoord(LinkCell* lc, const real_t rr[3])
{
   int t[3];
   for (int a = 0; a < 3; ++a)
   {
      t[a] = (int)floor((rr[a] - lc->localMin[a]) * lc->invBoxSize[a]);
      if (rr[a] < lc->localMax[a]) { if (t[a] == lc->gridSize[a]) t[a] = lc->gridSize[a] - 1; }
      else t[a] = lc->gridSize[a];
   }
   return getBoxFromTuple(lc, t[0], t[1], t[2]);
}

int maxOccupancy(LinkCell* lc)
{
   int localMax = 0, globalMax = 0;
   for (int b = 0; b < lc->nLocalBoxes; ++b) if (lc->nAtoms[b] > localMax) localMax = lc->nAtoms[b];
   profileStart(commReduceTimer);
   maxIntParallel(&localMax, &globalMax, 1);
   profileStop(commReduceTimer);
   return globalMax;
}
