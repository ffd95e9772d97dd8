// nl_blocks.cuh -- work blocks of the list-path force kernels: bx*by*bz link cells (x fastest).  The force kernels
// hand them to the SMs (force_nl.cuh, force_n3.cuh); the list build needs the same shape because pairs inside a
// block are stored once, on the row of the lower index (neighbor.cu).
#pragma once
#include "comd_dev.cuh"

struct NlBlocks
{
   int bx, by, bz;            // cells per work block
   int nbx, nby, nbz;         // work blocks per axis
   int total;                 // blocks this launch walks
   int phase;                 // 0 all blocks, 1 interior (no atom of the block has a ghost neighbour), 2 the boundary shell
   int chunk, nChunks;        // blocks per work chunk, chunks in this launch
   int oneShot;               // 1: a CTA works on one chunk and retires (grid = nChunks), so kernels of a higher-priority stream find SM slots
};
#define NLB_ALL 0
#define NLB_INTERIOR 1
#define NLB_BOUNDARY 2

// Blocks touching the outermost cell layer of the local domain form the boundary set; their atoms are the only
// ones with ghost neighbours, so everything else can be computed while the halo is still travelling.
__host__ __device__ inline int nlInteriorCount(const NlBlocks& B)
{
   return (B.nbx > 2 && B.nby > 2 && B.nbz > 2) ? (B.nbx - 2) * (B.nby - 2) * (B.nbz - 2) : 0;
}
// q-th block of the launch's set -> block coordinates
__device__ __forceinline__ void nlBlockCoord(const NlBlocks& B, int q, int& ix, int& iy, int& iz)
{
   if (B.phase == NLB_INTERIOR)
   {
      const int mx = B.nbx - 2, my = B.nby - 2;
      ix = 1 + q % mx; iy = 1 + (q / mx) % my; iz = 1 + q / (mx * my);
      return;
   }
   if (B.phase == NLB_ALL || nlInteriorCount(B) == 0)
   {
      ix = q % B.nbx; iy = (q / B.nbx) % B.nby; iz = q / (B.nbx * B.nby);
      return;
   }
   const int nA = 2 * B.nbx * B.nby, nB = (B.nbz - 2) * 2 * B.nbx;
   if (q < nA)
   {  // the two z faces, whole planes
      iz = q / (B.nbx * B.nby) ? B.nbz - 1 : 0;
      const int r = q % (B.nbx * B.nby);
      ix = r % B.nbx; iy = r / B.nbx;
   }
   else if (q < nA + nB)
   {  // y faces of the middle planes, whole rows
      q -= nA;
      iz = 1 + q / (2 * B.nbx);
      const int r = q % (2 * B.nbx);
      iy = r / B.nbx ? B.nby - 1 : 0; ix = r % B.nbx;
   }
   else
   {  // x faces of what is left
      q -= nA + nB;
      const int per = 2 * (B.nby - 2);
      iz = 1 + q / per;
      const int r = q % per;
      iy = 1 + r / 2; ix = (r & 1) ? B.nbx - 1 : 0;
   }
}

// cells per work block: the largest shape that still gives every SM several blocks.  4x4x4 with the once-per-pair
// kernels (76 % of the pairs of a copper crystal are inside such a block), 4x4x2 with the full-list kernels (the
// gathered positions of block + halo, ~90 KB, stay in L1).
static NlBlocks nlBlocksFor(const ComdDev* d)
{
   static const int shapes[][3] = { {4, 4, 4}, {4, 4, 2}, {4, 2, 2}, {2, 2, 2}, {2, 2, 1}, {2, 1, 1}, {1, 1, 1} };
   const DevGeom& g = d->g;
   NlBlocks B = {};
   for (int k = comdNlN3(d) ? 0 : 1; k < 7; ++k)
   {
      B.bx = shapes[k][0]; B.by = shapes[k][1]; B.bz = shapes[k][2];
      if (d->optNlBlock[0] > 0) { B.bx = d->optNlBlock[0]; B.by = d->optNlBlock[1]; B.bz = d->optNlBlock[2]; }
      B.nbx = (g.n[0] + B.bx - 1) / B.bx; B.nby = (g.n[1] + B.by - 1) / B.by; B.nbz = (g.n[2] + B.bz - 1) / B.bz;
      B.total = B.nbx * B.nby * B.nbz;
      if (d->optNlBlock[0] > 0 || B.total >= 148 * 8) break;
   }
   return B;
}
