// halo_runs.cuh -- the canonical "cell run" layout of a one-plane halo message, shared by the F'(rho) halo
// (force_eam.cu) and the ghost-position refresh of the neighbour-list path (neighbor.cu).
//
// A face's message is the concatenation of runs of consecutive cells, in an order both sides derive from the
// grid alone; inside a run atoms are in memory order (cell, then gid), which is identical on sender and receiver:
//   x faces: ny*nz runs of one cell      (send ix = 0 | nx-1, recv ix = -1 | nx,  iy,iz local)
//   y faces: nz runs of a whole x-row    (send iy = 0 | ny-1, recv iy = -1 | ny,  ix incl. halo, iz local)
//   z faces: one run = a whole xy-plane  (send iz = 0 | nz-1, recv iz = -1 | nz,  ix,iy incl. halo)
// (mkForceSendCellList / mkForceRecvCellList, haloExchange.c:1712-1801.)
#pragma once
#include "comd_dev.cuh"

struct FpFace { int firstCell, strideA, countA, strideB, countB, cellsPerRun, nRuns; };

__host__ __device__ inline FpFace fpFace(const DevGeom& g, int face, int isRecv)
{
   const int ax = face / 2, plus = face % 2;
   const int PX = g.P[0], PY = g.P[1];
   int coord;                          // unpadded coordinate of the plane along ax
   if (!isRecv) coord = plus ? g.n[ax] - 1 : 0; else coord = plus ? g.n[ax] : -1;
   FpFace f;
   if (ax == 0) { f.firstCell = cellIndex(g, coord, 0, 0); f.strideA = PX; f.countA = g.n[1]; f.strideB = PX * PY; f.countB = g.n[2]; f.cellsPerRun = 1; }
   else if (ax == 1) { f.firstCell = cellIndex(g, -1, coord, 0); f.strideA = 0; f.countA = 1; f.strideB = PX * PY; f.countB = g.n[2]; f.cellsPerRun = PX; }
   else { f.firstCell = cellIndex(g, -1, -1, coord); f.strideA = 0; f.countA = 1; f.strideB = 0; f.countB = 1; f.cellsPerRun = PX * PY; }
   f.nRuns = f.countA * f.countB;
   return f;
}

// message element m -> run by binary search in runOff, then a straight copy
__device__ __forceinline__ int fpLocate(const int* __restrict__ runOff, int nRuns, int m)
{
   int lo = 0, hi = nRuns;            // runOff[lo] <= m < runOff[hi]
   while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (runOff[mid] <= m) lo = mid; else hi = mid; }
   return lo;
}

static inline int haloRunStride(const DevGeom& g)
{
   int maxRuns = g.n[1] * g.n[2];
   if (g.n[2] > maxRuns) maxRuns = g.n[2];
   return maxRuns + 2;
}
