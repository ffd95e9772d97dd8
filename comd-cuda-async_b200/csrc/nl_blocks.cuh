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
   const int* order;          // traversal order of the launch's set: order[q] = linear block id (x fastest), or NULL for the index order
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
// q-th block of the launch's set -> block coordinates.  On the device every set but "all blocks in index order" comes with
// an order table (nlBlockOrder below: interior / boundary split, Hilbert traversal, largest-first), so the kernels carry one
// table read instead of the case analysis; the host enumerates the sets with nlBlockCoordOfSet.
__device__ __forceinline__ void nlBlockCoord(const NlBlocks& B, int q, int& ix, int& iy, int& iz)
{
   const int id = B.order ? __ldg(B.order + q) : q;
   ix = id % B.nbx; iy = (id / B.nbx) % B.nby; iz = id / (B.nbx * B.nby);
}
static inline void nlBlockCoordOfSet(const NlBlocks& B, int q, int& ix, int& iy, int& iz)
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
// kernels (76 % of the pairs of a copper crystal are inside such a block) where that leaves at least 8 blocks per CTA slot
// (2 x 148 slots): C5 on one GPU 12.45 against 12.85 ms/step with 4x4x2, but on eight GPUs (47^3 cells each, 5.8 blocks per
// slot) 1.84 against 1.74 -- the tail of the launch outweighs the pairs saved (profiles/r2_v3_block_shape_n8.txt,
// r2_v3_block_lanes_ab.txt).  4x4x2 with the full-list kernels (the gathered positions of block + halo, ~90 KB, stay in L1).
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
      if (d->optNlBlock[0] > 0 || B.total >= (k == 0 ? 148 * 16 : 148 * 8)) break;
   }
   return B;
}

// ---- Hilbert traversal of the work blocks (CoMD's -H / --hilbert, linkCells.c:28-63,151-174) -----------------------
// The reference renumbers its link cells along a 3-D Hilbert curve so that consecutive thread blocks work on cells
// that are neighbours in space (only for power-of-two grids).  Here the unit of scheduling is the work block, so the
// curve orders the blocks: chunks of consecutive blocks are handed to the SMs in curve order, and the blocks in
// flight at any time share their halo cells in L2.  Any grid size: the key is computed in the enclosing power-of-two
// cube (Skilling's transpose algorithm, "Programming the Hilbert curve", AIP Conf. Proc. 707, 2004) and the blocks of
// the launch's set are sorted by it.  Results do not depend on the order (per-atom sums are formed inside a block).
#include <algorithm>
#include <vector>
static inline unsigned long long nlHilbertKey(unsigned x, unsigned y, unsigned z, int bits)
{
   unsigned X[3] = { x, y, z };
   const unsigned M = 1u << (bits - 1);
   for (unsigned Q = M; Q > 1; Q >>= 1)                 // inverse undo of the excess work
   {
      const unsigned P = Q - 1;
      for (int i = 0; i < 3; ++i)
      {
         if (X[i] & Q) X[0] ^= P;
         else { const unsigned t = (X[0] ^ X[i]) & P; X[0] ^= t; X[i] ^= t; }
      }
   }
   for (int i = 1; i < 3; ++i) X[i] ^= X[i - 1];        // Gray encode
   unsigned t = 0;
   for (unsigned Q = M; Q > 1; Q >>= 1) if (X[2] & Q) t ^= Q - 1;
   for (int i = 0; i < 3; ++i) X[i] ^= t;
   unsigned long long key = 0;                          // interleave the transposed bits, most significant first
   for (int b = bits - 1; b >= 0; --b)
      for (int i = 0; i < 3; ++i) key = (key << 1) | ((X[i] >> b) & 1u);
   return key;
}

// The order table of one set as a host vector: mode 0 the set in its own order, 1 largest block first, 2 Hilbert curve.
// Pure host arithmetic (comd_block_order_table exposes it to the CPU tests: every set must list each of its blocks once).
static inline std::vector<int> nlBlockOrderTable(const int cellsN[3], int bx, int by, int bz, int phase, int mode)
{
   NlBlocks T = {};
   T.bx = bx; T.by = by; T.bz = bz;
   T.nbx = (cellsN[0] + bx - 1) / bx; T.nby = (cellsN[1] + by - 1) / by; T.nbz = (cellsN[2] + bz - 1) / bz;
   T.phase = phase;
   const int all = T.nbx * T.nby * T.nbz, interior = nlInteriorCount(T);
   const int n = phase == NLB_ALL ? all : (phase == NLB_INTERIOR ? interior : all - interior);
   std::vector<int> ids(n > 0 ? n : 0);
   if (n <= 0) return ids;
   int bits = 1;
   while ((1 << bits) < std::max(T.nbx, std::max(T.nby, T.nbz))) ++bits;
   std::vector<std::pair<unsigned long long, int>> keyed(n);
   for (int q = 0; q < n; ++q)
   {
      int ix, iy, iz;
      nlBlockCoordOfSet(T, q, ix, iy, iz);
      unsigned long long key = (unsigned long long)q;          // the set in its own order
      if (mode == 2) key = nlHilbertKey((unsigned)ix, (unsigned)iy, (unsigned)iz, bits);
      else if (mode == 1)
      {  // cells of the block, descending; index order among equals
         const int cx = std::min(bx, cellsN[0] - ix * bx), cy = std::min(by, cellsN[1] - iy * by), cz = std::min(bz, cellsN[2] - iz * bz);
         key = ((unsigned long long)(bx * by * bz - cx * cy * cz) << 32) | (unsigned)q;
      }
      keyed[q] = { key, ix + T.nbx * (iy + T.nby * iz) };
   }
   std::sort(keyed.begin(), keyed.end());
   for (int q = 0; q < n; ++q) ids[q] = keyed[q].second;
   return ids;
}

// order table of one phase set (device memory, cached in the handle until the block shape or the grid changes)
static int nlBlockOrder(ComdDev* d, NlBlocks& B, int phase, cudaStream_t s)
{
   B.order = nullptr;
   // Without -H: index order, except where a launch gives every CTA slot only a few blocks (small per-GPU domains): there the
   // set is walked largest block first, so the ragged blocks at the high faces (3/4, 9/16, 27/64 of a full one on a
   // 47-cell axis) fill the tail of the launch (longest-processing-time-first).
   const bool lpt = !d->optNlHilbert && B.total < 148 * 2 * 16 && (d->g.n[0] % B.bx || d->g.n[1] % B.by || d->g.n[2] % B.bz);
   const bool split = phase != NLB_ALL && nlInteriorCount(B) > 0;        // the interior / boundary sets are not index ranges
   if (!d->optNlHilbert && !lpt && !split) return COMD_OK;
   // one cached table per set, valid while the block shape, the grid and the kind of order stay the same
   const int mode = d->optNlHilbert ? 2 : (lpt ? 1 : 0);
   const int sig[7] = { B.bx, B.by, B.bz, B.nbx, B.nby, B.nbz, mode };
   bool same = true;
   for (int k = 0; k < 7; ++k) same = same && d->nlOrderSig[phase][k] == sig[k];
   if (!same)
   {
      if (d->nlOrder[phase]) { cudaFree(d->nlOrder[phase]); d->nlOrder[phase] = nullptr; }
      for (int k = 0; k < 7; ++k) d->nlOrderSig[phase][k] = sig[k];
   }
   if (!d->nlOrder[phase])
   {
      const std::vector<int> ids = nlBlockOrderTable(d->g.n, B.bx, B.by, B.bz, phase, mode);
      if (ids.empty()) return COMD_OK;
      CUDA_TRY(cudaMalloc(&d->nlOrder[phase], ids.size() * sizeof(int)));
      CUDA_TRY(cudaMemcpyAsync(d->nlOrder[phase], ids.data(), ids.size() * sizeof(int), cudaMemcpyHostToDevice, s));
      CUDA_TRY(cudaStreamSynchronize(s));            // ids goes out of scope; once per shape
   }
   B.order = d->nlOrder[phase];
   return COMD_OK;
}
