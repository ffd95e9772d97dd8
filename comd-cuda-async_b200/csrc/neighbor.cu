// neighbor.cu -- the optional neighbour-list path (sm_100a): list build over the 27-cell
// stencil with radius cutoff + skin, the max-displacement rebuild criterion, and the in-place
// ghost refresh that replaces re-binning on the steps between rebuilds.
//
// Reference behaviour being replaced (file:line into /root/reference/src-mpi):
//   buildNeighborListGpu / buildNeighborListKernel_warp   gpu_kernels.cu:1981-2019, 1494-1693
//   buildNeighborListCpu (list radius, lastR snapshot)    neighborList.c:101-191
//   neighborListUpdateRequired{Cpu,Gpu}                   neighborList.c:212-247, gpu_kernels.cu:1461-1480
//   redistributeAtomsGpuNL (rebuild-or-refresh decision)  timestep.c:278-354
//   unloadAtomsBuffer, in-place ghost update              haloExchange.c:1673-1683
// The list is FULL (every local atom lists all its neighbours, like the reference's GPU lists) for the full-list kernels
// (option nl_n3 0, LJ's default); for the once-per-pair kernels (nl_n3 1, EAM's default) a pair whose two atoms lie in the same
// work block is stored on the lower row only and tagged (force_n3.cuh), every other pair on both rows.  Either way the rows are
// row-major with a fixed stride, entries are indices into the sorted atom arrays (ghosts
// included) in ascending order, i.e. cell-major / gid-minor: summation order stays canonical.
#include "comd_dev.cuh"
#include "halo_runs.cuh"
#include "nl_blocks.cuh"
#include <cmath>

#define NL_G 8                  // lanes per atom in the build and force kernels

// ---- build -------------------------------------------------------------------------------
// One sweep over the 9 stencil rows, 8 lanes per atom, positions from the AoS gather copy (8 lanes
// read 256 contiguous bytes).  Pairs inside the cutoff go straight to the front of the row; pairs
// in the skin shell (rc, rc+skin] are parked in shared memory and appended behind them, both parts
// in ascending index order -- the force kernels then meet rounds that are all hits followed by
// rounds that are all misses instead of a 70 % mixture in every round.
#define NL_FARCAP 64            // skin-shell entries parked per atom
#define NL_U 4                  // candidate rounds in flight per trip
// Row layout: chunks of 32 entries, each stored transposed -- entry e = 8u + l of a chunk (round u, lane l of
// the force kernels) sits at 4l + u, so a lane reads the indices of four rounds with one 128-bit load.
// (G = lanes per atom of the force kernel that will walk the list: 8, or 4 for the once-per-pair kernels; chunk = 4 G entries)
__device__ __forceinline__ int nlSlot(int pos, int G) { return G == 8 ? ((pos & ~31) | ((pos & 7) << 2) | ((pos >> 3) & 3)) : ((pos & ~15) | ((pos & 3) << 2) | ((pos >> 2) & 3)); }
// N3 (bx > 0): a pair whose two atoms belong to the same work block of bx*by*bz cells is stored ONCE, on the row of
// the lower sorted index, tagged with N3_FLAG and the neighbour's (y,z) row inside the block (force_n3.cuh); the
// row of the higher index skips it -- without loading the position.  Pairs across a block face, and every pair with
// a ghost, stay on both rows as before.
#define N3_FLAG     0x80000000u
#define N3_ROWSHIFT 26
// N3: once-per-pair rule on (bx > 0); G: lanes per row of the force kernel that will walk the list (slot transposition).
// Template parameters, so that the full-list build (LJ's default) carries none of the block arithmetic: 7.96 -> see
// profiles/r2_experiments.md.
template <bool N3, int G>
__global__ void __launch_bounds__(256)
k_nl_build(DevGeom g, real_t rc2, real_t rl2, const real4* __restrict__ rq,
           const int* __restrict__ cell, const unsigned char* __restrict__ cellType, const int* __restrict__ cellStart,
           int* __restrict__ cnt, int* __restrict__ list, int* __restrict__ count, int stride, long rows,
           real_t* __restrict__ lx, real_t* __restrict__ ly, real_t* __restrict__ lz, unsigned long long* __restrict__ stat,
           int bx, int by, int bz)
{
   __shared__ int farBuf[8][4][NL_FARCAP];
   const unsigned full = 0xffffffffu;
   const int n = cnt[C_NATOMS];
   const int lane = threadIdx.x & 31, g4 = lane / NL_G, gl = lane % NL_G, w = threadIdx.x >> 5;
   const int warpsPerGrid = gridDim.x * (blockDim.x >> 5);
   const int rowY = g.P[0], rowZ = g.P[0] * g.P[1];
   const int nQuads = (n + 3) >> 2;
   int* myFar = farBuf[w][g4];
   for (int q = blockIdx.x * (blockDim.x >> 5) + w; q < nQuads; q += warpsPerGrid)
   {
      const int i = 4 * q + g4;
      int c = 0;
      bool act = false;
      if (i < n) { c = cell[i]; act = (cellType[c] & CT_LOCAL) != 0 && i < rows; }
      if (!__any_sync(full, act)) continue;
      real4 pi = make_real4(R(0.0), R(0.0), R(0.0), R(0.0));
      const int i0 = act ? i : 0;                      // a lane without a candidate re-reads its own atom
      if (act) pi = rq[i];
      // block membership of the 27 stencil cells relative to the atom's own cell (N3)
      int cx = 0, cy = 0, cz = 0, xSame = 0, ry = 0, rz = 0;      // ry, rz: the cell's (y, z) position inside its work block
      if (N3 && act)
      {
         cx = c % g.P[0] - 1; cy = (c / g.P[0]) % g.P[1] - 1; cz = c / rowZ - 1;
         ry = cy % by; rz = cz % bz;
         for (int dxc = -1; dxc <= 1; ++dxc)
         {
            const int x = cx + dxc;
            if (x >= 0 && x < g.n[0] && x / bx == cx / bx) xSame |= 1 << (dxc + 1);
         }
      }
      int nNear = 0, nFar = 0, nTwice = 0;
      int* row = list + (long)(act ? i : 0) * stride;
      #pragma unroll 1
      for (int r = 0; r < 9; ++r)
      {
         int jb = 0, je = 0;
         int sameLo = 0; unsigned sameLen = 0;           // [sameLo, sameLo + sameLen): the candidates of this row that share the atom's work block
         unsigned tag = 0; bool rowSame = false;
         if (act)
         {
            const int dyr = (r % 3) - 1, dzr = (r / 3) - 1;
            const int rc = c + dyr * rowY + dzr * rowZ;
            jb = cellStart[rc - 1]; je = cellStart[rc + 2];
            if (N3)
            {
               // same block in y and z: the neighbour row stays inside the domain and inside the block (no division per row)
               rowSame = (unsigned)(cy + dyr) < (unsigned)g.n[1] && (unsigned)(cz + dzr) < (unsigned)g.n[2]
                      && (unsigned)(ry + dyr) < (unsigned)by && (unsigned)(rz + dzr) < (unsigned)bz;
               if (rowSame)
               {
                  const int b1 = cellStart[rc], b2 = cellStart[rc + 1];
                  tag = N3_FLAG | (unsigned)((ry + dyr) + by * (rz + dzr)) << N3_ROWSHIFT;
                  // the in-block cells of a row are adjacent (the atom's own x cell is always one of them): one index range
                  sameLo = (xSame & 1) ? jb : ((xSame & 2) ? b1 : b2);
                  sameLen = (unsigned)(((xSame & 4) ? je : ((xSame & 2) ? b2 : b1)) - sameLo);
                  // Candidates that would only be skipped (same block, lower index: the pair lives on the other row) are
                  // cut out of the sweep where they form one end of the run: a lower stencil row of the same block keeps
                  // just its out-of-block cell, the atom's own row starts behind the atom when its x-1 cell is in the block.
                  if (r < 4)
                  {
                     if (xSame == 7) je = jb;
                     else if (xSame == 3) jb = b2;
                     else if (xSame == 6) je = b1;
                  }
                  else if (r == 4 && (xSame & 1)) jb = i + 1;
               }
            }
         }
         const int maxlen = __reduce_max_sync(full, je - jb);
         // four rounds of 8 candidates per trip: the four 32-byte position loads are issued before any of
         // the (ballot-serialised) bookkeeping, so a trip exposes one memory latency instead of four
         for (int base = 0; base < maxlen; base += NL_U * NL_G)
         {
            real4 pj[NL_U];
            bool ok[NL_U], both[NL_U];
            #pragma unroll
            for (int u = 0; u < NL_U; ++u)
            {
               const int j = jb + base + u * NL_G + gl;
               ok[u] = j < je && j != i;
               both[u] = N3 && (unsigned)(j - sameLo) < sameLen;   // same block iff the candidate's x cell is (row and block in y, z already are)
               if (both[u] && j < i) ok[u] = false;                // stored on row j
               pj[u] = rq[ok[u] ? j : i0];
            }
            #pragma unroll
            for (int u = 0; u < NL_U; ++u)
            {
               const int j = jb + base + u * NL_G + gl;
               const real_t dx = pi.x - pj[u].x, dy = pi.y - pj[u].y, dz = pi.z - pj[u].z;
               const real_t r2 = dx * dx + dy * dy + dz * dz;
               const bool near = ok[u] && r2 <= rc2;
               const bool far = ok[u] && !near && r2 <= rl2;
               const unsigned below = (1u << gl) - 1u, grp = (1u << NL_G) - 1u;
               const unsigned mN = (__ballot_sync(full, near) >> (g4 * NL_G)) & grp;
               const unsigned mF = (__ballot_sync(full, far) >> (g4 * NL_G)) & grp;
               const int ent = both[u] ? (int)((unsigned)j | tag) : j;
               if (near) { const int pos = nNear + __popc(mN & below); if (pos < stride) row[nlSlot(pos, G)] = ent; }
               if (far) { const int pos = nFar + __popc(mF & below); if (pos < NL_FARCAP) myFar[pos] = ent; }
               nNear += __popc(mN);
               nFar += __popc(mF);
               nTwice += (near || far) && both[u];
            }
         }
      }
      __syncwarp();
      const int nn = nNear + nFar;
      const bool over = nn > stride || nFar > NL_FARCAP;
      if (act && !over) for (int k = gl; k < nFar; k += NL_G) row[nlSlot(nNear + k, G)] = myFar[k];
      __syncwarp();
      if (act && gl == 0)
      {
         count[i] = over ? 0 : nn;
         if (over) atomicOr(&cnt[C_ERR], ERRB_NLFULL);
         lx[i] = pi.x; ly[i] = pi.y; lz[i] = pi.z;            // lastR snapshot (neighborList.c:176-183)
      }
      // statistics (pairs counted as a full list would hold them: an in-block pair is two entries' worth): one atomic per warp
      #pragma unroll
      for (int o = NL_G / 2; o > 0; o >>= 1) nTwice += __shfl_xor_sync(full, nTwice, o);
      int tot = (act && gl == 0) ? nn + nTwice : 0, mx = (act && gl == 0) ? nn : 0;
      #pragma unroll
      for (int o = 16; o > 0; o >>= 1) { tot += __shfl_xor_sync(full, tot, o); mx = max(mx, __shfl_xor_sync(full, mx, o)); }
      if (lane == 0) { atomicAdd(&stat[0], (unsigned long long)tot); atomicMax(&stat[1], (unsigned long long)mx); }
   }
}

// ---- rebuild criterion: any local atom moved more than skin/2 since the last build --------
__global__ void __launch_bounds__(256)
k_nl_check(real_t half2, const real_t* __restrict__ rx, const real_t* __restrict__ ry, const real_t* __restrict__ rz,
           const real_t* __restrict__ lx, const real_t* __restrict__ ly, const real_t* __restrict__ lz,
           const int* __restrict__ cell, const unsigned char* __restrict__ cellType, int* __restrict__ cnt)
{
   const int n = cnt[C_NATOMS];
   bool hit = false;
   for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
   {
      if (!(cellType[cell[i]] & CT_LOCAL)) continue;
      const real_t dx = lx[i] - rx[i], dy = ly[i] - ry[i], dz = lz[i] - rz[i];
      hit |= dx * dx + dy * dy + dz * dz > half2;
   }
   if (__any_sync(0xffffffffu, hit) && (threadIdx.x & 31) == 0) cnt[C_NLREBUILD] = 1;   // every writer stores the same value
}

// ---- ghost refresh: positions only, through the canonical cell-run layout of the F' halo (halo_runs.cuh) -------
// Between rebuilds the sorted arrays do not move, so every face's ghost cells hold exactly the atoms of the
// neighbour's outermost local plane (incl. the halo cells already filled by the earlier axis phases), in the same
// (cell, gid) order on both sides: the message is a straight concatenation of cell runs and needs no index lists.
// This also covers an atom that crossed a rank (or periodic) face at the last rebuild: it lives on as a ghost in
// the old owner's halo cell and as a local atom in the new owner's boundary cell -- two members of matching runs --
// so the old owner's copy follows the live atom.  (The first version replayed the send/recv lists of the rebuild's
// atom exchange; those listed the emigrant's stale halo-cell copy as the SOURCE and overwrote the live atom with it.
// The reference avoids the case by forcing a second re-bin + exchange whenever a rank received a new local atom,
// timestep.c:328-351; results are the same, the lists here simply never contain an emigrant.)
__global__ void __launch_bounds__(256)
k_nl_refresh_pack(DevGeom g, int axis, const int* __restrict__ runTab, int runStride,
                  const real_t* __restrict__ rx, const real_t* __restrict__ ry, const real_t* __restrict__ rz,
                  real_t* __restrict__ bufM, real_t* __restrict__ bufP, long msgCap, int* __restrict__ cnt,
                  real_t shMx, real_t shMy, real_t shMz, real_t shPx, real_t shPy, real_t shPz)
{
   const int d = blockIdx.y, face = 2 * axis + d;
   const FpFace F = fpFace(g, face, 0);
   const int* runStart = runTab + (long)(2 * face) * 2 * runStride;
   const int* runOff = runStart + runStride;
   const int total = runOff[F.nRuns];
   if (total > msgCap) { if (threadIdx.x == 0 && blockIdx.x == 0) atomicOr(&cnt[C_ERR], ERRB_CAPACITY); return; }
   real_t* out = d ? bufP : bufM;
   const real_t sx = d ? shPx : shMx, sy = d ? shPy : shMy, sz = d ? shPz : shMz;
   for (int m = blockIdx.x * blockDim.x + threadIdx.x; m < total; m += gridDim.x * blockDim.x)
   {
      const int k = fpLocate(runOff, F.nRuns, m);
      const int i = runStart[k] + (m - runOff[k]);
      out[3 * (long)m + 0] = rx[i] + sx;                            // same add as loadAtomsBuffer (haloExchange.c:1603-1605)
      out[3 * (long)m + 1] = ry[i] + sy;
      out[3 * (long)m + 2] = rz[i] + sz;
   }
}
__global__ void __launch_bounds__(256)
k_nl_refresh_unpack(DevGeom g, int axis, const int* __restrict__ runTab, int runStride,
                    real_t* __restrict__ rx, real_t* __restrict__ ry, real_t* __restrict__ rz,
                    const real_t* __restrict__ bufFromM, const real_t* __restrict__ bufFromP,
                    const int* __restrict__ nFromM, const int* __restrict__ nFromP, int* __restrict__ cnt, real4* __restrict__ rq)
{
   const int d = blockIdx.y, face = 2 * axis + d;
   const FpFace F = fpFace(g, face, 1);
   const int* runStart = runTab + (long)(2 * face + 1) * 2 * runStride;
   const int* runOff = runStart + runStride;
   const int total = runOff[F.nRuns];
   const int got = d ? *nFromP : *nFromM;
   if (got != total) { if (threadIdx.x == 0 && blockIdx.x == 0) atomicOr(&cnt[C_ERR], ERRB_FPMISMATCH); return; }
   const real_t* in = d ? bufFromP : bufFromM;
   for (int m = blockIdx.x * blockDim.x + threadIdx.x; m < total; m += gridDim.x * blockDim.x)
   {
      const int k = fpLocate(runOff, F.nRuns, m);
      const int i = runStart[k] + (m - runOff[k]);                  // a halo cell by construction: never a local atom
      const real_t x = in[3 * (long)m + 0], y = in[3 * (long)m + 1], z = in[3 * (long)m + 2];
      rx[i] = x; ry[i] = y; rz[i] = z;
      rq[i] = make_real4(x, y, z, R(0.0));
   }
}

// ------------------------------------------------------------------ host side
static int nlAllocate(ComdDev* d)
{
   // expected neighbours inside the list radius from the mean density of the local domain
   const DevGeom& g = d->g;
   double vol = 1.0;
   for (int a = 0; a < 3; ++a) vol *= (g.lMax[a] - g.lMin[a]);
   const int nLocal = d->hCnt[C_NLOCAL] > 0 ? d->hCnt[C_NLOCAL] : d->approxAtoms;
   const double rl = d->cutoff + d->skin;
   const double expect = 4.0 / 3.0 * 3.14159265358979323846 * rl * rl * rl * (double)nLocal / vol;
   int stride = (int)(expect * R(1.35)) + 16;
   stride = (stride + 31) / 32 * 32;                      // whole chunks (nlSlot)
   const long rows = d->cap;
   // Keep the first allocation: the row stride is sized once from the mean density (a row overflow is reported,
   // ERRB_NLFULL).  Re-deriving it at every rebuild made the stride depend on which counters happened to be
   // on the host (mid-redistribute they are not) and cost a 1.5 GB cudaFree + cudaMalloc inside a time step.
   if (d->nlList && d->nlRows >= rows && d->nlStride >= stride) return COMD_OK;   // (a larger cutoff + skin re-sizes)
   if (d->nlList) { cudaFree(d->nlList); cudaFree(d->nlCount); cudaFree(d->nlCountHit); for (int k = 0; k < 3; ++k) cudaFree(d->nlLastR[k]); }
   d->nlList = nullptr;
   CUDA_TRY(cudaMalloc(&d->nlList, (size_t)rows * stride * sizeof(int)));
   CUDA_TRY(cudaMalloc(&d->nlCount, rows * sizeof(int)));
   CUDA_TRY(cudaMemset(d->nlCount, 0, rows * sizeof(int)));
   CUDA_TRY(cudaMalloc(&d->nlCountHit, rows * sizeof(int)));
   CUDA_TRY(cudaMemset(d->nlCountHit, 0, rows * sizeof(int)));
   for (int k = 0; k < 3; ++k) CUDA_TRY(cudaMalloc(&d->nlLastR[k], rows * sizeof(real_t)));
   d->nlStride = stride; d->nlRows = rows;
   return COMD_OK;
}

int comdNeighborAllocate(ComdDev* d) { return nlAllocate(d); }      // the hit list of the link-cell EAM path lives in the same rows

int comdNeighborBuildIfNeeded(ComdDev* d)
{
   if (!d->useNL || d->nlValid) return COMD_OK;
   const double rl = d->cutoff + d->skin;
   for (int a = 0; a < 3; ++a)
      if (d->hostGeom.boxSize[a] < rl)
      { comdSetError("neighbour list: cell size %g < cutoff + skin %g on axis %d (size the link cells on cutoff+skin, CoMD.c:257-268)",
                     d->hostGeom.boxSize[a], rl, a); return COMD_E_ARG; }
   NVTX_RANGE("neighbour list build");
   int rc = nlAllocate(d);
   if (rc != COMD_OK) return rc;
   cudaStream_t s = d->stream;
   CUDA_TRY(cudaMemsetAsync(d->nlStat, 0, 2 * sizeof(unsigned long long), s));
   const int n = d->approxAtoms > 0 ? d->approxAtoms : (int)d->cap;
   // N3: the list is tied to the work-block shape of the force kernels (in-block pairs are stored once)
   const NlBlocks B = nlBlocksFor(d);
   const bool n3 = comdNlN3(d);
   if (n3 && d->cap > (1L << N3_ROWSHIFT)) { comdSetError("neighbour list: more than %ld atoms per GPU", 1L << N3_ROWSHIFT); return COMD_E_CAPACITY; }
   d->nlN3Shape[0] = n3 ? B.bx : 0; d->nlN3Shape[1] = n3 ? B.by : 0; d->nlN3Shape[2] = n3 ? B.bz : 0;
   d->nlLanes = n3 && d->optNlLanes != 8 ? 4 : 8;
#define NL_BUILD(N3F, GG) k_nl_build<N3F, GG><<<gridFor((n + 3) / 4, 8, 148 * 8), 256, 0, s>>>(d->g, d->cutoff * d->cutoff, rl * rl, d->A.rq, d->A.cell, d->cellType, \
      d->cellStart, d->cnt, d->nlList, d->nlCount, d->nlStride, d->nlRows, d->nlLastR[0], d->nlLastR[1], d->nlLastR[2], d->nlStat, \
      d->nlN3Shape[0], d->nlN3Shape[1], d->nlN3Shape[2])
   if (!n3) NL_BUILD(false, 8); else if (d->nlLanes == 4) NL_BUILD(true, 4); else NL_BUILD(true, 8);
#undef NL_BUILD
   LAUNCH_CHECK(d);
   d->nlValid = 1;
   d->nlBuilds++;
   return COMD_OK;
}

int comdNeighborAfterSort(ComdDev* d)
{
   d->runLayoutValid = 0;                          // new sort: new cell runs (halo_runs.cuh); computed once per rebuild
   return comdHaloRunLayout(d, d->stream);
}

// redistributeAtomsGpuNL (timestep.c:278-354): when no local atom has moved more than skin/2 the
// atoms stay where they are in memory and only the ghost positions are renewed (three axis
// phases, same atoms and order as the last full exchange); otherwise the caller re-bins.
int comdNeighborRefreshOrRebuild(ComdDev* d, int* didRefresh)
{
   cudaStream_t s = d->stream;
   *didRefresh = 0;
   if (d->checkFused) d->checkFused = 0;           // the drift kernel of this step evaluated the criterion (comd_core.cu, k_kick_drift)
   else
   {
      CUDA_TRY(cudaMemsetAsync(d->cnt + C_NLREBUILD, 0, sizeof(int), s));
      const double h = 0.5 * d->skin;
      k_nl_check<<<gridFor(d->approxAtoms, 256, 148 * 8), 256, 0, s>>>(h * h, d->A.r[0], d->A.r[1], d->A.r[2],
         d->nlLastR[0], d->nlLastR[1], d->nlLastR[2], d->A.cell, d->cellType, d->cnt); LAUNCH_CHECK(d);
   }
   int rc = comdAllreduceFlagDevice(d, d->cnt + C_NLREBUILD);     // "if one rebuilds, all rebuild"
   if (rc != COMD_OK) return rc;
   CUDA_TRY(cudaMemcpyAsync(d->hCnt + C_NLREBUILD, d->cnt + C_NLREBUILD, sizeof(int), cudaMemcpyDeviceToHost, s));
   if (d->specStep)
   {  // sync-free step (comd_dev_timestep): the host does not wait for the verdict here.  It enqueues the refresh and the
      // force kernels of the no-rebuild case -- the force kernels read the flag on the device and return at once if it is
      // set (PairParams::skipFlag) -- and looks at the flag after that, when the GPU has a whole step queued.
      // (inside a graph capture the record must be an external one to stay visible to cudaEventSynchronize)
      CUDA_TRY(cudaEventRecordWithFlags(d->evFlag, s, d->capturing ? cudaEventRecordExternal : cudaEventRecordDefault));
      PH_MARK(d, 1);
   }
   else
   {
      CUDA_TRY(cudaStreamSynchronize(s));
      PH_MARK(d, 1);
      if (d->hCnt[C_NLREBUILD]) return COMD_OK;    // caller runs the full re-bin + exchange
   }
   // More than one rank: the three exchange phases go to the high-priority exchange stream and the force
   // kernels start on the interior blocks meanwhile (force_nl.cuh); evRefresh marks "ghosts in place".
   const bool overlap = d->nRanks > 1 && d->optOverlap && d->inTimestep;   // only where the force call is known to follow
   if (overlap)
   {
      CUDA_TRY(cudaEventRecord(d->evStepBegin, s));
      CUDA_TRY(cudaStreamWaitEvent(d->commStream, d->evStepBegin, 0));
      d->xs = d->commStream;
   }
   cudaStream_t x = d->xs;
   const int runStride = haloRunStride(d->g);
   NVTX_RANGE("ghost refresh: pack / exchange / unpack x,y,z");
   for (int ax = 0; ax < 3; ++ax)
   {
      const int fM = 2 * ax, fP = fM + 1;
      dim3 grid(d->nRanks > 1 ? 32 : 148, 2);   // ranks > 1: few CTAs, they must fit beside an overlapped force launch
      k_nl_refresh_pack<<<grid, 256, 0, x>>>(d->g, ax, d->runOff, runStride, d->A.r[0], d->A.r[1], d->A.r[2],
         d->sendBuf[0], d->sendBuf[1], d->msgCap, d->cnt, d->shift[fM][0], d->shift[fM][1], d->shift[fM][2],
         d->shift[fP][0], d->shift[fP][1], d->shift[fP][2]); LAUNCH_CHECK(d);
      rc = comdExchangeRefreshPhase(d, ax);
      if (rc != COMD_OK) { d->xs = d->stream; return rc; }
      k_nl_refresh_unpack<<<grid, 256, 0, x>>>(d->g, ax, d->runOff, runStride, d->A.r[0], d->A.r[1], d->A.r[2],
         d->recvBuf[0], d->recvBuf[1], d->cnt + C_FPGOT0 + 2 * ax, d->cnt + C_FPGOT0 + 2 * ax + 1, d->cnt, d->A.rq); LAUNCH_CHECK(d);
   }
   if (overlap)
   {
      CUDA_TRY(cudaEventRecord(d->evRefresh, d->commStream));
      d->xs = d->stream;
      d->refreshInFlight = 1;
   }
   d->nlRefreshes++;
   *didRefresh = 1;
   return COMD_OK;
}

extern "C" int comd_dev_set_neighbor_list(ComdDev* d, int useNeighborList, double skinDistance)
{
   if (!d || skinDistance < R(0.0)) { comdSetError("comd_dev_set_neighbor_list: bad argument"); return COMD_E_ARG; }
   d->useNL = useNeighborList ? 1 : 0;
   d->skin = skinDistance;
   d->nlValid = 0; d->runLayoutValid = 0;
   return COMD_OK;
}

extern "C" int comd_dev_neighbor_stats(ComdDev* d, long* nPairsStored, int* maxPerAtom, int* nBuilds)
{
   if (!d) return COMD_E_ARG;
   CUDA_TRY(cudaSetDevice(d->device));
   unsigned long long st[2] = { 0, 0 };
   CUDA_TRY(cudaStreamSynchronize(d->stream));
   CUDA_TRY(cudaMemcpy(st, d->nlStat, sizeof st, cudaMemcpyDeviceToHost));
   if (nPairsStored) *nPairsStored = (long)st[0];
   if (maxPerAtom) *maxPerAtom = (int)st[1];
   if (nBuilds) *nBuilds = d->nlBuilds;
   return COMD_OK;
}

extern "C" int comd_dev_neighbor_counters(ComdDev* d, long* nBuilds, long* nRefreshes)
{
   if (!d) return COMD_E_ARG;
   if (nBuilds) *nBuilds = d->nlBuilds;
   if (nRefreshes) *nRefreshes = d->nlRefreshes;
   return COMD_OK;
}

extern "C" unsigned long long comd_hilbert_key(unsigned x, unsigned y, unsigned z, int bits)
{
   return bits >= 1 && bits <= 20 ? nlHilbertKey(x, y, z, bits) : 0ull;
}

// the order table the list kernels would walk for (cells, block shape, set, mode): needs no device (CPU tests)
extern "C" int comd_block_order_table(const int cells[3], int bx, int by, int bz, int phase, int mode, int capacity, int* out)
{
   if (!cells || !out || bx < 1 || by < 1 || bz < 1 || phase < 0 || phase > 2 || mode < 0 || mode > 2) return COMD_E_ARG;
   const std::vector<int> ids = nlBlockOrderTable(cells, bx, by, bz, phase, mode);
   if ((int)ids.size() > capacity) return COMD_E_CAPACITY;
   for (size_t k = 0; k < ids.size(); ++k) out[k] = ids[k];
   return (int)ids.size();
}
