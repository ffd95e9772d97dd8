// force_eam.cu -- three-phase EAM force (sm_100a) and the F'(rho) ghost fill.
//
// Replaces eamForce{1,2,3}Gpu and the EAM_Force_* kernels (gpu_kernels.cu:282-325,
// gpu_eam_thread_atom.h:35-287, gpu_eam_cta_cell.h:37-295) plus LoadForceBuffer /
// UnloadForceBuffer (gpu_redistribute.h:638-672); results must match the CPU
// eamForceCpuNL (eam.c:266-419) and its table lookup interpolate (eam.c:557-579):
//   pass 1  per pair r2 <= rc2 : r = sqrt(r2); (phi,phi') , (rho,.) ; f_i -= phi' dr/r ;
//           U_i += phi/2 ; rhobar_i += rho
//   pass 2  per local atom     : (F,F') = interp(F, rhobar_i) ; U_i += F ; dfEmbed_i = F'
//   halo    F' of ghost atoms  : three axis phases, plane by plane (haloExchange.c:1712-1860)
//   pass 3  per pair r2 <  rc2 : f_i -= (F'_i + F'_j) rho' dr/r
#include "force_n3.cuh"
#include "halo_runs.cuh"

// interpolate(), eam.c:557-579, incl. its one-sided clamps
__device__ __forceinline__ void tableLookup(const DevTable& t, real_t r, real_t& f, real_t& df)
{
   const real_t* __restrict__ tt = t.v;
   if (r < t.x0) r = t.x0;
   r = (r - t.x0) * t.invDx;
   int ii = (int)floor(r);
   if (ii > t.n) { ii = t.n; r = t.n / t.invDx; }
   r = r - floor(r);
   const real_t tm = __ldg(tt + ii - 1), t0 = __ldg(tt + ii), t1 = __ldg(tt + ii + 1), t2 = __ldg(tt + ii + 2);
   const real_t g1 = t1 - tm;
   const real_t g2 = t2 - t0;
   f = t0 + R(0.5) * r * (g1 + r * (t1 + tm - R(2.0) * t0));
   df = R(0.5) * (g1 + r * (g2 - g1)) * t.invDx;
}

__global__ void __launch_bounds__(128)
k_eam_pass1(DevGeom g, DevTable tPhi, DevTable tRho, DevTable tF, real_t rc2,
            const real_t* __restrict__ rx, const real_t* __restrict__ ry, const real_t* __restrict__ rz,
            const int* __restrict__ cell, const unsigned char* __restrict__ cellType,
            const int* __restrict__ cellStart, const int* __restrict__ cnt,
            real_t* __restrict__ fx, real_t* __restrict__ fy, real_t* __restrict__ fz, real_t* __restrict__ e,
            real_t* __restrict__ rhobar, real_t* __restrict__ dfEmbed)
{
   const int n = cnt[C_NATOMS];
   const int rowY = g.P[0], rowZ = g.P[0] * g.P[1];
   for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
   {
      const int c = cell[i];
      if (!(cellType[c] & CT_LOCAL)) continue;
      const real_t xi = rx[i], yi = ry[i], zi = rz[i];
      real_t ax = R(0.0), ay = R(0.0), az = R(0.0), u = R(0.0), rb = R(0.0);
      #pragma unroll 1
      for (int dz = -1; dz <= 1; ++dz)
      {
         #pragma unroll 1
         for (int dy = -1; dy <= 1; ++dy)
         {
            const int rc = c + dy * rowY + dz * rowZ;
            const int jb = cellStart[rc - 1], je = cellStart[rc + 2];
            for (int j = jb; j < je; ++j)
            {
               const real_t dx = xi - rx[j], dyy = yi - ry[j], dzz = zi - rz[j];
               const real_t r2 = dx * dx + dyy * dyy + dzz * dzz;
               if (r2 <= rc2 && j != i)
               {
                  const real_t r = sqrt(r2);
                  real_t phi, dphi, rho, drho;
                  tableLookup(tPhi, r, phi, dphi);
                  tableLookup(tRho, r, rho, drho);
                  const real_t s = dphi / r;
                  ax -= s * dx; ay -= s * dyy; az -= s * dzz;
                  u += R(0.5) * phi;
                  rb += rho;
               }
            }
         }
      }
      real_t fe, dfe;
      tableLookup(tF, rb, fe, dfe);
      fx[i] = ax; fy[i] = ay; fz[i] = az;
      e[i] = u + fe;
      rhobar[i] = rb;
      dfEmbed[i] = dfe;
   }
}

__global__ void __launch_bounds__(128)
k_eam_pass3(DevGeom g, DevTable tRho, real_t rc2,
            const real_t* __restrict__ rx, const real_t* __restrict__ ry, const real_t* __restrict__ rz,
            const int* __restrict__ cell, const unsigned char* __restrict__ cellType,
            const int* __restrict__ cellStart, const int* __restrict__ cnt, const real_t* __restrict__ dfEmbed,
            real_t* __restrict__ fx, real_t* __restrict__ fy, real_t* __restrict__ fz)
{
   const int n = cnt[C_NATOMS];
   const int rowY = g.P[0], rowZ = g.P[0] * g.P[1];
   for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
   {
      const int c = cell[i];
      if (!(cellType[c] & CT_LOCAL)) continue;
      const real_t xi = rx[i], yi = ry[i], zi = rz[i];
      const real_t dfi = dfEmbed[i];
      real_t ax = R(0.0), ay = R(0.0), az = R(0.0);
      #pragma unroll 1
      for (int dz = -1; dz <= 1; ++dz)
      {
         #pragma unroll 1
         for (int dy = -1; dy <= 1; ++dy)
         {
            const int rc = c + dy * rowY + dz * rowZ;
            const int jb = cellStart[rc - 1], je = cellStart[rc + 2];
            for (int j = jb; j < je; ++j)
            {
               const real_t dx = xi - rx[j], dyy = yi - ry[j], dzz = zi - rz[j];
               const real_t r2 = dx * dx + dyy * dyy + dzz * dzz;
               if (r2 < rc2 && j != i)
               {
                  const real_t r = sqrt(r2);
                  real_t rho, drho;
                  tableLookup(tRho, r, rho, drho);
                  const real_t s = (dfi + dfEmbed[j]) * drho / r;
                  ax -= s * dx; ay -= s * dyy; az -= s * dzz;
               }
            }
         }
      }
      fx[i] += ax; fy[i] += ay; fz[i] += az;
   }
}

// ------------------------------------------------------------------ F' halo (message layout: halo_runs.cuh)
// one block per (face, send|recv): run starts + exclusive offsets + total
__global__ void __launch_bounds__(1024)
k_fp_layout(DevGeom g, const int* __restrict__ cellStart, int* __restrict__ runTab, int runStride, int* __restrict__ cnt)
{
   __shared__ int warpSum[32];
   __shared__ int carry;
   const int face = blockIdx.x >> 1, isRecv = blockIdx.x & 1;
   const FpFace F = fpFace(g, face, isRecv);
   int* runStart = runTab + (long)blockIdx.x * 2 * runStride;
   int* runOff = runStart + runStride;
   const int t = threadIdx.x, lane = t & 31, w = t >> 5;
   if (t == 0) carry = 0;
   __syncthreads();
   for (int base = 0; base < F.nRuns; base += 1024)
   {
      const int k = base + t;
      int len = 0, b = 0;
      if (k < F.nRuns)
      {
         const int c0 = F.firstCell + (k % F.countA) * F.strideA + (k / F.countA) * F.strideB;
         b = cellStart[c0]; len = cellStart[c0 + F.cellsPerRun] - b;
      }
      int incl = len;
      #pragma unroll
      for (int o = 1; o < 32; o <<= 1) { int u = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += u; }
      if (lane == 31) warpSum[w] = incl;
      __syncthreads();
      if (w == 0)
      {
         int s = warpSum[lane], si = s;
         #pragma unroll
         for (int o = 1; o < 32; o <<= 1) { int u = __shfl_up_sync(0xffffffffu, si, o); if (lane >= o) si += u; }
         warpSum[lane] = si - s;
      }
      __syncthreads();
      const int excl = incl - len + warpSum[w] + carry;
      if (k < F.nRuns) { runStart[k] = b; runOff[k] = excl; }
      __syncthreads();
      if (t == 1023) carry = excl + len;
      __syncthreads();
   }
   if (t == 0) { runOff[F.nRuns] = carry; cnt[(isRecv ? C_FPRECV0 : C_FPSEND0) + face] = carry; }
}

__global__ void __launch_bounds__(256)
k_fp_pack(DevGeom g, int axis, const int* __restrict__ runTab, int runStride, const real_t* __restrict__ dfEmbed,
          real_t* __restrict__ bufM, real_t* __restrict__ bufP, long msgCap, int* __restrict__ cnt)
{
   const int d = blockIdx.y, face = 2 * axis + d;
   const FpFace F = fpFace(g, face, 0);
   const int* runStart = runTab + (long)(2 * face) * 2 * runStride;
   const int* runOff = runStart + runStride;
   const int total = runOff[F.nRuns];
   if (total > msgCap) { if (threadIdx.x == 0 && blockIdx.x == 0) atomicOr(&cnt[C_ERR], ERRB_CAPACITY); return; }
   real_t* out = d ? bufP : bufM;
   for (int m = blockIdx.x * blockDim.x + threadIdx.x; m < total; m += gridDim.x * blockDim.x)
   {
      const int k = fpLocate(runOff, F.nRuns, m);
      out[m] = dfEmbed[runStart[k] + (m - runOff[k])];
   }
}

__global__ void __launch_bounds__(256)
k_fp_unpack(DevGeom g, int axis, const int* __restrict__ runTab, int runStride, real_t* __restrict__ dfEmbed,
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
      const int a = runStart[k] + (m - runOff[k]);
      dfEmbed[a] = in[m];
      if (rq) rq[a].w = in[m];
   }
}

// run starts, offsets and message sizes of all six faces; in list mode the sorted arrays do not move between
// rebuilds, so the layout of the last sort stays valid (d->runLayoutValid) and is shared with the ghost refresh
int comdHaloRunLayout(ComdDev* d, cudaStream_t s)
{
   if (d->useNL && d->runLayoutValid) return COMD_OK;
   k_fp_layout<<<12, 1024, 0, s>>>(d->g, d->cellStart, d->runOff, haloRunStride(d->g), d->cnt); LAUNCH_CHECK(d);
   if (d->useNL) d->runLayoutValid = 1;
   return COMD_OK;
}

int comdFprimeHalo(ComdDev* d)
{
   NVTX_RANGE("F'(rho) halo: pack / exchange / unpack x,y,z");
   const DevGeom& g = d->g;
   cudaStream_t s = d->xs;
   const int runStride = haloRunStride(g);
   int rc = comdHaloRunLayout(d, s);
   if (rc != COMD_OK) return rc;
   for (int ax = 0; ax < 3; ++ax)
   {
      dim3 grid(d->nRanks > 1 ? 32 : 148, 2);   // ranks > 1: few CTAs, they must fit beside an overlapped force launch
      k_fp_pack<<<grid, 256, 0, s>>>(g, ax, d->runOff, runStride, d->dfEmbed, d->fpSend[0], d->fpSend[1], d->msgCap, d->cnt); LAUNCH_CHECK(d);
      rc = comdExchangeFprimePhase(d, ax);
      if (rc != COMD_OK) return rc;
      k_fp_unpack<<<grid, 256, 0, s>>>(g, ax, d->runOff, runStride, d->dfEmbed, d->fpRecv[0], d->fpRecv[1],
                                       d->cnt + C_FPGOT0 + 2 * ax, d->cnt + C_FPGOT0 + 2 * ax + 1, d->cnt, (d->useNL || (d->optForceVariant == 1 && d->optTileHits)) ? d->A.rq : nullptr); LAUNCH_CHECK(d);
   }
   return COMD_OK;
}

int comdForceEAM(ComdDev* d)
{
   const double rc2 = d->cutoff * d->cutoff;
   cudaStream_t s = d->stream;
   PairParams Q = {};
   Q.rc2 = rc2; Q.tPhi = d->tPhi; Q.tRho = d->tRho; Q.tF = d->tF; Q.tPR = d->tPR; Q.tW0 = d->tPR ? d->tW[0] : nullptr; Q.tW1 = d->tW[1]; Q.tDR = d->tDR;
   Q.rvPhi = d->rvPhi; Q.rvRho = d->rvRho; Q.r3 = d->r3;
   Q.skipFlag = d->specStep ? d->cnt + C_NLREBUILD : nullptr;
   {  // rhobar never exceeds the domain of F (beyond it the lookup clamps, eam.c:561-571): 2^53 quanta over that domain
      // leave a factor 4 to the 32-bit high word and let one pair contribute up to 1/4 of the domain
      const double rhoMax = fabs(d->tF.x0) + (d->tF.n + 1) / d->tF.invDx;
      int ex = 0; frexp(rhoMax, &ex);                   // rhoMax < 2^ex
      Q.n3ScaleRho = ldexp(1.0, 53 - ex);
   }
   if (d->useNL)
   {  // list path: rhobar, F' halo, then every pair force in pass 3 (force_nl.cuh); pair energies on request
      d->eValid = d->wantEnergy;
      const bool overlap = d->nRanks > 1 && d->optOverlap;
      int rc;
      nvtxRangePushA("EAM pass 1+2: rhobar, F, F'");
      const bool n3 = d->nlN3Shape[0] > 0;             // the list stores in-block pairs once (force_n3.cuh)
      if (d->refreshInFlight) { rc = n3 ? launchForceN3Split<PK_RHO>(d, Q, d->evRefresh) : launchForceNLSplit<PK_RHO>(d, Q, d->evRefresh); d->refreshInFlight = 0; }
      else rc = n3 ? launchForceN3<PK_RHO>(d, Q) : launchForceNL<PK_RHO>(d, Q);
      nvtxRangePop();
      if (rc != COMD_OK) return rc;
      PH_MARK(d, 3);
      NVTX_RANGE("EAM F' halo + pass 3");
      if (!overlap)
      {
         rc = comdFprimeHalo(d);
         if (rc != COMD_OK) return rc;
         PH_MARK(d, 4);
         if (n3) return d->wantEnergy ? launchForceN3<PK_EAMFE>(d, Q) : launchForceN3<PK_EAMF>(d, Q);
         return d->wantEnergy ? launchForceNL<PK_EAMFE>(d, Q) : launchForceNL<PK_EAMF>(d, Q);
      }
      // F'(rhobar) of the ghosts travels on the exchange stream while pass 3 works on the interior blocks
      CUDA_TRY(cudaEventRecord(d->evPass1, d->stream));
      CUDA_TRY(cudaStreamWaitEvent(d->commStream, d->evPass1, 0));
      d->xs = d->commStream;
      rc = comdFprimeHalo(d);
      d->xs = d->stream;
      if (rc != COMD_OK) return rc;
      CUDA_TRY(cudaEventRecord(d->evFp, d->commStream));
      PH_MARK(d, 4);
      if (n3) return d->wantEnergy ? launchForceN3Split<PK_EAMFE>(d, Q, d->evFp) : launchForceN3Split<PK_EAMF>(d, Q, d->evFp);
      return d->wantEnergy ? launchForceNLSplit<PK_EAMFE>(d, Q, d->evFp) : launchForceNLSplit<PK_EAMF>(d, Q, d->evFp);
   }
   d->eValid = 1;
   if (d->optForceVariant == 1 && d->optTileHits)
   {  // Link-cell path, hits written once: the tiled engine runs the 27-cell stencil for the density pass alone (TMA-staged rows,
      // FP32 filter, compaction, exact FP64 test) and writes every pair it found inside the cutoff to a per-step hit list; after the
      // F' halo the list kernel evaluates all pair forces from those hits instead of a second sweep over the 513 stencil candidates
      // per atom (round 1 ran the tiled engine twice: C3 2.29 -> see profiles/r2_experiments.md).  Same arithmetic as the list path
      // (eam.c:311-345 and :390-409 regrouped as PK_RHO / PK_EAMF), the cells are still re-binned every step.
      int rc = comdNeighborAllocate(d);
      if (rc != COMD_OK) return rc;
      d->eValid = d->wantEnergy;
      rc = launchForceTile<PK_RHOL>(d, Q, nullptr);
      if (rc != COMD_OK) return rc;
      rc = comdFprimeHalo(d);
      if (rc != COMD_OK) return rc;
      return d->wantEnergy ? launchForceNL<PK_EAMFE>(d, Q) : launchForceNL<PK_EAMF>(d, Q);
   }
   if (d->optForceVariant == 1)
   {
      int rc = launchForceTile<PK_EAM1>(d, Q, nullptr);
      if (rc != COMD_OK) return rc;
      rc = comdFprimeHalo(d);
      if (rc != COMD_OK) return rc;
      return launchForceTile<PK_EAM3>(d, Q, d->dfEmbed);
   }
   const int grid = gridFor(d->approxAtoms, 128, 148 * 64);
   k_eam_pass1<<<grid, 128, 0, s>>>(d->g, d->tPhi, d->tRho, d->tF, rc2, d->A.r[0], d->A.r[1], d->A.r[2], d->A.cell, d->cellType,
      d->cellStart, d->cnt, d->f[0], d->f[1], d->f[2], d->e, d->rhobar, d->dfEmbed); LAUNCH_CHECK(d);
   int rc = comdFprimeHalo(d);
   if (rc != COMD_OK) return rc;
   k_eam_pass3<<<grid, 128, 0, s>>>(d->g, d->tRho, rc2, d->A.r[0], d->A.r[1], d->A.r[2], d->A.cell, d->cellType,
      d->cellStart, d->cnt, d->dfEmbed, d->f[0], d->f[1], d->f[2]); LAUNCH_CHECK(d);
   return COMD_OK;
}
