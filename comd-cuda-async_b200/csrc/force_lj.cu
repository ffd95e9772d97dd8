// force_lj.cu -- Lennard-Jones pair force over the 27-cell stencil (sm_100a).
//
// Replaces ljForceGpu / LJ_Force_thread_atom / LJ_Force_cta_cell (gpu_kernels.cu:70-122,
// gpu_lj_thread_atom.h:31-143, gpu_lj_cta_cell.h:31-122); results must match the CPU
// ljForceCpuNL (ljForce.c:146-265): per pair with r2 <= rc2
//    ir2 = 1/r2; r6 = (s6*ir2)*(ir2*ir2); eLocal = r6*(r6-1) - eShift;
//    fr = -4*eps*r6*ir2*(12*r6-6); f_i -= dr*fr; U_i += eLocal       (U in units of 4 eps, halved at the end)
// Full-neighbour ("gather") form: every local atom accumulates its own force, so no
// atomics and a fixed summation order (cell-major, gid-minor).
#include "force_n3.cuh"

struct LjParams { real_t s6, eps, rc2, eShift; };

// One thread per atom; the 27-cell stencil is 9 contiguous atom runs thanks to the padded
// x-fastest cell order.  All lanes of a warp that share a cell read the same j -> one
// broadcast transaction per load.
__global__ void __launch_bounds__(128)
k_lj_force_rows(DevGeom g, LjParams P,
                const real_t* __restrict__ rx, const real_t* __restrict__ ry, const real_t* __restrict__ rz,
                const int* __restrict__ cell, const unsigned char* __restrict__ cellType,
                const int* __restrict__ cellStart, const int* __restrict__ cnt,
                real_t* __restrict__ fx, real_t* __restrict__ fy, real_t* __restrict__ fz, real_t* __restrict__ e)
{
   const int n = cnt[C_NATOMS];
   const int rowY = g.P[0], rowZ = g.P[0] * g.P[1];
   for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
   {
      const int c = cell[i];
      if (!(cellType[c] & CT_LOCAL)) continue;
      const real_t xi = rx[i], yi = ry[i], zi = rz[i];
      real_t ax = R(0.0), ay = R(0.0), az = R(0.0), u = R(0.0);
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
               if (r2 <= P.rc2 && j != i)
               {
                  const real_t ir2 = R(1.0) / r2;
                  const real_t r6 = (P.s6 * ir2) * (ir2 * ir2);
                  u += r6 * (r6 - R(1.0)) - P.eShift;
                  const real_t fr = -R(4.0) * P.eps * r6 * ir2 * (R(12.0) * r6 - R(6.0));
                  ax -= dx * fr; ay -= dyy * fr; az -= dzz * fr;
               }
            }
         }
      }
      fx[i] = ax; fy[i] = ay; fz[i] = az;
      e[i] = (R(0.5) * u) * (R(4.0) * P.eps);
   }
}

int comdForceLJ(ComdDev* d)
{
   LjParams P;
   const double s = d->ljSigma;
   P.s6 = s * s * s * s * s * s;
   P.eps = d->ljEps;
   P.rc2 = d->cutoff * d->cutoff;
   const real_t rc6 = P.s6 / (P.rc2 * P.rc2 * P.rc2);
   P.eShift = R(1.0) * rc6 * (rc6 - R(1.0));                 // POT_SHIFT 1.0, ljForce.c:83,173-174
   d->eValid = 1;
   if (d->useNL || d->optForceVariant == 1)
   {
      PairParams Q = {};
      Q.s6 = P.s6; Q.eps = P.eps; Q.eShift = P.eShift; Q.rc2 = P.rc2;
      Q.skipFlag = d->specStep ? d->cnt + C_NLREBUILD : nullptr;
      if (d->useNL && d->nlN3Shape[0] > 0)
      {  // in-block pairs once (force_n3.cuh); the per-atom energies only when somebody can read them
         d->eValid = d->wantEnergy;
         const bool split = d->refreshInFlight != 0;
         d->refreshInFlight = 0;
         if (d->wantEnergy) return split ? launchForceN3Split<PK_LJ>(d, Q, d->evRefresh) : launchForceN3<PK_LJ>(d, Q);
         return split ? launchForceN3Split<PK_LJF>(d, Q, d->evRefresh) : launchForceN3<PK_LJF>(d, Q);
      }
      if (d->useNL && d->refreshInFlight) { d->refreshInFlight = 0; return launchForceNLSplit<PK_LJ>(d, Q, d->evRefresh); }
      return d->useNL ? launchForceNL<PK_LJ>(d, Q) : launchForceTile<PK_LJ>(d, Q, nullptr);
   }
   k_lj_force_rows<<<gridFor(d->approxAtoms, 128, 148 * 64), 128, 0, d->stream>>>(d->g, P, d->A.r[0], d->A.r[1], d->A.r[2],
      d->A.cell, d->cellType, d->cellStart, d->cnt, d->f[0], d->f[1], d->f[2], d->e);
   LAUNCH_CHECK(d);
   return COMD_OK;
}
