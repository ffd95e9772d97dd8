// comd_core.cu -- lifecycle, atom I/O, link-cell binning (counting sort + scan), the atom
// halo pack/unpack kernels, velocity-Verlet updates and energy reductions for sm_100a.
//
// Reference behaviour being replaced (file:line into /root/reference/src-mpi):
//   updateLinkCellsCpu / updateLinkCellsGpu      linkCells.c:383-405, gpu_kernels.cu:470-504
//   getBoxFromCoord (ownership rule)             linkCells.c:448-480
//   loadAtomsBuffer / unloadAtomsBuffer          haloExchange.c:1576-1688
//   sortAtomsGpu (order inside boundary cells)   gpu_kernels.cu:1014-1043
//   advanceVelocity / advancePosition            timestep.c:135-180
//   kineticEnergy / computeEnergy                timestep.c:109-133, gpu_kernels.cu:1046-1059
#include "comd_dev.cuh"
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <algorithm>

// ------------------------------------------------------------------ errors
static thread_local char g_err[512] = "";
void comdSetError(const char* fmt, ...)
{
   va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof g_err, fmt, ap); va_end(ap);
}
extern "C" const char* comd_last_error(void) { return g_err; }
extern "C" const char* comd_version(void) { return "comd-cuda-async_b200 0.1 (sm_100a)"; }
extern "C" int comd_precision(void) { return (int)sizeof(real_t); }

// ------------------------------------------------------------------ small kernels
__global__ void k_set_int(int* p, int v, int n)
{
   for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) p[i] = v;
}

// start of a redistribute: the source pool is the current sorted array
__global__ void k_redist_begin(int* cnt)
{
   cnt[C_NSRC] = cnt[C_NATOMS];
   cnt[C_NSHELL] = 0;
   cnt[C_NLOCAL] = 0;
   for (int f = 0; f < 6; ++f) { cnt[C_SEND0 + f] = 0; cnt[C_RECV0 + f] = 0; }
}

// Re-bin: every atom that was local gets its new owner cell (linkCells.c:383-405 +
// 448-480); last step's ghosts are dropped (emptyHaloCells, linkCells.c:483-487).
// Atoms whose new cell lies in the 2-plane shell are remembered for the exchange.
__global__ void __launch_bounds__(256)
k_rebin(DevGeom g, const real_t* __restrict__ rx, const real_t* __restrict__ ry, const real_t* __restrict__ rz,
        const int* __restrict__ cellOld, const unsigned char* __restrict__ cellType,
        int* __restrict__ cellKey, int* __restrict__ cellCount, int* __restrict__ shell, int* __restrict__ cnt)
{
   const int n = cnt[C_NATOMS];
   const int lane = threadIdx.x & 31;
   const int stride = gridDim.x * blockDim.x;
   for (int base = blockIdx.x * blockDim.x + (threadIdx.x & ~31); base < n; base += stride)
   {
      const int i = base + lane;
      bool inShell = false;
      if (i < n)
      {
         int key = -1;
         if (cellType[cellOld[i]] & CT_LOCAL)
         {
            int err = 0;
            key = cellOfPos(g, rx[i], ry[i], rz[i], err);
            if (err) atomicOr(&cnt[C_ERR], ERRB_ESCAPE);
            atomicAdd(&cellCount[key], 1);
            inShell = cellType[key] & CT_SHELL;
         }
         cellKey[i] = key;
      }
      const unsigned m = __ballot_sync(0xffffffffu, inShell);
      if (m)
      {
         int b = 0;
         if (lane == 0) b = atomicAdd(&cnt[C_NSHELL], __popc(m));
         b = __shfl_sync(0xffffffffu, b, 0);
         if (inShell) shell[b + __popc(m & ((1u << lane) - 1))] = i;
      }
   }
}

// Pack the two faces of one axis: every shell atom whose (new) cell is in the two
// outermost planes incl. halo cells (mkAtomCellList, haloExchange.c:1543-1567) goes out,
// shifted by +-globalExtent across a periodic face (loadAtomsBuffer, :1576-1611).
__global__ void __launch_bounds__(256)
k_pack_atoms(int axis, const real_t* __restrict__ rx, const real_t* __restrict__ ry, const real_t* __restrict__ rz,
             const real_t* __restrict__ px, const real_t* __restrict__ py, const real_t* __restrict__ pz,
             const int* __restrict__ gid, const int* __restrict__ spec,
             const int* __restrict__ cellKey, const unsigned char* __restrict__ cellType,
             const int* __restrict__ shell, int* __restrict__ cnt,
             real_t* __restrict__ bufM, real_t* __restrict__ bufP, long msgCap,
             real_t shMx, real_t shMy, real_t shMz, real_t shPx, real_t shPy, real_t shPz)
{
   const int n = cnt[C_NSHELL];
   const int lane = threadIdx.x & 31;
   const int stride = gridDim.x * blockDim.x;
   const int fM = 2 * axis, fP = fM + 1;
   for (int base = blockIdx.x * blockDim.x + (threadIdx.x & ~31); base < n; base += stride)
   {
      const int s = base + lane;
      int i = -1; unsigned char t = 0;
      if (s < n) { i = shell[s]; t = cellType[cellKey[i]]; }
      #pragma unroll
      for (int d = 0; d < 2; ++d)
      {
         const int face = d ? fP : fM;
         const bool go = (t & CT_FACE(face)) != 0;
         const unsigned m = __ballot_sync(0xffffffffu, go);
         if (!m) continue;
         int b = 0;
         if (lane == 0) b = atomicAdd(&cnt[C_SEND0 + face], __popc(m));
         b = __shfl_sync(0xffffffffu, b, 0);
         if (go)
         {
            const long k = b + __popc(m & ((1u << lane) - 1));
            if (k >= msgCap) { atomicOr(&cnt[C_ERR], ERRB_CAPACITY); continue; }
            real_t* o = (d ? bufP : bufM) + k * COMD_MSG_REALS;
            msgPutId(o, gid[i], spec[i]);
            o += COMD_MSG_R0;
            o[0] = rx[i] + (d ? shPx : shMx);
            o[1] = ry[i] + (d ? shPy : shMy);
            o[2] = rz[i] + (d ? shPz : shMz);
            o[3] = px[i]; o[4] = py[i]; o[5] = pz[i];
         }
      }
   }
}

// Unpack one received message: putAtomInBox for every atom (unloadAtomsBuffer,
// haloExchange.c:1649-1688; linkCells.c:228-256): it joins the source pool with the cell its
// coordinates select, so ownership transfers by position alone.
__global__ void __launch_bounds__(256)
k_unpack_atoms(DevGeom g, const real_t* __restrict__ buf, const int* __restrict__ nRecvPtr,
               real_t* __restrict__ rx, real_t* __restrict__ ry, real_t* __restrict__ rz,
               real_t* __restrict__ px, real_t* __restrict__ py, real_t* __restrict__ pz,
               int* __restrict__ gid, int* __restrict__ spec,
               int* __restrict__ cellKey, int* __restrict__ cellCount, int* __restrict__ shell,
               int* __restrict__ cnt, long cap)
{
   const int n = *nRecvPtr;
   const int lane = threadIdx.x & 31;
   const int stride = gridDim.x * blockDim.x;
   for (int base = blockIdx.x * blockDim.x + (threadIdx.x & ~31); base < n; base += stride)
   {
      const int k = base + lane;
      const bool have = k < n;
      const unsigned m = __ballot_sync(0xffffffffu, have);
      int b = 0, bs = 0;
      if (lane == 0) { b = atomicAdd(&cnt[C_NSRC], __popc(m)); bs = atomicAdd(&cnt[C_NSHELL], __popc(m)); }
      b = __shfl_sync(0xffffffffu, b, 0);
      bs = __shfl_sync(0xffffffffu, bs, 0);
      if (!have) continue;
      const long j = b + lane;            // lanes with `have` are a prefix of the warp
      if (j >= cap) { atomicOr(&cnt[C_ERR], ERRB_CAPACITY); continue; }
      const real_t* in = buf + (long)k * COMD_MSG_REALS;
      int2 id;
      msgGetId(in, id.x, id.y);
      in += COMD_MSG_R0;
      const real_t x = in[0], y = in[1], z = in[2];
      int err = 0;
      const int c = cellOfPos(g, x, y, z, err);
      if (err) atomicOr(&cnt[C_ERR], ERRB_ESCAPE);
      gid[j] = id.x; spec[j] = id.y;
      rx[j] = x; ry[j] = y; rz[j] = z;
      px[j] = in[3]; py[j] = in[4]; pz[j] = in[5];
      cellKey[j] = c;
      atomicAdd(&cellCount[c], 1);
      shell[bs + lane] = (int)j;
   }
}

// ---- exclusive scan of the per-cell histogram -> cellStart (replaces the CUB call of
// gpu_scan.h:34-39).  Three small launches: tile scan, scan of tile sums, add back.
#define SCAN_TILE 2048
__global__ void __launch_bounds__(1024)
k_scan_tiles(const int* __restrict__ in, int* __restrict__ out, int* __restrict__ tileSum, int n)
{
   __shared__ int warpSum[32];
   const int t = threadIdx.x, lane = t & 31, w = t >> 5;
   const int i0 = blockIdx.x * SCAN_TILE + 2 * t;
   int a = i0 < n ? in[i0] : 0, b = i0 + 1 < n ? in[i0 + 1] : 0;
   int v = a + b, incl = v;
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
      if (lane == 31) tileSum[blockIdx.x] = si;
   }
   __syncthreads();
   const int excl = incl - v + warpSum[w];
   if (i0 < n) out[i0] = excl;
   if (i0 + 1 < n) out[i0 + 1] = excl + a;
}
__global__ void __launch_bounds__(1024)
k_scan_sums(int* __restrict__ tileSum, int nTiles, int* __restrict__ cellStart, int nCells, int* __restrict__ cnt)
{
   __shared__ int warpSum[32];
   __shared__ int carry;
   const int t = threadIdx.x, lane = t & 31, w = t >> 5;
   if (t == 0) carry = 0;
   __syncthreads();
   for (int base = 0; base < nTiles; base += 1024)
   {
      const int i = base + t;
      const int v = i < nTiles ? tileSum[i] : 0;
      int incl = v;
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
      const int excl = incl - v + warpSum[w] + carry;
      if (i < nTiles) tileSum[i] = excl;
      __syncthreads();
      if (t == 1023) carry = excl + v;
      __syncthreads();
   }
   if (t == 0) { cellStart[nCells] = carry; cnt[C_NATOMS] = carry; }
}
__global__ void k_scan_add(int* __restrict__ cellStart, const int* __restrict__ tileSum, int n)
{
   const int i = blockIdx.x * blockDim.x + threadIdx.x;
   if (i < n) cellStart[i] += tileSum[i / SCAN_TILE];
}

// counting-sort scatter: (gid, source index, cell) of every pool atom lands in its cell's
// segment; the order inside the segment is whatever the atomics give and is fixed next.
__global__ void __launch_bounds__(256)
k_scatter(const int* __restrict__ cellKey, const int* __restrict__ gid, const int* __restrict__ cellStart,
          int* __restrict__ cellFill, int* __restrict__ tmpSrc, int* __restrict__ tmpGid, int* __restrict__ tmpCell,
          int* __restrict__ cnt, long cap)
{
   const int n = cnt[C_NSRC];
   for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
   {
      const int c = cellKey[i];
      if (c < 0) continue;
      const long pos = cellStart[c] + atomicAdd(&cellFill[c], 1);
      if (pos >= cap) { atomicOr(&cnt[C_ERR], ERRB_CAPACITY); continue; }
      tmpSrc[pos] = i; tmpGid[pos] = gid[i]; tmpCell[pos] = c;
   }
}

// Canonical order + payload move: inside its cell every atom takes the rank of its gid
// (ascending), which makes the layout -- hence every summation order downstream --
// independent of atomics and identical on sender and receiver (what sortAtomsGpu /
// sortAtomsInCell aim for, gpu_kernels.cu:1014, haloExchange.c:1926-1981).
__global__ void __launch_bounds__(256)
k_rank_gather(DevGeom geo, const int* __restrict__ tmpSrc, const int* __restrict__ tmpGid, const int* __restrict__ tmpCell,
              const int* __restrict__ cellStart, const unsigned char* __restrict__ cellType,
              AtomArrays src, AtomArrays dst, int* __restrict__ cnt)
{
   const int n = cnt[C_NATOMS];
   const int lane = threadIdx.x & 31;
   const int stride = gridDim.x * blockDim.x;
   for (int base = blockIdx.x * blockDim.x + (threadIdx.x & ~31); base < n; base += stride)
   {
      const int t = base + lane;
      bool isLocal = false;
      if (t < n)
      {
         const int c = tmpCell[t], g = tmpGid[t];
         const int b = cellStart[c], e = cellStart[c + 1];
         int rank = 0;
         for (int u = b; u < e; ++u) { const int gu = tmpGid[u]; rank += (gu < g) || (gu == g && u < t); }
         const int d = b + rank, s = tmpSrc[t];
         const real_t x = src.r[0][s], y = src.r[1][s], z = src.r[2][s];
         dst.r[0][d] = x; dst.r[1][d] = y; dst.r[2][d] = z;
         dst.rf[d] = make_float4((float)(x - geo.lMin[0]), (float)(y - geo.lMin[1]), (float)(z - geo.lMin[2]), 0.f);
         dst.rq[d] = make_real4(x, y, z, R(0.0));
         dst.p[0][d] = src.p[0][s]; dst.p[1][d] = src.p[1][s]; dst.p[2][d] = src.p[2][s];
         dst.gid[d] = g; dst.spec[d] = src.spec[s]; dst.cell[d] = c;
         isLocal = cellType[c] & CT_LOCAL;
      }
      const unsigned m = __ballot_sync(0xffffffffu, isLocal);
      if (m && lane == 0) atomicAdd(&cnt[C_NLOCAL], __popc(m));
   }
}

// ---- integrator (timestep.c:143-176).  __dmul_rn/__dadd_rn keep the host's operation
// order (no FMA contraction) so positions, hence cell ids, track the CPU path bit by bit
// for as long as the forces do.
__global__ void __launch_bounds__(256)
k_advance_velocity(real_t dt, const int* __restrict__ cell, const unsigned char* __restrict__ cellType,
                   real_t* __restrict__ px, real_t* __restrict__ py, real_t* __restrict__ pz,
                   const real_t* __restrict__ fx, const real_t* __restrict__ fy, const real_t* __restrict__ fz,
                   const int* __restrict__ cnt)
{
   const int n = cnt[C_NATOMS];
   for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
   {
      if (!(cellType[cell[i]] & CT_LOCAL)) continue;
      px[i] = addRn(px[i], mulRn(dt, fx[i]));
      py[i] = addRn(py[i], mulRn(dt, fy[i]));
      pz[i] = addRn(pz[i], mulRn(dt, fz[i]));
   }
}
__global__ void __launch_bounds__(256)
k_advance_position(real_t dt, real_t invMass, const int* __restrict__ cell, const unsigned char* __restrict__ cellType,
                   real_t* __restrict__ rx, real_t* __restrict__ ry, real_t* __restrict__ rz,
                   const real_t* __restrict__ px, const real_t* __restrict__ py, const real_t* __restrict__ pz,
                   const int* __restrict__ cnt, real4* __restrict__ rq)
{
   const int n = cnt[C_NATOMS];
   for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
   {
      if (!(cellType[cell[i]] & CT_LOCAL)) continue;
      const real_t x = addRn(rx[i], mulRn(mulRn(dt, px[i]), invMass));
      const real_t y = addRn(ry[i], mulRn(mulRn(dt, py[i]), invMass));
      const real_t z = addRn(rz[i], mulRn(mulRn(dt, pz[i]), invMass));
      rx[i] = x; ry[i] = y; rz[i] = z;
      rq[i] = make_real4(x, y, z, R(0.0));
   }
}
// first half of a step in one pass: p += dt/2 f ; r += dt p / m
// First half of a step (timestep.c:62-68).  TWO: the second half-kick of the previous step is applied first --
// same f, two separately rounded additions, i.e. the bits advanceVelocity + advanceVelocity would give.
// CHECK: the neighbour list's rebuild criterion (any local atom further than skin/2 from the snapshot the
// list was built on, neighborList.c:212-234) is evaluated on the new position while it is in registers.
template <bool TWO, bool CHECK>
__global__ void __launch_bounds__(256)
k_kick_drift(real_t dtHalf, real_t dt, real_t invMass, const int* __restrict__ cell, const unsigned char* __restrict__ cellType,
             real_t* __restrict__ rx, real_t* __restrict__ ry, real_t* __restrict__ rz,
             real_t* __restrict__ px, real_t* __restrict__ py, real_t* __restrict__ pz,
             const real_t* __restrict__ fx, const real_t* __restrict__ fy, const real_t* __restrict__ fz,
             int* __restrict__ cnt, real4* __restrict__ rq,
             const real_t* __restrict__ lx, const real_t* __restrict__ ly, const real_t* __restrict__ lz, real_t half2)
{
   const int n = cnt[C_NATOMS];
   bool hit = false;
   for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
   {
      if (!(cellType[cell[i]] & CT_LOCAL)) continue;
      const real_t hx = mulRn(dtHalf, fx[i]), hy = mulRn(dtHalf, fy[i]), hz = mulRn(dtHalf, fz[i]);
      real_t a = px[i], b = py[i], c = pz[i];
      if (TWO) { a = addRn(a, hx); b = addRn(b, hy); c = addRn(c, hz); }
      a = addRn(a, hx); b = addRn(b, hy); c = addRn(c, hz);
      px[i] = a; py[i] = b; pz[i] = c;
      const real_t x = addRn(rx[i], mulRn(mulRn(dt, a), invMass));
      const real_t y = addRn(ry[i], mulRn(mulRn(dt, b), invMass));
      const real_t z = addRn(rz[i], mulRn(mulRn(dt, c), invMass));
      rx[i] = x; ry[i] = y; rz[i] = z;
      rq[i] = make_real4(x, y, z, R(0.0));
      if (CHECK)
      {
         const real_t dx = lx[i] - x, dy = ly[i] - y, dz = lz[i] - z;
         hit |= dx * dx + dy * dy + dz * dz > half2;
      }
   }
   if (CHECK && __any_sync(0xffffffffu, hit) && (threadIdx.x & 31) == 0) cnt[C_NLREBUILD] = 1;   // every writer stores the same value
}

// ---- energy: deterministic two-level reduction (fixed grid, fixed tree); replaces
// ReduceEnergy's atomicAdd(double) (gpu_reduce.h:32-98).
#define RED_BLOCKS 592
__global__ void __launch_bounds__(256)
k_energy_partial(real_t halfInvMass, const int* __restrict__ cell, const unsigned char* __restrict__ cellType,
                 const real_t* __restrict__ e, const real_t* __restrict__ px, const real_t* __restrict__ py,
                 const real_t* __restrict__ pz, const int* __restrict__ cnt, double* __restrict__ partial)
{
   __shared__ double sU[8], sK[8];
   const int n = cnt[C_NATOMS];
   double u = 0.0, k = 0.0;                        // sums in double in both builds
   for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
   {
      if (!(cellType[cell[i]] & CT_LOCAL)) continue;
      u += e[i];
      k += (px[i] * px[i] + py[i] * py[i] + pz[i] * pz[i]) * halfInvMass;
   }
   #pragma unroll
   for (int o = 16; o > 0; o >>= 1) { u += __shfl_down_sync(0xffffffffu, u, o); k += __shfl_down_sync(0xffffffffu, k, o); }
   const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
   if (lane == 0) { sU[w] = u; sK[w] = k; }
   __syncthreads();
   if (threadIdx.x == 0)
   {
      double a = 0.0, b = 0.0;
      for (int i = 0; i < 8; ++i) { a += sU[i]; b += sK[i]; }
      partial[2 * blockIdx.x] = a; partial[2 * blockIdx.x + 1] = b;
   }
}
__global__ void __launch_bounds__(1024)
k_energy_final(const double* __restrict__ partial, int nPart, double* __restrict__ out)
{
   __shared__ double sU[32], sK[32];
   double u = 0.0, k = 0.0;                        // sums in double in both builds
   for (int i = threadIdx.x; i < nPart; i += 1024) { u += partial[2 * i]; k += partial[2 * i + 1]; }
   #pragma unroll
   for (int o = 16; o > 0; o >>= 1) { u += __shfl_down_sync(0xffffffffu, u, o); k += __shfl_down_sync(0xffffffffu, k, o); }
   const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
   if (lane == 0) { sU[w] = u; sK[w] = k; }
   __syncthreads();
   if (threadIdx.x == 0)
   {
      double a = 0.0, b = 0.0;
      for (int i = 0; i < 32; ++i) { a += sU[i]; b += sK[i]; }
      out[0] = a; out[1] = b;
   }
}

// ------------------------------------------------------------------ host side
static int allocAtoms(AtomArrays& a, long cap)
{
   const long capP = cap + COMD_PAD;
   for (int k = 0; k < 3; ++k)
   {
      CUDA_TRY(cudaMalloc(&a.r[k], capP * sizeof(real_t)));
      CUDA_TRY(cudaMalloc(&a.p[k], capP * sizeof(real_t)));
      CUDA_TRY(cudaMemset(a.r[k], 0, capP * sizeof(real_t)));
   }
   CUDA_TRY(cudaMalloc(&a.gid, capP * sizeof(int)));
   CUDA_TRY(cudaMalloc(&a.spec, capP * sizeof(int)));
   CUDA_TRY(cudaMalloc(&a.cell, capP * sizeof(int)));
   CUDA_TRY(cudaMalloc(&a.rf, capP * sizeof(float4)));
   CUDA_TRY(cudaMemset(a.rf, 0, capP * sizeof(float4)));
   CUDA_TRY(cudaMalloc(&a.rq, capP * sizeof(real4)));
   CUDA_TRY(cudaMemset(a.rq, 0, capP * sizeof(real4)));
   return COMD_OK;
}
static void freeAtoms(AtomArrays& a)
{
   for (int k = 0; k < 3; ++k) { cudaFree(a.r[k]); cudaFree(a.p[k]); }
   cudaFree(a.gid); cudaFree(a.spec); cudaFree(a.cell); cudaFree(a.rf); cudaFree(a.rq);
}

extern "C" int comd_dev_create(ComdDev** out, int deviceId, const ComdGeometry* geom, long atomCapacity)
{
   if (!out || !geom || atomCapacity <= 0) { comdSetError("comd_dev_create: bad argument"); return COMD_E_ARG; }
   for (int a = 0; a < 3; ++a)
      if (geom->gridSize[a] < 2) { comdSetError("comd_dev_create: gridSize[%d]=%d < 2 (linkCells.c:181)", a, geom->gridSize[a]); return COMD_E_ARG; }
   int nDev = 0;
   if (cudaGetDeviceCount(&nDev) != cudaSuccess || nDev == 0)
   {
      comdSetError("comd_dev_create: no CUDA device (this library has no CPU fallback)");
      return COMD_E_CUDA;
   }
   if (deviceId < 0 || deviceId >= COMD_MAX_DEVICES) { comdSetError("comd_dev_create: device id %d outside 0..%d", deviceId, COMD_MAX_DEVICES - 1); return COMD_E_ARG; }
   CUDA_TRY(cudaSetDevice(deviceId));
   ComdDev* d = new ComdDev();
   memset(d, 0, sizeof(ComdDev));
   d->device = deviceId;
   d->hostGeom = *geom;
   DevGeom& g = d->g;
   for (int a = 0; a < 3; ++a)
   {
      g.n[a] = geom->gridSize[a]; g.P[a] = g.n[a] + 2;
      g.lMin[a] = geom->localMin[a]; g.lMax[a] = geom->localMax[a]; g.inv[a] = geom->invBoxSize[a];
   }
   g.nCells = g.P[0] * g.P[1] * g.P[2];
   d->cap = atomCapacity;
   d->rank = 0; d->nRanks = 1;
   // neighbours + periodic shifts exactly as initHaloExchange / initAtomHaloExchange
   // (haloExchange.c:756-770, 310-323; processorNum, decomposition.c:57-66)
   for (int f = 0; f < 6; ++f)
   {
      const int ax = f / 2, dir = (f % 2) ? +1 : -1;
      int pc[3] = { geom->procCoord[0], geom->procCoord[1], geom->procCoord[2] };
      pc[ax] = (pc[ax] + dir + geom->procGrid[ax]) % geom->procGrid[ax];
      d->nbrRank[f] = pc[0] + geom->procGrid[0] * (pc[1] + geom->procGrid[1] * pc[2]);
      for (int a = 0; a < 3; ++a) d->shift[f][a] = R(0.0);
      if (dir < 0 && geom->procCoord[ax] == 0) d->shift[f][ax] = +R(1.0) * geom->globalExtent[ax];
      if (dir > 0 && geom->procCoord[ax] == geom->procGrid[ax] - 1) d->shift[f][ax] = -R(1.0) * geom->globalExtent[ax];
   }
   d->rank = geom->procCoord[0] + geom->procGrid[0] * (geom->procCoord[1] + geom->procGrid[1] * geom->procCoord[2]);
   d->nRanks = geom->procGrid[0] * geom->procGrid[1] * geom->procGrid[2];

   CUDA_TRY(cudaStreamCreateWithFlags(&d->stream, cudaStreamNonBlocking));
   for (int k = 0; k < 256; ++k) CUDA_TRY(cudaEventCreate(&d->evForce[k]));
   CUDA_TRY(cudaEventCreate(&d->evT0)); CUDA_TRY(cudaEventCreate(&d->evT1));
   {
      int lo = 0, hi = 0;
      CUDA_TRY(cudaDeviceGetStreamPriorityRange(&lo, &hi));
      CUDA_TRY(cudaStreamCreateWithPriority(&d->commStream, cudaStreamNonBlocking, hi));
      CUDA_TRY(cudaEventCreateWithFlags(&d->evRefresh, cudaEventDisableTiming)); CUDA_TRY(cudaEventCreateWithFlags(&d->evPass1, cudaEventDisableTiming));
      CUDA_TRY(cudaEventCreateWithFlags(&d->evFp, cudaEventDisableTiming)); CUDA_TRY(cudaEventCreateWithFlags(&d->evStepBegin, cudaEventDisableTiming));
      CUDA_TRY(cudaStreamCreateWithFlags(&d->sideStream, cudaStreamNonBlocking));
      CUDA_TRY(cudaEventCreateWithFlags(&d->evSideBegin, cudaEventDisableTiming)); CUDA_TRY(cudaEventCreateWithFlags(&d->evSideDone, cudaEventDisableTiming));
      d->optOverlapMode = 1;
      CUDA_TRY(cudaEventCreateWithFlags(&d->evFlag, cudaEventDisableTiming));
      d->optNlSpeculate = 1;
      d->optTileHits = 1;
      d->optGraph = -1;
      d->xs = d->stream; d->optOverlap = 1; d->phStep = -1;
      d->e2eSpecArmed = 0;             // opt-in (option e2e_speculate): it only pays when the caller hands back bit-identical buffers
   }
   const long cap = d->cap;
   int rc;
   if ((rc = allocAtoms(d->A, cap)) != COMD_OK) return rc;
   if ((rc = allocAtoms(d->B, cap)) != COMD_OK) return rc;
   for (int k = 0; k < 3; ++k) { CUDA_TRY(cudaMalloc(&d->f[k], (cap + COMD_PAD) * sizeof(real_t))); CUDA_TRY(cudaMemset(d->f[k], 0, (cap + COMD_PAD) * sizeof(real_t))); }
   CUDA_TRY(cudaMalloc(&d->e, cap * sizeof(real_t))); CUDA_TRY(cudaMemset(d->e, 0, cap * sizeof(real_t)));
   CUDA_TRY(cudaMalloc(&d->cellStart, (g.nCells + 1) * sizeof(int)));
   CUDA_TRY(cudaMemset(d->cellStart, 0, (g.nCells + 1) * sizeof(int)));
   CUDA_TRY(cudaMalloc(&d->cellCount, g.nCells * sizeof(int)));
   CUDA_TRY(cudaMalloc(&d->cellFill, g.nCells * sizeof(int)));
   CUDA_TRY(cudaMalloc(&d->cellType, g.nCells));
   CUDA_TRY(cudaMalloc(&d->cellKey, cap * sizeof(int)));
   CUDA_TRY(cudaMalloc(&d->shell, cap * sizeof(int)));
   CUDA_TRY(cudaMalloc(&d->tmpSrc, cap * sizeof(int)));
   CUDA_TRY(cudaMalloc(&d->tmpGid, cap * sizeof(int)));
   CUDA_TRY(cudaMalloc(&d->tmpCell, cap * sizeof(int)));
   CUDA_TRY(cudaMalloc(&d->scanBlock, ((g.nCells + SCAN_TILE - 1) / SCAN_TILE + 1) * sizeof(int)));
   CUDA_TRY(cudaMalloc(&d->cnt, C_COUNT * sizeof(int)));
   CUDA_TRY(cudaMemset(d->cnt, 0, C_COUNT * sizeof(int)));
   CUDA_TRY(cudaMallocHost(&d->hCnt, C_COUNT * sizeof(int)));
   CUDA_TRY(cudaMalloc(&d->redPartial, 2 * RED_BLOCKS * sizeof(double)));
   CUDA_TRY(cudaMalloc(&d->redOut, 2 * sizeof(double)));
   CUDA_TRY(cudaMallocHost(&d->hRed, 2 * sizeof(double)));
   CUDA_TRY(cudaMalloc(&d->rhobar, (cap + COMD_PAD) * sizeof(real_t)));
   CUDA_TRY(cudaMalloc(&d->dfEmbed, (cap + COMD_PAD) * sizeof(real_t)));
   CUDA_TRY(cudaMemset(d->dfEmbed, 0, (cap + COMD_PAD) * sizeof(real_t)));

   // message capacity: the two largest planes of cells, generously populated
   // (bufCapacity, haloExchange.c:202-207, sized without MAXATOMS padding)
   long plane = 0;
   for (int a = 0; a < 3; ++a)
   {
      long p = (long)g.P[(a + 1) % 3] * g.P[(a + 2) % 3];
      if (p > plane) plane = p;
   }
   const double perCell = (double)cap / (double)g.nCells;
   d->msgCap = (long)(2 * plane * (perCell * R(1.5) + 8)) + 1024;
   if (d->msgCap > cap) d->msgCap = cap;
   for (int k = 0; k < 2; ++k)
   {
      CUDA_TRY(cudaMalloc(&d->sendBuf[k], d->msgCap * COMD_MSG_REALS * sizeof(real_t)));
      CUDA_TRY(cudaMalloc(&d->recvBuf[k], d->msgCap * COMD_MSG_REALS * sizeof(real_t)));
      CUDA_TRY(cudaMalloc(&d->fpSend[k], d->msgCap * sizeof(real_t)));
      CUDA_TRY(cudaMalloc(&d->fpRecv[k], d->msgCap * sizeof(real_t)));
   }
   CUDA_TRY(cudaMalloc(&d->runOff, (plane + 2) * 24 * sizeof(int)));
   CUDA_TRY(cudaMalloc(&d->nlStat, 2 * sizeof(unsigned long long)));
   CUDA_TRY(cudaMemset(d->nlStat, 0, 2 * sizeof(unsigned long long)));

   // cell classes
   std::vector<unsigned char> ct(g.nCells, 0);
   for (int z = 0; z < g.P[2]; ++z) for (int y = 0; y < g.P[1]; ++y) for (int x = 0; x < g.P[0]; ++x)
   {
      const int q[3] = { x, y, z };
      unsigned char t = 0;
      if (x >= 1 && x <= g.n[0] && y >= 1 && y <= g.n[1] && z >= 1 && z <= g.n[2]) t |= CT_LOCAL;
      for (int a = 0; a < 3; ++a)
      {
         if (q[a] <= 1) t |= CT_FACE(2 * a);
         if (q[a] >= g.P[a] - 2) t |= CT_FACE(2 * a + 1);
      }
      if (t & ~CT_LOCAL) t |= CT_SHELL;
      ct[x + g.P[0] * (y + g.P[1] * z)] = t;
   }
   CUDA_TRY(cudaMemcpy(d->cellType, ct.data(), g.nCells, cudaMemcpyHostToDevice));
   d->approxAtoms = 0;
   d->mass = R(1.0);
   d->optForceVariant = 1;
   d->optNlN3 = -1;
   if (const char* ev = getenv("COMD_NL_N3")) d->optNlN3 = atoi(ev) < 0 ? -1 : atoi(ev) != 0;
   d->optTileTx = 8;
   if (const char* ev = getenv("COMD_FORCE_VARIANT")) d->optForceVariant = atoi(ev);
   if (const char* ev = getenv("COMD_TILE_TX")) d->optTileTx = atoi(ev);
   *out = d;
   return COMD_OK;
}

extern "C" int comd_dev_destroy(ComdDev* d)
{
   if (!d) return COMD_OK;
   cudaSetDevice(d->device);
   cudaStreamSynchronize(d->stream);
   freeAtoms(d->A); freeAtoms(d->B);
   for (int k = 0; k < 3; ++k) { cudaFree(d->f[k]); cudaFree(d->nlLastR[k]); }
   cudaFree(d->e); cudaFree(d->cellStart); cudaFree(d->cellCount); cudaFree(d->cellFill); cudaFree(d->cellType);
   cudaFree(d->cellKey); cudaFree(d->shell); cudaFree(d->tmpSrc); cudaFree(d->tmpGid); cudaFree(d->tmpCell);
   cudaFree(d->scanBlock); cudaFree(d->cnt); cudaFreeHost(d->hCnt); cudaFree(d->redPartial); cudaFree(d->redOut);
   cudaFreeHost(d->hRed); cudaFree(d->rhobar); cudaFree(d->dfEmbed); cudaFree(d->tabMem); cudaFree(d->tPRmem); cudaFree(d->tW[0]); cudaFree(d->tW[1]); cudaFree(d->tDR); cudaFree(d->rvPhi); cudaFree(d->rvRho); cudaFree(d->r3); cudaFree(d->runOff);
   cudaFree(d->nlWork); cudaFree(d->nlList); cudaFree(d->nlCount); cudaFree(d->nlCountHit); cudaFree(d->nlStat);
   for (int k = 0; k < 2; ++k) { cudaFree(d->sendBuf[k]); cudaFree(d->recvBuf[k]); cudaFree(d->fpSend[k]); cudaFree(d->fpRecv[k]); }
   if (d->hStage) cudaFreeHost(d->hStage);
   cudaFree(d->rowTab); cudaFree(d->e2eU); cudaFree(d->e2eUgid); cudaFree(d->e2eFlag);
   for (int k = 0; k < 2; ++k) { cudaFree(d->e2eOut[k]); cudaFree(d->e2eOutGid[k]); }
   if (d->hE2eFlag) cudaFreeHost(d->hE2eFlag);
   for (int k = 0; k < 256; ++k) cudaEventDestroy(d->evForce[k]);
   cudaEventDestroy(d->evT0); cudaEventDestroy(d->evT1);
   if (d->optPhaseTiming && d->phaseSteps > 0 && d->rank == 0)
   {
      static const char* nm[6] = { "kick+drift+check+flag", "ghost refresh / rebuild", "pass 1 (or LJ force)", "F' halo", "pass 3", "kick" };
      fprintf(stderr, "[comd phase timing, rank 0, %ld steps] ", d->phaseSteps);
      const int n = (int)(d->phaseSteps < 1024 ? d->phaseSteps : 1024);
      for (int k = 0; k < 6; ++k)
      {  // median: the first NCCL calls and the rebuild steps are outliers by two orders of magnitude
         std::vector<float> v(d->phaseSample[k], d->phaseSample[k] + n);
         std::sort(v.begin(), v.end());
         fprintf(stderr, "%s median %.3f max %.3f ms | ", nm[k], v[n / 2], v[n - 1]);
      }
      fprintf(stderr, "\n[comd phase timing] last steps, phases in the order above:");
      for (long st = d->phaseSteps > 26 ? d->phaseSteps - 26 : 0; st < d->phaseSteps; ++st)
      {
         fprintf(stderr, "\n   step %3ld:", st);
         for (int k = 0; k < 6; ++k) fprintf(stderr, " %7.3f", d->phaseSample[k][st % 1024]);
      }
      fprintf(stderr, "\n");
   }
   for (int a = 0; a < 32; ++a) for (int b = 0; b < 8; ++b) if (d->evPh[a][b]) cudaEventDestroy(d->evPh[a][b]);
   cudaEventDestroy(d->evRefresh); cudaEventDestroy(d->evPass1); cudaEventDestroy(d->evFp); cudaEventDestroy(d->evStepBegin);
   cudaStreamDestroy(d->commStream);
   if (d->stepGraph) cudaGraphExecDestroy(d->stepGraph);
   for (int ph = 0; ph < 3; ++ph) if (d->nlOrder[ph]) cudaFree(d->nlOrder[ph]);
   cudaEventDestroy(d->evFlag); cudaEventDestroy(d->evSideBegin); cudaEventDestroy(d->evSideDone); cudaStreamDestroy(d->sideStream);
   if (d->copyStream)
   {
      cudaStreamDestroy(d->copyStream); cudaStreamDestroy(d->upStream);
      cudaEventDestroy(d->evCopy); cudaEventDestroy(d->evUpR); cudaEventDestroy(d->evUpP); cudaEventDestroy(d->evUp);
   }
   cudaStreamDestroy(d->stream);
   delete d;
   return COMD_OK;
}

extern "C" int comd_dev_set_mass(ComdDev* d, double m)
{
   if (d) d->stepGraphGen = -1;                     // the captured kick + drift kernel carries 1/m
   if (!d || !(m > R(0.0))) { comdSetError("comd_dev_set_mass: bad argument"); return COMD_E_ARG; }
   d->e2ePristine = 0;
   d->mass = m;
   return COMD_OK;
}

extern "C" int comd_dev_set_lj(ComdDev* d, double sigma, double epsilon, double cutoff)
{
   if (!d || !(cutoff > R(0.0))) { comdSetError("comd_dev_set_lj: bad argument"); return COMD_E_ARG; }
   for (int a = 0; a < 3; ++a)
      if (d->hostGeom.boxSize[a] < cutoff)
      { comdSetError("comd_dev_set_lj: cell size %g < cutoff %g on axis %d", d->hostGeom.boxSize[a], cutoff, a); return COMD_E_ARG; }
   d->potKind = 1; d->ljSigma = sigma; d->ljEps = epsilon; d->cutoff = cutoff;
   d->nlValid = 0; d->eValid = 0;                   // a list built for another cutoff is not this potential's list
   return COMD_OK;
}

extern "C" int comd_dev_set_eam(ComdDev* d, double cutoff, const ComdTable* phi, const ComdTable* rho, const ComdTable* f)
{
   if (!d || !phi || !rho || !f || !(cutoff > R(0.0))) { comdSetError("comd_dev_set_eam: bad argument"); return COMD_E_ARG; }
   for (int a = 0; a < 3; ++a)
      if (d->hostGeom.boxSize[a] < cutoff)
      { comdSetError("comd_dev_set_eam: cell size %g < cutoff %g on axis %d", d->hostGeom.boxSize[a], cutoff, a); return COMD_E_ARG; }
   CUDA_TRY(cudaSetDevice(d->device));
   const ComdTable* t[3] = { phi, rho, f };
   long total = 0;
   for (int k = 0; k < 3; ++k) total += t[k]->n + 3 + 1;
   if (d->tabMem) cudaFree(d->tabMem);
   CUDA_TRY(cudaMalloc(&d->tabMem, total * sizeof(real_t)));
   DevTable* out[3] = { &d->tPhi, &d->tRho, &d->tF };
   long off = 0;
   for (int k = 0; k < 3; ++k)
   {
      const int n = t[k]->n;
      std::vector<real_t> hv((size_t)n + 3);               // the padded table values[-1..n+1] in the device's precision
      for (int i = 0; i < n + 3; ++i) hv[i] = (real_t)t[k]->values[i - 1];
      CUDA_TRY(cudaMemcpy(d->tabMem + off, hv.data(), (n + 3) * sizeof(real_t), cudaMemcpyHostToDevice));
      out[k]->n = n; out[k]->x0 = t[k]->x0; out[k]->invDx = t[k]->invDx; out[k]->v = d->tabMem + off + 1;
      off += n + 3 + ((n + 3) & 1);      // keep every table 16-byte aligned
   }
   // phi and rho on one grid (true for funcfl and setfl files, eam.c:743-757,853-872):
   // interleave them so a pass-1 lookup is four 16-byte loads instead of eight 8-byte ones
   if (d->tPRmem) { cudaFree(d->tPRmem); d->tPRmem = nullptr; }
   d->tPR = nullptr;
   if (phi->n == rho->n && phi->x0 == rho->x0 && phi->invDx == rho->invDx)
   {
      const int n = phi->n;
      std::vector<real_t> h(2 * (size_t)(n + 3));
      for (int i = -1; i < n + 2; ++i) { h[2 * (i + 1)] = phi->values[i]; h[2 * (i + 1) + 1] = rho->values[i]; }
      CUDA_TRY(cudaMalloc(&d->tPRmem, h.size() * sizeof(real_t)));
      CUDA_TRY(cudaMemcpy(d->tPRmem, h.data(), h.size() * sizeof(real_t), cudaMemcpyHostToDevice));
      d->tPR = d->tPRmem + 1;
      // shifted copies: tW[0][k] = entry k-1, tW[1][k] = entry k; entries past n+1 repeat the last one
      for (int c = 0; c < 2; ++c)
      {
         if (d->tW[c]) { cudaFree(d->tW[c]); d->tW[c] = nullptr; }
         const int len = n + 8;
         std::vector<real_t> w(2 * (size_t)len);
         for (int k = 0; k < len; ++k)
         {
            int e = k - 1 + c; if (e > n + 1) e = n + 1;
            w[2 * k] = phi->values[e]; w[2 * k + 1] = rho->values[e];
         }
         CUDA_TRY(cudaMalloc(&d->tW[c], w.size() * sizeof(real_t)));
         CUDA_TRY(cudaMemcpy(d->tW[c], w.data(), w.size() * sizeof(real_t), cudaMemcpyHostToDevice));
      }
   }
   {  // pass-3 records: g1 = T[i+1]-T[i-1], q = (T[i+2]-T[i]) - g1, the two differences interpolate()
      // (eam.c:573-578) forms for the derivative, in the same IEEE operations
      if (d->tDR) { cudaFree(d->tDR); d->tDR = nullptr; }
      const int n = rho->n;
      std::vector<real_t> w(2 * (size_t)(n + 2));
      for (int i = 0; i <= n; ++i)
      {
         const real_t tm = rho->values[i - 1], t0 = rho->values[i], t1 = rho->values[i + 1];
         const real_t t2 = rho->values[i + 2 > n + 1 ? n + 1 : i + 2];
         const real_t g1 = t1 - tm, g2 = t2 - t0;
         w[2 * i] = g1; w[2 * i + 1] = g2 - g1;
      }
      CUDA_TRY(cudaMalloc(&d->tDR, w.size() * sizeof(real_t)));
      CUDA_TRY(cudaMemcpy(d->tDR, w.data(), w.size() * sizeof(real_t), cudaMemcpyHostToDevice));
   }
   {  // list path: per interval i the four numbers interpolate() (eam.c:573-578) forms from T[i-1..i+2], in the
      // same IEEE operations, as one 32-byte record: value = a + 0.5 t (g1 + t c2), slope = 0.5 (g1 + t q) invDx
      const ComdTable* tv[2] = { phi, rho };
      real4** dst[2] = { &d->rvPhi, &d->rvRho };
      std::vector<real_t> w3;
      const bool sameGrid = phi->n == rho->n && phi->x0 == rho->x0 && phi->invDx == rho->invDx;
      if (d->r3) { cudaFree(d->r3); d->r3 = nullptr; }
      if (sameGrid) w3.resize(4 * (size_t)(phi->n + 2));
      for (int k = 0; k < 2; ++k)
      {
         if (*dst[k]) { cudaFree(*dst[k]); *dst[k] = nullptr; }
         const int n = tv[k]->n;
         const double* T = tv[k]->values;
         std::vector<real_t> w(4 * (size_t)(n + 2));
         for (int i = 0; i <= n; ++i)
         {
            const real_t tm = T[i - 1], t0 = T[i], t1 = T[i + 1], t2 = T[i + 2 > n + 1 ? n + 1 : i + 2];
            const real_t g1 = t1 - tm, g2 = t2 - t0;
            w[4 * i] = t0; w[4 * i + 1] = g1; w[4 * i + 2] = t1 + tm - R(2.0) * t0; w[4 * i + 3] = g2 - g1;
            if (sameGrid) { w3[4 * i + 2 * k] = g1; w3[4 * i + 2 * k + 1] = g2 - g1; }
         }
         CUDA_TRY(cudaMalloc(dst[k], w.size() * sizeof(real_t)));
         CUDA_TRY(cudaMemcpy(*dst[k], w.data(), w.size() * sizeof(real_t), cudaMemcpyHostToDevice));
      }
      if (sameGrid)
      {
         CUDA_TRY(cudaMalloc(&d->r3, w3.size() * sizeof(real_t)));
         CUDA_TRY(cudaMemcpy(d->r3, w3.data(), w3.size() * sizeof(real_t), cudaMemcpyHostToDevice));
      }
   }
   d->potKind = 2; d->cutoff = cutoff; d->eValid = 0;
   d->nlValid = 0;
   return COMD_OK;
}

int comdCheckErrors(ComdDev* d, const char* where)
{
   // caller has synchronised and refreshed hCnt
   const int e = d->hCnt[C_ERR];
   if (!e) return COMD_OK;
   if (e & ERRB_CAPACITY) { comdSetError("%s: atom or message capacity exceeded (capacity %ld atoms)", where, d->cap); return COMD_E_CAPACITY; }
   if (e & ERRB_ESCAPE) { comdSetError("%s: an atom moved more than one link cell beyond the local domain in one step", where); return COMD_E_STATE; }
   if (e & ERRB_FPMISMATCH) { comdSetError("%s: F' halo message does not match the ghost population", where); return COMD_E_MISMATCH; }
   // the once-per-pair list kernels (force_n3.cuh) are the only list kernels with these two limits: say how to go on without them
   const char* n3Hint = comdNlN3(d) ? "; forces of this evaluation are invalid -- set option \"nl_n3\" to 0 (full-list kernels, no such limit) and upload the atoms again" : "";
   if (e & ERRB_TILE) { comdSetError("%s: a force tile / work block exceeds its shared-memory capacity (local density more than 1.3x the average)%s", where, n3Hint); return COMD_E_CAPACITY; }
   if (e & ERRB_FIXED) { comdSetError("%s: a pair contribution (> 32 eV/A: atoms closer than ~1.5 A?) or a per-atom sum (> 256 eV/A) left the fixed-point range of the in-block accumulators%s", where, n3Hint); return COMD_E_STATE; }
   if (e & ERRB_NLFULL) { comdSetError("%s: neighbour-list row overflow", where); return COMD_E_CAPACITY; }
   comdSetError("%s: device error bits 0x%x", where, e);
   return COMD_E_STATE;
}

static int syncCounters(ComdDev* d, const char* where)
{
   CUDA_TRY(cudaMemcpyAsync(d->hCnt, d->cnt, C_COUNT * sizeof(int), cudaMemcpyDeviceToHost, d->stream));
   CUDA_TRY(cudaStreamSynchronize(d->stream));
   d->approxAtoms = d->hCnt[C_NATOMS]; d->approxHasGhosts = 1;
   return comdCheckErrors(d, where);
}

extern "C" int comd_dev_sync(ComdDev* d)
{
   if (!d) return COMD_E_ARG;
   CUDA_TRY(cudaSetDevice(d->device));
   return syncCounters(d, "comd_dev_sync");
}

extern "C" int comd_dev_redistribute(ComdDev* d)
{
   if (!d) return COMD_E_ARG;
   d->e2ePristine = 0;
   CUDA_TRY(cudaSetDevice(d->device));
   const DevGeom& g = d->g;
   cudaStream_t s = d->stream;
   if (!(d->useNL && d->nlValid)) PH_MARK(d, 1);
   if (d->useNL && d->nlValid)
   {  // redistributeAtomsGpuNL (timestep.c:278-354): re-bin only when the list must be rebuilt,
      // otherwise refresh the ghost positions in place
      int didRefresh = 0;
      int rc = comdNeighborRefreshOrRebuild(d, &didRefresh);
      if (rc != COMD_OK || didRefresh) return rc;
   }
   NVTX_RANGE("redistribute: re-bin + atom halo + sort");
   const int nGuess = d->approxAtoms > 0 ? d->approxAtoms : (int)d->cap;
   CUDA_TRY(cudaMemsetAsync(d->cellCount, 0, g.nCells * sizeof(int), s));
   CUDA_TRY(cudaMemsetAsync(d->cellFill, 0, g.nCells * sizeof(int), s));
   k_redist_begin<<<1, 1, 0, s>>>(d->cnt); LAUNCH_CHECK(d);
   AtomArrays& A = d->A;
   k_rebin<<<gridFor(nGuess, 256), 256, 0, s>>>(g, A.r[0], A.r[1], A.r[2], A.cell, d->cellType, d->cellKey,
                                                 d->cellCount, d->shell, d->cnt); LAUNCH_CHECK(d);
   const int nShellGuess = nGuess;   // grid-stride: any grid is correct
   for (int ax = 0; ax < 3; ++ax)
   {
      const int fM = 2 * ax, fP = fM + 1;
      NVTX_RANGE(ax == 0 ? "atom halo x: pack / exchange / unpack" : (ax == 1 ? "atom halo y: pack / exchange / unpack" : "atom halo z: pack / exchange / unpack"));
      k_pack_atoms<<<gridFor(nShellGuess, 256, 148 * 4), 256, 0, s>>>(ax, A.r[0], A.r[1], A.r[2], A.p[0], A.p[1], A.p[2],
         A.gid, A.spec, d->cellKey, d->cellType, d->shell, d->cnt, d->sendBuf[0], d->sendBuf[1], d->msgCap,
         d->shift[fM][0], d->shift[fM][1], d->shift[fM][2], d->shift[fP][0], d->shift[fP][1], d->shift[fP][2]); LAUNCH_CHECK(d);
      int rc = comdExchangeAtomsPhase(d, ax);
      if (rc != COMD_OK) return rc;
      // unload order: the message that came across the - face first (haloExchange.c:1518-1519)
      for (int k = 0; k < 2; ++k)
      {
         k_unpack_atoms<<<gridFor(nShellGuess / 8 + 1, 256, 148 * 2), 256, 0, s>>>(g, d->recvBuf[k], d->cnt + C_RECV0 + fM + k,
            A.r[0], A.r[1], A.r[2], A.p[0], A.p[1], A.p[2], A.gid, A.spec, d->cellKey, d->cellCount, d->shell, d->cnt, d->cap);
         LAUNCH_CHECK(d);
      }
   }
   const int nTiles = (g.nCells + SCAN_TILE - 1) / SCAN_TILE;
   k_scan_tiles<<<nTiles, 1024, 0, s>>>(d->cellCount, d->cellStart, d->scanBlock, g.nCells); LAUNCH_CHECK(d);
   k_scan_sums<<<1, 1024, 0, s>>>(d->scanBlock, nTiles, d->cellStart, g.nCells, d->cnt); LAUNCH_CHECK(d);
   k_scan_add<<<(g.nCells + 255) / 256, 256, 0, s>>>(d->cellStart, d->scanBlock, g.nCells); LAUNCH_CHECK(d);
   k_scatter<<<gridFor(nGuess, 256), 256, 0, s>>>(d->cellKey, A.gid, d->cellStart, d->cellFill, d->tmpSrc, d->tmpGid,
                                                   d->tmpCell, d->cnt, d->cap); LAUNCH_CHECK(d);
   k_rank_gather<<<gridFor(nGuess, 256), 256, 0, s>>>(g, d->tmpSrc, d->tmpGid, d->tmpCell, d->cellStart, d->cellType,
                                                       d->A, d->B, d->cnt); LAUNCH_CHECK(d);
   AtomArrays t = d->A; d->A = d->B; d->B = t;
   d->nlValid = 0;
   d->runLayoutValid = 0;
   d->e2eOrderValid = 0;
   d->posDirty = 0;
   if (d->useNL) return comdNeighborAfterSort(d);
   return COMD_OK;
}

extern "C" int comd_dev_upload_atoms(ComdDev* d, int n, const int* gid, const int* species,
                                     const real_t* rx, const real_t* ry, const real_t* rz,
                                     const real_t* px, const real_t* py, const real_t* pz)
{
   if (!d || n < 0 || !gid || !rx || !ry || !rz) { comdSetError("comd_dev_upload_atoms: bad argument"); return COMD_E_ARG; }
   if (n > d->cap) { comdSetError("comd_dev_upload_atoms: %d atoms > capacity %ld", n, d->cap); return COMD_E_CAPACITY; }
   d->e2ePristine = 0;
   CUDA_TRY(cudaSetDevice(d->device));
   cudaStream_t s = d->stream;
   AtomArrays& A = d->A;
   const real_t* hr[3] = { rx, ry, rz };
   const real_t* hp[3] = { px, py, pz };
   for (int k = 0; k < 3; ++k)
   {
      CUDA_TRY(cudaMemcpyAsync(A.r[k], hr[k], n * sizeof(real_t), cudaMemcpyHostToDevice, s));
      if (hp[k]) CUDA_TRY(cudaMemcpyAsync(A.p[k], hp[k], n * sizeof(real_t), cudaMemcpyHostToDevice, s));
      else CUDA_TRY(cudaMemsetAsync(A.p[k], 0, n * sizeof(real_t), s));
      CUDA_TRY(cudaMemsetAsync(d->f[k], 0, d->cap * sizeof(real_t), s));
   }
   CUDA_TRY(cudaMemsetAsync(d->e, 0, d->cap * sizeof(real_t), s));
   CUDA_TRY(cudaMemcpyAsync(A.gid, gid, n * sizeof(int), cudaMemcpyHostToDevice, s));
   if (species) CUDA_TRY(cudaMemcpyAsync(A.spec, species, n * sizeof(int), cudaMemcpyHostToDevice, s));
   else CUDA_TRY(cudaMemsetAsync(A.spec, 0, n * sizeof(int), s));
   // mark everything "was local" so the re-bin keeps it: any local cell id will do
   k_set_int<<<gridFor(n, 256), 256, 0, s>>>(A.cell, cellIndex(d->g, 0, 0, 0), n); LAUNCH_CHECK(d);
   k_set_int<<<1, 1, 0, s>>>(d->cnt + C_NATOMS, n, 1); LAUNCH_CHECK(d);
   k_set_int<<<1, 1, 0, s>>>(d->cnt + C_ERR, 0, 1); LAUNCH_CHECK(d);
   d->approxAtoms = n; d->approxHasGhosts = 0;
   d->nlValid = 0; d->e2eOrderValid = 0;            // new atoms: never judge them against the old list's snapshot
   int rc = comd_dev_redistribute(d);
   if (rc != COMD_OK) return rc;
   return syncCounters(d, "comd_dev_upload_atoms");
}

extern "C" int comd_dev_num_local(ComdDev* d, int* nLocal)
{
   if (!d || !nLocal) return COMD_E_ARG;
   CUDA_TRY(cudaSetDevice(d->device));
   int rc = syncCounters(d, "comd_dev_num_local");
   *nLocal = d->hCnt[C_NLOCAL];
   return rc;
}
extern "C" int comd_dev_num_total(ComdDev* d, int* n)
{
   if (!d || !n) return COMD_E_ARG;
   CUDA_TRY(cudaSetDevice(d->device));
   int rc = syncCounters(d, "comd_dev_num_total");
   *n = d->hCnt[C_NATOMS];
   return rc;
}

static int downloadCommon(ComdDev* d, bool localOnly, int* nOut, int capacity, int* gid, int* species, int* refBox,
                          real_t* r[3], real_t* p[3], real_t* f[3], real_t* U)
{
   CUDA_TRY(cudaSetDevice(d->device));
   int rc = syncCounters(d, "comd_dev_download");
   if (rc != COMD_OK) return rc;
   const int n = d->hCnt[C_NATOMS];
   std::vector<int> cell(n), tmpI(n);
   std::vector<unsigned char> ct(d->g.nCells);
   std::vector<real_t> tmpD(n);
   CUDA_TRY(cudaMemcpy(cell.data(), d->A.cell, n * sizeof(int), cudaMemcpyDeviceToHost));
   CUDA_TRY(cudaMemcpy(ct.data(), d->cellType, d->g.nCells, cudaMemcpyDeviceToHost));
   std::vector<int> keep; keep.reserve(n);
   for (int i = 0; i < n; ++i) if (!localOnly || (ct[cell[i]] & CT_LOCAL)) keep.push_back(i);
   const int m = (int)keep.size();
   if (nOut) *nOut = m;
   if (m > capacity) { comdSetError("comd_dev_download: %d atoms > caller capacity %d", m, capacity); return COMD_E_CAPACITY; }
   if (refBox) for (int k = 0; k < m; ++k) refBox[k] = refBoxOfCell(d->g, cell[keep[k]]);
   auto pullI = [&](const int* src, int* dst) -> int {
      if (!dst) return COMD_OK;
      CUDA_TRY(cudaMemcpy(tmpI.data(), src, n * sizeof(int), cudaMemcpyDeviceToHost));
      for (int k = 0; k < m; ++k) dst[k] = tmpI[keep[k]];
      return COMD_OK; };
   auto pullD = [&](const real_t* src, real_t* dst) -> int {
      if (!dst) return COMD_OK;
      CUDA_TRY(cudaMemcpy(tmpD.data(), src, n * sizeof(real_t), cudaMemcpyDeviceToHost));
      for (int k = 0; k < m; ++k) dst[k] = tmpD[keep[k]];
      return COMD_OK; };
   if ((rc = pullI(d->A.gid, gid))) return rc;
   if ((rc = pullI(d->A.spec, species))) return rc;
   for (int k = 0; k < 3; ++k)
   {
      if (r && (rc = pullD(d->A.r[k], r[k]))) return rc;
      if (p && (rc = pullD(d->A.p[k], p[k]))) return rc;
      if (f && (rc = pullD(d->f[k], f[k]))) return rc;
   }
   if ((rc = pullD(d->e, U))) return rc;
   return COMD_OK;
}

extern "C" int comd_dev_download_atoms(ComdDev* d, int* nLocalOut, int capacity, int* gid, int* species, int* refBox,
                                       real_t* rx, real_t* ry, real_t* rz, real_t* px, real_t* py, real_t* pz,
                                       real_t* fx, real_t* fy, real_t* fz, real_t* U)
{
   if (!d) return COMD_E_ARG;
   real_t* r[3] = { rx, ry, rz }; real_t* p[3] = { px, py, pz }; real_t* f[3] = { fx, fy, fz };
   return downloadCommon(d, true, nLocalOut, capacity, gid, species, refBox, r, p, f, U);
}

extern "C" int comd_dev_download_all(ComdDev* d, int* nOut, int capacity, int* gid, int* refBox,
                                     real_t* rx, real_t* ry, real_t* rz)
{
   if (!d) return COMD_E_ARG;
   real_t* r[3] = { rx, ry, rz };
   return downloadCommon(d, false, nOut, capacity, gid, nullptr, refBox, r, nullptr, nullptr, nullptr);
}

extern "C" int comd_dev_cell_counts(ComdDev* d, int* nAtomsPerRefBox, int nTotalBoxes)
{
   if (!d || !nAtomsPerRefBox) return COMD_E_ARG;
   const DevGeom& g = d->g;
   const int expect = g.n[0] * g.n[1] * g.n[2] + 2 * ((g.n[0] + 2) * (g.n[1] + g.n[2] + 2) + g.n[1] * g.n[2]);
   if (nTotalBoxes != expect) { comdSetError("comd_dev_cell_counts: nTotalBoxes %d != %d", nTotalBoxes, expect); return COMD_E_ARG; }
   CUDA_TRY(cudaSetDevice(d->device));
   int rc = syncCounters(d, "comd_dev_cell_counts");
   if (rc != COMD_OK) return rc;
   std::vector<int> cs(g.nCells + 1);
   CUDA_TRY(cudaMemcpy(cs.data(), d->cellStart, (g.nCells + 1) * sizeof(int), cudaMemcpyDeviceToHost));
   for (int c = 0; c < g.nCells; ++c) nAtomsPerRefBox[refBoxOfCell(g, c)] = cs[c + 1] - cs[c];
   return COMD_OK;
}

extern "C" int comd_dev_halo_bytes(ComdDev* d, int kind, int face, long* sentBytes, long* receivedBytes)
{
   if (!d || face < 0 || face > 5 || kind < 0 || kind > 1) return COMD_E_ARG;
   CUDA_TRY(cudaSetDevice(d->device));
   int rc = syncCounters(d, "comd_dev_halo_bytes");
   const long per = kind == 0 ? COMD_MSG_REALS * (long)sizeof(real_t) : (long)sizeof(real_t);
   if (sentBytes) *sentBytes = per * d->hCnt[(kind == 0 ? C_SEND0 : C_FPSEND0) + face];
   if (receivedBytes) *receivedBytes = per * d->hCnt[(kind == 0 ? C_RECV0 : C_FPGOT0) + face];
   return rc;
}

extern "C" int comd_dev_advance_velocity(ComdDev* d, double dt)
{
   if (!d) return COMD_E_ARG;
   d->e2ePristine = 0;
   CUDA_TRY(cudaSetDevice(d->device));
   k_advance_velocity<<<gridFor(d->approxAtoms, 256), 256, 0, d->stream>>>(dt, d->A.cell, d->cellType, d->A.p[0], d->A.p[1], d->A.p[2],
      d->f[0], d->f[1], d->f[2], d->cnt); LAUNCH_CHECK(d);
   return COMD_OK;
}

extern "C" int comd_dev_advance_position(ComdDev* d, double dt)
{
   if (!d) return COMD_E_ARG;
   d->e2ePristine = 0;
   CUDA_TRY(cudaSetDevice(d->device));
   d->checkFused = 0;
   k_advance_position<<<gridFor(d->approxAtoms, 256), 256, 0, d->stream>>>(dt, R(1.0) / d->mass, d->A.cell, d->cellType,
      d->A.r[0], d->A.r[1], d->A.r[2], d->A.p[0], d->A.p[1], d->A.p[2], d->cnt, d->A.rq); LAUNCH_CHECK(d);
   d->posDirty = 1;
   return COMD_OK;
}

extern "C" int comd_dev_energy(ComdDev* d, double eLocal[2])
{
   if (!d || !eLocal) return COMD_E_ARG;
   if (d->potKind != 0 && !d->eValid) { comdSetError("comd_dev_energy: no force evaluation with energies since the atoms last moved"); return COMD_E_STATE; }
   CUDA_TRY(cudaSetDevice(d->device));
   k_energy_partial<<<RED_BLOCKS, 256, 0, d->stream>>>(R(0.5) / d->mass, d->A.cell, d->cellType, d->e, d->A.p[0], d->A.p[1], d->A.p[2],
      d->cnt, d->redPartial); LAUNCH_CHECK(d);
   k_energy_final<<<1, 1024, 0, d->stream>>>(d->redPartial, RED_BLOCKS, d->redOut); LAUNCH_CHECK(d);
   CUDA_TRY(cudaMemcpyAsync(d->hRed, d->redOut, 2 * sizeof(double), cudaMemcpyDeviceToHost, d->stream));
   int rc = syncCounters(d, "comd_dev_energy");
   eLocal[0] = d->hRed[0]; eLocal[1] = d->hRed[1];
   return rc;
}

// One force evaluation.  wantEnergy = 0 (inner steps of comd_dev_timestep, list path, EAM) skips the
// per-atom pair energies nobody can observe; the forces are the same bits either way (force_nl.cuh).
static int computeForce(ComdDev* d, int wantEnergy)
{
   if (d->potKind == 0) { comdSetError("comd_dev_compute_force: no potential set"); return COMD_E_STATE; }
   CUDA_TRY(cudaSetDevice(d->device));
   if (d->useNL) { int rc = comdNeighborBuildIfNeeded(d); if (rc != COMD_OK) return rc; }
   if (d->posDirty && !d->useNL && d->optForceVariant == 1) { int rc = comdRefreshShadow(d); if (rc != COMD_OK) return rc; }
   NVTX_RANGE("computeForce");
   const int slot = d->nForceEv % 128;
   if (!d->capturing) CUDA_TRY(cudaEventRecord(d->evForce[2 * slot], d->stream));     // (a replayed graph step carries no force timing)
   d->wantEnergy = wantEnergy;
   int rc = d->potKind == 1 ? comdForceLJ(d) : comdForceEAM(d);      // both set d->eValid
   if (rc != COMD_OK) return rc;
   if (d->potKind == 1 || !d->useNL) { PH_MARK(d, 3); PH_MARK(d, 4); }
   if (!d->capturing) { CUDA_TRY(cudaEventRecord(d->evForce[2 * slot + 1], d->stream)); d->nForceEv++; }
   return COMD_OK;
}

extern "C" int comd_dev_compute_force(ComdDev* d)
{
   if (!d) return COMD_E_ARG;
   d->e2ePristine = 0;
   return computeForce(d, 1);
}

// FP32 shadow positions after a drift that was not followed by a redistribute
__global__ void __launch_bounds__(256)
k_refresh_shadow(DevGeom g, const real_t* __restrict__ rx, const real_t* __restrict__ ry, const real_t* __restrict__ rz,
                 float4* __restrict__ rf, const int* __restrict__ cnt)
{
   const int n = cnt[C_NATOMS];
   for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
      rf[i] = make_float4((float)(rx[i] - g.lMin[0]), (float)(ry[i] - g.lMin[1]), (float)(rz[i] - g.lMin[2]), 0.f);
}
int comdRefreshShadow(ComdDev* d)
{
   k_refresh_shadow<<<gridFor(d->approxAtoms, 256), 256, 0, d->stream>>>(d->g, d->A.r[0], d->A.r[1], d->A.r[2], d->A.rf, d->cnt);
   LAUNCH_CHECK(d);
   d->posDirty = 0;
   return COMD_OK;
}

extern "C" int comd_dev_set_option(ComdDev* d, const char* key, long value)
{
   if (!d || !key) return COMD_E_ARG;
   d->stepGraphGen = -1;                            // a captured step holds launch shapes and arguments: capture again
   if (!strcmp(key, "force_variant")) { if (value < 0 || value > 1) { comdSetError("force_variant: 0 or 1"); return COMD_E_ARG; } d->optForceVariant = (int)value; return COMD_OK; }
   if (!strcmp(key, "nl_chunks_per_cta")) { if (value < 0 || value > 64) { comdSetError("nl_chunks_per_cta: 0..64"); return COMD_E_ARG; } d->optNlChunksPerCta = (int)value; return COMD_OK; }
   if (!strcmp(key, "nl_lanes")) { if (value != 4 && value != 8) { comdSetError("nl_lanes: 4 or 8"); return COMD_E_ARG; } d->optNlLanes = (int)value; d->nlValid = 0; return COMD_OK; }
   if (!strcmp(key, "nl_n3")) { d->optNlN3 = value < 0 ? -1 : value != 0; d->nlValid = 0; return COMD_OK; }
   if (!strcmp(key, "overlap")) { d->optOverlap = value != 0; return COMD_OK; }
   if (!strcmp(key, "graph_step")) { d->optGraph = value < 0 ? -1 : value != 0; return COMD_OK; }
   if (!strcmp(key, "nl_hilbert")) { d->optNlHilbert = value != 0; return COMD_OK; }
   if (!strcmp(key, "nl_speculate")) { d->optNlSpeculate = value != 0; return COMD_OK; }
   if (!strcmp(key, "overlap_mode")) { if (value < 0 || value > 1) { comdSetError("overlap_mode: 0 or 1"); return COMD_E_ARG; } d->optOverlapMode = (int)value; return COMD_OK; }
   if (!strcmp(key, "e2e_speculate")) { d->e2eSpecArmed = value != 0; return COMD_OK; }
   if (!strcmp(key, "overlap_sms")) { if (value < 0 || value > 148) { comdSetError("overlap_sms: 0..148"); return COMD_E_ARG; } d->optOverlapSms = (int)value; return COMD_OK; }
   if (!strcmp(key, "phase_timing"))
   {
      d->optPhaseTiming = value != 0;
      if (value) for (int a = 0; a < 32; ++a) for (int b = 0; b < 8; ++b) if (!d->evPh[a][b]) CUDA_TRY(cudaEventCreate(&d->evPh[a][b]));
      return COMD_OK;
   }
   if (!strcmp(key, "energy_every_step")) { d->optEnergyEveryStep = value != 0; return COMD_OK; }
   if (!strcmp(key, "nl_threads")) { if (value < 32 || value > 768 || value % 32) { comdSetError("nl_threads: multiple of 32 in 32..768"); return COMD_E_ARG; } d->optNlThreads = (int)value; return COMD_OK; }
   if (!strcmp(key, "nl_block")) { const int bx = (int)(value / 10000), by = (int)(value / 100 % 100), bz = (int)(value % 100);
      if (value != 0 && (bx < 1 || by < 1 || bz < 1 || by * bz > 32)) { comdSetError("nl_block: bx*10000 + by*100 + bz with by*bz <= 32, or 0 for auto"); return COMD_E_ARG; }
      d->optNlBlock[0] = bx; d->optNlBlock[1] = by; d->optNlBlock[2] = bz;
      if (comdNlN3(d)) d->nlValid = 0;              // in-block pairs are stored once: the list is tied to the block shape
      return COMD_OK; }
   if (!strcmp(key, "tile_hits")) { d->optTileHits = value != 0; return COMD_OK; }
   if (!strcmp(key, "tile_tx")) { if (value < 1 || value > 16) { comdSetError("tile_tx: 1..16"); return COMD_E_ARG; } d->optTileTx = (int)value; return COMD_OK; }
   comdSetError("comd_dev_set_option: unknown key '%s'", key);
   return COMD_E_ARG;
}

// first half of a step: p += dt/2 f ; r += dt p/m (timestep.c:62-68), gather copy refreshed
static int launchKickDrift(ComdDev* d, double dt, bool two)
{
   NVTX_RANGE("kick + drift");
   const bool check = d->useNL && d->nlValid;
   const real_t h = R(0.5) * d->skin;
   const int grid = gridFor(d->approxAtoms, 256);
#define KD_ARGS R(0.5) * dt, dt, R(1.0) / d->mass, d->A.cell, d->cellType, d->A.r[0], d->A.r[1], d->A.r[2], d->A.p[0], d->A.p[1], d->A.p[2], \
   d->f[0], d->f[1], d->f[2], d->cnt, d->A.rq, d->nlLastR[0], d->nlLastR[1], d->nlLastR[2], h * h
   if (check) CUDA_TRY(cudaMemsetAsync(d->cnt + C_NLREBUILD, 0, sizeof(int), d->stream));
   if (two && check) k_kick_drift<true, true><<<grid, 256, 0, d->stream>>>(KD_ARGS);
   else if (two) k_kick_drift<true, false><<<grid, 256, 0, d->stream>>>(KD_ARGS);
   else if (check) k_kick_drift<false, true><<<grid, 256, 0, d->stream>>>(KD_ARGS);
   else k_kick_drift<false, false><<<grid, 256, 0, d->stream>>>(KD_ARGS);
#undef KD_ARGS
   LAUNCH_CHECK(d);
   d->posDirty = 1;
   d->checkFused = check;                          // comdNeighborRefreshOrRebuild finds the flag already evaluated
   return COMD_OK;
}

// ---- CUDA-graph replay of the steady-state step (list path) -----------------------------------------------------
// Between two list builds every step without energies enqueues the same kernels with the same arguments: fused
// kick + kick + drift + displacement check, flag reduction and copy, ghost refresh, pass 1, F' halo, pass 3.  That
// sequence is captured once per list build (stream capture of the very code the eager step runs) and replayed with one
// cudaGraphLaunch; a later build updates the executable graph in place (same topology, new pointers and grids).  The
// sync-free protocol is unchanged: the flag's copy is followed by an external event record, the host looks at it after
// the launch, a stale list makes the captured force kernels return at once and the step is redone eagerly behind the
// rebuild.  Where the host enqueue time of a step exceeds its GPU time (small systems) this is the difference between a
// launch-bound and a GPU-bound step; large systems run eagerly (the host is a step ahead anyway) and keep the per-step
// force timing events.  Reference precedent: the -a 1 step that never returns to the host (timestep.c:257-265).
static bool graphWanted(const ComdDev* d)
{
   if (d->optGraph == 0 || d->optPhaseTiming) return false;
   if (d->optGraph < 0 && (d->nRanks > 1 || d->approxAtoms > 400000)) return false;
   if (d->nRanks > 1 && (d->fpCntValid & 7) != 7) return false;       // first exchange after a build still sizes its messages through the host
   return true;
}


static int captureStepGraph(ComdDev* d, double dt)
{
   const long launches0 = d->nLaunches, refreshes0 = d->nlRefreshes;
   const int posDirty0 = d->posDirty, checkFused0 = d->checkFused, refresh0 = d->refreshInFlight, eValid0 = d->eValid, wantE0 = d->wantEnergy;
   CUDA_TRY(cudaStreamBeginCapture(d->stream, cudaStreamCaptureModeThreadLocal));
   d->capturing = 1; d->specStep = 1;
   int rc = launchKickDrift(d, dt, true);
   if (rc == COMD_OK) rc = comd_dev_redistribute(d);
   if (rc == COMD_OK) rc = computeForce(d, 0);
   d->capturing = 0; d->specStep = 0;
   cudaGraph_t g = nullptr;
   const cudaError_t ce = cudaStreamEndCapture(d->stream, &g);
   d->stepGraphLaunches = d->nLaunches - launches0; d->stepGraphRefreshes = d->nlRefreshes - refreshes0;
   d->nLaunches = launches0; d->nlRefreshes = refreshes0;              // nothing ran yet: the replays count
   d->posDirty = posDirty0; d->checkFused = checkFused0; d->refreshInFlight = refresh0; d->eValid = eValid0; d->wantEnergy = wantE0;
   if (rc != COMD_OK || ce != cudaSuccess || !g)
   {
      if (g) cudaGraphDestroy(g);
      cudaGetLastError();
      if (rc == COMD_OK) { comdSetError("step graph: capture failed (%s)", cudaGetErrorString(ce)); rc = COMD_E_CUDA; }
      return rc;
   }
   bool ok = false;
   if (d->stepGraph)
   {
      cudaGraphExecUpdateResultInfo info;
      ok = cudaGraphExecUpdate(d->stepGraph, g, &info) == cudaSuccess;
      if (!ok) { cudaGetLastError(); cudaGraphExecDestroy(d->stepGraph); d->stepGraph = nullptr; }
   }
   if (!ok)
   {
      const cudaError_t ie = cudaGraphInstantiate(&d->stepGraph, g, 0);
      if (ie != cudaSuccess) { cudaGraphDestroy(g); d->stepGraph = nullptr; comdSetError("step graph: instantiate failed (%s)", cudaGetErrorString(ie)); return COMD_E_CUDA; }
   }
   cudaGraphDestroy(g);
   d->stepGraphGen = d->nlBuilds; d->stepGraphDt = dt;
   d->stepGraphOpts = d->optOverlap * 2 + d->optOverlapMode;
   return COMD_OK;
}

extern "C" int comd_dev_timestep(ComdDev* d, int nSteps, double dt)
{
   if (!d || nSteps < 0) return COMD_E_ARG;
   if (d->potKind == 0) { comdSetError("comd_dev_timestep: no potential set"); return COMD_E_STATE; }
   d->e2ePristine = 0;
   CUDA_TRY(cudaSetDevice(d->device));
   for (int it = 0; it < nSteps; ++it)
   {
      d->phStep = it; d->inTimestep = 1;
      PH_MARK(d, 0);
      const int wantE = it == nSteps - 1 || d->optEnergyEveryStep;
      int rc = COMD_OK;
      int spec = 0;
      bool replay = false;
      if (it > 0 && !wantE && d->useNL && d->nlValid && d->optNlSpeculate && graphWanted(d))
      {  // steady-state step: one graph launch (captured once per list build)
         if (!d->stepGraph || d->stepGraphGen != d->nlBuilds || d->stepGraphDt != dt || d->stepGraphOpts != d->optOverlap * 2 + d->optOverlapMode)
         {
            rc = captureStepGraph(d, dt);
            if (rc != COMD_OK) { d->optGraph = 0; cudaGetLastError(); }       // fall back to eager enqueueing for good
         }
         replay = rc == COMD_OK;
      }
      if (replay)
      {
         CUDA_TRY(cudaGraphLaunch(d->stepGraph, d->stream));
         d->nLaunches += d->stepGraphLaunches; d->nlRefreshes += d->stepGraphRefreshes; d->nGraphSteps++;
         d->posDirty = 1; d->checkFused = 0; d->refreshInFlight = 0; d->eValid = 0; d->wantEnergy = 0;
         spec = 1;
      }
      else
      {
         rc = launchKickDrift(d, dt, it > 0);     // it > 0: the previous step left its second half-kick to this kernel
         if (rc != COMD_OK) { d->inTimestep = 0; return rc; }
         // List mode with a valid list: the step is enqueued for the common case (no rebuild) without waiting for the
         // verdict of the displacement check; the host reads the flag once the force kernels are queued, i.e. while the
         // GPU is busy with them.  Where the reference needed its -a 1 two-stream step to stay ahead of the GPU
         // (timestep.c:257-265, eam.c:201-257).
         d->specStep = d->useNL && d->nlValid && d->optNlSpeculate;
         spec = d->specStep;
         rc = comd_dev_redistribute(d);            // marks 1 (decision known / flag on its way) and 2 (ghosts in place / enqueued)
         if (rc != COMD_OK) { d->inTimestep = 0; d->specStep = 0; return rc; }
         PH_MARK(d, 2);
         rc = computeForce(d, wantE);              // marks 3, 4
         d->specStep = 0;
         if (rc != COMD_OK) { d->inTimestep = 0; return rc; }
      }
      if (spec)
      {
         CUDA_TRY(cudaEventSynchronize(d->evFlag));
         if (d->hCnt[C_NLREBUILD])
         {  // some atom on some rank moved more than skin/2: the kernels above returned at once.  Now the real thing:
            // re-bin + atom exchange + sort + list build (redistributeAtomsGpuNL's rebuild branch, timestep.c:296-351)
            d->nlRefreshes--; d->nlSpecRedone++;
            if (!replay) d->nForceEv--;            // (a replayed step carries no timing events)
            d->nlValid = 0; d->refreshInFlight = 0;
            rc = comd_dev_redistribute(d);
            if (rc != COMD_OK) { d->inTimestep = 0; return rc; }
            rc = computeForce(d, wantE);
            if (rc != COMD_OK) { d->inTimestep = 0; return rc; }
         }
      }
      PH_MARK(d, 5);
      if (it == nSteps - 1)
      {
         rc = comd_dev_advance_velocity(d, R(0.5) * dt);
         if (rc != COMD_OK) { d->inTimestep = 0; return rc; }
      }
      PH_MARK(d, 6);
   }
   d->phStep = -1; d->inTimestep = 0;
   if (d->optPhaseTiming)
   {
      CUDA_TRY(cudaStreamSynchronize(d->stream));
      for (int it = 0; it < nSteps && it < 32; ++it)
      {
         for (int k = 0; k < 6; ++k)
         {
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, d->evPh[it][k], d->evPh[it][k + 1]) != cudaSuccess) ms = 0.f;
            d->phaseSample[k][d->phaseSteps % 1024] = ms;
         }
         d->phaseSteps++;
      }
      cudaGetLastError();
   }
   return COMD_OK;
}

extern "C" int comd_dev_step_stats(ComdDev* d, long* nGraphSteps, long* nRedoneSteps)
{
   if (!d) return COMD_E_ARG;
   if (nGraphSteps) *nGraphSteps = d->nGraphSteps;
   if (nRedoneSteps) *nRedoneSteps = d->nlSpecRedone;
   return COMD_OK;
}

extern "C" int comd_dev_kernel_stats(ComdDev* d, double* meanForceMs, long* nLaunches)
{
   if (!d) return COMD_E_ARG;
   CUDA_TRY(cudaSetDevice(d->device));
   if (meanForceMs)
   {
      // mean CUDA-event duration of the force calls since the previous query (ring of 128)
      CUDA_TRY(cudaStreamSynchronize(d->stream));
      const int n = d->nForceEv < 128 ? d->nForceEv : 128;
      double sum = 0.0;
      for (int k = 0; k < n; ++k)
      {
         float ms = 0.f;
         if (cudaEventElapsedTime(&ms, d->evForce[2 * k], d->evForce[2 * k + 1]) == cudaSuccess) sum += ms;
      }
      cudaGetLastError();
      if (n > 0) d->lastForceMs = sum / n;
      d->nForceEv = 0;
      *meanForceMs = d->lastForceMs;
   }
   if (nLaunches) *nLaunches = d->nLaunches;
   return COMD_OK;
}

// CUDA-event stopwatch on the library's own stream (bench.py times its K steps with it)
extern "C" int comd_dev_timer_start(ComdDev* d)
{
   if (!d) return COMD_E_ARG;
   CUDA_TRY(cudaSetDevice(d->device));
   CUDA_TRY(cudaStreamSynchronize(d->stream));
   CUDA_TRY(cudaEventRecord(d->evT0, d->stream));
   return COMD_OK;
}
extern "C" int comd_dev_timer_stop(ComdDev* d, double* ms)
{
   if (!d || !ms) return COMD_E_ARG;
   CUDA_TRY(cudaSetDevice(d->device));
   CUDA_TRY(cudaEventRecord(d->evT1, d->stream));
   CUDA_TRY(cudaEventSynchronize(d->evT1));
   float f = 0.f;
   CUDA_TRY(cudaEventElapsedTime(&f, d->evT0, d->evT1));
   *ms = f;
   return COMD_OK;
}

// ---- local atoms, compacted on the device -------------------------------------------------
// Local atoms of one (y,z) row of cells are one contiguous run of the sorted arrays; an exclusive
// scan over the ny*nz run lengths gives every run its place in a dense, ghost-free copy.
__global__ void __launch_bounds__(1024)
k_local_rows(DevGeom g, const int* __restrict__ cellStart, int* __restrict__ rowBeg, int* __restrict__ rowOff)
{
   __shared__ int warpSum[32];
   __shared__ int carry;
   const int nRows = g.n[1] * g.n[2];
   const int t = threadIdx.x, lane = t & 31, w = t >> 5;
   if (t == 0) carry = 0;
   __syncthreads();
   for (int base = 0; base < nRows; base += 1024)
   {
      const int k = base + t;
      int len = 0, b = 0;
      if (k < nRows)
      {
         const int c0 = cellIndex(g, 0, k % g.n[1], k / g.n[1]);
         b = cellStart[c0]; len = cellStart[c0 + g.n[0]] - b;
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
      if (k < nRows) { rowBeg[k] = b; rowOff[k] = excl; }
      __syncthreads();
      if (t == 1023) carry = excl + len;
      __syncthreads();
   }
   if (t == 0) rowOff[nRows] = carry;
}

struct CompactArrays { const real_t* in[10]; real_t* out[10]; const int* gidIn; int* gidOut; };
// arrays K0..K1-1 (0-2 r, 3-5 p, 6-8 f, 9 U) of the local atoms, row by row, into dense output arrays
template <int K0, int K1, bool GID>
__global__ void __launch_bounds__(256)
k_compact_local(CompactArrays A, const int* __restrict__ rowBeg, const int* __restrict__ rowOff, int nRows)
{
   for (int row = blockIdx.x; row < nRows; row += gridDim.x)
   {
      const int b = rowBeg[row], o = rowOff[row], len = rowOff[row + 1] - o;
      for (int t = threadIdx.x; t < len; t += blockDim.x)
      {
         #pragma unroll
         for (int k = K0; k < K1; ++k) A.out[k][o + t] = A.in[k][b + t];
         if (GID) A.gidOut[o + t] = A.gidIn[b + t];
      }
   }
}

// ---- e2e resume path: the caller hands back the atoms of the previous comd_dev_step_host call --------
// Local atom k of the compact (host) order lives at sorted index rowBeg[row] + (k - rowOff[row]).
// check: same gids in the same order, and no input position further than skin/2 from the snapshot the
// neighbour list was built on -> the sort and the list are still valid for these inputs.
__global__ void __launch_bounds__(256)
k_e2e_check(const int* __restrict__ rowBeg, const int* __restrict__ rowOff, int nRows, int n, real_t half2,
            const int* __restrict__ gidIn, const real_t* __restrict__ xin, const real_t* __restrict__ yin, const real_t* __restrict__ zin,
            const int* __restrict__ gid, const real_t* __restrict__ lx, const real_t* __restrict__ ly, const real_t* __restrict__ lz,
            int* __restrict__ flag)
{
   bool bad = false;
   if (blockIdx.x == 0 && threadIdx.x == 0 && rowOff[nRows] != n) bad = true;
   for (int row = blockIdx.x; row < nRows; row += gridDim.x)
   {
      const int b = rowBeg[row], o = rowOff[row], len = rowOff[row + 1] - o;
      for (int t = threadIdx.x; t < len; t += blockDim.x)
      {
         if (o + t >= n) { bad = true; break; }
         const real_t dx = xin[o + t] - lx[b + t], dy = yin[o + t] - ly[b + t], dz = zin[o + t] - lz[b + t];
         bad |= gidIn[o + t] != gid[b + t] || !(dx * dx + dy * dy + dz * dz <= half2);
      }
   }
   if (bad) *flag = 1;
}
struct ScatterArrays { const real_t* in[9]; real_t* out[9]; };
__global__ void __launch_bounds__(256)
k_e2e_scatter(ScatterArrays A, const int* __restrict__ rowBeg, const int* __restrict__ rowOff, int nRows, real4* __restrict__ rq)
{
   for (int row = blockIdx.x; row < nRows; row += gridDim.x)
   {
      const int b = rowBeg[row], o = rowOff[row], len = rowOff[row + 1] - o;
      for (int t = threadIdx.x; t < len; t += blockDim.x)
      {
         #pragma unroll
         for (int k = 0; k < 9; ++k) A.out[k][b + t] = A.in[k][o + t];
         rq[b + t] = make_real4(A.in[0][o + t], A.in[1][o + t], A.in[2][o + t], R(0.0));
      }
   }
}

static bool isPinned(const void* p)
{
   cudaPointerAttributes at;
   if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
   return at.type == cudaMemoryTypeHost;
}

// pinned host memory for callers that want their step_host buffers copied without staging
extern "C" int comd_host_alloc(void** p, size_t bytes)
{
   if (!p) return COMD_E_ARG;
   CUDA_TRY(cudaMallocHost(p, bytes));
   return COMD_OK;
}
extern "C" int comd_host_free(void* p)
{
   if (p) CUDA_TRY(cudaFreeHost(p));
   return COMD_OK;
}

// the inputs of a speculative call against what the previous call handed out, bit for bit
struct EqualArrays { const real_t* a[9]; const real_t* b[9]; };
__global__ void __launch_bounds__(256)
k_e2e_equal(EqualArrays Q, const int* __restrict__ ga, const int* __restrict__ gb, int n, int* __restrict__ flag)
{
   bool bad = false;
   for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x)
   {
      bad |= ga[t] != gb[t];
      #pragma unroll
      for (int k = 0; k < 9; ++k) bad |= realBits(Q.a[k][t]) != realBits(Q.b[k][t]);
   }
   if (bad) *flag = 1;
}

static bool rangesOverlap(const void* a, const void* b, size_t bytes)
{
   const char* x = static_cast<const char*>(a); const char* y = static_cast<const char*>(b);
   return x && y && x < y + bytes && y < x + bytes;
}

// End-to-end step through HOST buffers: one full timestep() whose inputs (gid, r, p, f at
// time t) start in host memory and whose results (gid, r, p, f, U at t+dt, device order, local
// atoms only) end there.  bench.py times this as `e2e`.  Buffers obtained from comd_host_alloc
// (pinned) are copied directly; pageable ones go through a pinned staging area.  The ghost-free
// result is compacted on the device, so only local atoms cross PCIe.
//
// Three ways through it, cheapest first:
//  * speculative (one rank, pinned buffers, the call follows another comd_dev_step_host call with nothing
//    in between): the device already holds what it handed out last time, so the step starts at once while
//    the inputs are still crossing PCIe on their own stream; a kernel then compares them bit for bit with
//    the previous results.  Equal -> the step that ran is the step that was asked for.  Different -> the
//    step is redone from the uploaded inputs (full path) and speculation is switched off for this handle.
//    An output buffer may be the input buffer of the same quantity (in-place update): each result is
//    copied out only after the upload of the array it may overwrite has finished.
//  * resume: same atoms in the same order and none further than skin/2 from the list's snapshot -> the
//    inputs are scattered to their rows, sort and list are kept.
//  * full: new atoms; re-bin, sort, rebuild.
static int stepHostImpl(ComdDev* d, int n, double dt, bool inputsOnDevice,
                        const int* gidIn, const real_t* const in[9],
                        int* nOut, int* gidOut, real_t* const outPtr[10], double eLocal[2])
{
   cudaStream_t s = d->stream;
   const long cap = d->cap;
   AtomArrays& A = d->A;
   real_t* devIn[9] = { A.r[0], A.r[1], A.r[2], A.p[0], A.p[1], A.p[2], d->f[0], d->f[1], d->f[2] };
   bool pinnedIn = inputsOnDevice || isPinned(gidIn);
   for (int k = 0; k < 9 && !inputsOnDevice; ++k) pinnedIn = pinnedIn && isPinned(in[k]);
   bool pinnedOut = !gidOut || isPinned(gidOut);
   for (int k = 0; k < 10; ++k) pinnedOut = pinnedOut && (!outPtr[k] || isPinned(outPtr[k]));
   if (!pinnedIn || !pinnedOut)
   {  // pinned staging: 11 slabs of `cap` doubles
      const long need = cap * 11;
      if (d->hStageDoubles < need)
      {
         if (d->hStage) cudaFreeHost(d->hStage);
         d->hStage = nullptr; d->hStageDoubles = 0;
         CUDA_TRY(cudaMallocHost(&d->hStage, need * sizeof(real_t)));
         d->hStageDoubles = need;
      }
   }
   real_t* st = d->hStage;
   const int nRows = d->g.n[1] * d->g.n[2];
   if (!d->rowTab) CUDA_TRY(cudaMalloc(&d->rowTab, (2 * (size_t)nRows + 2) * sizeof(int)));
   if (!d->e2eU)
   {
      CUDA_TRY(cudaMalloc(&d->e2eU, (size_t)cap * 9 * sizeof(real_t)));
      CUDA_TRY(cudaMalloc(&d->e2eUgid, (size_t)cap * sizeof(int)));
      for (int k = 0; k < 2; ++k)
      {
         CUDA_TRY(cudaMalloc(&d->e2eOut[k], (size_t)cap * 10 * sizeof(real_t)));
         CUDA_TRY(cudaMalloc(&d->e2eOutGid[k], (size_t)cap * sizeof(int)));
      }
      CUDA_TRY(cudaMalloc(&d->e2eFlag, sizeof(int)));
      CUDA_TRY(cudaMallocHost(&d->hE2eFlag, sizeof(int)));
      CUDA_TRY(cudaStreamCreateWithFlags(&d->copyStream, cudaStreamNonBlocking));
      CUDA_TRY(cudaStreamCreateWithFlags(&d->upStream, cudaStreamNonBlocking));
      CUDA_TRY(cudaEventCreateWithFlags(&d->evCopy, cudaEventDisableTiming));
      CUDA_TRY(cudaEventCreateWithFlags(&d->evUpR, cudaEventDisableTiming));
      CUDA_TRY(cudaEventCreateWithFlags(&d->evUpP, cudaEventDisableTiming));
      CUDA_TRY(cudaEventCreateWithFlags(&d->evUp, cudaEventDisableTiming));
   }
   int* rowBeg = d->rowTab; int* rowOff = d->rowTab + nRows;
   const int prev = d->e2eCur, nxt = prev ^ 1;
   real_t* scr[9]; real_t* cout[10]; EqualArrays Q;
   for (int k = 0; k < 9; ++k) { scr[k] = d->e2eU + k * cap; Q.a[k] = scr[k]; Q.b[k] = d->e2eOut[prev] + k * cap; }
   for (int k = 0; k < 10; ++k) cout[k] = d->e2eOut[nxt] + k * cap;
   int* ugid = d->e2eUgid; int* coutGid = d->e2eOutGid[nxt];

   // Resume path: the previous call left the atoms sorted, the list built and the local-row table of its
   // compaction in place.  If the caller hands the same atoms back in that order and none is further than
   // skin/2 from the list's snapshot, neither the sort nor the list has to be redone: the inputs are
   // scattered to their rows and the step starts from there.  All ranks must agree (the two paths issue
   // different collectives), hence the flag is summed over the ranks.
   const bool tryResume = !inputsOnDevice && d->useNL && d->nlValid && d->e2eOrderValid;
   bool cross = false;                             // an output buffer over the input buffer of another quantity
   if (!inputsOnDevice)
      for (int k = 0; k < 10; ++k)
      {
         for (int j = 0; j < 9; ++j) cross = cross || (j != k && rangesOverlap(outPtr[k], in[j], n * sizeof(real_t)));
         cross = cross || rangesOverlap(outPtr[k], gidIn, n * sizeof(int));
         if (k < 9) cross = cross || rangesOverlap(gidOut, in[k], n * sizeof(real_t));
      }
   const bool spec = tryResume && d->e2ePristine && d->e2eSpecArmed && d->nRanks == 1 && pinnedIn && pinnedOut && !cross && n == d->e2eLastM;
   cudaStream_t u = spec ? d->upStream : s;
   if (!inputsOnDevice)
   {
      if (spec) CUDA_TRY(cudaMemsetAsync(d->e2eFlag, 0, sizeof(int), u));
      {
         const int* src = gidIn;
         if (!pinnedIn) { memcpy(st + 9 * cap, gidIn, n * sizeof(int)); src = reinterpret_cast<const int*>(st + 9 * cap); }
         CUDA_TRY(cudaMemcpyAsync(ugid, src, n * sizeof(int), cudaMemcpyHostToDevice, u));
      }
      for (int k = 0; k < 9; ++k)
      {
         const real_t* src = in[k];
         if (!pinnedIn) { memcpy(st + k * cap, in[k], n * sizeof(real_t)); src = st + k * cap; }
         CUDA_TRY(cudaMemcpyAsync(scr[k], src, n * sizeof(real_t), cudaMemcpyHostToDevice, u));
         if (spec && k == 2) CUDA_TRY(cudaEventRecord(d->evUpR, u));
         if (spec && k == 5) CUDA_TRY(cudaEventRecord(d->evUpP, u));
         if (!spec && k == 2 && tryResume)
         {
            CUDA_TRY(cudaMemsetAsync(d->cnt + C_NLREBUILD, 0, sizeof(int), s));
            const real_t h = R(0.5) * d->skin;
            k_e2e_check<<<gridFor(nRows, 1, 148 * 8), 256, 0, s>>>(rowBeg, rowOff, nRows, n, h * h, ugid, scr[0], scr[1], scr[2],
               A.gid, d->nlLastR[0], d->nlLastR[1], d->nlLastR[2], d->cnt + C_NLREBUILD); LAUNCH_CHECK(d);
            int rcf = comdAllreduceFlagDevice(d, d->cnt + C_NLREBUILD);
            if (rcf != COMD_OK) return rcf;
            CUDA_TRY(cudaMemcpyAsync(d->hCnt + C_NLREBUILD, d->cnt + C_NLREBUILD, sizeof(int), cudaMemcpyDeviceToHost, s));
         }
      }
      if (spec)
      {
         k_e2e_equal<<<148 * 8, 256, 0, u>>>(Q, ugid, d->e2eOutGid[prev], n, d->e2eFlag); LAUNCH_CHECK(d);
         CUDA_TRY(cudaMemcpyAsync(d->hE2eFlag, d->e2eFlag, sizeof(int), cudaMemcpyDeviceToHost, u));
         CUDA_TRY(cudaEventRecord(d->evUp, u));
      }
   }
   bool resume = false;
   if (tryResume && !spec)
   {
      CUDA_TRY(cudaStreamSynchronize(s));
      resume = d->hCnt[C_NLREBUILD] == 0;
   }
   d->e2eOrderValid = 0;
   if (spec) { }                                   // the device holds these atoms already (checked by k_e2e_equal)
   else if (resume)
   {
      ScatterArrays W;
      for (int k = 0; k < 9; ++k) { W.in[k] = scr[k]; W.out[k] = devIn[k]; }
      k_e2e_scatter<<<gridFor(nRows, 1, 148 * 8), 256, 0, s>>>(W, rowBeg, rowOff, nRows, A.rq); LAUNCH_CHECK(d);
   }
   else
   {
      for (int k = 0; k < 9; ++k) CUDA_TRY(cudaMemcpyAsync(devIn[k], scr[k], n * sizeof(real_t), cudaMemcpyDeviceToDevice, s));
      CUDA_TRY(cudaMemcpyAsync(A.gid, ugid, n * sizeof(int), cudaMemcpyDeviceToDevice, s));
      CUDA_TRY(cudaMemsetAsync(A.spec, 0, n * sizeof(int), s));
      k_set_int<<<gridFor(n, 256), 256, 0, s>>>(A.cell, cellIndex(d->g, 0, 0, 0), n); LAUNCH_CHECK(d);
      k_set_int<<<1, 1, 0, s>>>(d->cnt + C_NATOMS, n, 1); LAUNCH_CHECK(d);
      d->approxAtoms = n; d->approxHasGhosts = 0;
      d->nlValid = 0;                              // new atoms: the list path must re-bin and rebuild
   }
   // ---- the step, in the pieces of comd_dev_timestep, so that results leave as soon as they are final:
   // positions and gids are final after the drift + redistribute and cross PCIe (copy stream) while the
   // force kernels run; momenta, forces and energies follow after the second kick.
   d->inTimestep = 1;
   int rc = launchKickDrift(d, dt, false);
   if (rc == COMD_OK) rc = comd_dev_redistribute(d);
   if (rc != COMD_OK) { d->inTimestep = 0; return rc; }
   const real_t* cin[10] = { d->A.r[0], d->A.r[1], d->A.r[2], d->A.p[0], d->A.p[1], d->A.p[2], d->f[0], d->f[1], d->f[2], d->e };
   k_local_rows<<<1, 1024, 0, s>>>(d->g, d->cellStart, rowBeg, rowOff); LAUNCH_CHECK(d);
   {
      CompactArrays C;
      for (int k = 0; k < 10; ++k) { C.in[k] = cin[k]; C.out[k] = cout[k]; }
      C.gidIn = d->A.gid; C.gidOut = coutGid;
      k_compact_local<0, 3, true><<<gridFor(nRows, 1, 148 * 8), 256, 0, s>>>(C, rowBeg, rowOff, nRows); LAUNCH_CHECK(d);
      CUDA_TRY(cudaEventRecord(d->evCopy, s));
      CUDA_TRY(cudaStreamWaitEvent(d->copyStream, d->evCopy, 0));
      // A result leaves only after the upload of the same quantity has finished.  In-place callers need that order
      // (the result lands on the input); for everybody else it keeps the two PCIe directions from running flat out
      // at the same time, which this platform does not reward: measured on C5 with separate result buffers and no
      // waits 53.5 ms per call, with the waits 42.3 (59.4 when the step waits for the whole upload first).
      if (spec) CUDA_TRY(cudaStreamWaitEvent(d->copyStream, d->evUpR, 0));
   }
   // The early copy needs the local count on the host before the step has finished: it is the input count on
   // one rank (nobody can leave) and on any rank whose atoms were resumed and not re-binned in this step.
   const bool earlyCopy = d->nRanks == 1 || (resume && d->nlValid);
   if (earlyCopy)
   {
      for (int k = 0; k < 3; ++k)
      {
         if (!outPtr[k]) continue;
         real_t* dst = pinnedOut ? outPtr[k] : st + k * cap;
         CUDA_TRY(cudaMemcpyAsync(dst, cout[k], n * sizeof(real_t), cudaMemcpyDeviceToHost, d->copyStream));
      }
      if (gidOut)
      {
         int* dst = pinnedOut ? gidOut : reinterpret_cast<int*>(st + 10 * cap);
         CUDA_TRY(cudaMemcpyAsync(dst, coutGid, n * sizeof(int), cudaMemcpyDeviceToHost, d->copyStream));
      }
   }
   rc = computeForce(d, 1);
   if (rc == COMD_OK) rc = comd_dev_advance_velocity(d, R(0.5) * dt);
   d->inTimestep = 0;
   if (rc != COMD_OK) return rc;
   {
      CompactArrays C;
      for (int k = 0; k < 10; ++k) { C.in[k] = cin[k]; C.out[k] = cout[k]; }
      C.gidIn = d->A.gid; C.gidOut = coutGid;
      k_compact_local<3, 10, false><<<gridFor(nRows, 1, 148 * 8), 256, 0, s>>>(C, rowBeg, rowOff, nRows); LAUNCH_CHECK(d);
   }
   rc = comd_dev_energy(d, eLocal);                // syncs, refreshes hCnt
   if (rc != COMD_OK) return rc;
   const int m = d->hCnt[C_NLOCAL];
   if (nOut) *nOut = m;
   if (earlyCopy && m != n) { comdSetError("comd_dev_step_host: %d local atoms after the step, %d on input", m, n); return COMD_E_STATE; }
   for (int k = earlyCopy ? 3 : 0; k < 10; ++k)
   {
      if (spec && k == 3) CUDA_TRY(cudaStreamWaitEvent(s, d->evUpP, 0));     // p leaves after p has arrived,
      if (spec && k == 6) CUDA_TRY(cudaStreamWaitEvent(s, d->evUp, 0));      // f and U after everything has
      if (!outPtr[k]) continue;
      real_t* dst = pinnedOut ? outPtr[k] : st + k * cap;
      CUDA_TRY(cudaMemcpyAsync(dst, cout[k], m * sizeof(real_t), cudaMemcpyDeviceToHost, s));
   }
   if (gidOut && !earlyCopy)
   {
      int* dst = pinnedOut ? gidOut : reinterpret_cast<int*>(st + 10 * cap);
      CUDA_TRY(cudaMemcpyAsync(dst, coutGid, m * sizeof(int), cudaMemcpyDeviceToHost, s));
   }
   CUDA_TRY(cudaStreamSynchronize(s));
   CUDA_TRY(cudaStreamSynchronize(d->copyStream));
   if (spec)
   {
      CUDA_TRY(cudaStreamSynchronize(d->upStream));
      if (*d->hE2eFlag != 0)
      {  // not the atoms the device held: the step that ran was the wrong one.  The inputs are complete in the
         // landing area (and possibly overwritten on the host by now): redo from there, never speculate again.
         d->e2eSpecArmed = 0;
         return stepHostImpl(d, n, dt, true, gidIn, in, nOut, gidOut, outPtr, eLocal);
      }
   }
   if (spec || resume) d->e2eResumes++;
   d->e2eCur = nxt; d->e2eLastM = m;
   d->e2eOrderValid = 1;                           // rowTab describes the order the caller now holds
   d->e2ePristine = 1;
   if (!pinnedOut)
   {
      for (int k = 0; k < 10; ++k) if (outPtr[k]) memcpy(outPtr[k], st + k * cap, m * sizeof(real_t));
      if (gidOut) memcpy(gidOut, st + 10 * cap, m * sizeof(int));
   }
   return COMD_OK;
}

extern "C" int comd_dev_step_host(ComdDev* d, int n, double dt,
                                  const int* gidIn, const real_t* rIn[3], const real_t* pIn[3], const real_t* fIn[3],
                                  int* nOut, int* gidOut, real_t* rOut[3], real_t* pOut[3], real_t* fOut[3], real_t* UOut,
                                  double eLocal[2])
{
   if (!d || n <= 0 || !gidIn || !rIn || !pIn || !fIn) { comdSetError("comd_dev_step_host: bad argument"); return COMD_E_ARG; }
   if (n > d->cap) { comdSetError("comd_dev_step_host: %d atoms > capacity %ld", n, d->cap); return COMD_E_CAPACITY; }
   if (d->potKind == 0) { comdSetError("comd_dev_step_host: no potential set"); return COMD_E_STATE; }
   CUDA_TRY(cudaSetDevice(d->device));
   const real_t* in[9] = { rIn[0], rIn[1], rIn[2], pIn[0], pIn[1], pIn[2], fIn[0], fIn[1], fIn[2] };
   real_t* outPtr[10] = { rOut ? rOut[0] : nullptr, rOut ? rOut[1] : nullptr, rOut ? rOut[2] : nullptr,
                          pOut ? pOut[0] : nullptr, pOut ? pOut[1] : nullptr, pOut ? pOut[2] : nullptr,
                          fOut ? fOut[0] : nullptr, fOut ? fOut[1] : nullptr, fOut ? fOut[2] : nullptr, UOut };
   return stepHostImpl(d, n, dt, false, gidIn, in, nOut, gidOut, outPtr, eLocal);
}

extern "C" int comd_dev_step_host_stats(ComdDev* d, long* nResumed)
{
   if (!d || !nResumed) return COMD_E_ARG;
   *nResumed = d->e2eResumes;
   return COMD_OK;
}

// ---- FP64 FMA-pipe peak, measured on this very GPU (the roofline denominator for the
// force kernels; MEASURED_PEAKS.json only carries HBM and bf16 figures).
__global__ void __launch_bounds__(256)
k_dfma_peak(double* out, int iters, double a, double b)
{
   double x0 = threadIdx.x * 1e-3, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
   for (int i = 0; i < iters; ++i)
   {
      x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
      x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
   }
   const double s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
   if (s == 12345.678) out[0] = s;        // never true; keeps the chain alive
}

__global__ void __launch_bounds__(256)
k_ffma_peak(double* out, int iters, float a, float b)
{
   float x0 = threadIdx.x * 1e-3f, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
   for (int i = 0; i < iters; ++i)
   {
      x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
      x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
   }
   const float s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
   if (s == 12345.678f) out[0] = s;
}

static int fmaPeak(ComdDev* d, double* tflops, bool single);
extern "C" int comd_dev_fp64_peak(ComdDev* d, double* tflops) { return fmaPeak(d, tflops, false); }
extern "C" int comd_dev_fp32_peak(ComdDev* d, double* tflops) { return fmaPeak(d, tflops, true); }

static int fmaPeak(ComdDev* d, double* tflops, bool single)
{
   if (!d || !tflops) return COMD_E_ARG;
   CUDA_TRY(cudaSetDevice(d->device));
   const int blocks = 148 * 8, threads = 256, iters = 1 << 15;
   cudaEvent_t a, b;
   CUDA_TRY(cudaEventCreate(&a)); CUDA_TRY(cudaEventCreate(&b));
   double best = 0.0;
   for (int rep = 0; rep < 6; ++rep)
   {
      CUDA_TRY(cudaEventRecord(a, d->stream));
      if (single) k_ffma_peak<<<blocks, threads, 0, d->stream>>>(d->redOut, iters, 0.999999f, 1e-9f);
      else k_dfma_peak<<<blocks, threads, 0, d->stream>>>(d->redOut, iters, 0.999999, 1e-9);
      LAUNCH_CHECK(d);
      CUDA_TRY(cudaEventRecord(b, d->stream));
      CUDA_TRY(cudaEventSynchronize(b));
      float ms = 0.f;
      CUDA_TRY(cudaEventElapsedTime(&ms, a, b));
      const double tf = 2.0 * 8.0 * (double)iters * blocks * threads / (ms * 1e-3) * 1e-12;
      if (rep > 0 && tf > best) best = tf;
   }
   cudaEventDestroy(a); cudaEventDestroy(b);
   *tflops = best;
   return COMD_OK;
}
