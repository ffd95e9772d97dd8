// force_tile.cuh -- the tiled pair-force engine shared by the LJ and EAM kernels (sm_100a).
//
// One CTA owns a run of TX consecutive local cells along x at fixed (y,z).  Thanks to the
// padded x-fastest cell order (comd_dev.cuh) the 27-cell stencils of all its atoms are covered
// by 9 contiguous atom runs (one per (dy,dz) row, TX+2 cells long).  The CTA
//   1. stages those 9 runs in shared memory with TMA bulk copies (cp.async.bulk, one
//      mbarrier): FP64 x/y/z (+ F' for EAM pass 3) and the FP32 float4 shadow positions;
//   2. FILTER: 8 lanes per atom scan the atom's 9 row segments in FP32 (FFMA pipe) against
//      a slightly enlarged cutoff and append survivors to lane-private queues -- no
//      FP64 work and no divergent pair maths in this loop;
//   3. PROCESS: the queues of the 8 lanes are compacted and each lane then evaluates one
//      surviving pair per round in FP64 (exact r2 re-test, so the FP32 filter never decides
//      a result): dense FP64-pipe work instead of the ~14 % lane utilisation a
//      thread-per-atom stencil loop gets;
//   4. reduces the 8 partial sums with xor shuffles (fixed order => bit-reproducible).
// Reference kernels replaced: LJ_Force_cta_cell (gpu_lj_cta_cell.h:31-122),
// EAM_Force_cta_cell / warp_atom (gpu_eam_cta_cell.h:37-295, gpu_eam_warp_atom.h:33-181).
#pragma once
#include "comd_dev.cuh"

#define FT_THREADS 256
#define FT_WARPS   (FT_THREADS / 32)
#define FT_TXMAX   16                 // cells per tile along x (upper bound)
#define FT_G       8                  // lanes per atom
#define FT_APW     (32 / FT_G)        // atoms per warp pass
#define FT_QCAP    24                 // lane-private queue depth
#define FT_CH      8                  // filter iterations between queue-room checks
#define FT_DCAP    (FT_QCAP * FT_G)   // dense queue depth per atom
#define FT_HDR     1024               // bytes reserved for TileHdr

struct TileCfg
{
   int   tilesX;        // tiles along x
   int   sMax;          // staged-atom capacity of the shared arrays
   float rcf2;          // FP32 filter threshold (cutoff + rounding margin)^2
};

struct TileHdr
{
   unsigned long long mbar;
   int rowGA[9];                      // global index (even) of the first staged atom of each row
   int rowS[9];                       // shared index of that atom
   int cs[9][FT_TXMAX + 3];           // shared index of every cell start of the row
   int iBeg, iEnd;                    // global range of the tile's own atoms
   int total;
   int rowDelta[9];                   // global index - shared index of the atoms of each row (PK_RHOL: hits leave as global indices)
};
static_assert(sizeof(TileHdr) <= FT_HDR, "TileHdr");

// kinds of pair work
#define PK_LJ   0
#define PK_EAM1 1
#define PK_EAM3 3
#define PK_RHOL 8                     // link-cell EAM pass 1 that also writes the step's hit list: rhobar + F, F' here, every pair
                                      // force in the list kernel k_force_nl<PK_EAMF|PK_EAMFE> that walks those hits (force_eam.cu)

struct PairParams
{
   // LJ
   real_t s6, eps, eShift;
   // both
   real_t rc2;
   // EAM
   DevTable tPhi, tRho, tF;
   const real2* tPR;                // interleaved {phi,rho} (same grid), entry 0; valid [-1, n+1]
   const real2* tW0; const real2* tW1;   // shifted copies of tPR for 256-bit window loads (comd_dev.cuh), or NULL
   const real2* tDR;                // pass 3: {g1, g2-g1} of rho per interval
   // list path (force_nl.cuh): 32-byte records, one sector per lookup
   const real4* rvPhi; const real4* rvRho; const real4* r3;
   double n3ScaleRho;                 // fixed-point scale of the rhobar accumulators (force_n3.cuh)
   const int* skipFlag;               // list kernels return at once when *skipFlag != 0 (speculatively enqueued step whose list turned out stale)
};

// ---- PTX helpers: mbarrier + 1-D bulk copy (TMA engine, UBLKCP in SASS) -----------------
__device__ __forceinline__ unsigned smemAddr(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbarInit(void* bar, int count)
{
   asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smemAddr(bar)), "r"(count));
   asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbarExpectTx(void* bar, unsigned bytes)
{
   asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smemAddr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbarWait(void* bar, unsigned phase)
{
   unsigned ok = 0;
   while (!ok)
      asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                   : "=r"(ok) : "r"(smemAddr(bar)), "r"(phase) : "memory");
}
__device__ __forceinline__ void bulkG2S(void* dst, const void* src, unsigned bytes, void* bar)
{
   asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                ::"r"(smemAddr(dst)), "l"(src), "r"(bytes), "r"(smemAddr(bar)) : "memory");
}

// one 256-bit read-only load (LDG.E.256 on sm_100a): two adjacent double2, 32-byte aligned
__device__ __forceinline__ void ldg256(const real2* p, real2& a, real2& b)
{
   real4 v;
   COMD_LDG_REAL4(".nc", v, p);
   a = make_real2(v.x, v.y); b = make_real2(v.z, v.w);
}

// interpolate() of eam.c:557-579 on four already-loaded stencil values
__device__ __forceinline__ void interpValue(real_t tm, real_t t0, real_t t1, real_t t, real_t& f)
{
   const real_t g1 = t1 - tm;
   f = t0 + R(0.5) * t * (g1 + t * (t1 + tm - R(2.0) * t0));
}
__device__ __forceinline__ void interpDeriv(real_t tm, real_t t0, real_t t1, real_t t2, real_t t, real_t invDx, real_t& df)
{
   const real_t g1 = t1 - tm;
   const real_t g2 = t2 - t0;
   df = R(0.5) * (g1 + t * (g2 - g1)) * invDx;
}
// index + fraction with the reference's one-sided clamps (eam.c:561-571)
__device__ __forceinline__ int tableIndex(const DevTable& T, real_t r, real_t& t)
{
   if (r < T.x0) r = T.x0;
   r = (r - T.x0) * T.invDx;
   int ii = realFloorInt(r);
   if (ii > T.n) { ii = T.n; r = T.n / T.invDx; }
   t = r - floor(r);
   return ii;
}
__device__ __forceinline__ void tableLookupFull(const DevTable& T, real_t r, real_t& f, real_t& df)
{
   real_t t;
   const int ii = tableIndex(T, r, t);
   const real_t* __restrict__ v = T.v + ii;
   const real_t tm = __ldg(v - 1), t0 = __ldg(v), t1 = __ldg(v + 1), t2 = __ldg(v + 2);
   interpValue(tm, t0, t1, t, f);
   interpDeriv(tm, t0, t1, t2, t, T.invDx, df);
}

// ---- one surviving pair, FP64 ---------------------------------------------------------
// acc layout: [0..2] force, [3] energy part, [4] rhobar (EAM pass 1)
template <int KIND>
__device__ __forceinline__ void pairEval(const PairParams& P, real_t dx, real_t dy, real_t dz, real_t r2,
                                         real_t dfi, real_t dfj, real_t* acc)
{
   if (KIND == PK_LJ)
   {  // ljForce.c:207-233
      const real_t ir2 = R(1.0) / r2;
      const real_t r6 = (P.s6 * ir2) * (ir2 * ir2);
      acc[3] += r6 * (r6 - R(1.0)) - P.eShift;
      const real_t fr = -R(4.0) * P.eps * r6 * ir2 * (R(12.0) * r6 - R(6.0));
      acc[0] -= dx * fr; acc[1] -= dy * fr; acc[2] -= dz * fr;
   }
   else if (KIND == PK_EAM1)
   {  // eam.c:311-345
      const real_t rinv = rsqrt(r2);
      const real_t r = r2 * rinv;
      real_t phi, dphi, rho;
      if (P.tW0)
      {
         real_t t;
         const int ii = tableIndex(P.tPhi, r, t);
         const real2* v = ((ii & 1) ? P.tW1 : P.tW0) + (ii & ~1);          // entries ii-1 .. ii+2, 32-byte aligned
         real2 vm, v0, v1, v2;
         ldg256(v, vm, v0); ldg256(v + 2, v1, v2);
         interpValue(vm.x, v0.x, v1.x, t, phi);
         interpDeriv(vm.x, v0.x, v1.x, v2.x, t, P.tPhi.invDx, dphi);
         interpValue(vm.y, v0.y, v1.y, t, rho);
      }
      else if (P.tPR)
      {
         real_t t;
         const int ii = tableIndex(P.tPhi, r, t);
         const real2* __restrict__ v = P.tPR + ii;
         const real2 vm = __ldg(v - 1), v0 = __ldg(v), v1 = __ldg(v + 1), v2 = __ldg(v + 2);
         interpValue(vm.x, v0.x, v1.x, t, phi);
         interpDeriv(vm.x, v0.x, v1.x, v2.x, t, P.tPhi.invDx, dphi);
         interpValue(vm.y, v0.y, v1.y, t, rho);
      }
      else
      {
         real_t drho;
         tableLookupFull(P.tPhi, r, phi, dphi);
         tableLookupFull(P.tRho, r, rho, drho);
      }
      const real_t s = dphi * rinv;
      acc[0] -= s * dx; acc[1] -= s * dy; acc[2] -= s * dz;
      acc[3] += R(0.5) * phi;
      acc[4] += rho;
   }
   else if (KIND == PK_RHOL)
   {  // eam.c:331-343, the density alone (value part of interpolate(), eam.c:557-579)
      const real_t r = r2 * rsqrt(r2);
      real_t t;
      const real_t* __restrict__ v = P.tRho.v + tableIndex(P.tRho, r, t);
      real_t rho;
      interpValue(__ldg(v - 1), __ldg(v), __ldg(v + 1), t, rho);
      acc[0] += rho;
   }
   else
   {  // eam.c:390-409
      const real_t rinv = rsqrt(r2);
      const real_t r = r2 * rinv;
      real_t t, drho;
      const int ii = tableIndex(P.tRho, r, t);
      if (P.tDR)
      {
         const real2 gq = __ldg(P.tDR + ii);
         drho = R(0.5) * (gq.x + t * gq.y) * P.tRho.invDx;
      }
      else
      {
         const real_t* __restrict__ v = P.tRho.v + ii;
         const real_t tm = __ldg(v - 1), t0 = __ldg(v), t1 = __ldg(v + 1), t2 = __ldg(v + 2);
         interpDeriv(tm, t0, t1, t2, t, P.tRho.invDx, drho);
      }
      const real_t s = (dfi + dfj) * drho * rinv;
      acc[0] -= s * dx; acc[1] -= s * dy; acc[2] -= s * dz;
   }
}

template <int KIND> struct PairTraits { static constexpr int NACC = KIND == PK_LJ ? 4 : (KIND == PK_EAM1 ? 5 : (KIND == PK_RHOL ? 1 : 3)); };

// hit list of PK_RHOL: the full-list row format of neighbor.cu (chunks of 32 entries, entry 8u+l stored at 4l+u)
struct TileHits { int* list; int* count; int* countHit; int stride; };
__device__ __forceinline__ int tileSlot(int pos) { return (pos & ~31) | ((pos & 7) << 2) | ((pos >> 3) & 3); }

// elements of a per-atom array in 16 bytes: TMA bulk copies start and end on such boundaries
#define FT_AL ((int)(16 / sizeof(real_t)))

// shared-memory bytes for a given staging capacity
template <int KIND>
__host__ __device__ inline size_t tileSmemBytes(int sMax)
{
   size_t b = FT_HDR;
   b += (size_t)sMax * sizeof(real_t) * (KIND == PK_EAM3 ? 4 : 3);      // x,y,z (+F')
   b += (size_t)sMax * 16;                                 // float4 shadow
   b += (size_t)FT_WARPS * FT_QCAP * 32 * 2;               // lane queues
   b += (size_t)FT_WARPS * FT_APW * FT_DCAP * 2;           // dense queues
   return b;
}

// ---- PROCESS: compact the 8 lane queues of every atom, then one pair per lane per round ----
template <int KIND>
__device__ __forceinline__ void processQueues(const PairParams& P, unsigned short* myQ, unsigned short* myD, int& qn,
                                              int gl, int si, real_t xi, real_t yi, real_t zi, real_t dfi,
                                              const real_t* sx, const real_t* sy, const real_t* sz, const real_t* sdf,
                                              real_t* acc, const TileHdr& H, int* hitRow, int hitStride, int& nHit)
{
   const unsigned full = 0xffffffffu;
   int incl = qn;
   #pragma unroll
   for (int o = 1; o < FT_G; o <<= 1) { const int u = __shfl_up_sync(full, incl, o, FT_G); if (gl >= o) incl += u; }
   const int total = __shfl_sync(full, incl, FT_G - 1, FT_G);
   const int w0 = incl - qn;
   for (int k = 0; k < qn; ++k) myD[w0 + k] = myQ[k * 32];
   __syncwarp();
   const int rounds = (__reduce_max_sync(full, total) + FT_G - 1) / FT_G;
   for (int rd = 0; rd < rounds; ++rd)
   {
      const int k = rd * FT_G + gl;
      bool hit = false; int jHit = 0;
      if (k < total)
      {
         const int j = myD[k];
         const real_t dx = xi - sx[j], dy = yi - sy[j], dz = zi - sz[j];
         const real_t r2 = dx * dx + dy * dy + dz * dz;
         const bool in = (KIND == PK_EAM3) ? (r2 < P.rc2) : (r2 <= P.rc2);     // eam.c:317 vs :397
         if (in && j != si) pairEval<KIND>(P, dx, dy, dz, r2, dfi, KIND == PK_EAM3 ? sdf[j] : R(0.0), acc);
         hit = in && j != si;
         jHit = j;
      }
      if (KIND == PK_RHOL)
      {  // the pair leaves as an entry of atom i's row: position among the 8 lanes of the atom by ballot, the neighbour as a
         // global index (the staged rows are straight copies: one offset per stencil row)
         const unsigned m = (__ballot_sync(full, hit) >> ((threadIdx.x & 31) / FT_G * FT_G)) & ((1u << FT_G) - 1u);
         if (hit)
         {
            int r = 0;
            #pragma unroll
            for (int q = 1; q < 9; ++q) r += jHit >= H.rowS[q];
            const int pos = nHit + __popc(m & ((1u << gl) - 1u));
            if (pos < hitStride) hitRow[tileSlot(pos)] = jHit + H.rowDelta[r];
         }
         nHit += __popc(m);
      }
   }
   __syncwarp();
   qn = 0;
}

template <int KIND>
__global__ void __launch_bounds__(FT_THREADS, 2)
k_force_tile(DevGeom g, TileCfg cfg, PairParams P,
             const real_t* __restrict__ rx, const real_t* __restrict__ ry, const real_t* __restrict__ rz,
             const float4* __restrict__ rf, const int* __restrict__ cell, const int* __restrict__ cellStart,
             const real_t* __restrict__ dfIn,
             real_t* __restrict__ fx, real_t* __restrict__ fy, real_t* __restrict__ fz, real_t* __restrict__ e,
             real_t* __restrict__ rhobar, real_t* __restrict__ dfOut, int* __restrict__ cnt, TileHits hits, real4* __restrict__ rqOut)
{
   extern __shared__ __align__(128) unsigned char smem[];
   constexpr int NACC = PairTraits<KIND>::NACC;
   const unsigned full = 0xffffffffu;
   TileHdr& H = *reinterpret_cast<TileHdr*>(smem);
   const int sMax = cfg.sMax;
   real_t* sx = reinterpret_cast<real_t*>(smem + FT_HDR);
   real_t* sy = sx + sMax;
   real_t* sz = sy + sMax;
   real_t* sdf = sz + sMax;
   unsigned char* q = reinterpret_cast<unsigned char*>(KIND == PK_EAM3 ? sdf + sMax : sdf);
   float4* sf = reinterpret_cast<float4*>(q); q += (size_t)sMax * 16;
   unsigned short* qLane = reinterpret_cast<unsigned short*>(q); q += FT_WARPS * FT_QCAP * 32 * 2;
   unsigned short* qDense = reinterpret_cast<unsigned short*>(q);

   const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
   int tile = blockIdx.x;
   const int tx = tile % cfg.tilesX; tile /= cfg.tilesX;
   const int iy = tile % g.n[1], iz = tile / g.n[1];
   const int cx0 = (int)((long)tx * g.n[0] / cfg.tilesX), cx1 = (int)((long)(tx + 1) * g.n[0] / cfg.tilesX);
   const int Tx = cx1 - cx0, nk = Tx + 3;

   if (tid == 0) mbarInit(&H.mbar, 1);
   for (int t = tid; t < 9 * nk; t += FT_THREADS)
   {
      const int r = t / nk, k = t - r * nk;
      H.cs[r][k] = cellStart[cellIndex(g, cx0 - 1, iy + (r % 3) - 1, iz + (r / 3) - 1) + k];
   }
   __syncthreads();
   if (warp == 0)
   {  // row layout + TMA issue: lane r < 9 owns row r
      int gs = 0, ge = 0;
      if (lane < 9) { gs = H.cs[lane][0]; ge = H.cs[lane][nk - 1]; }
      const int ga = gs & ~(FT_AL - 1), cntR = ((ge + FT_AL - 1) & ~(FT_AL - 1)) - ga;        // 16-byte aligned ends for the real_t arrays
      int incl = cntR;
      #pragma unroll
      for (int o = 1; o < 16; o <<= 1) { const int u = __shfl_up_sync(full, incl, o); if (lane >= o) incl += u; }
      const int excl = incl - cntR;
      const int total = __shfl_sync(full, incl, 8);
      if (lane < 9) { H.rowGA[lane] = ga; H.rowS[lane] = excl; H.rowDelta[lane] = ga - excl; }
      if (lane == 0) { H.total = total; H.iBeg = H.cs[4][1]; H.iEnd = H.cs[4][Tx + 1]; }
      if (total <= sMax)
      {
         constexpr unsigned perAtom = (unsigned)sizeof(real_t) * (KIND == PK_EAM3 ? 4 : 3) + 16u;
         if (lane == 0) mbarExpectTx(&H.mbar, (unsigned)total * perAtom);
         __syncwarp();
         if (lane < 9 && cntR > 0)
         {
            const unsigned nb = (unsigned)cntR * (unsigned)sizeof(real_t);
            bulkG2S(sx + excl, rx + ga, nb, &H.mbar);
            bulkG2S(sy + excl, ry + ga, nb, &H.mbar);
            bulkG2S(sz + excl, rz + ga, nb, &H.mbar);
            if (KIND == PK_EAM3) bulkG2S(sdf + excl, dfIn + ga, nb, &H.mbar);
            bulkG2S(sf + excl, rf + ga, (unsigned)cntR * 16u, &H.mbar);
         }
      }
   }
   __syncthreads();
   if (H.total > sMax)
   {
      if (tid == 0) atomicOr(&cnt[C_ERR], ERRB_TILE);
      return;
   }
   for (int t = tid; t < 9 * nk; t += FT_THREADS)
   {
      const int r = t / nk, k = t - r * nk;
      H.cs[r][k] = H.rowS[r] + (H.cs[r][k] - H.rowGA[r]);
   }
   __syncthreads();
   mbarWait(&H.mbar, 0);

   const int g4 = lane / FT_G, gl = lane % FT_G;
   const int iBeg = H.iBeg, nI = H.iEnd - iBeg;
   const int sI0 = H.rowS[4] + (iBeg - H.rowGA[4]);
   const int cellBase = cellIndex(g, cx0, iy, iz);
   unsigned short* myQ = qLane + warp * FT_QCAP * 32 + lane;
   unsigned short* myD = qDense + (warp * FT_APW + g4) * FT_DCAP;
   const float rcf2 = cfg.rcf2;

   for (int quad = warp; FT_APW * quad < nI; quad += FT_WARPS)
   {
      const int il = FT_APW * quad + g4;
      const bool act = il < nI;
      const int si = sI0 + il;
      int cxl = 0;
      float xif = 0.f, yif = 0.f, zif = 0.f;
      real_t xi = R(0.0), yi = R(0.0), zi = R(0.0), dfi = R(0.0);
      if (act)
      {
         cxl = cell[iBeg + il] - cellBase;
         const float4 v = sf[si];
         xif = v.x; yif = v.y; zif = v.z;
         xi = sx[si]; yi = sy[si]; zi = sz[si];
         if (KIND == PK_EAM3) dfi = sdf[si];
      }
      real_t acc[5] = { R(0.0), R(0.0), R(0.0), R(0.0), R(0.0) };
      int qn = 0, nHit = 0;
      int* hitRow = KIND == PK_RHOL ? hits.list + (long)(act ? iBeg + il : iBeg) * hits.stride : nullptr;
      #pragma unroll 1
      for (int r = 0; r < 9; ++r)
      {
         int jb = 0, je = 0;
         if (act) { jb = H.cs[r][cxl]; je = H.cs[r][cxl + 3]; }
         const int maxlen = __reduce_max_sync(full, je - jb);
         #pragma unroll 1
         for (int base = 0; base < maxlen; base += FT_G * FT_CH)
         {
            if (__any_sync(full, qn > FT_QCAP - FT_CH))
               processQueues<KIND>(P, myQ, myD, qn, gl, si, xi, yi, zi, dfi, sx, sy, sz, sdf, acc, H, hitRow, hits.stride, nHit);
            const int left = (maxlen - base + FT_G - 1) / FT_G;
            const int itEnd = left < FT_CH ? left : FT_CH;
            int k = jb + base + gl;
            for (int it = 0; it < itEnd; ++it, k += FT_G)
            {
               if (k < je)
               {
                  const float4 v = sf[k];
                  const float ddx = xif - v.x, ddy = yif - v.y, ddz = zif - v.z;
                  const float r2 = ddx * ddx + ddy * ddy + ddz * ddz;
                  if (r2 <= rcf2) { myQ[qn * 32] = (unsigned short)k; ++qn; }
               }
            }
         }
      }
      processQueues<KIND>(P, myQ, myD, qn, gl, si, xi, yi, zi, dfi, sx, sy, sz, sdf, acc, H, hitRow, hits.stride, nHit);
      #pragma unroll
      for (int a = 0; a < NACC; ++a)
      {
         #pragma unroll
         for (int o = FT_G / 2; o > 0; o >>= 1) acc[a] += __shfl_xor_sync(full, acc[a], o);
      }
      if (act && gl == 0)
      {
         const int i = iBeg + il;
         if (KIND == PK_LJ)
         {
            fx[i] = acc[0]; fy[i] = acc[1]; fz[i] = acc[2];
            e[i] = (R(0.5) * acc[3]) * (R(4.0) * P.eps);            // ljForce.c:256-261
         }
         else if (KIND == PK_EAM1)
         {
            real_t fe, dfe;
            tableLookupFull(P.tF, acc[4], fe, dfe);            // eam.c:352-371 (pass 2 fused)
            fx[i] = acc[0]; fy[i] = acc[1]; fz[i] = acc[2];
            e[i] = acc[3] + fe;
            rhobar[i] = acc[4];
            dfOut[i] = dfe;
         }
         else if (KIND == PK_RHOL)
         {
            real_t fe, dfe;
            tableLookupFull(P.tF, acc[0], fe, dfe);            // eam.c:352-371 (pass 2 fused)
            e[i] = fe;                                         // U_i = F(rhobar_i); the pair energies join in k_force_nl<PK_EAMFE>
            rhobar[i] = acc[0];
            dfOut[i] = dfe;
            rqOut[i].w = dfe;                                  // the copy the list kernel gathers
            hits.count[i] = nHit <= hits.stride ? nHit : 0;
            hits.countHit[i] = nHit <= hits.stride ? nHit : 0;
            if (nHit > hits.stride) atomicOr(&cnt[C_ERR], ERRB_NLFULL);
         }
         else
         {
            fx[i] += acc[0]; fy[i] += acc[1]; fz[i] += acc[2];
         }
      }
   }
}

// ---- host launcher ----------------------------------------------------------------------
#include <cmath>
template <int KIND>
static int launchForceTile(ComdDev* d, PairParams P, const real_t* dfIn)
{
   const DevGeom& g = d->g;
   int txT = d->optTileTx;
   if (txT > FT_TXMAX) txT = FT_TXMAX;
   if (txT > g.n[0]) txT = g.n[0];
   if (txT < 1) txT = 1;
   // mean population of a cell, from whichever count the host last saw (with or without ghosts)
   const double cellsCounted = d->approxHasGhosts ? (double)g.nCells : (double)g.n[0] * g.n[1] * g.n[2];
   const double perCell = (double)(d->approxAtoms > 0 ? d->approxAtoms : d->cap) / cellsCounted;
   int sMax; size_t bytes;
   for (;;)
   {  // staging capacity: the 9 rows at 1.25x the mean density + slack; two CTAs per SM if possible
      sMax = (int)(R(9.0) * (txT + 2) * perCell * R(1.25)) + 9 * 2 * (FT_AL - 1) + 96;
      sMax = (sMax + FT_AL - 1) & ~(FT_AL - 1);
      bytes = tileSmemBytes<KIND>(sMax);
      if (bytes <= 113 * 1024 || txT == 1) break;
      --txT;
   }
   if (bytes > 227 * 1024 || sMax > 65534)
   { comdSetError("force tile: %d staged atoms (%zu B) do not fit in shared memory", sMax, bytes); return COMD_E_CAPACITY; }
   TileCfg cfg;
   cfg.tilesX = (g.n[0] + txT - 1) / txT;
   cfg.sMax = sMax;
   double ext = 0.0;
   for (int a = 0; a < 3; ++a)
   {
      const double e = (g.lMax[a] - g.lMin[a]) + 2.0 / g.inv[a];
      if (e > ext) ext = e;
   }
   const double rc = sqrt(P.rc2);
   const double rcf = rc + 4.0 * sqrt(3.0) * ext * ldexp(1.0, -24);       // FP32 rounding of both shadow positions, doubled
   cfg.rcf2 = nextafterf((float)(rcf * rcf * (1.0 + 1e-6)), INFINITY);
   const int slot = KIND == PK_EAM3 ? 2 : (KIND == PK_RHOL ? 3 : KIND);
   TileHits hits = { d->nlList, d->nlCount, d->nlCountHit, d->nlStride };
   if ((size_t)d->tileSmemSet[slot] < bytes)
   {
      CUDA_TRY(cudaFuncSetAttribute(k_force_tile<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
      d->tileSmemSet[slot] = (int)bytes;
   }
   const int grid = cfg.tilesX * g.n[1] * g.n[2];
   k_force_tile<KIND><<<grid, FT_THREADS, bytes, d->stream>>>(g, cfg, P, d->A.r[0], d->A.r[1], d->A.r[2], d->A.rf, d->A.cell,
      d->cellStart, dfIn, d->f[0], d->f[1], d->f[2], d->e, d->rhobar, d->dfEmbed, d->cnt, hits, d->A.rq);
   LAUNCH_CHECK(d);
   return COMD_OK;
}
