// atoms_bench.cu -- what does the j-side scatter of a Newton-3 list kernel cost on the L1/shared data pipe?
// (sm_100a has a native shared-memory atomic add for 32-bit integers only: ATOMS.ADD; 64-bit integer and every
//  floating-point shared atomicAdd compile to an ATOMS.CAST.SPIN loop -- cuobjdump of this file shows it.)
// Patterns, one warp-instruction group per iteration, accumulators = NSLOT atoms x STRIDE words in shared memory:
//   M1  1 x ATOMS.ADD.32, 32 random slots
//   M2  2 x ATOMS.ADD.32 (hi/lo of one fixed-point value), 32 random slots            [pass 1: rho_j]
//   M3  6 x ATOMS.ADD.32 (hi/lo of x,y,z), 32 random slots                             [pass 3: f_j]
//   M4  6 x ATOMS.ADD.32, 4 groups of 8 lanes, each group two runs of 4 adjacent slots [sorted neighbours]
//   M5  3 x atomicAdd(u64) (CAS loop), 32 random slots
//   M6  3 x (DMUL + F2I.S64.F64 + split) only, no memory                               [conversion cost]
//   M7  LDG.256 gather (32 random sectors of a 70 KB window) + M3 in the same iteration [do they add up?]
//   M8  LDG.256 gather alone (reference, = P1 of gather_lines.cu)
//   M9  LDG.128 gather, 32 random lines;  M10  LDG.64 gather, 32 random lines
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/micro/atoms_bench.cu -o gpurun_out/atoms_bench
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1); } } while (0)
__device__ __forceinline__ unsigned mix(unsigned x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }
#define NSLOT 1200
#define STRIDE 7

__device__ __forceinline__ void addFix(unsigned* a, double v, double scale)
{
   const long long q = __double2ll_rn(v * scale);
   atomicAdd(a, (unsigned)(q >> 24));
   atomicAdd(a + 1, (unsigned)q & 0xffffffu);
}

template <int MODE>
__global__ void __launch_bounds__(1024, 1) k(const double4* __restrict__ tab, int span, int iters, double* out)
{
   extern __shared__ unsigned acc[];
   for (int t = threadIdx.x; t < NSLOT * STRIDE + 64; t += blockDim.x) acc[t] = 0;
   __syncthreads();
   const int lane = threadIdx.x & 31, warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
   double a0 = 0, a1 = 0;
   unsigned long long* acc64 = reinterpret_cast<unsigned long long*>(acc);
   #pragma unroll 2
   for (int it = 0; it < iters; ++it)
   {
      const unsigned sw = warp * 7919u + it * 104729u;
      int slot = mix(sw + lane * 31u) % NSLOT;
      if (MODE == 4) slot = (4 * (mix(sw + (lane >> 2) * 31u) % (NSLOT / 4)) + (lane & 3)) % NSLOT;
      const double v = 1e-3 * (double)(lane + 1) + 1e-6 * it;
      if (MODE == 1) atomicAdd(acc + slot * STRIDE, (unsigned)it);
      if (MODE == 2) addFix(acc + slot * STRIDE, v, 1125899906842624.0);
      if (MODE == 3 || MODE == 4 || MODE == 7)
      {
         addFix(acc + slot * STRIDE, v, 70368744177664.0);
         addFix(acc + slot * STRIDE + 2, -v, 70368744177664.0);
         addFix(acc + slot * STRIDE + 4, v * 0.5, 70368744177664.0);
      }
      if (MODE == 5)
      {
         atomicAdd(acc64 + slot * 3, (unsigned long long)__double2ll_rn(v * 70368744177664.0));
         atomicAdd(acc64 + slot * 3 + 1, (unsigned long long)__double2ll_rn(-v * 70368744177664.0));
         atomicAdd(acc64 + slot * 3 + 2, (unsigned long long)__double2ll_rn(0.5 * v * 70368744177664.0));
      }
      if (MODE == 6)
      {
         const long long q0 = __double2ll_rn((v + a0) * 70368744177664.0), q1 = __double2ll_rn((a0 - v) * 70368744177664.0), q2 = __double2ll_rn((v * 0.5 + a0) * 70368744177664.0);
         a0 += (double)(int)((q0 >> 24) + (q1 >> 24) + (q2 >> 24)) * 1e-30; a1 += (double)(int)(((unsigned)q0 & 0xffffffu) + ((unsigned)q1 & 0xffffffu) + ((unsigned)q2 & 0xffffffu)) * 1e-30;
      }
      if (MODE == 7 || MODE == 8)
      {
         const int idx = mix(sw + lane * 131u + 7u) % span;
         double4 w;
         asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(w.x), "=d"(w.y), "=d"(w.z), "=d"(w.w) : "l"(tab + idx));
         a0 += w.x + w.z; a1 += w.y + w.w;
      }
      if (MODE == 9)
      {
         const int idx = mix(sw + lane * 131u + 7u) % span;
         double2 w;
         asm volatile("ld.global.nc.v2.f64 {%0,%1}, [%2];" : "=d"(w.x), "=d"(w.y) : "l"(reinterpret_cast<const double2*>(tab + idx)));
         a0 += w.x; a1 += w.y;
      }
      if (MODE == 10)
      {
         const int idx = mix(sw + lane * 131u + 7u) % span;
         double w;
         asm volatile("ld.global.nc.f64 %0, [%1];" : "=d"(w) : "l"(reinterpret_cast<const double*>(tab + idx)));
         a0 += w;
      }
   }
   __syncthreads();
   double s = a0 + a1;
   for (int t = threadIdx.x; t < NSLOT * STRIDE; t += blockDim.x) s += acc[t];
   out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main()
{
   const int N = 2200;
   double4* d; CK(cudaMalloc(&d, N * sizeof(double4))); CK(cudaMemset(d, 0, N * sizeof(double4)));
   const int blocks = 148, threads = 1024, iters = 4096;
   double* out; CK(cudaMalloc(&out, blocks * threads * sizeof(double)));
   cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
   const char* names[11] = { "", "M1 1xATOMS.32 random", "M2 2xATOMS.32 random (rho_j)", "M3 6xATOMS.32 random (f_j)", "M4 6xATOMS.32 runs of 4", "M5 3x atomicAdd u64 CAS",
                             "M6 3x convert only", "M7 LDG.256 gather + M3", "M8 LDG.256 gather alone", "M9 LDG.128 32 lines", "M10 LDG.64 32 lines" };
   const size_t smem = (NSLOT * STRIDE + 64) * sizeof(unsigned);
#define RUN(M) CK(cudaFuncSetAttribute(k<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); k<M><<<blocks, threads, smem>>>(d, N, iters, out)
   for (int mode = 1; mode <= 10; ++mode)
   {
      float best = 1e30f;
      for (int rep = 0; rep < 4; ++rep)
      {
         cudaEventRecord(a);
         switch (mode) { case 1: RUN(1); break; case 2: RUN(2); break; case 3: RUN(3); break; case 4: RUN(4); break; case 5: RUN(5); break;
                         case 6: RUN(6); break; case 7: RUN(7); break; case 8: RUN(8); break; case 9: RUN(9); break; case 10: RUN(10); break; }
         cudaEventRecord(b); CK(cudaEventSynchronize(b));
         float ms; cudaEventElapsedTime(&ms, a, b); if (rep && ms < best) best = ms;
      }
      const double loads = (double)blocks * threads / 32 * iters;
      printf("%-32s %8.3f ms  %6.2f clk per warp-iteration per SM (at 1.965 GHz)\n", names[mode], best, best * 1e-3 * 1.965e9 * 148 / loads);
   }
   CK(cudaGetLastError());
   return 0;
}
