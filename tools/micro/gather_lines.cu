// gather_lines.cu -- what does an LDG.E.256 gather cost in the L1 data pipe: per lane, per distinct 32-byte
// sector, or per distinct 128-byte line?  Patterns over a 70 KB hot window (always L1-resident):
//   P1  32 lanes -> 32 random sectors (random lines)                 [table lookups]
//   P2  8 groups of 4 lanes -> 8 random lines, 4 distinct sectors in each
//   P3  8 groups of 4 lanes -> 8 random sectors, every sector read by 4 lanes (broadcast) [4 atoms sharing 8 neighbours]
//   P4  4 groups of 8 lanes -> per group ~4.4 lines / 8 sectors (two runs of 4 adjacent sectors) [today's position gathers]
//   P5  32 lanes -> 1 sector (full broadcast)
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/micro/gather_lines.cu -o gpurun_out/gather_lines
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1); } } while (0)
__device__ __forceinline__ unsigned mix(unsigned x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }

template <int MODE>
__global__ void __launch_bounds__(256) k(const double4* __restrict__ tab, int span, int iters, double* out)
{
   const int lane = threadIdx.x & 31, warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
   double a0 = 0, a1 = 0;
   #pragma unroll 4
   for (int it = 0; it < iters; ++it)
   {
      const unsigned sw = warp * 7919u + it * 104729u;
      int idx;
      if (MODE == 1) idx = mix(sw + lane * 31u) % span;
      else if (MODE == 2) idx = 4 * (mix(sw + (lane >> 2) * 31u) % (span / 4)) + (lane & 3);
      else if (MODE == 3) idx = mix(sw + (lane >> 2) * 31u) % span;
      else if (MODE == 4) idx = (4 * (mix(sw + (lane >> 2) * 31u) % (span / 4)) + (lane & 3) + (mix(sw + (lane >> 3)) & 3)) % span;
      else idx = mix(sw) % span;
      double4 v;
      asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w) : "l"(tab + idx));
      a0 += v.x + v.z; a1 += v.y + v.w;
   }
   out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1;
}

int main()
{
   const int N = 2200;
   double4* d; CK(cudaMalloc(&d, N * sizeof(double4))); CK(cudaMemset(d, 0, N * sizeof(double4)));
   const int blocks = 148 * 4, threads = 256, iters = 4096;
   double* out; CK(cudaMalloc(&out, blocks * threads * sizeof(double)));
   cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
   const char* names[6] = { "", "P1 32 random sectors", "P2 8 lines x 4 sectors", "P3 8 sectors x 4 lanes", "P4 ~4.4 lines per 8 lanes", "P5 one sector" };
   for (int mode = 1; mode <= 5; ++mode)
   {
      float best = 1e30f;
      for (int rep = 0; rep < 4; ++rep)
      {
         cudaEventRecord(a);
         if (mode == 1) k<1><<<blocks, threads>>>(d, N, iters, out);
         if (mode == 2) k<2><<<blocks, threads>>>(d, N, iters, out);
         if (mode == 3) k<3><<<blocks, threads>>>(d, N, iters, out);
         if (mode == 4) k<4><<<blocks, threads>>>(d, N, iters, out);
         if (mode == 5) k<5><<<blocks, threads>>>(d, N, iters, out);
         cudaEventRecord(b); CK(cudaEventSynchronize(b));
         float ms; cudaEventElapsedTime(&ms, a, b); if (rep && ms < best) best = ms;
      }
      const double loads = (double)blocks * threads / 32 * iters;
      printf("%-28s %8.3f ms  %6.2f clk per warp-load per SM (at 1.965 GHz)\n", names[mode], best, best * 1e-3 * 1.965e9 * 148 / loads);
   }
   CK(cudaGetLastError());
   return 0;
}
