// gather_bench.cu -- how fast can one SM gather random 32-byte records?  LDG.256 vs 2x LDG.128 vs
// texture fetches (int4 texels) vs shared memory (2x LDS.128).  Decides where the EAM tables live.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/micro/gather_bench.cu -o gpurun_out/gather_bench
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ unsigned lcg(unsigned& s) { s = s * 1664525u + 1013904223u; return s >> 8; }

// hot window: `span` records starting at `base`; every lane draws an independent index each iteration
template <int MODE>
__global__ void __launch_bounds__(256) k_gather(const double4* __restrict__ tab, cudaTextureObject_t tex, int base, int span, int iters, double* out)
{
   extern __shared__ double4 sm[];
   if (MODE == 3) { for (int i = threadIdx.x; i < span; i += blockDim.x) sm[i] = tab[base + i]; __syncthreads(); }
   unsigned s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
   double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
   #pragma unroll 4
   for (int it = 0; it < iters; ++it)
   {
      const int idx = (int)(lcg(s) % (unsigned)span);
      double4 v;
      if (MODE == 0) asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w) : "l"(tab + base + idx));
      else if (MODE == 1)
      {
         const double2* p = reinterpret_cast<const double2*>(tab + base + idx);
         const double2 u = __ldg(p), w = __ldg(p + 1);
         v = make_double4(u.x, u.y, w.x, w.y);
      }
      else if (MODE == 2)
      {
         const int4 u = tex1Dfetch<int4>(tex, 2 * (base + idx)), w = tex1Dfetch<int4>(tex, 2 * (base + idx) + 1);
         v = make_double4(__hiloint2double(u.y, u.x), __hiloint2double(u.w, u.z), __hiloint2double(w.y, w.x), __hiloint2double(w.w, w.z));
      }
      else v = sm[idx];
      a0 += v.x; a1 += v.y; a2 += v.z; a3 += v.w;
   }
   out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3;
}

int main()
{
   const int N = 10003;
   std::vector<double4> h(N);
   for (int i = 0; i < N; ++i) h[i] = make_double4(i, 2 * i, 3 * i, 4 * i);
   double4* d; CK(cudaMalloc(&d, N * sizeof(double4))); CK(cudaMemcpy(d, h.data(), N * sizeof(double4), cudaMemcpyHostToDevice));
   cudaResourceDesc rd = {}; rd.resType = cudaResourceTypeLinear; rd.res.linear.devPtr = d; rd.res.linear.sizeInBytes = N * sizeof(double4);
   rd.res.linear.desc = cudaCreateChannelDesc<int4>();
   cudaTextureDesc td = {}; td.readMode = cudaReadModeElementType;
   cudaTextureObject_t tex; CK(cudaCreateTextureObject(&tex, &rd, &td, nullptr));
   const int blocks = 148 * 4, threads = 256, iters = 4096;
   double* out; CK(cudaMalloc(&out, blocks * threads * sizeof(double)));
   cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
   const char* names[4] = { "LDG.256", "2xLDG.128", "2xTEX int4", "smem 2xLDS.128" };
   const int spans[3] = { 256, 2200, 6400 };              // 8 KB, 70 KB (hot set of 4 shells), 205 KB
   for (int sp = 0; sp < 3; ++sp)
      for (int mode = 0; mode < 4; ++mode)
      {
         const int span = spans[sp];
         size_t smem = mode == 3 ? span * sizeof(double4) : 0;
         if (smem > 200 * 1024) continue;
         float best = 1e30f;
         for (int rep = 0; rep < 4; ++rep)
         {
            cudaEventRecord(a);
            if (mode == 0) k_gather<0><<<blocks, threads, 0>>>(d, tex, 3600, span, iters, out);
            if (mode == 1) k_gather<1><<<blocks, threads, 0>>>(d, tex, 3600, span, iters, out);
            if (mode == 2) k_gather<2><<<blocks, threads, 0>>>(d, tex, 3600, span, iters, out);
            if (mode == 3) { CK(cudaFuncSetAttribute(k_gather<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); k_gather<3><<<blocks, threads, smem>>>(d, tex, 3600, span, iters, out); }
            cudaEventRecord(b); CK(cudaEventSynchronize(b));
            float ms; cudaEventElapsedTime(&ms, a, b); if (rep && ms < best) best = ms;
         }
         const double lookups = (double)blocks * threads * iters;
         printf("span %5d recs (%4d KB) %-16s %8.3f ms  %6.2f G lookups/s  %5.2f clk/warp-lookup/SM\n", span, span * 32 / 1024, names[mode], best,
                lookups / best * 1e-6, best * 1e-3 * 1.965e9 * 148 / (lookups / 32));
      }
   CK(cudaGetLastError());
   return 0;
}
