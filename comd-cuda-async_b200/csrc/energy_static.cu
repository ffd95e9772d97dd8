// energy_static.cu -- comd_energy_sums_static: the potential + kinetic energy sums of a set of atoms, computed
// without a single device allocation.  No counterpart in the reference.
//
// Why it exists.  comd_dev_energy() reduces in two levels through a per-handle buffer of block partials that
// comd_dev_create() got from cudaMalloc.  Two kinds of caller cannot go that way:
//   * an analysis hook that a host program calls from inside its time-step loop while halo traffic and force
//     kernels are in flight on other streams: cudaMalloc / cudaFree wait for the whole device, so an allocation
//     there stalls exactly the overlap the step was arranged for;
//   * a host that has already handed all of HBM to its own arena or caching allocator (the reference grabs its
//     per-rank slabs once in AllocateGpu; a torch process does the same through its caching allocator): a
//     cudaMalloc issued by a library underneath it can simply fail, and the host has no way to lend the library
//     memory through the handle-based interface.
// Such a caller already owns the per-atom arrays and a place for the two results.  What it lacks is the scratch
// between the two reduction levels, so that scratch lives here, in a file-scope __device__ array that is part of
// the library image: every GPU that loads the library carries its 256 KB from module load until the process exits.
//
// Limits and sharing rules (also in include/comd_b200.h):
//   * at most COMD_ENERGY_STATIC_MAX_ATOMS atoms per call (one partial pair per 4096-atom tile, 16384 tiles);
//     a larger count is refused with COMD_E_ARG before anything is launched;
//   * within one process there is ONE array per GPU and every call on that GPU uses it.  Calls that follow each
//     other on one stream are fine.  Calls whose kernels may be on the GPU at the same time (issued to different
//     streams, from one host thread or several) race on it and return garbage; order them with an event.
//     Another process has its own copy of the module and so its own array;
//   * the array keeps whatever the previous call left.  Nothing in it has to start at zero: a call writes
//     every slot it later reads (tiles 0 .. nTiles-1) and reads no other, so stale partials of a longer
//     earlier call are never picked up;
//   * the first call on each GPU looks the array's device address up (cudaGetSymbolAddress, which also makes the
//     runtime load this module on that GPU).  Make that first call during set-up, outside any stream capture.
#include <mutex>
#include "comd_dev.cuh"

#define ES_TILE        4096                         // atoms reduced by one CTA
#define ES_THREADS     256
#define ES_MAX_TILES   16384                        // 16384 * 4096 = COMD_ENERGY_STATIC_MAX_ATOMS
#define ES_MAX_DEVICES 64

static_assert((long)ES_TILE * ES_MAX_TILES == COMD_ENERGY_STATIC_MAX_ATOMS, "header and scratch size disagree");

// The scratch: (U, K) partial of every tile.  In host code below, `g_esPartial` is only a handle for the runtime's
// symbol table -- NOT a device address; the address a kernel may be given comes from cudaGetSymbolAddress.
static __device__ __align__(256) double g_esPartial[2 * ES_MAX_TILES];

// Level 1: CTA t sums atoms [t*ES_TILE, min(n, (t+1)*ES_TILE)).  Same per-atom expressions and the same
// double accumulators as k_energy_partial (comd_core.cu); the tiling is fixed by n alone, so the result does not
// depend on the grid the launch happens to get.
__global__ void __launch_bounds__(ES_THREADS)
k_es_partial(long n, real_t halfInvMass, const real_t* __restrict__ e, const real_t* __restrict__ px,
             const real_t* __restrict__ py, const real_t* __restrict__ pz, double* __restrict__ partial)
{
   __shared__ double sU[ES_THREADS / 32], sK[ES_THREADS / 32];
   const long lo = (long)blockIdx.x * ES_TILE;
   const long hi = lo + ES_TILE < n ? lo + ES_TILE : n;
   double u = 0.0, k = 0.0;
   for (long i = lo + threadIdx.x; i < hi; i += ES_THREADS)
   {
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
      for (int i = 0; i < ES_THREADS / 32; ++i) { a += sU[i]; b += sK[i]; }
      partial[2 * blockIdx.x] = a; partial[2 * blockIdx.x + 1] = b;
   }
}

// Level 2: one CTA folds the nTiles partials this call wrote (and only those) in a fixed order.
__global__ void __launch_bounds__(1024)
k_es_final(const double* __restrict__ partial, int nTiles, double* __restrict__ out)
{
   __shared__ double sU[32], sK[32];
   double u = 0.0, k = 0.0;
   for (int i = threadIdx.x; i < nTiles; i += 1024) { u += partial[2 * i]; k += partial[2 * i + 1]; }
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

// Device address of g_esPartial on each GPU, looked up once per GPU with that GPU current.
static double*    g_esAddr[ES_MAX_DEVICES];
static std::mutex g_esLock;

static int esScratch(int device, double** out)
{
   std::lock_guard<std::mutex> hold(g_esLock);
   if (!g_esAddr[device])
   {
      double* p = nullptr;
      CUDA_TRY(cudaGetSymbolAddress((void**)&p, g_esPartial));
      if (!p) { comdSetError("comd_energy_sums_static: no device address for the scratch array on GPU %d", device); return COMD_E_CUDA; }
      g_esAddr[device] = p;
   }
   *out = g_esAddr[device];
   return COMD_OK;
}

extern "C" int comd_energy_sums_static(int deviceId, void* stream, long nAtoms, double mass,
                                       const comd_real* U, const comd_real* px, const comd_real* py, const comd_real* pz,
                                       double* eOut)
{
   if (nAtoms < 0 || !(mass > 0.0) || !eOut || (nAtoms > 0 && (!U || !px || !py || !pz)))
   { comdSetError("comd_energy_sums_static: bad argument (nAtoms %ld, mass %g, or a null array)", nAtoms, mass); return COMD_E_ARG; }
   if (nAtoms > COMD_ENERGY_STATIC_MAX_ATOMS)
   {
      comdSetError("comd_energy_sums_static: %ld atoms exceed the %ld the built-in scratch array holds partials for; "
                   "split the call or use comd_dev_energy", nAtoms, (long)COMD_ENERGY_STATIC_MAX_ATOMS);
      return COMD_E_ARG;
   }
   if (deviceId < 0 || deviceId >= ES_MAX_DEVICES) { comdSetError("comd_energy_sums_static: device id %d outside 0..%d", deviceId, ES_MAX_DEVICES - 1); return COMD_E_ARG; }
   CUDA_TRY(cudaSetDevice(deviceId));
   double* partial = nullptr;
   int rc = esScratch(deviceId, &partial);
   if (rc != COMD_OK) return rc;
   cudaStream_t s = (cudaStream_t)stream;
   const int nTiles = (int)((nAtoms + ES_TILE - 1) / ES_TILE);
   if (nTiles > 0)
   {
      k_es_partial<<<nTiles, ES_THREADS, 0, s>>>(nAtoms, (real_t)(R(0.5) / mass), U, px, py, pz, partial);
      CUDA_TRY(cudaGetLastError());
   }
   k_es_final<<<1, 1024, 0, s>>>(partial, nTiles, eOut);
   CUDA_TRY(cudaGetLastError());
   return COMD_OK;
}
