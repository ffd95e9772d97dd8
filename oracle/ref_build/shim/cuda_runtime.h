/* Host-only stand-in for <cuda_runtime.h>, used ONLY to compile the reference's
 * CPU path (-m cpu_nl) as the parity oracle.  "Device" memory is plain host
 * memory; streams/events are inert.  Test infrastructure, not product code. */
#ifndef ORACLE_SHIM_CUDA_RUNTIME_H
#define ORACLE_SHIM_CUDA_RUNTIME_H
#include <stdlib.h>
#include <string.h>
#include <stddef.h>
typedef int cudaError_t;
typedef struct CUstream_st* cudaStream_t;
typedef struct CUevent_st* cudaEvent_t;
enum { cudaSuccess = 0 };
enum cudaMemcpyKind { cudaMemcpyHostToHost = 0, cudaMemcpyHostToDevice = 1,
                      cudaMemcpyDeviceToHost = 2, cudaMemcpyDeviceToDevice = 3 };
struct cudaDeviceProp { char name[256]; int multiProcessorCount; size_t totalGlobalMem;
                        int major, minor; int clockRate; int memoryClockRate; int memoryBusWidth; int ECCEnabled; };
typedef struct cudaDeviceProp cudaDeviceProp;
static inline cudaError_t cudaMalloc(void** p, size_t n) { *p = calloc(1, n ? n : 1); return 0; }
static inline cudaError_t cudaMallocHost(void** p, size_t n) { *p = calloc(1, n ? n : 1); return 0; }
#ifdef REF_BIND_B200
/* CoMD-ref-b200 runs the reference's GPU set-up and tear-down (AllocateGpu / DestroyGpu): DestroyGpu hands cudaFree pointers it
 * never allocated (spline coefficients of an unused option, gpu_utility.c:330-334) -- an error code from the real runtime, a crash
 * from free().  The process ends right after, so nothing is released here. */
static inline cudaError_t cudaFree(void* p) { (void)p; return 0; }
#else
static inline cudaError_t cudaFree(void* p) { free(p); return 0; }
#endif
static inline cudaError_t cudaFreeHost(void* p) { free(p); return 0; }
static inline cudaError_t cudaMemcpy(void* d, const void* s, size_t n, int k) { (void)k; memmove(d, s, n); return 0; }
static inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, int k, cudaStream_t st) { (void)k; (void)st; memmove(d, s, n); return 0; }
static inline cudaError_t cudaMemset(void* d, int v, size_t n) { memset(d, v, n); return 0; }
static inline cudaError_t cudaMemsetAsync(void* d, int v, size_t n, cudaStream_t st) { (void)st; memset(d, v, n); return 0; }
static inline cudaError_t cudaHostRegister(void* p, size_t n, unsigned f) { (void)p; (void)n; (void)f; return 0; }
static inline cudaError_t cudaHostUnregister(void* p) { (void)p; return 0; }
static inline cudaError_t cudaDeviceSynchronize(void) { return 0; }
static inline cudaError_t cudaStreamSynchronize(cudaStream_t s) { (void)s; return 0; }
static inline cudaError_t cudaStreamCreate(cudaStream_t* s) { *s = 0; return 0; }
static inline cudaError_t cudaStreamCreateWithPriority(cudaStream_t* s, unsigned f, int pr) { (void)f; (void)pr; *s = 0; return 0; }
static inline cudaError_t cudaStreamDestroy(cudaStream_t s) { (void)s; return 0; }
static inline cudaError_t cudaDeviceGetStreamPriorityRange(int* lo, int* hi) { *lo = 0; *hi = 0; return 0; }
static inline cudaError_t cudaEventCreate(cudaEvent_t* e) { *e = 0; return 0; }
static inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t* e, unsigned f) { (void)f; *e = 0; return 0; }
static inline cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t s) { (void)e; (void)s; return 0; }
static inline cudaError_t cudaEventSynchronize(cudaEvent_t e) { (void)e; return 0; }
static inline cudaError_t cudaEventDestroy(cudaEvent_t e) { (void)e; return 0; }
static inline cudaError_t cudaStreamWaitEvent(cudaStream_t s, cudaEvent_t e, unsigned f) { (void)s; (void)e; (void)f; return 0; }
static inline cudaError_t cudaGetLastError(void) { return 0; }
static inline const char* cudaGetErrorString(cudaError_t e) { (void)e; return "oracle shim: no device"; }
static inline cudaError_t cudaGetDeviceCount(int* n) { *n = 1; return 0; }
static inline cudaError_t cudaGetDevice(int* d) { *d = 0; return 0; }
static inline cudaError_t cudaSetDevice(int d) { (void)d; return 0; }
static inline cudaError_t cudaDeviceReset(void) { return 0; }
static inline cudaError_t cudaGetDeviceProperties(struct cudaDeviceProp* p, int d) {
   (void)d; memset(p, 0, sizeof(*p)); strcpy(p->name, "host (oracle shim)"); p->multiProcessorCount = 1;
   p->clockRate = 1; p->memoryClockRate = 1; p->memoryBusWidth = 64; return 0; }
#define cudaStreamNonBlocking 0x01
#define cudaHostRegisterDefault 0
#endif
