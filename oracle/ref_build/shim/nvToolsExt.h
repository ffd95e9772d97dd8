/* inert NVTX stand-in for the oracle build */
#ifndef ORACLE_SHIM_NVTX_H
#define ORACLE_SHIM_NVTX_H
#include <stdint.h>
#define NVTX_VERSION 1
#define NVTX_EVENT_ATTRIB_STRUCT_SIZE 0
#define NVTX_COLOR_ARGB 1
#define NVTX_MESSAGE_TYPE_ASCII 1
typedef struct { int version, size, colorType; uint32_t color; int messageType; struct { const char* ascii; } message; } nvtxEventAttributes_t;
static inline int nvtxRangePushEx(const nvtxEventAttributes_t* a) { (void)a; return 0; }
static inline int nvtxRangePushA(const char* s) { (void)s; return 0; }
static inline int nvtxRangePop(void) { return 0; }
#endif
