#ifndef ORACLE_SHIM_CUDA_H
#define ORACLE_SHIM_CUDA_H
#define CUDA_VERSION 12090
#include "cuda_runtime.h"
#endif
