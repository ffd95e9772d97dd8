/* Link-time stand-ins for the symbols the reference's host C leaves undefined
 * once gpu_kernels.cu and comm.cc are dropped.  Only the
 * CPU path (-m cpu_nl) is ever run through this build, so every device entry
 * point aborts if reached.  Test infrastructure only. */
#include <stdio.h>
#include <stdlib.h>

#define DEAD(name) void name(void) { fprintf(stderr, "oracle/_ref: device entry point %s reached on the CPU oracle\n", #name); abort(); }
#define NOOP(name) void name(void) { }
#define ZERO(name) int name(void) { return 0; }

/* the libmp layer (comm.h): "not in use" is the answer the reference's host code branches on */
ZERO(comm_use_comm) ZERO(comm_use_async) ZERO(comm_use_gpu_comm) ZERO(comm_use_gdrdma)
DEAD(comm_flush) DEAD(comm_irecv) DEAD(comm_isend) DEAD(comm_isend_on_stream) DEAD(comm_progress)
DEAD(comm_send_ready) DEAD(comm_send_ready_on_stream) DEAD(comm_wait_all) DEAD(comm_wait_all_on_stream)
DEAD(comm_wait_ready_on_stream) DEAD(comm_init) DEAD(comm_finalize) ZERO(comm_select_device)   /* -DDO_MPI main asks it for a device id (CoMD.c:112) */
DEAD(exchangeDataForceGpu_KI)

#ifndef REF_BIND_B200      /* CoMD-ref-b200 takes the device entry points from ref_bind_b200.c (libcomd_b200.so behind them) */
NOOP(emptyNeighborListGpu) NOOP(initHashTableGpu) NOOP(emptyHashTableGpu)

DEAD(advancePositionGpu) DEAD(advanceVelocityGpu) DEAD(buildAtomListGpu) DEAD(buildNeighborListGpu)
DEAD(compactCellsGpu) DEAD(computeEnergy)
DEAD(eamForce1Gpu) DEAD(eamForce1GpuAsync) DEAD(eamForce2Gpu) DEAD(eamForce2GpuAsync)
DEAD(eamForce3Gpu) DEAD(eamForce3GpuAsync) DEAD(getAtomMsgSoAPtr) DEAD(ljForceGpu)
DEAD(loadAtomsBufferFromGpu_Async) DEAD(loadAtomsBufferFromGpu_Comm) DEAD(loadForceBufferFromGpu)
DEAD(loadForceBufferFromGpu_Async) DEAD(loadForceBufferFromGpu_Comm)
DEAD(neighborListUpdateRequiredGpu) DEAD(pairlistUpdateRequiredGpu) DEAD(sortAtomsGpu)
DEAD(unloadAtomsBufferToGpu) DEAD(unloadAtomsBufferToGpu_Async) DEAD(unloadAtomsBufferToGpu_Comm)
DEAD(unloadForceBufferToGpu) DEAD(unloadForceBufferToGpu_Async) DEAD(unloadForceBufferToGpu_Comm)
DEAD(unloadForceScanCells) DEAD(updateLinkCellsGpu) DEAD(updateNeighborsGpu) DEAD(updateNeighborsGpuAsync)
#endif

int mpi_comm_rank = 0;
