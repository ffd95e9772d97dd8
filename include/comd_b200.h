/* comd_b200.h -- C ABI of the B200-native device layer for CoMD's short-range force path.
 *
 * This is the drop-in boundary: the set of entry points the reference's host C
 * (timestep.c, ljForce.c, eam.c, haloExchange.c, CoMD.c) reaches its GPU through.
 * The reference declares them in src-mpi/gpu_kernels.h:13-92 and gpu_utility.h:55-69
 * and passes a ~600-byte SimGpu struct of raw device pointers BY VALUE
 * (gpu_types.h:159-190).  Here every call takes one opaque per-rank handle instead;
 * each declaration below names the reference interface it replaces.
 * INTEGRATION.md shows the one-line rebinding of each reference call site.
 *
 * Conventions
 *   - plain C linkage, plain pointers and sizes; no C++/torch types.
 *   - every function returns 0 on success, a negative COMD_E_* code on failure;
 *     comd_last_error() returns a human-readable message for the calling thread.
 *     (The reference prints and exit(-1)s on any CUDA error, gpu_utility.h:71-84;
 *     a library must not, so the error is returned and the host decides.)
 *   - one host thread per rank, one rank per GPU (CoMD.c:105-117).
 *   - all work is enqueued on the handle's own stream; calls that return values to
 *     the host synchronise that stream, the others do not.
 *   - atom data are structure-of-arrays, FP64 (COMD_DOUBLE; -DCOMD_SINGLE builds
 *     switch comd_real to float like mytype.h:8-20).
 *   - there is no CPU fallback: without a CUDA device every call fails with
 *     COMD_E_CUDA.
 */
#ifndef COMD_B200_H
#define COMD_B200_H

#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

/* Per-atom arrays (r, p, f, U) cross this interface as comd_real, the reference's real_t (mytype.h:11-24):
 * double in libcomd_b200.so, float in libcomd_b200_f32.so (built with -DCOMD_SINGLE, the reference's
 * "DOUBLE_PRECISION = OFF").  Scalars, geometry, tables and energy sums stay double in both; the library
 * rounds tables to comd_real when it uploads them.  comd_precision() tells which one was loaded. */
#ifdef COMD_SINGLE
typedef float comd_real;
#else
typedef double comd_real;
#endif

#define COMD_OK            0
#define COMD_E_CUDA       -1   /* a CUDA runtime call failed (or no device)          */
#define COMD_E_ARG        -2   /* bad argument                                       */
#define COMD_E_CAPACITY   -3   /* atom / message capacity exceeded                   */
#define COMD_E_STATE      -4   /* call order violated (e.g. force before potential)  */
#define COMD_E_COMM       -5   /* NCCL failure                                       */
#define COMD_E_MISMATCH   -6   /* F' halo: ghost-cell population differs from sender */

typedef struct ComdDev ComdDev;      /* opaque; replaces SimGpu (gpu_types.h:159-190) */

/* Local link-cell geometry + domain decomposition of one rank.
 * Filled by the host exactly as initLinkCells (linkCells.c:122-182) and
 * initDecomposition (decomposition.c:18-50) do; the device never re-derives
 * boxSize/invBoxSize so the ownership arithmetic stays bit-identical. */
typedef struct ComdGeometry
{
   int    gridSize[3];       /* local cells per axis (LinkCell.gridSize)           */
   double localMin[3];       /* LinkCell.localMin                                   */
   double localMax[3];       /* LinkCell.localMax                                   */
   double boxSize[3];        /* LinkCell.boxSize                                    */
   double invBoxSize[3];     /* LinkCell.invBoxSize                                 */
   double globalExtent[3];   /* Domain.globalExtent                                 */
   int    procGrid[3];       /* Domain.procGrid                                     */
   int    procCoord[3];      /* Domain.procCoord                                    */
} ComdGeometry;

/* Uniform-grid table as the reference stores it (InterpolationObject,
 * CoMDTypes.h:169-176): values points at entry 0 of an array valid on [-1, n+1]
 * (padding of initInterpolationObject, eam.c:496-519). */
typedef struct ComdTable
{
   int           n;
   double        x0;
   double        invDx;
   const double* values;
} ComdTable;

/* ---- lifecycle: SetupGpu / AllocateGpu / DestroyGpu (gpu_utility.h:55-62) -------- */
int  comd_dev_create(ComdDev** out, int deviceId, const ComdGeometry* geom, long atomCapacity);
int  comd_dev_destroy(ComdDev* dev);
const char* comd_last_error(void);
const char* comd_version(void);
int  comd_precision(void);                  /* sizeof(comd_real): 8 or 4 */

/* ---- potential + species: the potential block of CopyDataToGpu (gpu_utility.c:432-600) */
int  comd_dev_set_mass(ComdDev* dev, double massInternalUnits);
int  comd_dev_set_lj(ComdDev* dev, double sigma, double epsilon, double cutoff);              /* LjPotentialGpu */
int  comd_dev_set_eam(ComdDev* dev, double cutoff,
                      const ComdTable* phi, const ComdTable* rho, const ComdTable* f);         /* EamPotentialGpu */

/* ---- atoms in / out: CopyDataToGpu (gpu_utility.c:432), GetLocalAtomsFromGpu (:602) --
 * Upload takes the rank's local atoms in any order (host pointers, SoA).  It bins
 * them, builds the ghost shell and sorts (i.e. it ends with comd_dev_redistribute).
 * Download returns LOCAL atoms in device order (cell-major in the reference's own box
 * numbering is NOT implied; key by gid).  Any output pointer may be NULL. */
int  comd_dev_upload_atoms(ComdDev* dev, int nLocal, const int* gid, const int* species,
                           const comd_real* rx, const comd_real* ry, const comd_real* rz,
                           const comd_real* px, const comd_real* py, const comd_real* pz);
int  comd_dev_download_atoms(ComdDev* dev, int* nLocalOut, int capacity,
                             int* gid, int* species, int* refBox,
                             comd_real* rx, comd_real* ry, comd_real* rz,
                             comd_real* px, comd_real* py, comd_real* pz,
                             comd_real* fx, comd_real* fy, comd_real* fz, comd_real* U);
int  comd_dev_num_local(ComdDev* dev, int* nLocal);        /* sumAtoms input (CoMD.c:442-457) */
int  comd_dev_num_total(ComdDev* dev, int* nLocalPlusGhost);
/* per-cell populations in the REFERENCE's box numbering (getBoxFromTuple,
 * linkCells.c:299-346), nTotalBoxes entries: what CoMD.c:445 meant to copy back. */
int  comd_dev_cell_counts(ComdDev* dev, int* nAtomsPerRefBox, int nTotalBoxes);
/* all atoms incl. ghosts, for the halo parity tests */
int  comd_dev_download_all(ComdDev* dev, int* nOut, int capacity, int* gid, int* refBox,
                           comd_real* rx, comd_real* ry, comd_real* rz);

/* ---- binning + atom halo: updateLinkCellsGpu (gpu_kernels.cu:470), haloExchange of
 * atoms (haloExchange.c:732; pack/unpack gpu_kernels.cu:507-638), buildAtomListGpu
 * (:554), sortAtomsGpu (:1014): one call = redistributeAtoms (timestep.c:212-276). */
int  comd_dev_redistribute(ComdDev* dev);

/* ---- forces: ljForceGpu (gpu_kernels.cu:70), eamForce{1,2,3}Gpu (:282-325) with the
 * F'(rho) halo exchange between pass 2 and 3 (eam.c:241). = pot->force(sim). */
int  comd_dev_compute_force(ComdDev* dev);

/* ---- integrator: advanceVelocityGpu / advancePositionGpu (gpu_kernels.cu:327-349) */
int  comd_dev_advance_velocity(ComdDev* dev, double dt);
int  comd_dev_advance_position(ComdDev* dev, double dt);
/* computeEnergy (gpu_kernels.cu:1046): eLocal[0] = potential, eLocal[1] = kinetic (rank-local sums) */
int  comd_dev_energy(ComdDev* dev, double eLocal[2]);
/* nSteps of timestep() (timestep.c:48-100) without returning to the host in between:
 * kick, drift, redistribute, force, kick.  Leaves r, p, f and the per-atom energies U at
 * t + nSteps*dt.  On the neighbour-list path the EAM pair energies, which only the caller can
 * observe, are evaluated at the last of the nSteps force calls (the forces of every step come
 * from the same instructions either way); option "energy_every_step" restores eam.c's schedule. */
int  comd_dev_timestep(ComdDev* dev, int nSteps, double dt);

/* ---- neighbour list (optional path): buildNeighborListGpu (gpu_kernels.cu:1981),
 * neighborListUpdateRequiredGpu (:1461).  skin = relativeSkin * cutoff (CoMD.c:257-264).
 * useNeighborList != 0 switches force + redistribute to the list variants. */
int  comd_dev_set_neighbor_list(ComdDev* dev, int useNeighborList, double skinDistance);
int  comd_dev_neighbor_stats(ComdDev* dev, long* nPairsStored, int* maxPerAtom, int* nBuilds);
/* list builds and in-place ghost refreshes since creation (the two outcomes of redistributeAtomsGpuNL,
 * timestep.c:278-354); bench.py reports how many of each fell into its timed window */
int  comd_dev_neighbor_counters(ComdDev* dev, long* nBuilds, long* nRefreshes);

/* ---- multi-GPU: replaces the MPI wrappers of parallel.h:17-63 for the halo path and
 * the libmp layer of comm.h:40-74.  The 128-byte id is made on rank 0 and handed to
 * the other ranks by whatever launcher is in use (torchrun store, a file, a pipe). */
#define COMD_COMM_ID_BYTES 128
int  comd_comm_unique_id(unsigned char id[COMD_COMM_ID_BYTES]);
int  comd_comm_init(ComdDev* dev, const unsigned char id[COMD_COMM_ID_BYTES], int rank, int nRanks);
int  comd_comm_allreduce_sum(ComdDev* dev, double* inout, int n);   /* addRealParallel, parallel.c:130 */
int  comd_comm_allreduce_sum_int(ComdDev* dev, long* inout, int n); /* addIntParallel,  parallel.c:120 */
int  comd_comm_barrier(ComdDev* dev);                               /* barrierParallel, parallel.c:93  */

/* ---- tuning knobs (no reference counterpart; defaults are the measured-best settings):
 *   "force_variant"      link-cell path: 1 = tiled TMA-staged filter/process kernels (default), 0 = thread-per-atom
 *                        stencil loop (the shape of LJ_Force_thread_atom, gpu_lj_thread_atom.h:31-143)
 *   "tile_hits"          link-cell EAM path: 1 (default) the tiled density pass writes the step's hit list and the list kernel
 *                        evaluates the pair forces from it; 0: the tiled engine sweeps the stencil twice (round 1)
 *   "tile_tx"            link-cell path: target cells per force tile along x (1..16)
 *   "energy_every_step"  list path: per-atom EAM pair energies at every step of comd_dev_timestep (default 0)
 *   "nl_n3"              list path: 1 = pairs inside a work block are stored and evaluated once, for both
 *                        atoms (half-list form of neighborList.c:123,142 / eam.c:311-345 inside the block, shared-memory
 *                        fixed-point accumulators); 0 = full list, every pair met from both ends; -1 (default) = 1 for EAM,
 *                        0 for LJ (whose pair term is cheaper than the scatter).  Invalidates the list.
 *   "nl_lanes"           once-per-pair kernels: lanes that walk one atom's list row, 4 (default) or 8.  Invalidates the list.
 *   "nl_block"           list path: cells per work block as bx*10000 + by*100 + bz (by*bz <= 32), 0 = auto
 *   "nl_threads"         list path: cap on the CTA size of the force kernels (multiple of 32)
 *   "nl_chunks_per_cta"  list path: work chunks per persistent CTA handed out through the global counter (default 32)
 *   "overlap"            list path, N > 1: halo exchange overlapped with the interior blocks (default 1)
 *   "graph_step"         list path inside comd_dev_timestep: steps between two list builds that need no energies are replayed
 *                        from a CUDA graph captured once per build (one launch per step).  -1 (default): on one rank with up
 *                        to 400 000 atoms, where a step is launch-bound; 0 off; 1 on wherever the step qualifies
 *   "nl_hilbert"         list path: work blocks handed to the SMs along a 3-D Hilbert curve instead of x-fastest index order
 *                        (CoMD's -H / --hilbert, linkCells.c:28-63,151-174; any grid size; results unchanged; default 0)
 *   "nl_speculate"       list path inside comd_dev_timestep: 1 (default) enqueue each step for the no-rebuild case without waiting
 *                        for the displacement check (the force kernels return at once if the list turns out stale, the
 *                        host then re-bins and repeats them); 0: wait for the verdict every step (round 1)
 *   "overlap_mode"       1 (default): interior blocks as one-shot CTAs on all SMs, boundary shell on a side stream that waits
 *                        for the exchange only; 0: persistent interior CTAs on "overlap_sms" SMs, boundary launch behind them
 *   "overlap_sms"        overlap_mode 0: SMs an overlapped interior launch may fill (default 140 of 148)
 *   "phase_timing"       per-phase CUDA-event timing of comd_dev_timestep, reported on stderr by comd_dev_destroy
 *   "e2e_speculate"      0 (default): comd_dev_step_host waits for its upload before the step starts.  1: a call that
 *                        follows another one may start from the atoms the device still holds while its inputs are
 *                        uploaded (see there) -- pays only when the caller hands back bit-identical buffers */
int  comd_dev_set_option(ComdDev* dev, const char* key, long value);

/* Bytes of the last halo message packed for / received across `face` (x-,x+,y-,y+,z-,z+): kind 0 = atom exchange
 * (56-byte AtomMsg per atom, haloExchange.h:32-38), kind 1 = F'(rho) exchange (8 bytes per atom).  What the
 * reference's loadBuffer returns and unloadBuffer is handed (haloExchange.h:84,98); the buffers themselves stay in HBM. */
int  comd_dev_halo_bytes(ComdDev* dev, int kind, int face, long* sentBytes, long* receivedBytes);

/* ---- measurement hooks (bench.py / tests) ---------------------------------------- */
int  comd_dev_sync(ComdDev* dev);
/* Mean CUDA-event duration (ms) of the force calls (comd_dev_compute_force, also inside
 * comd_dev_timestep) since the previous query, and the kernels launched since creation. */
int  comd_dev_kernel_stats(ComdDev* dev, double* meanForceMs, long* nLaunches);
/* CUDA-event stopwatch on the handle's stream: start syncs then records, stop records and syncs. */
int  comd_dev_timer_start(ComdDev* dev);
int  comd_dev_timer_stop(ComdDev* dev, double* ms);
/* FP64 FMA-pipe peak of this GPU (DFMA microbenchmark, TFLOP/s): the roofline denominator
 * for the force kernels; the driver's MEASURED_PEAKS.json has no FP64 entry. */
/* steps of comd_dev_timestep replayed from the CUDA graph (option "graph_step"), and speculatively enqueued steps that were
 * repeated behind a list rebuild (option "nl_speculate") */
int  comd_dev_step_stats(ComdDev* dev, long* nGraphSteps, long* nRedoneSteps);
int  comd_dev_fp64_peak(ComdDev* dev, double* tflops);
int  comd_dev_fp32_peak(ComdDev* dev, double* tflops);    /* FFMA twin: the roofline denominator of the -DCOMD_SINGLE build */
/* position of work block (x, y, z) on the 3-D Hilbert curve of order `bits` (the key option "nl_hilbert" sorts by; needs no device) */
unsigned long long comd_hilbert_key(unsigned x, unsigned y, unsigned z, int bits);
/* the order in which the list kernels walk the work blocks (bx*by*bz link cells each) of a domain of cells[3] link cells: set 0 all
 * blocks, 1 interior (no ghost neighbour), 2 boundary shell; mode 0 the set's own order, 1 largest block first, 2 Hilbert curve.
 * Writes linear block ids (x fastest) to out, returns their number or a negative COMD_E_*; needs no device */
int  comd_block_order_table(const int cells[3], int bx, int by, int bz, int set, int mode, int capacity, int* out);
/* End-to-end step through HOST buffers (bench.py `e2e`): one full timestep() whose inputs
 * (gid, r, p, f at time t; SoA, any order, all local to this rank) start in host memory and
 * whose results (gid, r, p, f, U at t+dt, in device order, *nOut of them) end there.
 * Buffers from comd_host_alloc (pinned) are copied directly, pageable ones through pinned staging
 * inside the call; the ghost-free result is compacted on the device before it crosses PCIe. */
int  comd_host_alloc(void** p, size_t bytes);      /* cudaMallocHost; any rank, any time after comd_dev_create */
int  comd_host_free(void* p);
int  comd_dev_step_host(ComdDev* dev, int nLocal, double dt,
                        const int* gidIn, const comd_real* rIn[3], const comd_real* pIn[3], const comd_real* fIn[3],
                        int* nOut, int* gidOut, comd_real* rOut[3], comd_real* pOut[3], comd_real* fOut[3], comd_real* UOut,
                        double eLocal[2]);
/* A caller that hands the atoms of one comd_dev_step_host call to the next (same gids, same order, no
 * position further than skin/2 from the neighbour list's snapshot) is served without re-sorting and
 * without rebuilding the list; anything else takes the full upload path.  *nResumed counts the former.
 * Every call waits for its upload before the step starts.  Only with option "e2e_speculate" 1 (off by default), on one
 * rank, with pinned buffers and no other call on the handle in between, does such a call start early: from the atoms the
 * device handed out last time, while the uploaded inputs are compared with them bit for bit; if they differ the step is
 * redone from the upload and the handle stops speculating.  A result buffer may be the input buffer of the same quantity
 * (in-place update); a result buffer over the input of ANOTHER quantity switches the early start off. */
int  comd_dev_step_host_stats(ComdDev* dev, long* nResumed);

/* ---- energy sums without any device allocation.  No counterpart in the reference (its ReduceEnergy,
 * gpu_reduce.h:32-98, always works on the slabs AllocateGpu made).
 * Every other entry point needs a ComdDev, and comd_dev_create() builds one out of cudaMalloc calls.  That shuts
 * out callers that must not, or cannot, allocate on the device: a diagnostics hook invoked inside a running
 * time-step loop (cudaMalloc and cudaFree wait for the whole GPU and would serialise the halo/compute overlap),
 * and a host that has already given all of HBM to its own arena or caching allocator, underneath which a
 * library's cudaMalloc may just fail.  Such a caller owns its per-atom arrays; this call needs nothing else.
 *   U, px, py, pz : DEVICE pointers, nAtoms entries each -- per-atom potential energy and momentum of the atoms
 *                   to be summed (pass local atoms only; there is no ghost filter here)
 *   eOut          : DEVICE pointer to 2 doubles: eOut[0] = sum U, eOut[1] = sum |p|^2 / (2 mass), the two
 *                   values comd_dev_energy() returns, summed in double in both precision builds
 *   stream        : a cudaStream_t (NULL = the default stream).  The two kernels are enqueued there and the
 *                   call returns without waiting; eOut is valid once that stream has reached them.
 * The block partials between the two reduction levels sit in a fixed 256 KB __device__ array inside the library
 * (each GPU that loads the library holds it until the process ends), sized for COMD_ENERGY_STATIC_MAX_ATOMS
 * atoms: a larger nAtoms returns COMD_E_ARG and launches nothing.  The call itself uses no cudaMalloc, cudaFree
 * or other allocating routine.  All calls of a process on one GPU share that array: back-to-back calls on one
 * stream are fine, calls that can overlap on the GPU (different streams) race and must be ordered by the caller.
 * Nothing in the array needs clearing between calls -- each call overwrites every slot it reads.
 * The FIRST call on each GPU also resolves the array's device address (cudaGetSymbolAddress, which loads the
 * module there): make that call once during set-up, before any stream capture begins. */
#define COMD_ENERGY_STATIC_MAX_ATOMS 67108864L      /* 16384 tiles of 4096 atoms */
int  comd_energy_sums_static(int deviceId, void* stream, long nAtoms, double mass,
                             const comd_real* U, const comd_real* px, const comd_real* py, const comd_real* pz,
                             double* eOut);

#ifdef __cplusplus
}
#endif
#endif /* COMD_B200_H */
