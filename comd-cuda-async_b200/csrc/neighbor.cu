// neighbor.cu -- optional neighbour-list path (buildNeighborListGpu, gpu_kernels.cu:1981;
// rebuild criterion neighborList.c:212-247).  Filled in after the link-cell path.
#include "comd_dev.cuh"

int comdNeighborBuildIfNeeded(ComdDev* d) { (void)d; return COMD_OK; }

extern "C" int comd_dev_set_neighbor_list(ComdDev* d, int useNeighborList, double skinDistance)
{
   if (!d) return COMD_E_ARG;
   if (useNeighborList) { comdSetError("comd_dev_set_neighbor_list: neighbour-list path not available in this build"); return COMD_E_STATE; }
   d->useNL = 0; d->skin = skinDistance;
   return COMD_OK;
}

extern "C" int comd_dev_neighbor_stats(ComdDev* d, long* nPairsStored, int* maxPerAtom, int* nBuilds)
{
   if (!d) return COMD_E_ARG;
   if (nPairsStored) *nPairsStored = 0;
   if (maxPerAtom) *maxPerAtom = 0;
   if (nBuilds) *nBuilds = d->nlBuilds;
   return COMD_OK;
}
