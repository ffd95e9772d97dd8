/* host_dump.c -- test tool (CPU only): run the host layer's initial-state path exactly as
 * initSimulation does up to (not including) the device, and dump it for tests/test_host_layer.py.
 *   host_dump OUT.bin <CoMD options>      multi-rank: start it under comd-launch, one file per rank
 * Layout: int32 nLocal, nGlobal, grid[3], nTotalBoxes; double cutoff, mass, lat, boxSize[3], invBoxSize[3],
 *         localMin[3], localMax[3]; then gid[n] (int32), box[n] (int32), r[3][n], p[3][n] (real_t).
 *         With --doeam the three tables follow, phi, rho, F: int32 n; double x0, invDx; real_t values[-1 .. n+1].
 */
#include "comd_host.h"
#include <stdlib.h>
#include <string.h>

int main(int argc, char** argv)
{
   initParallel(&argc, &argv);
   if (argc < 2) { fprintf(stderr, "usage: host_dump OUT.bin [CoMD options]\n"); return 2; }
   char path[4096];
   snprintf(path, sizeof path, getNRanks() > 1 ? "%s.%d" : "%s", argv[1], getMyRank());
   Command cmd = parseCommandLine(argc - 1, argv + 1);
   SimFlat* sim = calloc(1, sizeof *sim);
   sim->pot = cmd.doeam ? initEamPot(cmd.potDir, cmd.potName, cmd.potType) : initLjPot();
   const real_t lat = cmd.lat < 0.0 ? sim->pot->lat : cmd.lat;
   sim->species = calloc(1, sizeof(SpeciesData));
   sim->species->mass = sim->pot->mass;
   real3 ext = { cmd.nx * lat, cmd.ny * lat, cmd.nz * lat };
   sim->domain = initDecomposition(cmd.xproc, cmd.yproc, cmd.zproc, ext);
   const int useNL = strstr(cmd.method, "_nl") != NULL;
   const real_t skin = useNL ? sim->pot->cutoff * cmd.relativeSkinDistance : 0.0;
   sim->boxes = initLinkCells(sim->domain, sim->pot->cutoff + skin, cmd.doHilbert);
   sim->atoms = initAtoms(sim->boxes, skin);
   createFccLattice(cmd.nx, cmd.ny, cmd.nz, lat, sim);
   setTemperature(sim, cmd.temperature);
   randomDisplacements(sim, cmd.initialDelta);

   FILE* fp = fopen(path, "wb");
   if (!fp) { perror(path); return 1; }
   const Atoms* a = sim->atoms;
   const LinkCell* b = sim->boxes;
   int hdr[6] = { a->nLocal, a->nGlobal, b->gridSize[0], b->gridSize[1], b->gridSize[2], b->nTotalBoxes };
   double dh[15] = { sim->pot->cutoff, sim->pot->mass, lat, b->boxSize[0], b->boxSize[1], b->boxSize[2], b->invBoxSize[0], b->invBoxSize[1],
                     b->invBoxSize[2], b->localMin[0], b->localMin[1], b->localMin[2], b->localMax[0], b->localMax[1], b->localMax[2] };
   fwrite(hdr, sizeof hdr, 1, fp);
   fwrite(dh, sizeof dh, 1, fp);
   fwrite(a->gid, sizeof(int), a->nLocal, fp);
   for (int i = 0; i < a->nLocal; ++i)
   {
      const real_t rr[3] = { a->r[0][i], a->r[1][i], a->r[2][i] };
      const int box = getBoxFromCoord(sim->boxes, rr);
      fwrite(&box, sizeof box, 1, fp);
   }
   for (int k = 0; k < 3; ++k) fwrite(a->r[k], sizeof(real_t), a->nLocal, fp);
   for (int k = 0; k < 3; ++k) fwrite(a->p[k], sizeof(real_t), a->nLocal, fp);
   if (cmd.doeam)
   {  /* what eamReadFuncfl / eamReadSetfl + initInterpolationObject left in the potential (host/potentials.c) */
      const EamPotential* e = (const EamPotential*)sim->pot;
      const InterpolationObject* tab[3] = { e->phi, e->rho, e->f };
      for (int k = 0; k < 3; ++k)
      {
         const double xd[2] = { tab[k]->x0, tab[k]->invDx };
         fwrite(&tab[k]->n, sizeof(int), 1, fp);
         fwrite(xd, sizeof xd, 1, fp);
         fwrite(tab[k]->values - 1, sizeof(real_t), tab[k]->n + 3, fp);
      }
   }
   fclose(fp);
   if (printRank()) printf("host_dump: %d atoms (%d global), eKinetic/atom %.15g\n", a->nLocal, a->nGlobal, sim->eKinetic / a->nGlobal);
   destroyParallel();
   return 0;
}
