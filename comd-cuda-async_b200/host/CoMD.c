/* CoMD.c -- the driver: same command line, same screen and YAML output, same main loop as the
 * reference's CoMD.c:86-187 (main), :200-327 (initSimulation), :413-549 (validate, sumAtoms,
 * printThings, printSimulationDataYaml), :552-613 (sanityChecks).  Everything between the print
 * statements runs on the GPU through the C ABI of include/comd_b200.h.
 */
#include "comd_host.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <strings.h>

extern void comdTimerSyncHook(ComdDev* dev);

static BasePotential* initPotential(int doeam, const char* potDir, const char* potName, const char* potType)
{
   return doeam ? initEamPot(potDir, potName, potType) : initLjPot();
}

static SpeciesData* initSpecies(BasePotential* pot)
{
   SpeciesData* sp = calloc(1, sizeof *sp);
   strcpy(sp->name, pot->name);
   sp->atomicNo = pot->atomicNo;
   sp->mass = pot->mass;
   return sp;
}

/* AllocateGpu + CopyDataToGpu (gpu_utility.c:165-282,432-600): one device context per rank */
static void createDevice(SimFlat* s, int doeam)
{
   ComdGeometry g;
   for (int a = 0; a < 3; ++a)
   {
      g.gridSize[a] = s->boxes->gridSize[a];
      g.localMin[a] = s->boxes->localMin[a];   g.localMax[a] = s->boxes->localMax[a];
      g.boxSize[a] = s->boxes->boxSize[a];     g.invBoxSize[a] = s->boxes->invBoxSize[a];
      g.globalExtent[a] = s->domain->globalExtent[a];
      g.procGrid[a] = s->domain->procGrid[a];  g.procCoord[a] = s->domain->procCoord[a];
   }
   /* the pool holds locals + last step's ghosts + this step's ghosts while redistributing */
   double shell = 1.0;
   for (int a = 0; a < 3; ++a) shell *= (g.gridSize[a] + 2.0) / g.gridSize[a];
   const long cap = (long)(s->atoms->nLocal * (2.0 * shell - 1.0) * 1.15) + 65536;
   DEV_TRY(comd_dev_create(&s->dev, localDeviceId(), &g, cap));
   DEV_TRY(comd_dev_set_mass(s->dev, s->pot->mass));
   if (doeam)
   {
      EamPotential* e = (EamPotential*)s->pot;
      /* ComdTable carries doubles in either precision: a float table widens exactly (the device layer of the
       * single-precision build rounds back to the same floats) */
      InterpolationObject* src[3] = { e->phi, e->rho, e->f };
      ComdTable t[3];
      double* wide[3];
      for (int k = 0; k < 3; ++k)
      {
         wide[k] = malloc((src[k]->n + 3) * sizeof(double));
         for (int i = -1; i <= src[k]->n + 1; ++i) wide[k][i + 1] = src[k]->values[i];
         t[k].n = src[k]->n; t[k].x0 = src[k]->x0; t[k].invDx = src[k]->invDx; t[k].values = wide[k] + 1;
      }
      DEV_TRY(comd_dev_set_eam(s->dev, e->cutoff, &t[0], &t[1], &t[2]));
      for (int k = 0; k < 3; ++k) free(wide[k]);
   }
   else
   {
      LjPotential* lj = (LjPotential*)s->pot;
      DEV_TRY(comd_dev_set_lj(s->dev, lj->sigma, lj->epsilon, lj->cutoff));
   }
   const int useNL = s->method == THREAD_ATOM_NL || s->method == WARP_ATOM_NL;
   DEV_TRY(comd_dev_set_neighbor_list(s->dev, useNL, s->skinDistance));
   if (s->useHilbert) DEV_TRY(comd_dev_set_option(s->dev, "nl_hilbert", 1));    /* -H: Hilbert traversal of the work blocks */
   if (getNRanks() > 1)
   {
      unsigned char id[COMD_COMM_ID_BYTES];
      DEV_TRY(parallelCommId(id));
      DEV_TRY(comd_comm_init(s->dev, id, getMyRank(), getNRanks()));
   }
   comdTimerSyncHook(s->dev);
}

SimFlat* initSimulation(Command cmd)
{
   SimFlat* sim = calloc(1, sizeof *sim);
   sim->nSteps = cmd.nSteps;
   sim->printRate = cmd.printRate;
   sim->dt = cmd.dt;
   sim->gpuAsync = cmd.gpuAsync;
   sim->gpuProfile = cmd.gpuProfile;
   sim->ljInterpolation = cmd.ljInterpolation;
   sim->spline = cmd.spline;
   sim->usePairlist = cmd.usePairlist;
   sim->useHilbert = cmd.doHilbert;
   if (sim->gpuProfile) sim->nSteps = 0;         /* profile mode: one force call, no steps (CoMD.c:215-218) */

   if (!strcmp(cmd.method, "thread_atom")) sim->method = THREAD_ATOM;
   else if (!strcmp(cmd.method, "warp_atom")) sim->method = WARP_ATOM;
   else if (!strcmp(cmd.method, "warp_atom_nl")) sim->method = WARP_ATOM_NL;
   else if (!strcmp(cmd.method, "cta_cell")) sim->method = CTA_CELL;
   else if (!strcmp(cmd.method, "thread_atom_nl")) sim->method = THREAD_ATOM_NL;
   else if (!strcmp(cmd.method, "cpu_nl"))
   {
      if (printRank()) printf("Error: -m cpu_nl is the reference's CPU path; this build has no CPU force path (see oracle/ for the checker)\n");
      exit(-1);
   }
   else
   {
      if (printRank()) printf("Error: You have to specify a valid method: -m [thread_atom,thread_atom_nl,warp_atom,warp_atom_nl,cta_cell,cpu_nl]\n");
      exit(-1);
   }
   const int useNL = sim->method == THREAD_ATOM_NL || sim->method == WARP_ATOM_NL;
   if (printRank())
   {  /* the reference's method names select two very different kernels here: say which, and what it costs */
      if (useNL) printf("Force path: neighbour list (-m %s): once-per-pair list kernels, the fast path of this build\n", cmd.method);
      else printf("Force path: link cells (-m %s): 27-cell stencil over TMA-staged tiles, re-binned every step; expect ~2.3x the step time of "
                  "-m thread_atom_nl (measured on B200, EAM: C3 1.77 vs 0.78 ms/step, C5 30 vs 12 ms/step)\n", cmd.method);
   }
   if ((cmd.ljInterpolation || cmd.spline || cmd.usePairlist) && printRank())
      printf("Note: -I / -P / -L select experimental kernels of the reference that this build does not carry; ignored.\n");

   sim->pot = initPotential(cmd.doeam, cmd.potDir, cmd.potName, cmd.potType);
   real_t latticeConstant = cmd.lat;
   if (cmd.lat < 0.0) latticeConstant = sim->pot->lat;
   sanityChecks(cmd, sim->pot->cutoff, latticeConstant, sim->pot->latticeType);
   sim->species = initSpecies(sim->pot);

   real3 globalExtent = { cmd.nx * latticeConstant, cmd.ny * latticeConstant, cmd.nz * latticeConstant };
   sim->domain = initDecomposition(cmd.xproc, cmd.yproc, cmd.zproc, globalExtent);

   real_t skinDistance = 0.0;
   if (useNL)
   {
      skinDistance = sim->pot->cutoff * cmd.relativeSkinDistance;
      if (printRank()) printf("Skin-Distance: %f\n", skinDistance);
   }
   sim->skinDistance = skinDistance;
   sim->boxes = initLinkCells(sim->domain, sim->pot->cutoff + skinDistance, cmd.doHilbert);
   sim->atoms = initAtoms(sim->boxes, skinDistance);

   createFccLattice(cmd.nx, cmd.ny, cmd.nz, latticeConstant, sim);
   setTemperature(sim, cmd.temperature);
   randomDisplacements(sim, cmd.initialDelta);

   sim->atomExchange = initAtomHaloExchange(sim->domain, sim->boxes);
   if (cmd.doeam) ((EamPotential*)sim->pot)->forceExchange = initForceHaloExchange(sim->domain, sim->boxes, 1);

   createDevice(sim, cmd.doeam);
   /* upload = CopyDataToGpu + the first redistributeAtoms (bin, ghost shell, canonical sort) */
   startTimer(redistributeTimer);
   Atoms* a = sim->atoms;
   DEV_TRY(comd_dev_upload_atoms(sim->dev, a->nLocal, a->gid, a->iSpecies, a->r[0], a->r[1], a->r[2], a->p[0], a->p[1], a->p[2]));
   stopTimer(redistributeTimer);

   startTimer(computeForceTimer);
   computeForce(sim);
   stopTimer(computeForceTimer);
   kineticEnergyGpu(sim);
   return sim;
}

void destroySimulation(SimFlat** ps)
{
   if (!ps || !*ps) return;
   SimFlat* s = *ps;
   comdTimerSyncHook(NULL);
   if (s->dev) comd_dev_destroy(s->dev);
   if (s->pot) s->pot->destroy(&s->pot);
   destroyLinkCells(&s->boxes);
   destroyAtoms(s->atoms);
   destroyHaloExchange(&s->atomExchange);
   free(s->species);
   free(s->domain);
   free(s);
   *ps = NULL;
}

/* sumAtoms (CoMD.c:442-457).  The reference sums a host array its GPU path never refreshes
 * (the copy-back is commented out at :445); here the per-cell populations come from the device. */
void sumAtoms(SimFlat* s)
{
   DEV_TRY(comd_dev_cell_counts(s->dev, s->boxes->nAtoms, s->boxes->nTotalBoxes));
   s->atoms->nLocal = 0;
   for (int b = 0; b < s->boxes->nLocalBoxes; ++b) s->atoms->nLocal += s->boxes->nAtoms[b];
   startTimer(commReduceTimer);
   addIntParallel(&s->atoms->nLocal, &s->atoms->nGlobal, 1);
   stopTimer(commReduceTimer);
}

void printThings(SimFlat* s, int iStep, double elapsedTime)
{
   static int iStepPrev = -1;
   static int firstCall = 1;
   const int nEval = iStep - iStepPrev;          /* 1 for the zeroth step */
   iStepPrev = iStep;
   if (!printRank()) return;
   if (firstCall)
   {
      firstCall = 0;
      fprintf(screenOut,
              "#                                                                                         Performance\n"
              "#  Loop   Time(fs)       Total Energy   Potential Energy     Kinetic Energy  Temperature   (us/atom)     # Atoms\n");
      fflush(screenOut);
   }
   const real_t time = iStep * s->dt;
   const real_t eTotal = (s->ePotential + s->eKinetic) / s->atoms->nGlobal;
   const real_t eK = s->eKinetic / s->atoms->nGlobal;
   const real_t eU = s->ePotential / s->atoms->nGlobal;
   const real_t Temp = (s->eKinetic / s->atoms->nGlobal) / (kB_eV * 1.5);
   const double timePerAtom = 1.0e6 * elapsedTime / (double)(nEval * s->atoms->nLocal);
   fprintf(screenOut, " %6d %10.2f %18.12f %18.12f %18.12f %12.4f %10.4f %12d\n", iStep, time, eTotal, eU, eK, Temp, timePerAtom,
           s->atoms->nGlobal);
}

void printSimulationDataYaml(FILE* file, SimFlat* s)
{
   const int maxOcc = maxOccupancy(s->boxes);     /* collective: all ranks */
   if (!printRank() || !file) return;
   const Domain* d = s->domain;
   const LinkCell* b = s->boxes;
   fprintf(file, "Simulation data: \n");
   fprintf(file, "  Total atoms        : %d\n", s->atoms->nGlobal);
   fprintf(file, "  Min global bounds  : [ %14.10f, %14.10f, %14.10f ]\n", d->globalMin[0], d->globalMin[1], d->globalMin[2]);
   fprintf(file, "  Max global bounds  : [ %14.10f, %14.10f, %14.10f ]\n", d->globalMax[0], d->globalMax[1], d->globalMax[2]);
   printSeparator(file);
   fprintf(file, "Decomposition data: \n");
   fprintf(file, "  Processors         : %6d,%6d,%6d\n", d->procGrid[0], d->procGrid[1], d->procGrid[2]);
   fprintf(file, "  Local boxes        : %6d,%6d,%6d = %8d\n", b->gridSize[0], b->gridSize[1], b->gridSize[2],
           b->gridSize[0] * b->gridSize[1] * b->gridSize[2]);
   fprintf(file, "  Box size           : [ %14.10f, %14.10f, %14.10f ]\n", b->boxSize[0], b->boxSize[1], b->boxSize[2]);
   fprintf(file, "  Box factor         : [ %14.10f, %14.10f, %14.10f ] \n", b->boxSize[0] / s->pot->cutoff, b->boxSize[1] / s->pot->cutoff,
           b->boxSize[2] / s->pot->cutoff);
   fprintf(file, "  Max Link Cell Occupancy: %d (dense storage, no MAXATOMS)\n", maxOcc);
   printSeparator(file);
   fprintf(file, "Potential data: \n");
   s->pot->print(file, s->pot);
   const int perAtomSize = 10 * sizeof(real_t) + 2 * sizeof(int);
   const float totalMemLocal = (float)perAtomSize * s->atoms->nLocal / 1024 / 1024;
   const float totalMemGlobal = (float)perAtomSize * s->atoms->nGlobal / 1024 / 1024;
   int nAll = 0;
   for (int k = 0; k < b->nTotalBoxes; ++k) nAll += b->nAtoms[k];
   printSeparator(file);
   fprintf(file, "Memory data: \n");
   fprintf(file, "  Intrinsic atom footprint = %4d B/atom \n", perAtomSize);
   fprintf(file, "  Total atom footprint     = %7.3f MB (%6.2f MB/node)\n", totalMemGlobal, totalMemLocal);
   fprintf(file, "  Link cell atom footprint = %7.3f MB/node\n", totalMemLocal);
   fprintf(file, "  Link cell atom footprint = %7.3f MB/node (including halo cell data\n", (float)perAtomSize * nAll / 1024 / 1024);
   fflush(file);
}

void sanityChecks(Command cmd, double cutoff, double latticeConst, char latticeType[8])
{
   int failCode = 0;
   if (cmd.xproc * cmd.yproc * cmd.zproc != getNRanks())
   {
      failCode |= 1;
      if (printRank()) fprintf(screenOut, "\nNumber of MPI ranks must match xproc * yproc * zproc\n");
   }
   const double minx = 2 * cutoff * cmd.xproc, miny = 2 * cutoff * cmd.yproc, minz = 2 * cutoff * cmd.zproc;
   if (cmd.nx * latticeConst < minx || cmd.ny * latticeConst < miny || cmd.nz * latticeConst < minz)
   {
      failCode |= 2;
      if (printRank())
         fprintf(screenOut, "\nSimulation too small.\n"
                 "  Increase the number of unit cells to make the simulation\n"
                 "  at least (%3.2f, %3.2f. %3.2f) Ansgstroms in size\n", minx, miny, minz);
   }
   if (strcasecmp(latticeType, "FCC") != 0)
   {
      failCode |= 4;
      if (printRank()) fprintf(screenOut, "\nOnly FCC Lattice type supported, not %s. Fatal Error.\n", latticeType);
   }
   int checkCode = failCode;
   bcastParallel(&checkCode, sizeof(int), 0);
   if (checkCode != failCode) failCode |= 8;      /* ranks disagree: cannot happen with identical command lines */
   if (failCode != 0) exit(failCode);
}

Validate* initValidate(SimFlat* sim)
{
   sumAtoms(sim);
   Validate* val = calloc(1, sizeof *val);
   val->eTot0 = (sim->ePotential + sim->eKinetic) / sim->atoms->nGlobal;
   val->nAtoms0 = sim->atoms->nGlobal;
   if (printRank())
   {
      fprintf(screenOut, "\n");
      printSeparator(screenOut);
      fprintf(screenOut, "Initial energy : %14.12f, atom count : %d \n", val->eTot0, val->nAtoms0);
      fprintf(screenOut, "\n");
   }
   return val;
}

void validateResult(const Validate* val, SimFlat* sim)
{
   if (!printRank()) return;
   const real_t eFinal = (sim->ePotential + sim->eKinetic) / sim->atoms->nGlobal;
   const int nAtomsDelta = sim->atoms->nGlobal - val->nAtoms0;
   fprintf(screenOut, "\n\nSimulation Validation:\n");
   fprintf(screenOut, "  Initial energy  : %14.12f\n", val->eTot0);
   fprintf(screenOut, "  Final energy    : %14.12f\n", eFinal);
   fprintf(screenOut, "  eFinal/eInitial : %f\n", eFinal / val->eTot0);
   if (nAtomsDelta == 0) fprintf(screenOut, "  Final atom count : %d, no atoms lost\n", sim->atoms->nGlobal);
   else
   {
      fprintf(screenOut, "#############################\n");
      fprintf(screenOut, "# WARNING: %6d atoms lost #\n", nAtomsDelta);
      fprintf(screenOut, "#############################\n");
   }
}

#ifndef COMD_NO_MAIN
int main(int argc, char** argv)
{
   initParallel(&argc, &argv);
   profileStart(totalTimer);
   yamlBegin();
   timestampBarrier("Starting Initialization\n");
   yamlAppInfo(yamlFile);
   yamlAppInfo(screenOut);

   Command cmd = parseCommandLine(argc, argv);
   printCmdYaml(yamlFile, &cmd);
   printCmdYaml(screenOut, &cmd);

   SimFlat* sim = initSimulation(cmd);
   sumAtoms(sim);
   printSimulationDataYaml(yamlFile, sim);
   printSimulationDataYaml(screenOut, sim);

   Validate* validate = initValidate(sim);
   timestampBarrier("Initialization Finished\n");
   timestampBarrier("Starting simulation\n");

   const int nSteps = sim->nSteps, printRate = sim->printRate;
   int iStep = 0;
   profileStart(loopTimer);
   for (; iStep < nSteps;)
   {
      startTimer(commReduceTimer);
      sumAtoms(sim);
      stopTimer(commReduceTimer);
      printThings(sim, iStep, getElapsedTime(timestepTimer));
      startTimer(timestepTimer);
      timestep(sim, printRate, sim->dt);
      stopTimer(timestepTimer);
      iStep += printRate;
   }
   profileStop(loopTimer);

   sumAtoms(sim);
   printThings(sim, iStep, getElapsedTime(timestepTimer));
   timestampBarrier("Ending simulation\n");

   validateResult(validate, sim);
   profileStop(totalTimer);
   {  /* algorithmic flop per atom-step: neighbours of an FCC site inside the cutoff x (27 LJ | 93 EAM) + integrator (+ embedding) */
      const real_t lat = cmd.lat < 0.0 ? sim->pot->lat : cmd.lat, rc = sim->pot->cutoff;
      int nNbr = 0;
      const int m = (int)(rc / lat) + 2;
      for (int i = -2 * m; i <= 2 * m; ++i) for (int j = -2 * m; j <= 2 * m; ++j) for (int k = -2 * m; k <= 2 * m; ++k)
      {  /* FCC sites = half-integer triples (in units of lat/2) with an even coordinate sum */
         if ((i + j + k) & 1 || (i == 0 && j == 0 && k == 0)) continue;
         const real_t r2 = 0.25 * lat * lat * (real_t)(i * i + j * j + k * k);
         if (r2 <= rc * rc) ++nNbr;
      }
      double peak = 0.0;
#ifdef COMD_SINGLE
      DEV_TRY(comd_dev_fp32_peak(sim->dev, &peak));      /* real_t = float: the FFMA pipe is the one the forces run on */
#else
      DEV_TRY(comd_dev_fp64_peak(sim->dev, &peak));
#endif
      setRooflineInfo(cmd.doeam ? 93.0 * nNbr + 42.0 : 27.0 * nNbr + 22.0, peak, getNRanks());
   }
   printPerformanceResults(sim->atoms->nGlobal, sim->printRate);
   printPerformanceResultsYaml(yamlFile);

   destroySimulation(&sim);
   free(validate);
   yamlEnd();
   timestampBarrier("CoMD Ending\n");
   destroyParallel();
   return 0;
}
#endif
