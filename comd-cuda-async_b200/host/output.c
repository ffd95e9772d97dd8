/* output.c -- wall-clock timers and the YAML / screen reports, in the reference's formats
 * (performanceTimers.c:55-68,127-335; yamlOutput.c:45-128).  Device work is asynchronous, so
 * profileStop() of a timer that brackets device calls first drains the device stream
 * (comdTimerSyncHook): kernel time lands in the timer that launched it, not in whichever timer
 * next happens to synchronise (the reference's K20c logs book force time under "eamHalo").
 */
#define _GNU_SOURCE
#include "comd_host.h"
#include <inttypes.h>
#include <math.h>
#include <string.h>
#include <sys/utsname.h>
#include <time.h>
#include <unistd.h>

/* the pipe the roofline lines are quoted against: CoMD.c measures the FFMA peak in the real_t = float build, the DFMA peak otherwise */
#ifdef COMD_SINGLE
#define COMD_FP_NAME  "FP32"
#define COMD_FMA_NAME "FFMA"
#else
#define COMD_FP_NAME  "FP64"
#define COMD_FMA_NAME "DFMA"
#endif

#ifndef COMD_VARIANT
#define COMD_VARIANT "CoMD-b200"
#endif
#ifndef COMD_VERSION
#define COMD_VERSION "1.1-b200"
#endif

FILE* yamlFile = NULL;
static ComdDev* g_timerDev = NULL;
void comdTimerSyncHook(ComdDev* dev) { g_timerDev = dev; }

static const char* timerName[numberOfTimers] = {
   "total", "loop", "timestep", "  position", "  velocity", "  neighborList", "  redistribute", "    atomHalo",
   "  force", "    eamHalo", "commHalo", "commReduce" };

typedef struct
{
   double start, total, elapsed;
   uint64_t count;
   int minRank, maxRank;
   double minValue, maxValue, average, stdev;
} Timer;
static Timer perfTimer[numberOfTimers];
static struct { double atomRate, atomAllRate, atomsPerUSec, flopPerAtomStep, fp64PeakTflops; int nRanks; } perfGlobal;

/* Figure of merit beside the timers: the algorithmic work of one atom-step (full-neighbour form, FMA = 2,
 * div = sqrt = 1: 27 flop per LJ neighbour, 93 per EAM neighbour, + integrator / embedding) and the FP64 FMA-pipe peak
 * measured on this GPU; printPerformanceResults turns the atom rate into a roofline fraction. */
void setRooflineInfo(double flopPerAtomStep, double fp64PeakTflopsPerGpu, int nRanks)
{
   perfGlobal.flopPerAtomStep = flopPerAtomStep; perfGlobal.fp64PeakTflops = fp64PeakTflopsPerGpu; perfGlobal.nRanks = nRanks;
}

static double now(void)
{
   struct timespec t;
   clock_gettime(CLOCK_MONOTONIC, &t);
   return t.tv_sec + 1e-9 * t.tv_nsec;
}

void profileStart(enum TimerHandle h) { perfTimer[h].start = now(); }

void profileStop(enum TimerHandle h)
{
   if (g_timerDev && h >= timestepTimer && h <= eamHaloTimer) comd_dev_sync(g_timerDev);
   const double dt = now() - perfTimer[h].start;
   perfTimer[h].count += 1;
   perfTimer[h].total += dt;
   perfTimer[h].elapsed += dt;
}

/* seconds accumulated since the previous call (CoMD.c:150 prints it per interval) */
double getElapsedTime(enum TimerHandle h)
{
   const double e = perfTimer[h].elapsed;
   perfTimer[h].elapsed = 0.0;
   return e;
}

static void timerStats(void)
{
   double sendBuf[numberOfTimers], recvBuf[numberOfTimers];
   RankReduceData in[numberOfTimers], lo[numberOfTimers], hi[numberOfTimers];
   for (int i = 0; i < numberOfTimers; ++i) { sendBuf[i] = perfTimer[i].total; in[i].val = perfTimer[i].total; in[i].rank = getMyRank(); }
   addDoubleParallel(sendBuf, recvBuf, numberOfTimers);
   minRankDoubleParallel(in, lo, numberOfTimers);
   maxRankDoubleParallel(in, hi, numberOfTimers);
   for (int i = 0; i < numberOfTimers; ++i)
   {
      perfTimer[i].average = recvBuf[i] / (double)getNRanks();
      perfTimer[i].minValue = lo[i].val; perfTimer[i].minRank = lo[i].rank;
      perfTimer[i].maxValue = hi[i].val; perfTimer[i].maxRank = hi[i].rank;
      const double d = perfTimer[i].total - perfTimer[i].average;
      sendBuf[i] = d * d;
   }
   addDoubleParallel(sendBuf, recvBuf, numberOfTimers);
   for (int i = 0; i < numberOfTimers; ++i) perfTimer[i].stdev = sqrt(recvBuf[i] / (double)getNRanks());
}

void printPerformanceResults(int nGlobalAtoms, int printRate)
{
   timerStats();
   if (!printRank()) return;
   const double loopTime = perfTimer[loopTimer].total;
   fprintf(screenOut, "\n\nTimings for Rank %d\n", getMyRank());
   fprintf(screenOut, "        Timer        # Calls    Avg/Call (s)   Total (s)    %% Loop\n");
   fprintf(screenOut, "___________________________________________________________________\n");
   for (int i = 0; i < numberOfTimers; ++i)
   {
      const Timer* t = &perfTimer[i];
      if (t->count > 0)
         fprintf(screenOut, "%-16s%12" PRIu64 "     %8.4f      %8.4f    %8.2f\n", timerName[i], t->count,
                 t->total / (double)t->count, t->total, t->total / loopTime * 100.0);
   }
   fprintf(screenOut, "\nTiming Statistics Across %d Ranks:\n", getNRanks());
   fprintf(screenOut, "        Timer        Rank: Min(s)       Rank: Max(s)      Avg(s)    Stdev(s)\n");
   fprintf(screenOut, "_____________________________________________________________________________\n");
   for (int i = 0; i < numberOfTimers; ++i)
   {
      const Timer* t = &perfTimer[i];
      if (t->count > 0)
         fprintf(screenOut, "%-16s%6d:%10.4f  %6d:%10.4f  %10.4f  %10.4f\n", timerName[i], t->minRank, t->minValue,
                 t->maxRank, t->maxValue, t->average, t->stdev);
   }
   const double atomsPerTask = nGlobalAtoms / (real_t)getNRanks();
   const Timer* ts = &perfTimer[timestepTimer];
   perfGlobal.atomRate = ts->average * 1e6 / (atomsPerTask * ts->count * printRate);
   perfGlobal.atomAllRate = ts->average * 1e6 / ((double)nGlobalAtoms * ts->count * printRate);
   perfGlobal.atomsPerUSec = 1.0 / perfGlobal.atomAllRate;
   const char* bar = "---------------------------------------------------";
   fprintf(screenOut, "\n%s\n Average atom update rate:     %6.2f us/atom/task\n%s\n\n", bar, perfGlobal.atomRate, bar);
   fprintf(screenOut, "\n%s\n Average all atom update rate: %6.2f us/atom\n%s\n\n", bar, perfGlobal.atomAllRate, bar);
   fprintf(screenOut, "\n%s\n Average atom rate:            %6.2f atoms/us\n%s\n\n", bar, perfGlobal.atomsPerUSec, bar);
   /* the headline metric of BASELINE.json in its own unit */
   fprintf(screenOut, " Atom-steps per second:        %.6e\n\n", perfGlobal.atomsPerUSec * 1e6);
   if (perfGlobal.fp64PeakTflops > 0.0)
   {
      const double tf = perfGlobal.atomsPerUSec * 1e6 * perfGlobal.flopPerAtomStep * 1e-12;
      fprintf(screenOut, " %s roofline:                %.3f TFLOP/s algorithmic (%.0f flop/atom-step) of %.2f TFLOP/s measured %s peak on %d GPU%s = %.4f\n\n",
              COMD_FP_NAME, tf, perfGlobal.flopPerAtomStep, perfGlobal.fp64PeakTflops * perfGlobal.nRanks, COMD_FMA_NAME, perfGlobal.nRanks, perfGlobal.nRanks > 1 ? "s" : "",
              tf / (perfGlobal.fp64PeakTflops * perfGlobal.nRanks));
   }
}

void printPerformanceResultsYaml(FILE* file)
{
   if (!printRank() || !file) return;
   const double loopTime = perfTimer[loopTimer].total;
   fprintf(file, "\nPerformance Results:\n");
   fprintf(file, "  TotalRanks: %d\n", getNRanks());
   fprintf(file, "  ReportingTimeUnits: seconds\n");
   fprintf(file, "Performance Results For Rank %d:\n", getMyRank());
   for (int i = 0; i < numberOfTimers; ++i)
   {
      const Timer* t = &perfTimer[i];
      if (t->count == 0) continue;
      fprintf(file, "  Timer: %s\n", timerName[i]);
      fprintf(file, "    CallCount: %" PRIu64 "\n", t->count);
      fprintf(file, "    AvgPerCall: %8.4f\n", t->total / (double)t->count);
      fprintf(file, "    Total:      %8.4f\n", t->total);
      fprintf(file, "    PercentLoop: %8.2f\n", t->total / loopTime * 100);
   }
   fprintf(file, "Performance Results Across Ranks:\n");
   for (int i = 0; i < numberOfTimers; ++i)
   {
      const Timer* t = &perfTimer[i];
      if (t->count == 0) continue;
      fprintf(file, "  Timer: %s\n", timerName[i]);
      fprintf(file, "    MinRank: %d\n", t->minRank);
      fprintf(file, "    MinTime: %8.4f\n", t->minValue);
      fprintf(file, "    MaxRank: %d\n", t->maxRank);
      fprintf(file, "    MaxTime: %8.4f\n", t->maxValue);
      fprintf(file, "    AvgTime: %8.4f\n", t->average);
      fprintf(file, "    StdevTime: %8.4f\n", t->stdev);
   }
   fprintf(file, "Performance Global Update Rates:\n");
   fprintf(file, "  AtomUpdateRate:\n    AverageRate: %6.2f\n    Units: us/atom/task\n", perfGlobal.atomRate);
   fprintf(file, "  AllAtomUpdateRate:\n    AverageRate: %6.2f\n    Units: us/atom\n", perfGlobal.atomAllRate);
   fprintf(file, "  AtomRate:\n    AverageRate: %6.2f\n    Units: atoms/us\n", perfGlobal.atomsPerUSec);
   fprintf(file, "  AtomStepsPerSecond: %.6e\n", perfGlobal.atomsPerUSec * 1e6);
   if (perfGlobal.fp64PeakTflops > 0.0)
   {
      const double tf = perfGlobal.atomsPerUSec * 1e6 * perfGlobal.flopPerAtomStep * 1e-12;
      fprintf(file, "  Roofline" COMD_FP_NAME ":\n    AlgorithmicFlopPerAtomStep: %.0f\n    AchievedTFlops: %.4f\n    MeasuredPeakTFlopsPerGpu: %.3f\n    Gpus: %d\n    Fraction: %.4f\n",
              perfGlobal.flopPerAtomStep, tf, perfGlobal.fp64PeakTflops, perfGlobal.nRanks, tf / (perfGlobal.fp64PeakTflops * perfGlobal.nRanks));
   }
   fprintf(file, "\n");
}

/* ---- YAML file + application banner ---- */
void printSeparator(FILE* file) { fprintf(file, "\n"); }

void yamlBegin(void)
{
   if (!printRank()) return;
   time_t raw; time(&raw);
   struct tm* t = localtime(&raw);
   char name[128];
   snprintf(name, sizeof name, "%s.%04d:%02d:%02d-%02d:%02d:%02d.yaml", COMD_VARIANT, t->tm_year + 1900, t->tm_mon + 1, t->tm_mday,
            t->tm_hour, t->tm_min, t->tm_sec);
   yamlFile = fopen(name, "w");
   if (!yamlFile) yamlFile = fopen("/dev/null", "w");
}

void yamlEnd(void)
{
   if (!printRank() || !yamlFile) return;
   fclose(yamlFile);
   yamlFile = NULL;
}

void yamlAppInfo(FILE* file)
{
   if (!printRank() || !file) return;
   struct utsname u; memset(&u, 0, sizeof u); uname(&u);
   printSeparator(file);
   fprintf(file, "Mini-Application Name    : %s\n", COMD_VARIANT);
   fprintf(file, "Mini-Application Version : %s\n", COMD_VERSION);
   fprintf(file, "Platform:\n");
   fprintf(file, "  hostname: %s\n", u.nodename);
   fprintf(file, "  kernel name: '%s'\n", u.sysname);
   fprintf(file, "  kernel release: '%s'\n", u.release);
   fprintf(file, "  processor: '%s'\n", u.machine);
   fprintf(file, "Build:\n");
   fprintf(file, "  CC: '%s'\n", "gcc (host C11) + nvcc -gencode arch=compute_100a,code=sm_100a (device layer)");
   fprintf(file, "  compiler version: '%s'\n", __VERSION__);
   fprintf(file, "  device layer: '%s'\n", comd_version());
   fprintf(file, "  using MPI: false (NCCL over NVLink + shared-memory host collectives)\n");
   fprintf(file, "  Threading: none\n");
   fprintf(file, "  Double Precision: %s\n", sizeof(real_t) == sizeof(double) ? "true" : "false");
   time_t raw; time(&raw);
   struct tm* t = localtime(&raw);
   fprintf(file, "Run Date/Time: %04d-%02d-%02d, %02d:%02d:%02d\n", t->tm_year + 1900, t->tm_mon + 1, t->tm_mday, t->tm_hour,
           t->tm_min, t->tm_sec);
   fprintf(file, "\n");
   fflush(file);
}
