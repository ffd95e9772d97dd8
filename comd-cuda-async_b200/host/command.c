/* command.c -- the command line of CoMD: the same 26 options, letters, defaults and YAML echo as
 * mycommand.c:187-320 (note -s is gpuProfile and -H is hilbert there, whatever its doc table says).
 * One table drives getopt_long, the help text and the defaults.
 *
 * --method keeps the reference's names.  thread_atom / warp_atom / cta_cell select the link-cell
 * force path (re-bin every step); thread_atom_nl / warp_atom_nl select the neighbour-list path
 * (skin = relativeSkinDistance * cutoff); cpu_nl is refused: this build has no CPU force path
 * (the reference's CPU path is the oracle under oracle/, not a product feature).
 */
#include "comd_host.h"
#include <getopt.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>

typedef struct { const char* name; char letter; int hasArg; char type; size_t offset; size_t size; const char* help; } OptionRow;
#define ROW(name, letter, hasArg, type, field, help) { name, letter, hasArg, type, offsetof(Command, field), sizeof(((Command*)0)->field), help }

static const OptionRow table[] = {
   { "help", 'h', 0, 'i', 0, 0, "print this message" },
   ROW("potDir", 'd', 1, 's', potDir, "potential directory"),
   ROW("potName", 'p', 1, 's', potName, "potential name"),
   ROW("potType", 't', 1, 's', potType, "potential type (funcfl or setfl)"),
   ROW("doeam", 'e', 0, 'i', doeam, "compute eam potentials"),
   ROW("nx", 'x', 1, 'i', nx, "number of unit cells in x"),
   ROW("ny", 'y', 1, 'i', ny, "number of unit cells in y"),
   ROW("nz", 'z', 1, 'i', nz, "number of unit cells in z"),
   ROW("xproc", 'i', 1, 'i', xproc, "processors in x direction"),
   ROW("yproc", 'j', 1, 'i', yproc, "processors in y direction"),
   ROW("zproc", 'k', 1, 'i', zproc, "processors in z direction"),
   ROW("nSteps", 'N', 1, 'i', nSteps, "number of time steps"),
   ROW("printRate", 'n', 1, 'i', printRate, "number of steps between output"),
   ROW("dt", 'D', 1, 'd', dt, "time step (in fs)"),
   ROW("lat", 'l', 1, 'd', lat, "lattice parameter (Angstroms)"),
   ROW("temp", 'T', 1, 'd', temperature, "initial temperature (K)"),
   ROW("delta", 'r', 1, 'd', initialDelta, "initial delta (Angstroms)"),
   ROW("hilbert", 'H', 0, 'i', doHilbert, "space-filling curve for the traversal of cells"),
   ROW("skinDistance", 'S', 1, 'd', relativeSkinDistance, "skinDistance (relative to cutoff (default: 0.1))"),
   ROW("method", 'm', 1, 's', method, "thread_atom,warp_atom,warp_atom_nl,cta_cell,thread_atom_nl,cpu_nl"),
   ROW("gpuAsync", 'a', 1, 'i', gpuAsync, "communicaton hiding optimization using streams"),
   ROW("gpuProfile", 's', 0, 'i', gpuProfile, "profiling mode: reboxing disabled, single kernel run"),
   ROW("ljInterpolation", 'I', 0, 'i', ljInterpolation, "compute Lennard-Jones potential using interpolation (gpu only)"),
   ROW("spline", 'P', 0, 'i', spline, "use splines for interpolation (gpu only)"),
   ROW("usePairlist", 'L', 0, 'i', usePairlist, "use pairlists for cta_cell method in Lennard-Jones computation"),
};
enum { nRows = sizeof table / sizeof table[0] };

static void printArgs(void)
{
   size_t longest = 1;
   for (int k = 0; k < nRows; ++k) if (strlen(table[k].name) > longest) longest = strlen(table[k].name);
   fprintf(screenOut, "\n  Arguments are: \n");
   for (int k = 0; k < nRows; ++k)
      fprintf(screenOut, "   --%-*s -%c  arg=%1d type=%c  %s\n", (int)longest, table[k].name, table[k].letter, table[k].hasArg, table[k].type,
              table[k].help);
   fprintf(screenOut, "\n\n");
}

Command parseCommandLine(int argc, char** argv)
{
   Command cmd;
   memset(&cmd, 0, sizeof cmd);
   strcpy(cmd.potDir, "pots");
   strcpy(cmd.potType, "funcfl");
   cmd.nx = cmd.ny = cmd.nz = 20;
   cmd.xproc = cmd.yproc = cmd.zproc = 1;
   cmd.nSteps = 100;
   cmd.printRate = 10;
   cmd.dt = 1.0;
   cmd.lat = -1.0;
   cmd.temperature = 600.0;
   cmd.initialDelta = 0.0;
   cmd.relativeSkinDistance = 0.1;
   strcpy(cmd.method, "thread_atom");

   struct option longOpts[nRows + 1];
   char shortOpts[2 * nRows + 1];
   int so = 0;
   for (int k = 0; k < nRows; ++k)
   {
      longOpts[k].name = table[k].name;
      longOpts[k].has_arg = table[k].hasArg ? required_argument : no_argument;
      longOpts[k].flag = NULL;
      longOpts[k].val = table[k].letter;
      shortOpts[so++] = table[k].letter;
      if (table[k].hasArg) shortOpts[so++] = ':';
   }
   memset(&longOpts[nRows], 0, sizeof longOpts[0]);
   shortOpts[so] = 0;

   int help = 0, c;
   optind = 1;
   while ((c = getopt_long(argc, argv, shortOpts, longOpts, NULL)) != -1)
   {
      const OptionRow* row = NULL;
      for (int k = 0; k < nRows; ++k) if (table[k].letter == c) row = &table[k];
      if (!row) { help = 1; continue; }              /* unknown option: getopt already complained */
      if (row->letter == 'h') { help = 1; continue; }
      char* dst = (char*)&cmd + row->offset;
      if (!row->hasArg) *(int*)dst = 1;
      else if (row->type == 'i') *(int*)dst = atoi(optarg);
      else if (row->type == 'd') *(double*)dst = atof(optarg);
      else { strncpy(dst, optarg, row->size - 1); dst[row->size - 1] = 0; }
   }
   if (strlen(cmd.potName) == 0)
   {
      if (strcmp(cmd.potType, "setfl") == 0) strcpy(cmd.potName, "Cu01.eam.alloy");
      if (strcmp(cmd.potType, "funcfl") == 0) strcpy(cmd.potName, "Cu_u6.eam");
   }
   if (help)
   {
      if (printRank()) printArgs();
      exit(2);
   }
   return cmd;
}

void printCmdYaml(FILE* file, Command* cmd)
{
   if (!printRank() || !file) return;
   fprintf(file,
           "Command Line Parameters:\n"
           "  doeam: %d\n"
           "  potDir: %s\n"
           "  potName: %s\n"
           "  potType: %s\n"
           "  nx: %d\n"
           "  ny: %d\n"
           "  nz: %d\n"
           "  xproc: %d\n"
           "  yproc: %d\n"
           "  zproc: %d\n"
           "  Lattice constant: %g Angstroms\n"
           "  nSteps: %d\n"
           "  printRate: %d\n"
           "  Time step: %g fs\n"
           "  Initial Temperature: %g K\n"
           "  Initial Delta: %g Angstroms\n\n"
           "  GPU async opt: %d\n"
           "  GPU profiling mode: %d\n"
           "  GPU method: %s\n"
           "  Space-filling (Hilbert): %d\n"
           "\n",
           cmd->doeam, cmd->potDir, cmd->potName, cmd->potType, cmd->nx, cmd->ny, cmd->nz, cmd->xproc, cmd->yproc, cmd->zproc,
           cmd->lat, cmd->nSteps, cmd->printRate, cmd->dt, cmd->temperature, cmd->initialDelta, cmd->gpuAsync, cmd->gpuProfile,
           cmd->method, cmd->doHilbert);
   fflush(file);
}
