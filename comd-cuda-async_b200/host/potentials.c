/* potentials.c -- the BasePotential "classes": Lennard-Jones and tabulated EAM.
 *
 * Constructors, print and destroy follow ljForce.c:102-134 and eam.c:141-172,420-443; the readers
 * restate eamReadSetfl (eam.c:680-757) and eamReadFuncfl (:802-872) and the table padding of
 * initInterpolationObject (:496-519) so that the tables handed to the device are the reference's,
 * value for value.  `force` forwards to the device layer: there is no CPU force loop in this build.
 * Differences: every rank reads the (small) file itself, so eamBcastPotential has nothing to do;
 * the LJ cutoff multiple is a named constant (2.5 sigma as BASELINE.json states; the fork under
 * /root/reference hard-codes 5 sigma at ljForce.c:114), overridable with the
 * environment variable COMD_LJ_CUTOFF_SIGMAS; setfl files get their element name from line 4
 * (the reference leaves pot->name uninitialised on that path).
 */
#include "comd_host.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define LJ_CUTOFF_SIGMAS 2.5

/* ---------------------------------------------------------------- Lennard-Jones */
static int ljForce(SimFlat* s)
{
   DEV_TRY(comd_dev_compute_force(s->dev));
   s->hostAtomsStale = 1;
   return 0;
}

static void ljPrint(FILE* file, BasePotential* pot)
{
   LjPotential* lj = (LjPotential*)pot;
   fprintf(file, "  Potential type   : Lennard-Jones\n");
   fprintf(file, "  Species name     : %s\n", lj->name);
   fprintf(file, "  Atomic number    : %d\n", lj->atomicNo);
   fprintf(file, "  Mass             : %lg amu\n", lj->mass / amuToInternalMass);
   fprintf(file, "  Lattice Type     : %s\n", lj->latticeType);
   fprintf(file, "  Lattice spacing  : %lg Angstroms\n", lj->lat);
   fprintf(file, "  Cutoff           : %lg Angstroms\n", lj->cutoff);
   fprintf(file, "  Epsilon          : %lg eV\n", lj->epsilon);
   fprintf(file, "  Sigma            : %lg Angstroms\n", lj->sigma);
}

static void ljDestroy(BasePotential** pot)
{
   if (!pot || !*pot) return;
   free(*pot);
   *pot = NULL;
}

BasePotential* initLjPot(void)
{
   LjPotential* lj = calloc(1, sizeof *lj);
   lj->force = ljForce;
   lj->print = ljPrint;
   lj->destroy = ljDestroy;
   lj->sigma = 2.315;
   lj->epsilon = 0.167;
   lj->mass = 63.55 * amuToInternalMass;
   lj->lat = 3.615;
   strcpy(lj->latticeType, "FCC");
   double mult = LJ_CUTOFF_SIGMAS;
   const char* ev = getenv("COMD_LJ_CUTOFF_SIGMAS");
   if (ev && atof(ev) > 0.0) mult = atof(ev);
   lj->cutoff = mult * lj->sigma;
   strcpy(lj->name, "Cu");
   lj->atomicNo = 29;
   return (BasePotential*)lj;
}

/* ---------------------------------------------------------------- EAM tables */
static InterpolationObject* makeTable(int n, real_t x0, real_t dx, const real_t* data)
{
   InterpolationObject* t = calloc(1, sizeof *t);
   real_t* store = calloc(n + 3, sizeof(real_t));
   t->values = store + 1;                       /* valid on [-1, n+1] */
   t->n = n;
   t->invDx = 1.0 / dx;
   t->x0 = x0;
   for (int i = 0; i < n; ++i) t->values[i] = data[i];
   t->values[-1] = t->values[0];
   t->values[n + 1] = t->values[n] = t->values[n - 1];
   return t;
}

static void freeTable(InterpolationObject** t)
{
   if (!t || !*t) return;
   free((*t)->values - 1);
   free(*t);
   *t = NULL;
}

/* the reference's lookup (eam.c:557-579), kept on the host for diagnostics and tests only */
void interpolate(InterpolationObject* table, real_t r, real_t* f, real_t* df)
{
   const real_t* tt = table->values;
   if (r < table->x0) r = table->x0;
   r = (r - table->x0) * table->invDx;
   int ii = (int)floor(r);
   if (ii > table->n) { ii = table->n; r = table->n / table->invDx; }
   r = r - floor(r);
   const real_t g1 = tt[ii + 1] - tt[ii - 1];
   const real_t g2 = tt[ii + 2] - tt[ii];
   *f = tt[ii] + 0.5 * r * (g1 + r * (tt[ii + 1] + tt[ii - 1] - 2.0 * tt[ii]));
   *df = 0.5 * (g1 + r * (g2 - g1)) * table->invDx;
}

static FILE* openPot(const char* who, const char* dir, const char* name)
{
   char path[4096];
   snprintf(path, sizeof path, "%s/%s", dir, name);
   FILE* fp = fopen(path, "r");
   if (!fp)
   {
      fprintf(stderr, "%s: potential file %s not found (--potDir; the NIST Cu tables are input data, see README)\n", who, path);
      exit(-1);
   }
   return fp;
}

static void readNumbers(FILE* fp, real_t* buf, int n, const char* who)
{
   for (int i = 0; i < n; ++i)
      if (fscanf(fp, FMT1, buf + i) != 1) { fprintf(stderr, "%s: short table (entry %d of %d)\n", who, i, n); exit(-1); }
}

static void readSetfl(EamPotential* pot, const char* dir, const char* name)
{
   FILE* fp = openPot("eamReadSetfl", dir, name);
   char line[4096];
   for (int k = 0; k < 4; ++k) if (!fgets(line, sizeof line, fp)) line[0] = 0;     /* 3 comment lines + element line */
   int nElems = 0; char elem[64] = "";
   sscanf(line, "%d %63s", &nElems, elem);
   if (nElems != 1) { fprintf(stderr, "eamReadSetfl: CoMD 1.1 does not support alloys and cannot read setfl files with multiple species\n"); exit(-1); }
   strncpy(pot->name, elem, 2); pot->name[2] = 0;
   int nRho = 0, nR = 0; double dRho = 0, dR = 0, cutoff = 0;      /* doubles in either precision (eam.c:707,834) */
   if (!fgets(line, sizeof line, fp)) line[0] = 0;
   sscanf(line, "%d %le %d %le %le", &nRho, &dRho, &nR, &dR, &cutoff);
   pot->cutoff = cutoff;
   int atomicNo = 0; double mass = 0, lat = 0; char lattice[8] = "";
   if (!fgets(line, sizeof line, fp)) line[0] = 0;
   sscanf(line, "%d %le %le %7s", &atomicNo, &mass, &lat, lattice);
   pot->atomicNo = atomicNo; pot->lat = lat; pot->mass = mass * amuToInternalMass;
   strcpy(pot->latticeType, lattice);
   real_t* buf = malloc((nRho > nR ? nRho : nR) * sizeof(real_t));
   const real_t x0 = 0.0;
   readNumbers(fp, buf, nRho, "eamReadSetfl");
   pot->f = makeTable(nRho, x0, dRho, buf);
   readNumbers(fp, buf, nR, "eamReadSetfl");
   pot->rho = makeTable(nR, x0, dR, buf);
   readNumbers(fp, buf, nR, "eamReadSetfl");                 /* the file tabulates r*phi(r) */
   for (int i = 1; i < nR; ++i) { const real_t r = x0 + i * dR; buf[i] /= r; }
   buf[0] = buf[1] + (buf[1] - buf[2]);
   pot->phi = makeTable(nR, x0, dR, buf);
   free(buf);
   fclose(fp);
}

static void readFuncfl(EamPotential* pot, const char* dir, const char* name)
{
   FILE* fp = openPot("eamReadFuncfl", dir, name);
   char line[4096], elem[64] = "";
   if (!fgets(line, sizeof line, fp)) line[0] = 0;
   sscanf(line, "%63s", elem);
   strncpy(pot->name, elem, 2); pot->name[2] = 0;
   int atomicNo = 0; double mass = 0, lat = 0; char lattice[8] = "";
   if (!fgets(line, sizeof line, fp)) line[0] = 0;
   sscanf(line, "%d %le %le %7s", &atomicNo, &mass, &lat, lattice);
   pot->atomicNo = atomicNo; pot->lat = lat; pot->mass = mass * amuToInternalMass;
   strcpy(pot->latticeType, lattice);
   int nRho = 0, nR = 0; double dRho = 0, dR = 0, cutoff = 0;      /* doubles in either precision (eam.c:707,834) */
   if (!fgets(line, sizeof line, fp)) line[0] = 0;
   sscanf(line, "%d %le %d %le %le", &nRho, &dRho, &nR, &dR, &cutoff);
   pot->cutoff = cutoff;
   const real_t x0 = 0.0;
   real_t* buf = malloc((nRho > nR ? nRho : nR) * sizeof(real_t));
   readNumbers(fp, buf, nRho, "eamReadFuncfl");
   pot->f = makeTable(nRho, x0, dRho, buf);
   readNumbers(fp, buf, nR, "eamReadFuncfl");                /* Z(r): phi = Z^2/r in Hartree*Bohr */
   for (int i = 1; i < nR; ++i)
   {
      const real_t r = x0 + i * dR;
      buf[i] *= buf[i] / r;
      buf[i] *= hartreeToEv * bohrToAngs;
   }
   buf[0] = buf[1] + (buf[1] - buf[2]);
   pot->phi = makeTable(nR, x0, dR, buf);
   readNumbers(fp, buf, nR, "eamReadFuncfl");
   pot->rho = makeTable(nR, x0, dR, buf);
   free(buf);
   fclose(fp);
}

static int eamForce(SimFlat* s)
{
   /* passes 1+2, the F'(rho) halo, pass 3 -- one device call (eam.c:196-264) */
   profileStart(eamHaloTimer); profileStop(eamHaloTimer);
   DEV_TRY(comd_dev_compute_force(s->dev));
   s->hostAtomsStale = 1;
   return 0;
}

static void eamPrint(FILE* file, BasePotential* pot)
{
   EamPotential* e = (EamPotential*)pot;
   fprintf(file, "  Potential type  : EAM\n");
   fprintf(file, "  Species name    : %s\n", e->name);
   fprintf(file, "  Atomic number   : %d\n", e->atomicNo);
   fprintf(file, "  Mass            : %lg amu\n", e->mass / amuToInternalMass);
   fprintf(file, "  Lattice type    : %s\n", e->latticeType);
   fprintf(file, "  Lattice spacing : %lg Angstroms\n", e->lat);
   fprintf(file, "  Cutoff          : %lg Angstroms\n", e->cutoff);
}

static void eamDestroy(BasePotential** pPot)
{
   if (!pPot || !*pPot) return;
   EamPotential* e = (EamPotential*)*pPot;
   freeTable(&e->phi); freeTable(&e->rho); freeTable(&e->f);
   destroyHaloExchange(&e->forceExchange);
   free(e);
   *pPot = NULL;
}

void eamBcastPotential(EamPotential* pot) { (void)pot; }    /* every rank read the file itself */

BasePotential* initEamPot(const char* dir, const char* file, const char* type)
{
   EamPotential* e = calloc(1, sizeof *e);
   e->force = eamForce;
   e->print = eamPrint;
   e->destroy = eamDestroy;
   if (strcmp(type, "setfl") == 0) readSetfl(e, dir, file);
   else if (strcmp(type, "funcfl") == 0) readFuncfl(e, dir, file);
   else
   {
      fprintf(stderr, "initEamPot: Potential type %s not supported. Fatal Error.\n", type);
      exit(-1);
   }
   eamBcastPotential(e);
   return (BasePotential*)e;
}
