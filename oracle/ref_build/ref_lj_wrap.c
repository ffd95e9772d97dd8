/* Compiles the reference ljForce.c where it lies and makes its hard-coded LJ
 * cutoff multiple (5*sigma, ljForce.c:114) selectable at run time; the default
 * is BASELINE.json's 2.5*sigma.  Nothing from the reference is copied here. */
#define initLjPot refInitLjPotAsShipped
#include "ljForce.c"
#undef initLjPot

static double g_ljCutoffMult = 2.5;
void ref_set_lj_cutoff_mult(double m) { g_ljCutoffMult = m; }
double ref_get_lj_cutoff_mult(void) { return g_ljCutoffMult; }

BasePotential* initLjPot(void)
{
   LjPotential* pot = (LjPotential*)refInitLjPotAsShipped();
   pot->cutoff = g_ljCutoffMult * pot->sigma;
   return (BasePotential*)pot;
}
