/* Compiles the reference eam.c where it lies and exports its file-static
 * table lookup so the restated oracle and the CUDA path can be checked against
 * the reference's own interpolate() (eam.c:557-579) and table construction
 * (eam.c:496-519, 680-872). */
#include "eam.c"

/* which: 0 = phi, 1 = rho, 2 = F */
static InterpolationObject* refTable(BasePotential* bp, int which)
{
   EamPotential* pot = (EamPotential*)bp;
   return which == 0 ? pot->phi : (which == 1 ? pot->rho : pot->f);
}
void ref_eam_interpolate(BasePotential* bp, int which, int n, const double* x, double* f, double* df)
{
   InterpolationObject* t = refTable(bp, which);
   for (int i = 0; i < n; ++i) { real_t a, b; interpolate(t, x[i], &a, &b); f[i] = a; df[i] = b; }
}
int ref_eam_table_n(BasePotential* bp, int which) { return refTable(bp, which)->n; }
void ref_eam_table(BasePotential* bp, int which, double* x0, double* invDx, double* valuesPadded /* n+3, from index -1 */)
{
   InterpolationObject* t = refTable(bp, which);
   *x0 = t->x0; *invDx = t->invDx;
   for (int i = -1; i < t->n + 2; ++i) valuesPadded[i + 1] = t->values[i];
}
