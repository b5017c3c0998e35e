/* oracle/shim: link-time stand-ins for the GSL entry points declared in
 * gsl/*.h.  Nothing on the graded path reaches them (HumanEyeCamera only,
 * SURVEY.md F12); OmniCamera's constructor calls gsl_rng_alloc and ignores the
 * result.  TEST INFRASTRUCTURE ONLY. */
#include <stdio.h>
#include <stdlib.h>
#include "gsl/gsl_randist.h"
#include "gsl/gsl_errno.h"
#include "gsl/gsl_roots.h"
static void die(const char *n) { fprintf(stderr, "oracle shim: %s reached (GSL absent)\n", n); abort(); }
const gsl_rng_type *gsl_rng_default = 0;
const gsl_root_fsolver_type *gsl_root_fsolver_brent = 0;
const gsl_rng_type *gsl_rng_env_setup(void) { return 0; }
gsl_rng *gsl_rng_alloc(const gsl_rng_type *T) { (void)T; return 0; }
void gsl_ran_bivariate_gaussian(const gsl_rng *r, double a, double b, double c, double *x, double *y)
{ (void)r; (void)a; (void)b; (void)c; (void)x; (void)y; die("gsl_ran_bivariate_gaussian"); }
gsl_error_handler_t *gsl_set_error_handler_off(void) { return 0; }
gsl_root_fsolver *gsl_root_fsolver_alloc(const gsl_root_fsolver_type *T) { (void)T; die("gsl_root_fsolver_alloc"); return 0; }
void gsl_root_fsolver_free(gsl_root_fsolver *s) { (void)s; }
int gsl_root_fsolver_set(gsl_root_fsolver *s, gsl_function *f, double a, double b) { (void)s; (void)f; (void)a; (void)b; die("gsl_root_fsolver_set"); return -1; }
int gsl_root_fsolver_iterate(gsl_root_fsolver *s) { (void)s; die("gsl_root_fsolver_iterate"); return -1; }
const char *gsl_root_fsolver_name(const gsl_root_fsolver *s) { (void)s; return "stub"; }
double gsl_root_fsolver_root(const gsl_root_fsolver *s) { (void)s; return 0; }
double gsl_root_fsolver_x_lower(const gsl_root_fsolver *s) { (void)s; return 0; }
double gsl_root_fsolver_x_upper(const gsl_root_fsolver *s) { (void)s; return 0; }
int gsl_root_test_interval(double a, double b, double c, double d) { (void)a; (void)b; (void)c; (void)d; return 0; }
