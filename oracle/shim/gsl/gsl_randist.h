/* oracle/shim: declaration-only stand-in for the GSL headers that
 * /root/reference/src/pbrt/cameras.h:27 and cameras.cpp:33-34 include.
 * GSL is not vendored by the reference and not installed here.  Only
 * HumanEyeCamera (out of scope) calls these at run time;
 * the definitions in gsl_stubs.c abort if they are ever reached.
 * TEST INFRASTRUCTURE ONLY - never linked into the product library. */
#ifndef ORACLE_SHIM_GSL_RANDIST_H
#define ORACLE_SHIM_GSL_RANDIST_H
#ifdef __cplusplus
extern "C" {
#endif
typedef struct gsl_rng_type_s gsl_rng_type;
typedef struct gsl_rng_s gsl_rng;
extern const gsl_rng_type *gsl_rng_default;
const gsl_rng_type *gsl_rng_env_setup(void);
gsl_rng *gsl_rng_alloc(const gsl_rng_type *T);
void gsl_ran_bivariate_gaussian(const gsl_rng *r, double sigma_x, double sigma_y,
                                double rho, double *x, double *y);
#ifdef __cplusplus
}
#endif
#endif
