/* oracle/shim: see gsl_randist.h.  TEST INFRASTRUCTURE ONLY. */
#ifndef ORACLE_SHIM_GSL_ROOTS_H
#define ORACLE_SHIM_GSL_ROOTS_H
#ifdef __cplusplus
extern "C" {
#endif
typedef struct gsl_function_struct {
    double (*function)(double x, void *params);
    void *params;
} gsl_function;
typedef struct gsl_root_fsolver_type_s gsl_root_fsolver_type;
typedef struct gsl_root_fsolver_s gsl_root_fsolver;
extern const gsl_root_fsolver_type *gsl_root_fsolver_brent;
gsl_root_fsolver *gsl_root_fsolver_alloc(const gsl_root_fsolver_type *T);
void gsl_root_fsolver_free(gsl_root_fsolver *s);
int gsl_root_fsolver_set(gsl_root_fsolver *s, gsl_function *f, double x_lower, double x_upper);
int gsl_root_fsolver_iterate(gsl_root_fsolver *s);
const char *gsl_root_fsolver_name(const gsl_root_fsolver *s);
double gsl_root_fsolver_root(const gsl_root_fsolver *s);
double gsl_root_fsolver_x_lower(const gsl_root_fsolver *s);
double gsl_root_fsolver_x_upper(const gsl_root_fsolver *s);
int gsl_root_test_interval(double x_lower, double x_upper, double epsabs, double epsrel);
#ifdef __cplusplus
}
#endif
#endif
