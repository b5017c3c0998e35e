/* oracle/shim: see gsl_randist.h.  TEST INFRASTRUCTURE ONLY. */
#ifndef ORACLE_SHIM_GSL_ERRNO_H
#define ORACLE_SHIM_GSL_ERRNO_H
#ifdef __cplusplus
extern "C" {
#endif
enum { GSL_SUCCESS = 0, GSL_FAILURE = -1, GSL_CONTINUE = -2, GSL_EINVAL = 4 };
typedef void gsl_error_handler_t(const char *reason, const char *file, int line, int gsl_errno);
gsl_error_handler_t *gsl_set_error_handler_off(void);
#ifdef __cplusplus
}
#endif
#endif
