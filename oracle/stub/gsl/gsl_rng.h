/* Empty stand-in so the reference's gfl_utils.h (only its MAX macro is used on
 * the fast-binless path) can be included without GSL installed. Test
 * infrastructure only. */
#ifndef ORACLE_STUB_GSL_RNG_H
#define ORACLE_STUB_GSL_RNG_H
typedef struct gsl_rng gsl_rng;
#endif
