/* Empty stand-in (see gsl_rng.h). */
#ifndef ORACLE_STUB_GSL_RANDIST_H
#define ORACLE_STUB_GSL_RANDIST_H
#endif
