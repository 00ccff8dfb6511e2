/* Empty stand-in (see gsl_rng.h). */
#ifndef ORACLE_STUB_GSL_SF_H
#define ORACLE_STUB_GSL_SF_H
#endif
