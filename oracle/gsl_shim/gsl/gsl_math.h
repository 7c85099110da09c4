/* oracle/gsl_shim/gsl/gsl_math.h -- TEST INFRASTRUCTURE (oracle). Not part of the product path.
 * The reference includes this header (Parameters.cpp:8-10) but uses nothing from it. */
#ifndef GSL_SHIM_gsl_math_H
#define GSL_SHIM_gsl_math_H
#include <cmath>
#endif
